"""engine.graphed(): the CUDA-graph replay of a whole training step must train exactly like the eager loop."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _train(graphed: bool):
    from b200ml.models import mnist_mlp
    from b200ml.ops.loss import cross_entropy
    from b200ml.parallel.engine import DataParallelEngine
    torch.manual_seed(0)
    dev = torch.device("cuda")
    model = mnist_mlp().to(dev).train()
    eng = DataParallelEngine(model, optimizer="adam", lr=1e-3)

    def one_step(x, y):
        loss = cross_entropy(model(x), y)
        eng.zero_grad()
        loss.backward()
        eng.step()
        return loss.detach()

    step = eng.graphed(one_step) if graphed else one_step
    g = torch.Generator().manual_seed(1)
    losses = []
    for i in range(8):
        n = 32 if i != 5 else 7                      # one short batch: a second signature, stays eager
        x = torch.randn(n, 784, generator=g).to(dev)
        y = torch.randint(0, 10, (n,), generator=g).to(dev)
        losses.append(float(step(x, y)))
    w = torch.cat([p.detach().reshape(-1).clone() for p in model.parameters()])
    replays = step.replays if graphed else 0
    eng.close()
    return w, losses, replays


def test_graphed_step_matches_eager():
    w_e, l_e, _ = _train(False)
    w_g, l_g, replays = _train(True)
    assert replays == 5                              # 7 full batches: 2 eager warm-ups, then capture + 5 replays
    assert torch.equal(w_e, w_g), (w_e - w_g).abs().max()
    assert l_e == l_g


def test_model_is_picklable_after_training_with_graphs():
    """TorchDistributor returns rank 0's value through cloudpickle: after engine.close() a trained Transformer must not
    hold CUDA graphs, streams or hooks (this broke `pytorch_machine_translator.py --conf spark.executor.instances=2`)."""
    import cloudpickle
    from b200ml.models.transformer import Transformer
    from b200ml.ops.loss import linear_cross_entropy
    from b200ml.parallel.engine import DataParallelEngine
    torch.manual_seed(0)
    dev = torch.device("cuda")
    S = 64
    model = Transformer(128, 256, 2, 0.1, 1, S, 300, 200, 300).to(dev).train()
    eng = DataParallelEngine(model, optimizer="adam", lr=1e-3)
    la = (torch.tril(torch.ones(S, S, device=dev)) == 0)[None, None]

    def one_step(src, trg):
        eng.zero_grad()
        h = model(src, trg, (src != 0)[:, None, None, :], la, la, return_hidden=True)
        loss = linear_cross_entropy(h, model.linear.weight, model.linear.bias, trg, ignore_index=0)
        loss.backward()
        eng.step()
        return loss.detach()

    step = eng.graphed(one_step)
    for _ in range(4):
        step(torch.randint(1, 200, (4, S), device=dev), torch.randint(1, 300, (4, S), device=dev))
    assert step.replays >= 1
    eng.close()
    blob = cloudpickle.dumps(model)
    clone = cloudpickle.loads(blob)
    assert set(clone.state_dict()) == set(model.state_dict())
