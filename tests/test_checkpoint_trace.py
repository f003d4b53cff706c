"""Checkpoint/resume and tracing helpers (CPU)."""
import torch


def _train(model, opt, steps, seed):
    from b200ml.ops.loss import cross_entropy
    g = torch.Generator().manual_seed(seed)
    for _ in range(steps):
        x = torch.randn(16, 4, generator=g)
        y = torch.randint(0, 3, (16,), generator=g)
        opt.zero_grad()
        cross_entropy(model(x), y).backward()
        opt.step()


def test_checkpoint_resume_reproduces_uninterrupted_run(tmp_path):
    from b200ml.models import Multilayer_perceptor
    from b200ml.parallel.engine import DataParallelEngine
    from b200ml.utils.checkpoint import load_checkpoint, save_checkpoint
    torch.manual_seed(0)
    ref = Multilayer_perceptor()
    init = {k: v.clone() for k, v in ref.state_dict().items()}
    e_ref = DataParallelEngine(ref, optimizer="adam", lr=0.01)
    _train(ref, e_ref, 3, seed=1)
    _train(ref, e_ref, 3, seed=2)

    a = Multilayer_perceptor()
    a.load_state_dict(init)
    e_a = DataParallelEngine(a, optimizer="adam", lr=0.01)
    _train(a, e_a, 3, seed=1)
    save_checkpoint(str(tmp_path / "ck.pt"), a, e_a, extra={"epoch": 1})
    sd = torch.load(tmp_path / "ck.pt", weights_only=False)["model"]
    assert list(sd.keys()) == list(init.keys()) and all(v.dtype == torch.float32 for v in sd.values())

    b = Multilayer_perceptor()
    e_b = DataParallelEngine(b, optimizer="adam", lr=0.01)
    extra = load_checkpoint(str(tmp_path / "ck.pt"), b, e_b)
    assert extra == {"epoch": 1} and int(e_b.step_tensor) == 3
    _train(b, e_b, 3, seed=2)
    for k in init:
        assert torch.allclose(b.state_dict()[k], ref.state_dict()[k], atol=1e-6), k


def test_phase_timer_and_launch_counter():
    from b200ml.utils.trace import PhaseTimer, kernel_launches, nvtx_range
    t = PhaseTimer("cpu")
    with t.phase("fwd"), nvtx_range("x"):
        sum(range(1000))
    with t.phase("fwd"):
        pass
    s = t.summary()
    assert "fwd" in s and s["fwd"] >= 0 and isinstance(kernel_launches(), int)
