"""Diagnostic (not a benchmark): why do the two arms of bench.py print different `final_loss` values?

Trains the headline translator config for a few dozen steps on the batches bench.py uses, four ways, and prints the loss
curves side by side:

  ref        the UNMODIFIED reference transformer.py (baseline/_ref), bf16 autocast, torch.optim.Adam, its own init
  own        this framework, its own random init (what bench.py does)
  own@ref    this framework, weights loaded from the reference model's state_dict (same init as `ref`)
  ref0/own0  both with dropout off (reference in eval() mode - its SentenceEmbedding hard-codes Dropout(0.1)), same init

    python bench/loss_curve_check.py [--steps 45]
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
from torch import nn  # noqa: E402

import bench as B  # noqa: E402


class _Vocab:
    def __init__(self, n):
        self.n = n

    def __len__(self):
        return self.n


def build_ref(dev, p):
    sys.path.insert(0, os.path.join(ROOT, "baseline", "_ref"))
    import transformer as ref_transformer

    c = B.CFG
    torch.manual_seed(1234)
    return ref_transformer.Transformer(c["d_model"], c["ffn_hidden"], c["num_heads"], p, c["num_layers"],
                                       c["max_sequence_length"], c["v_de"], _Vocab(c["v_en"]), _Vocab(c["v_de"])).to(dev)


def train_ref(model, batches, steps, dev, train_mode=True):
    c = B.CFG
    S = c["max_sequence_length"]
    model.train(train_mode)
    opt = torch.optim.Adam(model.parameters(), lr=c["lr"])
    loss_fn = nn.CrossEntropyLoss(ignore_index=0, reduction="none")
    la = (torch.tril(torch.ones(S, S)) == 0).unsqueeze(0).unsqueeze(0).to(dev)
    out = []
    for i in range(steps):
        s, t = batches[i % len(batches)]
        src, trg = s.to(dev), t.to(dev)
        pad = (src != 0).unsqueeze(1).unsqueeze(2)
        with torch.autocast("cuda", dtype=torch.bfloat16):
            pred = model(src, trg, pad, la, la)
        loss = loss_fn(pred.view(-1, pred.size(-1)).float(), trg.view(-1))
        valid = trg.view(-1) != 0
        loss = loss[valid].sum() / valid.sum()
        opt.zero_grad()
        loss.backward()
        opt.step()
        out.append(float(loss))
    return out


def train_own(batches, steps, dev, p, state_dict=None):
    import b200ml  # noqa: F401
    from b200ml.models.transformer import Transformer
    from b200ml.ops.loss import linear_cross_entropy
    from b200ml.parallel.engine import DataParallelEngine
    from b200ml.utils.rng import manual_seed

    c = B.CFG
    S = c["max_sequence_length"]
    manual_seed(1234, dev)
    model = Transformer(c["d_model"], c["ffn_hidden"], c["num_heads"], p, c["num_layers"], S, c["v_de"], c["v_en"],
                        c["v_de"]).to(dev)
    if state_dict is not None:
        model.load_state_dict(state_dict, strict=True)
    else:
        train_own.init_std = {k: round(float(v.float().std()), 4) for k, v in model.state_dict().items() if v.dim() >= 2}
    model.train()
    eng = DataParallelEngine(model, optimizer="adam", lr=c["lr"])
    la = (torch.tril(torch.ones(S, S, device=dev)) == 0)[None, None]
    out = []
    for i in range(steps):
        s, t = batches[i % len(batches)]
        src, trg = s.to(dev), t.to(dev)
        eng.begin_step()
        h = model(src, trg, (src != 0)[:, None, None, :], la, la, return_hidden=True)
        loss = linear_cross_entropy(h, model.linear.weight, model.linear.bias, trg, ignore_index=0)
        loss.backward()
        eng.finish_step()
        out.append(float(loss))
    eng.close()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=45)
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    c = B.CFG
    batches = B.make_batches(25, c["batch_per_gpu"], c["max_sequence_length"], c["v_en"], c["v_de"], seed=100)
    res = {}
    rm = build_ref(dev, c["drop_prob"])
    sd = {k: v.detach().clone() for k, v in rm.state_dict().items()}
    stats = {k: (float(v.float().std()), tuple(v.shape)) for k, v in sd.items() if v.dim() >= 2}
    res["ref"] = train_ref(rm, batches, a.steps, dev)
    res["own"] = train_own(batches, a.steps, dev, c["drop_prob"])
    res["own@ref"] = train_own(batches, a.steps, dev, c["drop_prob"], sd)
    rm0 = build_ref(dev, 0.0)
    rm0.load_state_dict(sd)
    res["ref0"] = train_ref(rm0, batches, a.steps, dev, train_mode=False)
    res["own0"] = train_own(batches, a.steps, dev, 0.0, sd)
    pick = sorted(set(list(range(0, a.steps, 4)) + [a.steps - 1]))
    print("step      " + " ".join(f"{i:7d}" for i in pick))
    for k, v in res.items():
        print(f"{k:9s} " + " ".join(f"{v[i]:7.3f}" for i in pick))
    print(json.dumps({"loss_curves": {k: [round(x, 4) for x in v] for k, v in res.items()},
                      "ref_init_std": {k: round(s, 4) for k, (s, _) in stats.items()},
                      "own_init_std": getattr(train_own, "init_std", {})}))


if __name__ == "__main__":
    main()
