"""Checkpoint / resume (SURVEY §5.4 — absent from the reference, whose ``model = distributor.run(train_func)`` hints at
the intent).  The model part is the plain ``state_dict()`` with the reference's keys, fp32, loadable by ``torch.load`` +
``load_state_dict`` into the reference classes; the optimizer part is the engine's (or FlatOptimizer's) step counter and
Adam moments in parameter order, so it survives a different bucket layout / world size."""
from __future__ import annotations

import os

import torch


def _per_param_state(engine_or_opt):
    flat = engine_or_opt.flat
    out = {}
    if getattr(engine_or_opt, "m", None) is not None:
        # multi-GPU engine: only a slice's owner keeps its Adam moments current -> gather them first (collective)
        if hasattr(engine_or_opt, "gather_moments"):
            m, v = engine_or_opt.gather_moments()
        else:
            m, v = engine_or_opt.m, engine_or_opt.v
        for i, p in enumerate(flat.params):
            o, n = flat.slot(i)
            out[i] = {"m": m[o:o + n].detach().cpu().clone(), "v": v[o:o + n].detach().cpu().clone()}
    return out


def save_checkpoint(path: str, model, engine_or_opt=None, extra: dict | None = None) -> None:
    """Call on EVERY rank: gathering the Adam moments from the slice owners is a collective (the fp32 master pull is
    one-sided); only rank 0 writes.  The file holds full per-parameter state, so it can be resumed at another world
    size or bucket layout: ``load_checkpoint`` gives every rank the full moments, which covers whatever it will own."""
    if engine_or_opt is not None and hasattr(engine_or_opt, "sync_master"):
        engine_or_opt.sync_master()
    rank = int(os.environ.get("RANK", "0"))
    blob = {"model": {k: v.detach().cpu() for k, v in model.state_dict().items()}, "extra": extra or {}}
    if engine_or_opt is not None:
        names = {id(p): n for n, p in model.named_parameters()}
        blob["optimizer"] = {"step": int(engine_or_opt.step_tensor.item()), "kind": engine_or_opt.kind,
                             "order": [names.get(id(p)) for p in engine_or_opt.flat.params],
                             "state": _per_param_state(engine_or_opt)}
    if rank == 0:
        os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
        tmp = path + ".tmp"
        torch.save(blob, tmp)
        os.replace(tmp, path)


def load_checkpoint(path: str, model, engine_or_opt=None, map_location="cpu") -> dict:
    blob = torch.load(path, map_location=map_location, weights_only=False)
    model.load_state_dict(blob["model"], strict=True)
    if engine_or_opt is not None:
        flat = engine_or_opt.flat
        flat.refresh_working_copy()           # parameters are views into the flat master buffer: re-cast to bf16
        opt = blob.get("optimizer")
        if opt:
            engine_or_opt.step_tensor.fill_(opt["step"])
            names = {n: p for n, p in model.named_parameters()}
            index = {id(p): i for i, p in enumerate(flat.params)}
            for j, name in enumerate(opt["order"]):
                st = opt["state"].get(j)
                if st is None or name not in names:
                    continue
                i = index[id(names[name])]
                o, n = flat.slot(i)
                engine_or_opt.m[o:o + n].copy_(st["m"])
                engine_or_opt.v[o:o + n].copy_(st["v"])
    return blob.get("extra", {})
