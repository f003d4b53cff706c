#!/usr/bin/env python
"""Build libb200ml.so in-tree with nvcc for sm_100a (cross-compiles without a GPU).

    python build.py            # incremental (per-file object cache under build/)
    python build.py --force    # rebuild everything
    python build.py --sass     # also dump SASS mnemonic summaries into profiles/sass/

The library is plain CUDA/C++ behind a C ABI (include/b200ml.h) and is loaded with ctypes, so a build takes
seconds per file and the resulting .so travels to the GPU box with the repo snapshot.
"""
from __future__ import annotations

import argparse
import concurrent.futures as cf
import hashlib
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "machine_learning---apache-spark_b200")
OUT = os.path.join(PKG, "lib", "libb200ml.so")
BUILD = os.path.join(ROOT, "build")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
NVCC_FLAGS = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr",
              "-Xptxas", "-v", "-I", os.path.join(ROOT, "include")]


def nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def sources() -> list[str]:
    srcs = []
    for sub in ("csrc/kernels", "csrc/comm", "csrc/runtime"):
        d = os.path.join(ROOT, sub)
        if not os.path.isdir(d):
            continue
        for f in sorted(os.listdir(d)):
            if f.endswith((".cu", ".cpp")):
                srcs.append(os.path.join(d, f))
    return srcs


def _dep_hash(src: str) -> str:
    h = hashlib.sha256()
    h.update(" ".join(NVCC_FLAGS + ARCH).encode())
    deps = [src]
    for sub in ("include", "csrc/common"):
        d = os.path.join(ROOT, sub)
        deps += [os.path.join(d, f) for f in sorted(os.listdir(d))]
    for p in deps:
        with open(p, "rb") as fh:
            h.update(fh.read())
    return h.hexdigest()[:16]


def compile_one(src: str, force: bool) -> tuple[str, str]:
    os.makedirs(BUILD, exist_ok=True)
    tag = _dep_hash(src)
    base = os.path.relpath(src, ROOT).replace(os.sep, "_")
    obj = os.path.join(BUILD, f"{base}.{tag}.o")
    log = ""
    if force or not os.path.exists(obj):
        for old in os.listdir(BUILD):
            if old.startswith(base + ".") and (old.endswith(".o") or old.endswith(".o.log")):
                os.remove(os.path.join(BUILD, old))
        cmd = [nvcc(), *ARCH, *NVCC_FLAGS, "-x", "cu", "-c", src, "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        log = r.stdout + r.stderr
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src}:\n{log}")
        with open(obj + ".log", "w") as fh:
            fh.write(log)
    return obj, log


def build(force: bool = False, verbose: bool = False) -> str:
    srcs = sources()
    objs = []
    with cf.ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 4)) as ex:
        for obj, log in ex.map(lambda s: compile_one(s, force), srcs):
            objs.append(obj)
            if verbose and log:
                print(log)
    newest = max(os.path.getmtime(o) for o in objs)
    if force or not os.path.exists(OUT) or os.path.getmtime(OUT) < newest:
        os.makedirs(os.path.dirname(OUT), exist_ok=True)
        cmd = [nvcc(), *ARCH, "-shared", "-Xcompiler", "-fPIC", "-o", OUT, *objs, "-lcudart"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("link failed:\n" + r.stdout + r.stderr)
    return OUT


def dump_sass() -> None:
    """Mnemonic histogram per kernel — the evidence that tcgen05/TMA are really in the binary."""
    outdir = os.path.join(ROOT, "profiles", "sass")
    os.makedirs(outdir, exist_ok=True)
    r = subprocess.run(["cuobjdump", "-sass", OUT], capture_output=True, text=True)
    cur, hist = None, {}
    for line in r.stdout.splitlines():
        s = line.strip()
        if s.startswith("Function :"):
            cur = s.split(":", 1)[1].strip()
            hist[cur] = {}
        elif cur and s.startswith("/*") and "*/" in s:
            body = s.split("*/", 1)[1].strip()
            if not body or body.startswith("/*"):
                continue
            tok = body.split()[0]
            if tok.startswith("@"):
                tok = body.split()[1] if len(body.split()) > 1 else tok
            mn = tok.rstrip(";").split(".")[0]
            hist[cur][mn] = hist[cur].get(mn, 0) + 1
    keys = ("UTCHMMA", "UTCQMMA", "UTCBAR", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "SYNCS", "HMMA", "MULTIMEM",
            "RED", "ATOM")
    with open(os.path.join(outdir, "summary.txt"), "w") as fh:
        fh.write("# cuobjdump -sass libb200ml.so : per-kernel counts of the Blackwell-specific mnemonics\n")
        for fn, h in sorted(hist.items()):
            total = sum(h.values())
            picks = {k: v for k, v in h.items() if any(k.startswith(p) for p in keys)}
            fh.write(f"{fn}\n    instructions={total} " + " ".join(f"{k}={v}" for k, v in sorted(picks.items())) + "\n")
    print("wrote", os.path.join(outdir, "summary.txt"))


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--sass", action="store_true")
    ap.add_argument("-v", "--verbose", action="store_true")
    a = ap.parse_args()
    try:
        out = build(a.force, a.verbose)
    except RuntimeError as e:
        print(e, file=sys.stderr)
        sys.exit(1)
    print("built", out)
    if a.sass:
        dump_sass()
