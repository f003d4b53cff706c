"""SASS excerpts of the shipped kernels (CPU only: cuobjdump on the built library) -> profiles/sass/*.sass.

For every kernel that uses the Blackwell paths, the instructions around its tensor-core issue loop (UTC*MMA), its TMEM
reads (LDTM), its TMA / bulk copies (UTMALDG, UTMASTG, UBLKCP), its mbarrier traffic (SYNCS) and - for the all-reduce -
its multimem / peer accesses are kept with a few lines of context, so that the listing shows the real inner loops, not
just mnemonic counts (those are in summary.txt, `python build.py --sass`).

    python bench/sass_excerpts.py
"""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "machine_learning---apache-spark_b200", "lib", "libb200ml.so")
OUT = os.path.join(ROOT, "profiles", "sass")
KEYS = ("UTCHMMA", "UTCQMMA", "UTCBAR", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UBLKCP", "UTMAPF", "MULTIMEM", "SYNCS",
        "DFMA", "REDUX")
WANT = ("gemm_bf16_kernel", "attn_fwd_kernel", "attn_bwd_fused_kernel", "conv_igemm_kernel", "conv_wgrad_igemm_kernel",
        "allreduce_bulk_kernel", "allreduce_nvls_kernel", "allreduce_update_kernel", "xent_rows_kernel",
        "mlpc_dgemm_kernel", "lstm_rec_fwd_kernel", "lstm2_fwd_kernel", "lstm2_bwd_kernel", "ln_bwd_kernel",
        "embedding_bwd_sorted_kernel", "optim_step_kernel")
CONTEXT, MAX_WINDOWS, MAX_LINES = 4, 10, 170


def demangle(names):
    r = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True)
    return r.stdout.splitlines()


def main():
    os.makedirs(OUT, exist_ok=True)
    txt = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    funcs, cur = {}, None
    for line in txt.splitlines():
        s = line.strip()
        if s.startswith("Function :"):
            cur = s.split(":", 1)[1].strip()
            funcs[cur] = []
        elif cur and s.startswith("/*") and "*/" in s:
            body = s.split("*/", 1)[1].strip()
            if body and not body.startswith("/*"):
                funcs[cur].append(re.sub(r"\s*/\*.*$", "", body).rstrip(" ;") + " ;")
    names = list(funcs)
    pretty = dict(zip(names, demangle(names)))
    written = {}
    for mangled, ins in funcs.items():
        name = pretty[mangled]
        base = next((w for w in WANT if w in name), None)
        if base is None:
            continue
        tmpl = re.search(base + r"<([^>]*)>", name)
        tag = base + ("_" + re.sub(r"[^0-9A-Za-z]+", "_", tmpl.group(1)).strip("_") if tmpl else "")
        hits = [i for i, l in enumerate(ins) if any(k in l for k in KEYS)]
        if not hits:
            continue
        counts = {}
        for l in ins:
            for k in KEYS:
                if k in l:
                    counts[k] = counts.get(k, 0) + 1
        windows, last_end = [], -1
        for i in hits:
            lo, hi = max(i - CONTEXT, 0), min(i + CONTEXT + 1, len(ins))
            if lo <= last_end and windows:
                windows[-1][1] = hi
            else:
                windows.append([lo, hi])
            last_end = hi
        # prefer windows that contain the tensor-core / TMEM instructions, keep program order
        def score(w):
            seg = ins[w[0]:w[1]]
            return sum(3 for l in seg if "UTC" in l and "MMA" in l) + sum(2 for l in seg if "LDTM" in l or "MULTIMEM" in l) + \
                sum(1 for l in seg if "UBLKCP" in l or "UTMA" in l or "DFMA" in l)
        keep = sorted(sorted(windows, key=score, reverse=True)[:MAX_WINDOWS])
        lines = [f"// {name}", f"// {len(ins)} SASS instructions; mnemonic counts: " +
                 ", ".join(f"{k} {v}" for k, v in sorted(counts.items())),
                 "// excerpts (instruction index: text), '...' marks skipped instructions", ""]
        n = 0
        for lo, hi in keep:
            lines.append("        ...")
            for i in range(lo, hi):
                lines.append(f"{i:6d}: {ins[i]}")
                n += 1
            if n > MAX_LINES:
                break
        path = os.path.join(OUT, tag[:80] + ".sass")
        if tag in written and written[tag] >= len(ins):
            continue
        written[tag] = len(ins)
        with open(path, "w") as fh:
            fh.write("\n".join(lines) + "\n")
    print(f"wrote {len(written)} excerpts to {OUT}")


if __name__ == "__main__":
    sys.exit(main())
