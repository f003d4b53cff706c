"""Turn `ncu --page raw --csv` output (one row per profiled launch) into the per-kernel roofline table kept under
profiles/: launches of the same kernel are merged, and each kernel is placed against the MEASURED peaks
(MEASURED_PEAKS.json: HBM copy bandwidth, cuBLAS bf16 throughput).

    python bench/ncu_table.py gpurun_out/step_raw.csv "Transformer-MT step" > profiles/ncu_transformer_step.md
"""
import csv
import json
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def fnum(v):
    try:
        return float(v.replace(",", ""))
    except (ValueError, AttributeError):
        return 0.0


def short(name):
    name = re.sub(r"\(.*", "", name)
    name = name.replace("void ", "").replace("b200::", "")
    return name[:70]


def main(path, title):
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
    hbm = peaks.get("hbm_gbs", 6650.0)
    rows = list(csv.reader(open(path, newline="")))
    # skip ncu banner lines before the header
    start = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    hdr, units = rows[start], rows[start + 1]
    col = {k: i for i, k in enumerate(hdr)}

    def get(r, k):
        return r[col[k]] if k in col and col[k] < len(r) else ""

    def unit(k):
        return units[col[k]] if k in col else ""

    agg = {}
    for r in rows[start + 2:]:
        if len(r) < len(hdr) // 2:
            continue
        name = short(get(r, "Kernel Name"))
        t = fnum(get(r, "gpu__time_duration.sum"))
        tu = unit("gpu__time_duration.sum")
        t_us = t / 1e3 if tu in ("ns", "nsecond") else (t if tu in ("us", "usecond") else t * 1e3 if tu in ("ms", "msecond") else t)
        rd, wr = fnum(get(r, "dram__bytes_read.sum")), fnum(get(r, "dram__bytes_write.sum"))
        scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        rd *= scale.get(unit("dram__bytes_read.sum"), 1)
        wr *= scale.get(unit("dram__bytes_write.sum"), 1)
        a = agg.setdefault(name, {"n": 0, "us": 0.0, "rd": 0.0, "wr": 0.0, "tensor": [], "dram": [], "regs": "", "grid": "",
                                  "smem": "", "warps": []})
        a["n"] += 1
        a["us"] += t_us
        a["rd"] += rd
        a["wr"] += wr
        a["tensor"].append(fnum(get(r, "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active")))
        a["dram"].append(fnum(get(r, "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed")))
        a["warps"].append(fnum(get(r, "sm__warps_active.avg.pct_of_peak_sustained_active")))
        a["regs"] = get(r, "launch__registers_per_thread")
        a["grid"] = get(r, "launch__grid_size")
        a["smem"] = get(r, "launch__shared_mem_per_block_dynamic")
    print(f"# ncu --set full, {title}\n")
    print(f"One profiled step (eager, kernels serialised, side streams and PDL off; ncu replays each launch, caches are "
          f"cold for the first pass: compare shares, not absolutes).  `HBM %` = (dram read + write) / time against the "
          f"measured copy bandwidth of {hbm:.0f} GB/s (MEASURED_PEAKS.json); `tensor %` = "
          f"sm__pipe_tensor_cycles_active of peak; `dram %` = gpu__dram_throughput of peak as ncu reports it.\n")
    print("| kernel | launches | us (sum) | DRAM read MB | DRAM write MB | HBM % of measured | dram % (ncu) | tensor % | warps active % | regs | grid (last) | dyn smem (last) |")
    print("|---|---:|---:|---:|---:|---:|---:|---:|---:|---:|---:|---:|")
    tot = 0.0
    for name, a in sorted(agg.items(), key=lambda kv: -kv[1]["us"]):
        tot += a["us"]
        gbs = (a["rd"] + a["wr"]) / (a["us"] * 1e-6) / 1e9 if a["us"] > 0 else 0.0
        avg = lambda v: sum(v) / max(len(v), 1)  # noqa: E731
        print(f"| `{name}` | {a['n']} | {a['us']:.1f} | {a['rd'] / 1e6:.1f} | {a['wr'] / 1e6:.1f} | {100 * gbs / hbm:.0f} | "
              f"{avg(a['dram']):.0f} | {avg(a['tensor']):.0f} | {avg(a['warps']):.0f} | {a['regs']} | {a['grid']} | {a['smem']} |")
    print(f"\nsum of kernel time: {tot:.0f} us")


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else "step")
