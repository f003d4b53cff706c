"""Reads an `ncu --page source --csv` export (one or more kernels) and prints, per kernel, the stall-reason totals and the
instructions with the most warp-stall samples together with their two main stall reasons.

    ncu -i rep.ncu-rep --page source --csv > src.csv ; python bench/ncu_source_top.py src.csv [N]
"""
import csv
import sys


def main():
    path = sys.argv[1]
    top_n = int(sys.argv[2]) if len(sys.argv) > 2 else 25
    rows = list(csv.reader(open(path)))
    i = 0
    while i < len(rows):
        if rows[i] and rows[i][0] == "Kernel Name":
            name = rows[i][1].split("(")[0]
            hdr = rows[i + 1]
            ix = {n: k for k, n in enumerate(hdr)}
            j = i + 2
            data = []
            while j < len(rows) and not (rows[j] and rows[j][0] == "Kernel Name"):
                if len(rows[j]) == len(hdr):
                    data.append(rows[j])
                j += 1
            report(name, hdr, ix, data, top_n)
            i = j
        else:
            i += 1


def report(name, hdr, ix, data, top_n):
    stalls = [n for n in hdr if n.startswith("stall_") and "Not Issued" not in n]
    total = sum(int(r[ix["# Samples"]] or 0) for r in data)
    tot = {n: sum(int(r[ix[n]] or 0) for r in data) for n in stalls}
    print(f"== {name}: {total} samples")
    print("   " + ", ".join(f"{n[6:]} {v}" for n, v in sorted(tot.items(), key=lambda x: -x[1])[:7]))
    top = sorted(data, key=lambda r: -int(r[ix["# Samples"]] or 0))[:top_n]
    for r in sorted(top, key=lambda r: int(r[0], 16) if r[0].startswith("0x") else int(r[0])):
        st = sorted(((n[6:], int(r[ix[n]] or 0)) for n in stalls), key=lambda x: -x[1])[:2]
        print(f"   {r[0][-5:]} {r[ix['# Samples']]:>5} {r[ix['Instructions Executed']]:>7}  {r[1][:64]:<64} {st}")


if __name__ == "__main__":
    main()
