"""Top stalled SASS instructions of one kernel from `ncu -i x.ncu-rep --page source --csv ...` output.

    ncu -i rep --page source --csv --kernel-name regex:NAME --launch-skip N --launch-count 1 > src.csv
    python tools/ncu_stalls.py src.csv [top]
"""
import csv
import sys


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    top_n = int(sys.argv[2]) if len(sys.argv) > 2 else 40
    hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[hi]
    print(rows[0][1] if rows[0] and rows[0][0] == "Kernel Name" else "")
    data = [r for r in rows[hi + 1:] if len(r) == len(hdr) and r[0] != "Address"]
    si, src, ie = hdr.index("# Samples"), hdr.index("Source"), hdr.index("Instructions Executed")
    stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
    tot = sum(int(r[si]) for r in data)
    agg = {}
    for r in data:
        for i in stall_cols:
            agg[hdr[i]] = agg.get(hdr[i], 0) + int(r[i])
    print("total samples", tot, " by reason:", ", ".join(f"{k[6:]}={100 * v / max(tot, 1):.0f}%" for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]))
    for r in sorted(data, key=lambda r: -int(r[si]))[:top_n]:
        st = sorted(((int(r[i]), hdr[i][6:]) for i in stall_cols), reverse=True)[:2]
        print(f"{int(r[si]):6d} {100 * int(r[si]) / max(tot, 1):5.1f}% exec={r[ie]:>8s} {r[src].strip()[:66]:66s} {st}")


if __name__ == "__main__":
    main()
