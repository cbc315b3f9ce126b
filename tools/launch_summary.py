"""Summarise an ncu launch list (--metrics gpu__time_duration.sum --csv): time share per kernel name.

    python tools/launch_summary.py gpurun_out/x_launches.csv [out.txt]
"""
import csv
import re
import sys
from collections import OrderedDict


def main():
    rows = []
    with open(sys.argv[1]) as f:
        lines = [l for l in f if l.startswith('"')]
    rd = csv.reader(lines)
    hdr = next(rd)
    kn, mv, mu = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg = OrderedDict()
    total = 0.0
    n = 0
    for r in rd:
        v = float(r[mv].replace(",", ""))
        us = v * {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(r[mu], 1e-3)
        name = re.sub(r"\(.*", "", r[kn]).replace("srn::", "").replace("__nv_bfloat16", "bf16").replace("void ", "")
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += us
        total += us
        n += 1
    out = [f"# {sys.argv[1]}: {n} launches, {total / 1e3:.3f} ms of kernel time "
           "(ncu per-launch durations are cold-cache and serialised: compare SHARES, not absolutes)"]
    for name, (cnt, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        out.append(f"{name:60s} x{cnt:5d} {us / 1e3:9.3f} ms {100 * us / total:5.1f}%  avg {us / cnt:7.1f} us")
    text = "\n".join(out)
    print(text)
    if len(sys.argv) > 2:
        open(sys.argv[2], "w").write(text + "\n")


if __name__ == "__main__":
    main()
