"""Summarise an ncu launch list (--metrics gpu__time_duration.sum[,dram__bytes_read.sum,dram__bytes_write.sum] --csv):
time share (and DRAM bytes) per kernel name.

    python tools/launch_summary.py gpurun_out/x_launches.csv [out.txt]
"""
import csv
import re
import sys
from collections import OrderedDict

UNIT = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6, "byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def main():
    with open(sys.argv[1]) as f:
        lines = [l for l in f if l.startswith('"')]
    rd = csv.reader(lines)
    hdr = next(rd)
    ki, kn, mn, mv, mu = (hdr.index(k) for k in ("ID", "Kernel Name", "Metric Name", "Metric Value", "Metric Unit"))
    agg = OrderedDict()
    tot = {"t": 0.0, "r": 0.0, "w": 0.0}
    ids = set()
    for r in rd:
        v = float(r[mv].replace(",", "")) * UNIT.get(r[mu], 1.0)
        name = re.sub(r"\(.*", "", r[kn]).replace("srn::", "").replace("__nv_bfloat16", "bf16").replace("void ", "")
        a = agg.setdefault(name, {"n": set(), "t": 0.0, "r": 0.0, "w": 0.0})
        a["n"].add(r[ki])
        ids.add(r[ki])
        key = "t" if "time_duration" in r[mn] else ("r" if "bytes_read" in r[mn] else "w")
        a[key] += v
        tot[key] += v
    out = [f"# {sys.argv[1]}: {len(ids)} launches, {tot['t'] / 1e3:.3f} ms of kernel time, DRAM read {tot['r'] / 1e9:.3f} GB, "
           f"write {tot['w'] / 1e9:.3f} GB (ncu per-launch durations are cold-cache and serialised: compare SHARES, not absolutes)"]
    for name, a in sorted(agg.items(), key=lambda kv: -kv[1]["t"]):
        n = len(a["n"])
        out.append(f"{name[:58]:58s} x{n:5d} {a['t'] / 1e3:9.3f} ms {100 * a['t'] / max(tot['t'], 1e-9):5.1f}%  avg {a['t'] / n:7.1f} us"
                   f"  dram R {a['r'] / 1e6:9.1f} MB  W {a['w'] / 1e6:9.1f} MB")
    text = "\n".join(out)
    print(text)
    if len(sys.argv) > 2:
        open(sys.argv[2], "w").write(text + "\n")


if __name__ == "__main__":
    main()
