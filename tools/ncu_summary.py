"""Summarise an `ncu --page raw --csv` export: one line per profiled launch with the numbers the roofline uses.

    python tools/ncu_summary.py gpurun_out/x_raw.csv [out.txt]
"""
import csv
import sys

COLS = [
    ("dur_us", "gpu__time_duration.sum", 1e-3),
    ("tensor%", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", 1),
    ("dramR_MB", "dram__bytes_read.sum", None),
    ("dramW_MB", "dram__bytes_write.sum", None),
    ("dram%", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", 1),
    ("lts%", "lts__throughput.avg.pct_of_peak_sustained_elapsed", 1),
    ("sm%", "sm__throughput.avg.pct_of_peak_sustained_elapsed", 1),
    ("warps%", "sm__warps_active.avg.pct_of_peak_sustained_active", 1),
    ("regs", "launch__registers_per_thread", 1),
    ("grid", "launch__grid_size", 1),
]


def find(hdr, name):
    for i, h in enumerate(hdr):
        if h == name:
            return i
    for i, h in enumerate(hdr):
        if h.endswith("." + name):
            return i
    return None


def to_bytes(v, unit):
    v = float(v.replace(",", ""))
    u = unit.lower()
    return v * {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}.get(u, 1)


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    hdr, units, data = rows[0], rows[1], rows[2:]
    kn = find(hdr, "Kernel Name")
    idx = {label: find(hdr, name) for label, name, _ in COLS}
    lines = [f"# {sys.argv[1]}: {len(data)} profiled launches (ncu --set full, --clock-control none; cold-cache, serialised)"]
    lines.append(f"{'kernel':58s} " + " ".join(f"{l:>9s}" for l, _, _ in COLS))
    for r in data:
        name = r[kn]
        name = name.replace("srn::", "").replace("__nv_bfloat16", "bf16")
        out = []
        for label, _, scale in COLS:
            i = idx[label]
            if i is None or r[i] == "":
                out.append(f"{'-':>9s}")
                continue
            if scale is None:
                out.append(f"{to_bytes(r[i], units[i]) / 1e6:9.1f}")
            else:
                v = float(r[i].replace(",", ""))
                if label == "dur_us":
                    u = units[i].lower()
                    v = v * {"ns": 1e-3, "us": 1, "ms": 1e3, "s": 1e6}.get(u, 1e-3)
                out.append(f"{v:9.1f}")
        lines.append(f"{name[:58]:58s} " + " ".join(out))
    text = "\n".join(lines)
    print(text)
    if len(sys.argv) > 2:
        open(sys.argv[2], "w").write(text + "\n")


if __name__ == "__main__":
    main()
