"""profiles/traffic.json from an ncu launch list (csv with gpu__time_duration.sum, dram__bytes_read.sum,
dram__bytes_write.sum per launch): per-kernel average DRAM bytes per launch, read by bench.py for roofline.traffic.

    python tools/make_traffic.py profiles/r01_v6_launches.csv profiles/traffic.json
"""
import csv
import json
import re
import sys

UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1.0, "ms": 1e3}


def short(name: str) -> str:
    """'void conv_umma_kernel<256, 128, false>(…)' -> 'conv_umma_kernel<256>' (the names bench.py's per-kernel table uses)."""
    m = re.search(r"(\w+_kernel)(?:<([^>(]*)>)?", name)
    if not m:
        return name.split("(")[0]
    base, targs = m.group(1), m.group(2)
    if base.startswith("conv_umma") and targs:
        return f"{base}<{targs.split(',')[0].strip()}>"
    return base


def main(src, dst):
    """ncu --csv without --page raw: one row per (launch ID, metric)."""
    rows = [r for r in csv.reader(l for l in open(src) if l.startswith('"'))]
    col = {n: i for i, n in enumerate(rows[0])}
    per_launch = {}
    for r in rows[1:]:
        e = per_launch.setdefault(r[col["ID"]], {"k": short(r[col["Kernel Name"]]), "bytes": 0.0, "us": 0.0, "tensor": 0.0})
        name, unit = r[col["Metric Name"]], r[col["Metric Unit"]]
        val = float(r[col["Metric Value"]].replace(",", "") or 0)
        if name in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            e["bytes"] += val * UNIT[unit]
        elif name == "gpu__time_duration.sum":
            e["us"] += val * UNIT[unit]
        elif name.startswith("sm__pipe_tensor"):
            e["tensor"] = val
    out = {}
    for e in per_launch.values():
        o = out.setdefault(e["k"], {"launches": 0, "bytes": 0.0, "us": 0.0, "tensor_pct": 0.0})
        o["launches"] += 1
        o["bytes"] += e["bytes"]
        o["us"] += e["us"]
        o["tensor_pct"] += e["tensor"]
    res = {k: {"launches": e["launches"], "dram_bytes_per_launch": e["bytes"] / e["launches"],
               "avg_us_under_ncu": e["us"] / e["launches"], "tensor_pipe_active_pct": e["tensor_pct"] / e["launches"]}
           for k, e in out.items()}
    json.dump(res, open(dst, "w"), indent=1)
    tot = sum(e["us"] for e in out.values())
    for k, e in sorted(out.items(), key=lambda kv: -kv[1]["us"]):
        print(f"{k:32s} launches {e['launches']:4d}  share {100 * e['us'] / tot:5.1f}%  avg {e['us'] / e['launches']:8.1f} us  "
              f"DRAM {e['bytes'] / e['launches'] / 1e6:8.1f} MB  tensor {e['tensor_pct'] / e['launches']:5.1f}%")


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
