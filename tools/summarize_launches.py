"""Aggregate an `ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --csv`
launch list into a per-kernel share table (markdown).  usage: summarize_launches.py launches.csv [title]"""
import csv
import re
import sys
from collections import defaultdict


def short(name):
    name = re.sub(r"^void ", "", name)
    name = re.sub(r"\(.*$", "", name)
    name = name.replace("paddle_lite_b200::", "").replace("(anonymous namespace)::", "")
    return name[:70]


def main():
    path = sys.argv[1]
    title = sys.argv[2] if len(sys.argv) > 2 else path
    rows = []
    with open(path, newline="") as f:
        lines = [ln for ln in f if ln.startswith('"')]
    rd = csv.DictReader(lines)
    per_id = defaultdict(dict)
    names = {}
    for r in rd:
        per_id[r["ID"]][r["Metric Name"]] = (float(r["Metric Value"].replace(",", "")), r["Metric Unit"])
        names[r["ID"]] = r["Kernel Name"]
    agg = defaultdict(lambda: [0, 0.0, 0.0, 0.0])
    for i, m in per_id.items():
        t, unit = m.get("gpu__time_duration.sum", (0.0, "ns"))
        scale = {"ns": 1e-3, "us": 1.0, "usecond": 1.0, "nsecond": 1e-3, "ms": 1e3, "msecond": 1e3}.get(unit, 1e-3)
        a = agg[short(names[i])]
        a[0] += 1
        a[1] += t * scale
        a[2] += m.get("dram__bytes_read.sum", (0.0, ""))[0]
        a[3] += m.get("dram__bytes_write.sum", (0.0, ""))[0]
    total = sum(a[1] for a in agg.values())
    print(f"## {title}\n")
    print(f"{len(per_id)} launches, {total / 1e3:.3f} ms total (ncu-serialised, cold cache: shares matter, not absolutes)\n")
    print("| kernel | launches | total us | share | avg us | dram read MB | dram write MB |")
    print("|---|---:|---:|---:|---:|---:|---:|")
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"| `{k}` | {a[0]} | {a[1]:.1f} | {100 * a[1] / total:.1f}% | {a[1] / a[0]:.1f} | {a[2] / 1e6:.1f} | {a[3] / 1e6:.1f} |")


if __name__ == "__main__":
    main()
