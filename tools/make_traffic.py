"""Turn an ncu launch list (gpu__time_duration + dram bytes, csv) of `bench.py` into
  profiles/traffic.json   { "<net>_b<batch>": {"igemm_dram_bytes_per_launch": ..., ...} }
which bench.py reads for roofline.traffic.   usage: make_traffic.py launches.csv net batch igemm_launches_per_step"""
import csv
import json
import os
import sys
from collections import defaultdict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main():
    path, net, batch, per_step = sys.argv[1], sys.argv[2], int(sys.argv[3]), int(sys.argv[4])
    with open(path, newline="") as f:
        lines = [ln for ln in f if ln.startswith('"')]
    per_id = defaultdict(dict)
    names = {}
    order = []
    for r in csv.DictReader(lines):
        if r["ID"] not in names:
            order.append(r["ID"])
        names[r["ID"]] = r["Kernel Name"]
        per_id[r["ID"]][r["Metric Name"]] = float(r["Metric Value"].replace(",", ""))
    ig = [i for i in order if "igemm_i8_kernel" in names[i]]
    whole = len(ig) // per_step * per_step
    ig = ig[len(ig) - whole:]  # whole steps only, the last ones (steady state)
    rd = sum(per_id[i].get("dram__bytes_read.sum", 0.0) for i in ig)
    wr = sum(per_id[i].get("dram__bytes_write.sum", 0.0) for i in ig)
    out_path = os.path.join(ROOT, "profiles", "traffic.json")
    tj = json.load(open(out_path)) if os.path.exists(out_path) else {}
    tj[f"{net}_b{batch}"] = {
        "igemm_dram_bytes_per_launch": (rd + wr) / max(len(ig), 1),
        "igemm_dram_read_bytes_per_step": rd / max(whole // per_step, 1),
        "igemm_dram_write_bytes_per_step": wr / max(whole // per_step, 1),
        "igemm_launches_counted": len(ig), "steps_counted": whole // per_step,
        "source": os.path.basename(path),
    }
    json.dump(tj, open(out_path, "w"), indent=1, sort_keys=True)
    print(json.dumps(tj[f"{net}_b{batch}"], indent=1))


if __name__ == "__main__":
    main()
