"""Per-step DRAM bytes from ncu launch lists of `bench.py --no-extras` (tools/gpu_l2_traffic.sh): the last two whole
steps (57 launches each for ResNet-50: quantise, 53 convs, 2 pools, fc), per kernel family."""
import csv
import sys
from collections import defaultdict

PER_STEP = 57


def load(path):
    lines = [ln for ln in open(path, newline="") if ln.startswith('"')]
    per_id, names, order = defaultdict(dict), {}, []
    for r in csv.DictReader(lines):
        if r["ID"] not in names:
            order.append(r["ID"])
        names[r["ID"]] = r["Kernel Name"]
        per_id[r["ID"]][r["Metric Name"]] = float(r["Metric Value"].replace(",", ""))
    return [(names[i], per_id[i]) for i in order]


for path in sys.argv[1:]:
    k = load(path)
    # the bench's steps are the launches whose family sequence repeats; take the last 2 * PER_STEP launches of the
    # conv-net kernels (the tail of the run is the timed graph replay)
    net = [(n, m) for n, m in k if any(t in n for t in ("igemm_i8", "pool", "quant", "gemv", "dw3x3"))]
    tail = net[-2 * PER_STEP:]
    rd = sum(m.get("dram__bytes_read.sum", 0.0) for _, m in tail) / 2
    wr = sum(m.get("dram__bytes_write.sum", 0.0) for _, m in tail) / 2
    us = sum(m.get("gpu__time_duration.sum", 0.0) for _, m in tail) / 2 / 1e3
    print(f"{path}: {len(k)} launches profiled; per step: DRAM read {rd / 1e9:.3f} GB, write {wr / 1e9:.3f} GB, "
          f"sum of serialised kernel times {us:.0f} us")
