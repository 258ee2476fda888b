"""Key metrics of every kernel in a .ncu-rep as a markdown table (reads `ncu -i … --page raw --csv`)."""
import csv
import io
import subprocess
import sys

KEYS = [
    ("gpu__time_duration.sum", "time"),
    ("dram__bytes_read.sum", "dram rd"),
    ("dram__bytes_write.sum", "dram wr"),
    ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram %"),
    ("lts__t_bytes.sum", "L2 bytes"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM %"),
    ("sm__inst_executed_pipe_tensor_op_gen5.sum", "tcgen05 inst"),
    ("sm__pipe_tensor_subpipe_gen5_cycles_active.avg.pct_of_peak_sustained_active", "tensor pipe %"),
    ("sm__pipe_tensor_op_gen5_cycles_active.avg.pct_of_peak_sustained_active", "tensor pipe %"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue %"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "occupancy %"),
    ("launch__registers_per_thread", "regs"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
]


def main():
    rep = sys.argv[1]
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr_i = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    hdr, units, data = rows[hdr_i], rows[hdr_i + 1], rows[hdr_i + 2:]
    col = {h: i for i, h in enumerate(hdr)}
    tensor_cols = [h for h in hdr if ("imma" in h or "pipe_tensor_cycles_active.avg" in h or "mem_tensor_cycles_active.avg" in h)]
    print(f"## {rep.split('/')[-1]}\n")
    for r in data:
        if len(r) < len(hdr):
            continue
        print(f"### `{r[col['Kernel Name']][:90]}`  grid {r[col.get('Grid Size', 0)]} block {r[col.get('Block Size', 0)]}\n")
        print("| metric | value | unit |\n|---|---:|---|")
        seen = set()
        for k, _ in KEYS:
            if k in col and k not in seen:
                seen.add(k)
                print(f"| {k} | {r[col[k]]} | {units[col[k]]} |")
        for k in tensor_cols:
            if k not in seen:
                seen.add(k)
                print(f"| {k} | {r[col[k]]} | {units[col[k]]} |")
        print()


if __name__ == "__main__":
    main()
