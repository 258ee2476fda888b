"""One markdown table of key ncu metrics, one column per selected launch of a .ncu-rep.
usage: ncu_layers_table.py rep.ncu-rep idx:label [idx:label ...]"""
import csv
import io
import subprocess
import sys

KEYS = [
    ("gpu__time_duration.sum", "duration"),
    ("sm__cycles_elapsed.max", "SM cycles"),
    ("dram__bytes_read.sum", "DRAM read"),
    ("dram__bytes_write.sum", "DRAM write"),
    ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "DRAM throughput % of peak"),
    ("lts__t_bytes.sum", "L2 bytes"),
    ("lts__t_sector_hit_rate.pct", "L2 hit rate"),
    ("sm__inst_executed_pipe_tensor_op_gen5.sum", "tcgen05 instructions"),
    ("sm__pipe_tensor_subpipe_gen5_cycles_active.avg.pct_of_peak_sustained_active", "tensor pipe active %"),
    ("TPC.TriageCompute.sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed", "tensor pipe active % (realtime)"),
    ("sm__mem_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "TMEM/tensor-memory active %"),
    ("smsp__inst_executed.sum", "warp instructions"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue slots busy %"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved occupancy %"),
    ("launch__registers_per_thread", "registers / thread"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
    ("l1tex__data_bank_conflicts_pipe_lsu.sum", "smem bank conflicts (LSU)"),
]


def main():
    rep = sys.argv[1]
    sel = [(int(a.split(":")[0]), a.split(":", 1)[1]) for a in sys.argv[2:]]
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr_i = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    hdr, units, data = rows[hdr_i], rows[hdr_i + 1], rows[hdr_i + 2:]
    col = {h: i for i, h in enumerate(hdr)}
    print("| metric | unit | " + " | ".join(lbl for _, lbl in sel) + " |")
    print("|---|---|" + "---:|" * len(sel))
    print("| kernel | | " + " | ".join("`" + data[i][col["Kernel Name"]].split("(")[0].replace("void ", "") + "`" for i, _ in sel) + " |")
    for k, name in KEYS:
        if k in col:
            print(f"| {name} | {units[col[k]]} | " + " | ".join(data[i][col[k]] for i, _ in sel) + " |")
    # top stall reasons per launch (warp-state samples)
    stall = [h for h in hdr if h.startswith("smsp__pcsamp_warps_issue_stalled_") and not h.endswith("_not_issued")]
    if stall:
        print("\nTop warp-stall reasons (pc samples, share of all samples):\n")
        for i, lbl in sel:
            vals = []
            for h in stall:
                try:
                    vals.append((float(data[i][col[h]].replace(",", "")), h.replace("smsp__pcsamp_warps_issue_stalled_", "")))
                except ValueError:
                    pass
            tot = sum(v for v, _ in vals) or 1.0
            top = sorted(vals, reverse=True)[:5]
            print(f"* {lbl}: " + ", ".join(f"{n} {100 * v / tot:.0f}%" for v, n in top))


if __name__ == "__main__":
    main()
