"""One line per kernel launch of a .ncu-rep raw csv page: duration, DRAM bytes and % of peak, L2 %, tensor-pipe
active share, issue-slot use.   usage: ncu_brief.py report.ncu-rep  (or a csv made by `ncu -i rep --page raw --csv`)"""
import csv
import io
import subprocess
import sys

COLS = [("gpu__time_duration.sum", "us"), ("dram__bytes_read.sum", "rdMB"), ("dram__bytes_write.sum", "wrMB"),
        ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
        ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "l2%"),
        ("lts__t_sector_hit_rate.pct", "l2hit%"),
        ("l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "l1%"),
        ("sm__cycles_elapsed.max", "cyc"),
        ("TPC.TriageCompute.sm__pipe_tensor_subpipe_imma_cycles_active_realtime.avg", "imma_cyc"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%"),
        ("smsp__inst_executed.sum", "inst")]


def main():
    src = sys.argv[1]
    if src.endswith(".csv"):
        text = open(src).read()
    else:
        text = subprocess.run(["ncu", "-i", src, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(text)))
    hi = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    hdr, units, data = rows[hi], rows[hi + 1], rows[hi + 2:]
    col = {h: i for i, h in enumerate(hdr)}
    print("| # | kernel | grid | " + " | ".join(n for _, n in COLS) + " | tensor% |")
    print("|---|---|---|" + "---:|" * (len(COLS) + 1))
    for k, r in enumerate(data):
        if len(r) < len(hdr):
            continue
        vals = []
        for c, _ in COLS:
            v = r[col[c]] if c in col else ""
            try:
                f = float(v.replace(",", ""))
                if units[col[c]] == "byte":
                    f /= 1e6
                elif units[col[c]] == "Kbyte":
                    f /= 1e3
                elif units[col[c]] == "Gbyte":
                    f *= 1e3
                elif units[col[c]] in ("ns", "nsecond"):
                    f /= 1e3
                elif units[col[c]] in ("ms", "msecond"):
                    f *= 1e3
                v = f"{f:.1f}" if f < 1e6 else f"{f:.3g}"
            except ValueError:
                pass
            vals.append(v)
        try:
            tp = 100.0 * float(r[col[COLS[8][0]]].replace(",", "")) / float(r[col[COLS[7][0]]].replace(",", ""))
            tps = f"{tp:.1f}"
        except Exception:
            tps = ""
        name = r[col["Kernel Name"]].replace("void ", "").replace("pl::", "")[:34]
        print(f"| {k} | `{name}` | {r[col['Grid Size']]} | " + " | ".join(vals) + f" | {tps} |")


if __name__ == "__main__":
    main()
