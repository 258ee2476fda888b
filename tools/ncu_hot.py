"""Top stall lines of one kernel launch from `ncu --page source --csv` (SASS view).
usage: ncu_hot.py report.ncu-rep launch_index [top_n]"""
import csv
import subprocess
import sys

rep, idx = sys.argv[1], int(sys.argv[2])
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--launch-skip", str(idx), "--launch-count", "1"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
col = {h: i for i, h in enumerate(hdr)}
data = [r for r in rows[hi + 1:] if len(r) >= len(hdr) and r[0] != "Address"]
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
tot = sum(int(r[col["# Samples"]] or 0) for r in data)
print(rows[0][1][:100], "total samples", tot)
agg = {h: 0 for h in stalls}
for r in data:
    for h in stalls:
        agg[h] += int(r[col[h]] or 0)
print("stall totals:", {k: v for k, v in sorted(agg.items(), key=lambda kv: -kv[1]) if v > tot * 0.01})
ranked = sorted(enumerate(data), key=lambda kv: -int(kv[1][col["# Samples"]] or 0))[:top]
for i, r in sorted(ranked):
    n = int(r[col["# Samples"]] or 0)
    why = sorted(((int(r[col[h]] or 0), h[6:]) for h in stalls), reverse=True)[:2]
    print(f"{i:5d} {100.0 * n / tot:5.1f}%  {r[col['Source']][:90]:90s} {why[0][1]}:{why[0][0]} {why[1][1]}:{why[1][0]}")
