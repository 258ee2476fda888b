"""Per-layer table of a bench.py --breakdown file against both roofs.  usage: show_breakdown.py bd.json [min_ms]"""
import json
import sys

HBM, TENSOR = 6532.9, 4500.0  # GB/s (MEASURED_PEAKS.json), int8 TOP/s (tools/mma_probe.py)
t = [r for r in json.load(open(sys.argv[1])) if "ms" in r]  # (the last entry describes the timing source)
min_ms = float(sys.argv[2]) if len(sys.argv) > 2 else 0.0
tot = sum(r["ms"] for r in t)
fl_tot = 0.0
groups = {}
for r in t:
    th = r["mbytes"] / HBM * 1e-3 * 1e3 / 1e3 if r["mbytes"] else 0.0
    th = r["mbytes"] / (HBM * 1e3) * 1e3 / 1e3 * 1e3 / 1e3 if False else r["mbytes"] * 1e6 / (HBM * 1e9) * 1e3
    tt = r["gops"] * 1e9 / (TENSOR * 1e12) * 1e3 if r["gops"] else 0.0
    fl = max(th, tt)
    fl_tot += fl
    key = (r["kind"], tuple(r["in"]), tuple(r["out"]), r["k"], r["stride"])
    g = groups.setdefault(key, [0, 0.0, 0.0, r])
    g[0] += 1
    g[1] += r["ms"]
    g[2] += fl
print(f"total eager {tot:.4f} ms, sum of per-layer floors {fl_tot:.4f} ms ({fl_tot / tot:.2f})")
for key, (n, ms, fl, r) in sorted(groups.items(), key=lambda kv: -kv[1][1]):
    if ms < min_ms:
        continue
    print(f"{n:2d}x {r['kind']:8s} in={r['in']} out={r['out']} k={r['k']} s={r['stride']}: {ms:.4f} ms total "
          f"({ms / n * 1e3:.1f} us each, {100 * ms / tot:.1f}%)  floor {fl / n * 1e3:.1f} us  frac {fl / ms:.2f}  "
          f"tops {r['tops'] or 0:.0f} gbs {r['gbs'] or 0:.0f}")
