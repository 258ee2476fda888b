"""Top stall locations from `ncu --page source --csv` output (helper for reading profiles here)."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 30
hdr = None
data = []
for r in rows:
    if r and r[0] == "Address":
        hdr = r
        continue
    if hdr and len(r) == len(hdr):
        data.append(r)
iS, iSrc, iAddr = hdr.index("# Samples"), hdr.index("Source"), hdr.index("Address")
iEx = hdr.index("Instructions Executed")


def num(x):
    try:
        return int(float(x))
    except ValueError:
        return 0


tot = sum(num(r[iS]) for r in data)
print("total samples", tot, "rows", len(data))
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
agg = {h: sum(num(r[hdr.index(h)]) for r in data) for h in stalls}
print("stall totals:", dict(sorted(agg.items(), key=lambda kv: -kv[1])[:8]))
for r in sorted(data, key=lambda r: -num(r[iS]))[:n]:
    st = {h: num(r[hdr.index(h)]) for h in stalls}
    st = {k: v for k, v in st.items() if v > 0}
    print(r[iAddr][-5:], r[iS].rjust(6), r[iEx].rjust(9), r[iSrc][:100], dict(sorted(st.items(), key=lambda kv: -kv[1])[:3]))
