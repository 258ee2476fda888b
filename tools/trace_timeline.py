"""Print the PL_DBG_TRACE timeline of CTA 0 (igemm_i8_kernel) as per-tile deltas.
usage: trace_timeline.py trace.txt [first_tile] [n_tiles]"""
import sys

NAMES = {
    0: ["top", "wait-empty(A)", "loads issued", "wait-rempty/last-empty"],
    1: ["top", "wait-tempty", "wait-full(first)", "wait-full(last)", "commits", "wait-full(2nd)"],
    2: ["top", "wait-tfull", "body done", "arrived", "pre-bar", "post-bar", "end"],
    3: ["top", "wait-tfull", "body done", "arrived", "pre-bar", "post-bar", "end"],
}
ROLE = {0: "producer", 1: "mma", 2: "epi(first warp)", 3: "epi(last warp)"}


def main():
    path = sys.argv[1]
    t0 = int(sys.argv[2]) if len(sys.argv) > 2 else 8
    nt = int(sys.argv[3]) if len(sys.argv) > 3 else 6
    rows = {}
    for ln in open(path):
        if ln.startswith("#"):
            print(ln.strip())
            continue
        v = [int(x) for x in ln.split()]
        rows[(v[0], v[1])] = v[2:]
    base = min(x for v in rows.values() for x in v if x > 0)
    for r in range(4):
        print(f"--- {ROLE[r]}")
        prev_top = None
        for t in range(t0, t0 + nt):
            ev = rows.get((r, t))
            if not ev or ev[0] == 0:
                continue
            s = f"tile {t:3d} top@{ev[0] - base:8d}"
            if prev_top is not None:
                s += f" (period {ev[0] - prev_top:5d})"
            prev_top = ev[0]
            for i, nm in enumerate(NAMES[r]):
                if i and ev[i]:
                    s += f" | {nm} +{ev[i] - ev[0]}"
            print(s)
    # steady-state period per role
    for r in range(4):
        tops = [rows[(r, t)][0] for t in range(64) if rows.get((r, t)) and rows[(r, t)][0]]
        if len(tops) > 12:
            mid = tops[4:-4]
            print(f"{ROLE[r]}: mean period {(mid[-1] - mid[0]) / (len(mid) - 1):.0f} cycles over {len(mid)} tiles; first top @{tops[0] - base}, last @{tops[-1] - base}")


if __name__ == "__main__":
    main()
