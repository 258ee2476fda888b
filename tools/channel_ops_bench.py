"""HBM roofline of the int8 NHWC channel-movement kernels (concat / split / shuffle_channel / relu).
Algorithmic bytes: every element read once and written once (2 B per element).  Working sets are
larger than the 126 MB L2 (or rotate over several buffers) so that DRAM is what is timed."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402


def time_us(fn, iters=30, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) * 1e3 / iters


def main():
    pl = ge.load_package().capi
    peak = None
    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs")
    except Exception:
        pass
    cases = [("resnet-sized 256x56x56", (256, 56, 56), 128), ("256x28x28", (256, 28, 28), 256),
             ("humanseg b32 24x24", (32, 24, 24), 36), ("humanseg b32 12x12", (32, 12, 12), 72)]
    for name, pix, half in cases:
        g = torch.Generator(device="cuda").manual_seed(1)
        nset = max(1, int(300e6 // (pix[0] * pix[1] * pix[2] * half * 4)) + 1)  # rotate > L2
        sets = []
        for _ in range(min(nset, 8)):
            a = torch.randint(-128, 128, pix + (half,), dtype=torch.int8, device="cuda", generator=g)
            b = torch.randint(-128, 128, pix + (half,), dtype=torch.int8, device="cuda", generator=g)
            sets.append((a, b, torch.empty(pix + (2 * half,), dtype=torch.int8, device="cuda"),
                         torch.empty(pix + (2 * half,), dtype=torch.int8, device="cuda")))
        n = pix[0] * pix[1] * pix[2] * 2 * half
        state = {"i": 0}

        def nxt():
            state["i"] = (state["i"] + 1) % len(sets)
            return sets[state["i"]]

        def f_concat():
            a, b, w, _ = nxt()
            pl.concat([a, b], y=w)

        def f_split():
            a, b, w, _ = nxt()
            pl.split(w, [half, half], ys=[a, b])

        def f_shuffle():
            _, _, w, o = nxt()
            pl.shuffle_channel(w, 2, y=o)

        def f_relu():
            _, _, w, o = nxt()
            pl.relu_i8(w, 127, y=o)

        for op, fn in (("concat", f_concat), ("split", f_split), ("shuffle_channel g=2", f_shuffle), ("relu", f_relu)):
            us = time_us(fn)
            gbs = 2.0 * n / us / 1e3
            print(json.dumps({"op": op, "shape": name, "channels": 2 * half, "MB_moved": round(2 * n / 1e6, 1),
                              "us": round(us, 2), "GB/s": round(gbs, 1),
                              "frac_hbm_peak": None if not peak else round(gbs / peak, 3)}), flush=True)


if __name__ == "__main__":
    main()
