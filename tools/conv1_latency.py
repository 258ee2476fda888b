"""BASELINE config 1: single int8 conv2d 3x3 s1 p1, 1x64x56x56 -> 64 ch, per-channel scales, int8 out:
latency of one launch (CUDA events over back-to-back launches, the reference's own convention
lite/tests/math/conv_int8_compute_test.cc:35-36: warm-up 10, repeats 100), checked against the oracle;
plus the batched variant (N = 256) for TOPS."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as ge  # noqa: E402
import orc  # noqa: E402

pl = ge.load_package().capi


def run(N):
    rng = np.random.default_rng(1234)
    C = K = 64
    x = rng.integers(-127, 128, (N, 56, 56, C), dtype=np.int8)
    w = rng.integers(-127, 128, (K, 3, 3, C), dtype=np.int8)
    ws = ((1 + np.arange(K) % 7) / (7 * 127)).astype(np.float32)
    bias = rng.uniform(-1, 1, K).astype(np.float32)
    sf, bf, _ = orc.fold_scales(ws, K, 1 / 127, C * 9 / 127 / 4, orc.OUT_INT8, bias)
    xd, wp = torch.from_numpy(x).cuda(), pl.pack_conv_weight(torch.from_numpy(w).cuda())
    sfd, bfd = torch.from_numpy(sf).cuda(), torch.from_numpy(bf).cuda()
    y = torch.empty((N, 56, 56, K), dtype=torch.int8, device="cuda")
    fn = lambda: pl.conv2d(xd, wp, K, 3, 3, sfd, bfd, (1, 1), (1, 1, 1, 1), act=pl.ACT_RELU, y=y)
    fn()
    torch.cuda.synchronize()
    ok = None
    if N == 1:
        want = orc.conv2d(x, w, sf, bf, (1, 1), (1, 1, 1, 1), act=orc.ACT_RELU)
        ok = bool(np.array_equal(y.cpu().numpy(), want))
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(100):
            fn()
    for _ in range(10):
        g.replay()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    g.replay()
    b.record()
    torch.cuda.synchronize()
    us = a.elapsed_time(b) * 1e3 / 100
    ops = 2.0 * N * 56 * 56 * K * 9 * C
    byts = N * 56 * 56 * (C + K) + K * 9 * C + 8 * K
    print(json.dumps({"config": f"conv2d 3x3 s1 p1 {N}x56x56x64 -> 64 int8 relu", "us_per_launch": round(us, 2),
                      "tops": round(ops / us / 1e6, 1), "gbs": round(byts / us / 1e3, 1), "bit_exact_vs_oracle": ok,
                      "timing": "100 launches in one CUDA graph, CUDA events"}))


if __name__ == "__main__":
    run(1)
    run(256)
