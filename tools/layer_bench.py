"""Per-layer micro-benchmark of the C-ABI kernels at BASELINE shapes (CUDA events,
L2-cold: each iteration uses a different buffer set when the tensors are < 256 MB).
   python tools/layer_bench.py [name ...] [--iters N] [--batch B]"""
import argparse
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

pl = ge.load_package().capi

LAYERS = {
    # name: (H, W, C, K, k, stride, pad, act, residual)
    "stem7x7": (224, 224, 3, 64, 7, 2, 3, "relu", False),
    "stem3x3": (224, 224, 3, 32, 3, 2, 1, "relu", False),
    "stem7x7_s2d": (224, 224, 3, 64, 7, 2, 3, "relu", "s2d"),
    "stem3x3_s2d": (224, 224, 3, 32, 3, 2, 1, "relu", "s2d"),
    "r2_1x1_64_256": (56, 56, 64, 256, 1, 1, 0, "none", False),
    "r2_1x1_64_256_res": (56, 56, 64, 256, 1, 1, 0, "relu", True),
    "r2_1x1_64_64": (56, 56, 64, 64, 1, 1, 0, "relu", False),
    "r2_1x1_256_64": (56, 56, 256, 64, 1, 1, 0, "relu", False),
    "r2_3x3_64": (56, 56, 64, 64, 3, 1, 1, "relu", False),
    "r3_3x3_128": (28, 28, 128, 128, 3, 1, 1, "relu", False),
    "r3_1x1_128_512_res": (28, 28, 128, 512, 1, 1, 0, "relu", True),
    "r4_3x3_256": (14, 14, 256, 256, 3, 1, 1, "relu", False),
    "r4_1x1_256_1024_res": (14, 14, 256, 1024, 1, 1, 0, "relu", True),
    "r4_1x1_1024_256": (14, 14, 1024, 256, 1, 1, 0, "relu", False),
    "r5_3x3_512": (7, 7, 512, 512, 3, 1, 1, "relu", False),
    "r3_1x1_512_128": (28, 28, 512, 128, 1, 1, 0, "relu", False),
    "r3_1x1_256_128": (56, 56, 256, 128, 1, 1, 0, "relu", False),
    "r5_1x1_2048_512": (7, 7, 2048, 512, 1, 1, 0, "relu", False),
    "r3_3x3s2_128": (56, 56, 128, 128, 3, 2, 1, "relu", False),
    "r4_3x3s2_256": (28, 28, 256, 256, 3, 2, 1, "relu", False),
    "m_pw_112_32_64": (112, 112, 32, 64, 1, 1, 0, "relu", False),
    "m_pw_14_512_512": (14, 14, 512, 512, 1, 1, 0, "relu", False),
    "r5_1x1_512_2048_res": (7, 7, 512, 2048, 1, 1, 0, "relu", True),
    "dw_112_32_s1": (112, 112, 32, 32, 3, 1, 1, "relu", "dw"),
    "dw_112_64_s2": (112, 112, 64, 64, 3, 2, 1, "relu", "dw"),
    "dw_56_128_s1": (56, 56, 128, 128, 3, 1, 1, "relu", "dw"),
    "dw_14_512_s1": (14, 14, 512, 512, 3, 1, 1, "relu", "dw"),
}
ACT = {"none": pl.ACT_NONE, "relu": pl.ACT_RELU}


def bench(name, batch, iters, nsets):
    H, W, C, K, k, st, pad, act, res = LAYERS[name]
    dev = "cuda"
    ld = 4 if C <= 4 else (C + 15) // 16 * 16
    P = (H + 2 * pad - k) // st + 1
    g = torch.Generator(device=dev).manual_seed(1)
    xs = [torch.randint(-127, 128, (batch, H, W, ld), dtype=torch.int8, device=dev, generator=g) for _ in range(nsets)]
    ys = [torch.empty((batch, P, P, K), dtype=torch.int8, device=dev) for _ in range(nsets)]
    sc = torch.full((K,), 1e-3, device=dev)
    bs = torch.zeros((K,), device=dev)
    dw = res == "dw"
    s2d = res == "s2d"
    thin = (not dw) and (not s2d) and C <= 4
    if s2d:
        geo = pl.s2d_geometry(batch, H, W, C, k, k, (st, st), (pad,) * 4)
        xf = torch.rand((batch, C, H, W), device=dev, generator=g) * 2 - 1
        xs = [pl.quantize_nchw_to_s2d(xf, 1 / 127, geo) for _ in range(nsets)]
        w = pl.pack_s2d_conv_weight(torch.randint(-127, 128, (K, k, k, C), dtype=torch.int8, device=dev, generator=g), geo)
    elif dw:
        w = torch.randint(-127, 128, (K, k, k), dtype=torch.int8, device=dev, generator=g)
    elif thin:
        w = pl.pack_thin_conv_weight(torch.randint(-127, 128, (K, k, k, C), dtype=torch.int8, device=dev, generator=g))
    else:
        w = pl.pack_conv_weight(torch.randint(-127, 128, (K, k, k, C), dtype=torch.int8, device=dev, generator=g))
    rs = [torch.randint(-127, 128, (batch, P, P, K), dtype=torch.int8, device=dev, generator=g) for _ in range(nsets)] if res is True else None

    def run(i):
        j = i % nsets
        if s2d:
            pl.s2d_conv2d(xs[j], w, batch, H, W, C, K, k, k, sc, bs, (st, st), (pad,) * 4, act=ACT[act], y=ys[j])
        elif dw:
            pl.dwconv2d(xs[j], w, sc, bs, k, k, (st, st), (pad,) * 4, act=ACT[act], y=ys[j], Cc=C)
        elif thin:
            pl.thin_conv2d(xs[j], w, K, k, k, C, sc, bs, (st, st), (pad,) * 4, act=ACT[act], y=ys[j])
        else:
            pl.conv2d(xs[j], w, K, k, k, sc, bs, (st, st), (pad,) * 4, act=ACT[act], y=ys[j], Cc=C,
                      residual=None if rs is None else rs[j], res_scale=0.5)

    for i in range(3):
        run(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(iters):
        run(i)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    cg = 1 if dw else C
    ops = 2.0 * batch * P * P * K * k * k * cg
    byts = batch * (H * W * C + P * P * K * (2 if res is True else 1)) + K * k * k * cg
    plan = ""
    if not dw:
        q = pl.last_igemm_plan()
        plan = (f" plan: mode {q['mode']} BN {q['BN']} BK {q['BK']} stages {q['stages']} kg {q.get('kgroup', 1)} pairs {q['cta_pairs']}"
                f" resident {q['filter_resident']} grid {q['grid']}")
    print(f"{name:24s} b{batch} ms={ms:.4f} TOPS={ops / ms / 1e9:8.1f} GB/s={byts / ms / 1e6:8.1f} "
          f"(bytes {byts / 1e6:.1f} MB){plan}", flush=True)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("names", nargs="*")
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--nsets", type=int, default=2)
    a = ap.parse_args()
    for n in (a.names or list(LAYERS)):
        bench(n, a.batch, a.iters, a.nsets)
