"""BASELINE config 3 (humanseg_shfv75_0119_quant, the int8 ShuffleNetV2 segmentation model shipped in the
reference): every conv2d / depthwise_conv2d of its `__model__` (62 layers, channels 3/9/18/36/72/144, fixture
tests/golden/humanseg_layers.json made by tools/extract_model_layers.py) run through the C-ABI with random int8
data and per-channel scales: bit-exact check against the oracle at batch 2, timing at batch 32.
The non-int8 ops of the model (conv2d_transpose, concat/split/shuffle, softmax ...) are SURVEY 8(f) rows.
  python tools/humanseg_layers.py [--batch 32]"""
import argparse
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


def rup(a, b):
    return (a + b - 1) // b * b


def layer_list():
    return json.load(open(os.path.join(ROOT, "tests", "golden", "humanseg_layers.json")))["layers"]


def make_case(l, batch, seed):
    """Random data for one layer: returns (runner, oracle_fn, macs, bytes)."""
    rng = np.random.default_rng(seed)
    _, C, H, W = l["in_dims"]
    K, Cg, R, S = l["filter_dims"]
    st, pd = tuple(l["strides"]), l["paddings"]
    pads = (pd[0], pd[0], pd[1], pd[1])
    dw = l["type"] == "depthwise_conv2d"
    x = rng.integers(-127, 128, (batch, H, W, C), dtype=np.int8)
    w = rng.integers(-127, 128, (K, R, S, Cg), dtype=np.int8)
    ws = ((1 + np.arange(K) % 7) / (7 * 127)).astype(np.float32)
    bias = rng.uniform(-1, 1, K).astype(np.float32)
    sf, bf, _ = orc.fold_scales(ws, K, 1 / 127, Cg * R * S / 127 / 4, orc.OUT_INT8, bias)
    sfd, bfd = torch.from_numpy(sf).cuda(), torch.from_numpy(bf).cuda()
    P = (H + 2 * pd[0] - R) // st[0] + 1
    Q = (W + 2 * pd[1] - S) // st[1] + 1
    ldy = rup(K, 16)
    y = torch.zeros((batch, P, Q, ldy), dtype=torch.int8, device="cuda")
    if dw:
        ld = rup(C, 16)
        xd = torch.zeros((batch, H, W, ld), dtype=torch.int8, device="cuda")
        xd[..., :C] = torch.from_numpy(x).cuda()
        wd = torch.from_numpy(np.ascontiguousarray(w.reshape(K, R, S))).cuda()
        run = lambda: pl.dwconv2d(xd, wd, sfd, bfd, R, S, st, pads, act=pl.ACT_RELU, y=y, Cc=C)
        want = lambda: orc.conv2d(x, w, sf, bf, st, pads, groups=C, act=orc.ACT_RELU)
    elif C <= 4:
        # stride-2 thin stem: fp32 NCHW in, space-to-depth quantiser + s2d conv (x above is replaced by what the
        # quantiser must hold so that the oracle sees the same int8 values)
        g = pl.s2d_geometry(batch, H, W, C, R, S, st, pads)
        assert g.supported
        xf = rng.uniform(-1, 1, (batch, C, H, W)).astype(np.float32)
        x = orc.nchw_to_nhwc(orc.quantize(xf, 1 / 127))
        xs = pl.quantize_nchw_to_s2d(torch.from_numpy(xf).cuda(), 1 / 127, g)
        wp = pl.pack_s2d_conv_weight(torch.from_numpy(w).cuda(), g)
        run = lambda: pl.s2d_conv2d(xs, wp, batch, H, W, C, K, R, S, sfd, bfd, st, pads, act=pl.ACT_RELU, y=y, ldy=ldy)
        want = lambda: orc.conv2d(x, w, sf, bf, st, pads, act=orc.ACT_RELU)
    else:
        ld = rup(C, 16)
        xd = torch.zeros((batch, H, W, ld), dtype=torch.int8, device="cuda")
        xd[..., :C] = torch.from_numpy(x).cuda()
        wp = pl.pack_conv_weight(torch.from_numpy(w).cuda())
        run = lambda: pl.conv2d(xd, wp, K, R, S, sfd, bfd, st, pads, act=pl.ACT_RELU, y=y, Cc=C, ldy=ldy)
        want = lambda: orc.conv2d(x, w, sf, bf, st, pads, act=orc.ACT_RELU)
    macs = batch * P * Q * K * R * S * Cg
    byts = batch * (H * W * C + P * Q * K) + K * R * S * Cg + 8 * K
    return run, want, y, K, macs, byts


def check(batch=2):
    """Every distinct layer shape, bit-exact against the oracle."""
    seen, n = set(), 0
    for i, l in enumerate(layer_list()):
        key = json.dumps([l[k] for k in ("type", "in_dims", "filter_dims", "strides", "paddings", "groups")])
        if key in seen:
            continue
        seen.add(key)
        run, want, y, K, _, _ = make_case(l, batch, 100 + i)
        run()
        torch.cuda.synchronize()
        got = y.cpu().numpy()[..., :K]
        assert np.array_equal(got, want()), f"humanseg layer {i} {l['type']} {l['in_dims']} {l['filter_dims']} differs"
        assert not y.cpu().numpy()[..., K:].any(), "pad channels must be written as zeros"
        n += 1
    return n


def bench(batch=32):
    tot_ms = tot_macs = tot_bytes = 0.0
    rows = []
    for i, l in enumerate(layer_list()):
        run, _, _, _, macs, byts = make_case(l, batch, 100 + i)
        for _ in range(3):
            run()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for _ in range(20):
                run()
        g.replay()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        g.replay()
        b.record()
        torch.cuda.synchronize()
        ms = a.elapsed_time(b) / 20
        rows.append((l["type"], l["in_dims"][1:], l["filter_dims"], l["strides"], ms, byts / ms / 1e6))
        tot_ms += ms
        tot_macs += macs
        tot_bytes += byts
    return {"batch": batch, "layers": len(rows), "sum_ms": round(tot_ms, 4), "gmac": round(tot_macs / 1e9, 3),
            "algorithmic_mbytes": round(tot_bytes / 1e6, 1), "gbs": round(tot_bytes / tot_ms / 1e6, 1),
            "tops": round(2 * tot_macs / tot_ms / 1e9, 1), "images_per_s_conv_layers_only": round(batch / (tot_ms * 1e-3), 0),
            "timing": "each layer: 20 back-to-back launches in a CUDA graph, CUDA events (L2-warm: the tensors are < 126 MB)"}, rows


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=32)
    a = ap.parse_args()
    n = check()
    summary, rows = bench(a.batch)
    summary["distinct_shapes_bit_exact_vs_oracle"] = n
    for r in rows:
        print(f"{r[0]:17s} in={r[1]} w={r[2]} s={r[3]} us={r[4] * 1e3:.2f} GB/s={r[5]:.0f}")
    print(json.dumps(summary))
