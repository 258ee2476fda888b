"""pl_set_tile_order is a scheduling hint only: conv (plain GEMM / im2col / halo / resident-filter / space-to-depth
paths, CTA pairs), depthwise, gemm and pooling give the oracle's bytes whether they walk their tile list
first-to-last or last-to-first."""
import numpy as np
import pytest
import torch

import orc
from test_gpu_parity import conv_inputs, dev, host, pad_c, rup, run_conv

pytestmark = pytest.mark.gpu


@pytest.fixture
def reversed_order(pl):
    pl.set_tile_order(True)
    yield
    pl.set_tile_order(False)


CASES = [
    # N, H, W, C, K, R, S, stride, pads, dil
    (6, 56, 56, 64, 64, 3, 3, 1, (1, 1, 1, 1), 1),      # halo mode, resident 3x3 filter, many tiles per CTA
    (8, 28, 28, 128, 512, 1, 1, 1, (0, 0, 0, 0), 1),    # plain GEMM, two N tiles, resident filter slice per CTA
    (9, 28, 28, 64, 256, 1, 1, 1, (0, 0, 0, 0), 1),     # odd tile count
    (4, 28, 28, 128, 128, 3, 3, 2, (1, 1, 1, 1), 1),    # im2col TMA
    (3, 14, 14, 256, 512, 3, 3, 1, (1, 1, 1, 1), 1),    # halo, two channel blocks, two N tiles (streamed filter)
    (2, 7, 7, 512, 1000, 1, 1, 1, (0, 0, 0, 0), 1),     # ragged N
]


@pytest.mark.parametrize("case", CASES)
@pytest.mark.parametrize("act", ["relu", "none"])
def test_conv_reversed_matches_oracle(pl, reversed_order, case, act):
    N, H, W, Cc, K, R, S, st, pads, dl = case
    x, w, ws, bias, in_scale, out_scale = conv_inputs(case, 11)
    sf, bf, _ = orc.fold_scales(ws, K, in_scale, out_scale, orc.OUT_INT8, bias)
    a_o, a_p = (orc.ACT_RELU, pl.ACT_RELU) if act == "relu" else (orc.ACT_NONE, pl.ACT_NONE)
    want = orc.conv2d(x, w, sf, bf, (st, st), pads, (dl, dl), act=a_o)
    got = run_conv(pl, case, x, w, sf, bf, a_p, 0.0, pl.OUT_INT8, pl.ROUND_ARM)
    assert np.array_equal(got[..., :K], want)


def test_conv_residual_reversed_matches_oracle(pl, reversed_order):
    case = (8, 28, 28, 128, 512, 1, 1, 1, (0, 0, 0, 0), 1)
    N, H, W, Cc, K, R, S, st, pads, dl = case
    x, w, ws, bias, in_scale, out_scale = conv_inputs(case, 5)
    rng = np.random.default_rng(6)
    res = rng.integers(-127, 128, (N, H, W, K), dtype=np.int8)
    sf, bf, _ = orc.fold_scales(ws, K, in_scale, out_scale, orc.OUT_INT8, bias)
    want = orc.conv2d(x, w, sf, bf, (1, 1), pads, act=orc.ACT_RELU, residual=res, res_scale=np.float32(0.37))
    y = pl.conv2d(dev(x), pl.pack_conv_weight(dev(w)), K, 1, 1, dev(sf), dev(bf), (1, 1), pads, act=pl.ACT_RELU,
                  residual=dev(res), res_scale=0.37)
    assert np.array_equal(host(y), want)


def test_cta_pairs_reversed_matches_forward(pl, reversed_order, monkeypatch):
    monkeypatch.setenv("PL_CTA2", "2")
    monkeypatch.setenv("PL_FORCE_BN", "128")  # (the small problem would otherwise take tiles too narrow for pairs)
    pl.refresh_knobs()
    try:
        case = (4, 14, 14, 256, 256, 3, 3, 1, (1, 1, 1, 1), 1)
        N, H, W, Cc, K, R, S, st, pads, dl = case
        x, w, ws, bias, in_scale, out_scale = conv_inputs(case, 2)
        sf, bf, _ = orc.fold_scales(ws, K, in_scale, out_scale, orc.OUT_INT8, bias)
        want = orc.conv2d(x, w, sf, bf, (1, 1), pads, act=orc.ACT_RELU)
        got = run_conv(pl, case, x, w, sf, bf, pl.ACT_RELU, 0.0, pl.OUT_INT8, pl.ROUND_ARM)
        assert pl.last_igemm_plan()["cta_pairs"] == 1
        assert np.array_equal(got, want)
    finally:
        monkeypatch.undo()
        pl.refresh_knobs()


def test_s2d_stem_reversed_matches_oracle(pl, reversed_order):
    rng = np.random.default_rng(3)
    N, H = 3, 64
    xf = rng.uniform(-1, 1, (N, 3, H, H)).astype(np.float32)
    w = rng.integers(-127, 128, (64, 7, 7, 3), dtype=np.int8)
    ws = ((rng.random(64, dtype=np.float32) + 0.5) / 127).astype(np.float32)
    bias = rng.standard_normal(64).astype(np.float32)
    sf, bf, _ = orc.fold_scales(ws, 64, 1 / 127, 3 * 49 / 127 / 4, orc.OUT_INT8, bias)
    geo = pl.s2d_geometry(N, H, H, 3, 7, 7, (2, 2), (3, 3, 3, 3))
    assert geo.supported
    want = orc.conv2d(orc.nchw_to_nhwc(orc.quantize(xf, 1 / 127)), w, sf, bf, (2, 2), (3, 3, 3, 3), act=orc.ACT_RELU)
    xs = pl.quantize_nchw_to_s2d(dev(xf), 1 / 127, geo)
    got = pl.s2d_conv2d(xs, pl.pack_s2d_conv_weight(dev(w), geo), N, H, H, 3, 64, 7, 7, dev(sf), dev(bf), (2, 2),
                        (3, 3, 3, 3), act=pl.ACT_RELU)
    assert np.array_equal(host(got), want)


@pytest.mark.parametrize("stride", [1, 2])
def test_depthwise_reversed_matches_oracle(pl, reversed_order, stride):
    rng = np.random.default_rng(9)
    N, H, C = 5, 56, 128
    x = rng.integers(-127, 128, (N, H, H, C), dtype=np.int8)
    w = rng.integers(-127, 128, (C, 3, 3), dtype=np.int8)
    ws = ((rng.random(C, dtype=np.float32) + 0.5) / 127).astype(np.float32)
    bias = rng.standard_normal(C).astype(np.float32)
    sf, bf, _ = orc.fold_scales(ws, C, 1 / 127, 9 / 127 / 2, orc.OUT_INT8, bias)
    want = orc.conv2d(x, w.reshape(C, 3, 3, 1), sf, bf, (stride, stride), (1, 1, 1, 1), groups=C, act=orc.ACT_RELU)
    got = pl.dwconv2d(dev(x), dev(w), dev(sf), dev(bf), 3, 3, (stride, stride), (1, 1, 1, 1), act=pl.ACT_RELU)
    assert np.array_equal(host(got), want)


def test_pool_and_gemm_reversed_match_forward(pl):
    rng = np.random.default_rng(4)
    x = dev(rng.integers(-127, 128, (4, 56, 56, 64), dtype=np.int8))
    A = dev(rng.integers(-127, 128, (700, 512), dtype=np.int8))
    Bt = dev(rng.integers(-127, 128, (384, 512), dtype=np.int8))
    s = dev(np.full(384, 1e-3, np.float32))
    b = dev(np.zeros(384, np.float32))
    outs = []
    for rev in (False, True):
        pl.set_tile_order(rev)
        try:
            outs.append((host(pl.pool2d(x, (3, 3), (2, 2), (1, 1, 1, 1), pl.POOL_MAX)),
                         host(pl.pool2d(x, (7, 7), (7, 7), (0, 0, 0, 0), pl.POOL_AVG)),
                         host(pl.gemm(A, Bt, s, b, out=pl.OUT_INT8)), host(pl.gemm(A, Bt, s, b, out=pl.OUT_FP32))))
        finally:
            pl.set_tile_order(False)
    for f, r in zip(*outs):
        assert np.array_equal(f, r)


@pytest.mark.parametrize("shape", [(3, 57, 61, 64), (2, 56, 56, 32), (1, 9, 7, 16), (2, 112, 112, 64)])
@pytest.mark.parametrize("pad", [0, 1])
@pytest.mark.parametrize("rev", [False, True])
def test_max_pool_3x3s2_row_pairs_match_oracle(pl, shape, pad, rev):
    """the 3x3 / s2 max pool computes two output rows per thread (pool3x3s2_max_pair_kernel): odd and even output heights,
    with and without padding, both tile orders"""
    N, H, W, C = shape
    rng = np.random.default_rng(H + W + C + pad)
    x = rng.integers(-127, 128, (N, H, W, C), dtype=np.int8)
    want = orc.pool2d(x, (3, 3), (2, 2), (pad,) * 4, orc.POOL_MAX)
    pl.set_tile_order(rev)
    try:
        got = host(pl.pool2d(dev(x), (3, 3), (2, 2), (pad,) * 4, pl.POOL_MAX))
    finally:
        pl.set_tile_order(False)
    assert np.array_equal(got, want)
