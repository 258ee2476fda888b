"""GPU parity of the CTA-pair launches (tcgen05 cta_group::2, igemm_i8.cu template flag CTA2): every case is
forced onto pairs (PL_CTA2=2, tile width through PL_FORCE_BN so that the problems stay small enough for the
oracle), checked to have taken that path (pl_last_igemm_plan) and compared bit for bit with the oracle.
Covers the three A-operand modes (plain GEMM / 1x1, im2col TMA, halo reuse), every epilogue variant, odd
block counts (the pair's second CTA computes a block that does not exist), ragged M / N / K, tile pairs and
the banded tile order on top of CTA pairs."""
import numpy as np
import pytest
import torch

import orc
from test_gpu_parity import conv_inputs, dev, host, pad_c, rup

pytestmark = pytest.mark.gpu


@pytest.fixture
def knobs(pl, monkeypatch):
    """set PL_* developer knobs for one test; the library re-reads them on refresh"""
    def set_(**kw):
        for k, v in kw.items():
            monkeypatch.setenv(k, str(v))
        pl.refresh_knobs()
    yield set_
    monkeypatch.undo()
    pl.refresh_knobs()


def gemm_case(pl, M, N, K, out, act, seed=3):
    rng = np.random.default_rng(seed)
    A = rng.integers(-127, 128, (M, K), dtype=np.int8)
    Bt = rng.integers(-127, 128, (N, K), dtype=np.int8)
    scale = ((1 + np.arange(N) % 5) / 5 * 40.0 / (5376.0 * K ** 0.5)).astype(np.float32)  # int8 range well used
    bias = rng.uniform(-1, 1, N).astype(np.float32)
    a = {"none": orc.ACT_NONE, "relu": orc.ACT_RELU, "relu6": orc.ACT_RELU6}[act]
    arg = 60.0 if act == "relu6" else 0.0
    if out == "int32":
        want = orc.gemm(A, Bt, None, None, b_is_nk=True, out=orc.OUT_INT32)
        got = pl.gemm(dev(A), dev(Bt), None, None, out=pl.OUT_INT32)
    else:
        o = orc.OUT_INT8 if out == "int8" else orc.OUT_FP32
        s = scale
        want = orc.gemm(A, Bt, s, bias, b_is_nk=True, act=a, act_arg=arg, out=o)
        got = pl.gemm(dev(A), dev(Bt), dev(s), dev(bias), a, arg, pl.OUT_INT8 if out == "int8" else pl.OUT_FP32)
    return host(got), want


GEMM_SHAPES = [
    # M, N, K, forced BN
    (256, 256, 512, 256),    # one pair, one tile
    (384, 256, 512, 128),    # odd block count: the last pair's second CTA has no block
    (300, 208, 528, 256),    # ragged M, N and K
    (1300, 320, 1024, 64),   # 11 blocks x 5 n-tiles, ragged everywhere
    (2048, 512, 256, 128),   # more pair tiles than clusters on a 2-tile-per-cluster grid
    (40000, 128, 128, 128),  # every cluster walks several pair tiles (ring wrap-around, accumulator ring)
]


@pytest.mark.parametrize("shape", GEMM_SHAPES)
@pytest.mark.parametrize("out,act", [("int8", "relu"), ("int8", "none"), ("int8", "relu6"), ("fp32", "none"),
                                     ("fp32", "relu"), ("int32", "none")])
def test_gemm_pairs(pl, knobs, shape, out, act):
    M, N, K, bn = shape
    if M > 4096 and out != "int8":
        pytest.skip("large-M case only on the staged int8 path")
    knobs(PL_CTA2=2, PL_FORCE_BN=bn, PL_NO_GEMV=1)
    got, want = gemm_case(pl, M, N, K, out, act)
    plan = pl.last_igemm_plan()
    assert plan["cta_pairs"] == 1 and plan["BN"] == bn, plan
    assert np.array_equal(got, want)
    if out == "int8":
        assert len(np.unique(want)) > 8


def test_gemm_pairs_match_single_cta(pl, knobs):
    """same launch with and without pairs: identical bytes (fp32 out, so every accumulator bit shows)"""
    knobs(PL_CTA2=2, PL_FORCE_BN=128)
    a, want = gemm_case(pl, 1000, 384, 1536, "fp32", "none", seed=9)
    assert pl.last_igemm_plan()["cta_pairs"] == 1
    knobs(PL_CTA2=0, PL_FORCE_BN=128)
    b, _ = gemm_case(pl, 1000, 384, 1536, "fp32", "none", seed=9)
    assert pl.last_igemm_plan()["cta_pairs"] == 0
    assert np.array_equal(a, b) and np.array_equal(a, want)


def test_gemm_pairs_of_tile_pairs(pl, knobs):
    """256-row tile pairs (two accumulators per stage) on top of CTA pairs: a 512 x 256 cluster tile"""
    knobs(PL_CTA2=2, PL_FORCE_BN=256, PL_MPAIR_FORCE=1, PL_CTA2_MPAIR_MIN_K=1)
    for M in (1024, 1300):
        got, want = gemm_case(pl, M, 512, 384, "int8", "relu", seed=M)
        plan = pl.last_igemm_plan()
        assert plan["cta_pairs"] == 1 and plan["tile_pairs"] == 1, plan
        assert np.array_equal(got, want)


def test_gemm_pairs_banded_order(pl, knobs):
    knobs(PL_CTA2=2, PL_FORCE_BN=64, PL_GROUP_M=2)
    got, want = gemm_case(pl, 1300, 640, 256, "int8", "none", seed=5)
    plan = pl.last_igemm_plan()
    assert plan["cta_pairs"] == 1 and plan["band_rows"] == 2, plan
    assert np.array_equal(got, want)


CONV_CASES = [
    # N, H, W, C, K, R, S, stride, pads, dil, forced BN, expected mode
    (2, 20, 20, 64, 128, 3, 3, 2, (1, 1, 1, 1), 1, 128, 1),     # im2col, ragged last block
    (9, 9, 7, 32, 64, 3, 3, 2, (1, 1, 1, 1), 1, 64, 1),          # tiny: a single pair
    (5, 17, 17, 96, 192, 5, 5, 1, (4, 4, 4, 4), 2, 64, 1),       # dilation, C = 96 (K-block tail), odd block count
    (3, 14, 14, 64, 64, 3, 3, 1, (1, 1, 1, 1), 1, 64, 3),        # halo reuse, 2 bands per image, 3 pairs
    (3, 28, 28, 128, 256, 3, 3, 1, (1, 1, 1, 1), 1, 128, 3),     # halo, 2 n-tiles, odd block count
    (1, 14, 14, 256, 256, 3, 3, 1, (1, 1, 1, 1), 1, 256, 3),     # ResNet res4 3x3: 2 channel blocks
    (5, 14, 14, 512, 128, 1, 1, 1, (0, 0, 0, 0), 1, 128, 0),     # 1x1 = plain GEMM path, 8 blocks (7.66 ragged)
]


@pytest.mark.parametrize("case", CONV_CASES)
@pytest.mark.parametrize("out,act", [("int8", "relu"), ("int8", "none"), ("fp32", "none")])
def test_conv_pairs(pl, knobs, case, out, act):
    *geo, bn, mode = case
    N, H, W, Cc, K, R, S, st, pads, dl = geo
    knobs(PL_CTA2=2, PL_FORCE_BN=bn)
    x, w, ws, bias, si, so = conv_inputs(tuple(geo), 41)
    a = {"none": orc.ACT_NONE, "relu": orc.ACT_RELU}[act]
    o = orc.OUT_INT8 if out == "int8" else orc.OUT_FP32
    sf, bf, _ = orc.fold_scales(ws, K, si, so, o, bias, a, 0.0)
    want = orc.conv2d(x, w, sf, bf, (st, st), pads, (dl, dl), 1, a, 0.0, o)
    xd = dev(pad_c(x, rup(Cc, 16)))
    wp = pl.pack_conv_weight(dev(w))
    got = host(pl.conv2d(xd, wp, K, R, S, dev(sf), dev(bf), (st, st), pads, (dl, dl), a, 0.0,
                         pl.OUT_INT8 if out == "int8" else pl.OUT_FP32, Cc=Cc))
    plan = pl.last_igemm_plan()
    assert plan["cta_pairs"] == 1 and plan["mode"] == mode, plan
    assert np.array_equal(got, want)


@pytest.mark.parametrize("case", [CONV_CASES[3], CONV_CASES[6], (3, 14, 14, 256, 1024, 1, 1, 1, (0, 0, 0, 0), 1, 256, 0)])
def test_conv_pairs_fused_residual(pl, knobs, case):
    *geo, bn, mode = case
    N, H, W, Cc, K, R, S, st, pads, dl = geo
    knobs(PL_CTA2=2, PL_FORCE_BN=bn)
    x, w, ws, bias, si, so = conv_inputs(tuple(geo), 43)
    sf, bf, _ = orc.fold_scales(ws, K, si, so, orc.OUT_INT8, bias, orc.ACT_RELU, 0.0)
    P = (H + pads[0] + pads[1] - (dl * (R - 1) + 1)) // st + 1
    Q = (W + pads[2] + pads[3] - (dl * (S - 1) + 1)) // st + 1
    res = np.random.default_rng(1).integers(-127, 128, (N, P, Q, K), dtype=np.int8)
    rs = np.float32(0.37)
    want = orc.conv2d(x, w, sf, bf, (st, st), pads, (dl, dl), 1, orc.ACT_RELU, 0.0, orc.OUT_INT8,
                      residual=res, res_scale=rs)
    xd = dev(pad_c(x, rup(Cc, 16)))
    wp = pl.pack_conv_weight(dev(w))
    got = host(pl.conv2d(xd, wp, K, R, S, dev(sf), dev(bf), (st, st), pads, (dl, dl), pl.ACT_RELU, 0.0, pl.OUT_INT8,
                         Cc=Cc, residual=dev(res), res_scale=float(rs)))
    plan = pl.last_igemm_plan()
    assert plan["cta_pairs"] == 1 and plan["epi"] == 3, plan
    assert np.array_equal(got, want)


def test_halo_pairs_with_resident_half_filter(pl, knobs):
    """PL_CTA2_HRES: a halo launch whose filter fits in shared memory runs as pairs, each CTA keeping half of it"""
    knobs(PL_CTA2_HRES=2, PL_FORCE_BN=128)
    geo = (3, 28, 28, 128, 128, 3, 3, 1, (1, 1, 1, 1), 1)
    N, H, W, Cc, K, R, S, st, pads, dl = geo
    x, w, ws, bias, si, so = conv_inputs(geo, 47)
    sf, bf, _ = orc.fold_scales(ws, K, si, so, orc.OUT_INT8, bias, orc.ACT_RELU, 0.0)
    want = orc.conv2d(x, w, sf, bf, (st, st), pads, (dl, dl), 1, orc.ACT_RELU, 0.0, orc.OUT_INT8)
    got = host(pl.conv2d(dev(x), pl.pack_conv_weight(dev(w)), K, R, S, dev(sf), dev(bf), (st, st), pads, (dl, dl),
                         pl.ACT_RELU, 0.0, pl.OUT_INT8, Cc=Cc))
    plan = pl.last_igemm_plan()
    assert plan["cta_pairs"] == 1 and plan["filter_resident"] == 1 and plan["mode"] == 3, plan
    assert np.array_equal(got, want)


@pytest.mark.parametrize("kg", [0, 2, 4])
def test_kblock_groups(pl, knobs, kg):
    """PL_KGROUP: several ring slots under one full barrier / one commit (off by default: measured slower), with a
    short last group, a resident and a streamed filter, single CTAs and pairs"""
    for cta2, bn, (M, N, K) in ((0, 128, (1000, 256, 640)), (2, 128, (1000, 256, 896)), (0, 64, (40000, 64, 384))):
        knobs(PL_KGROUP=kg, PL_CTA2=cta2, PL_FORCE_BN=bn, PL_NO_GEMV=1)
        got, want = gemm_case(pl, M, N, K, "int8", "relu", seed=kg + M)
        plan = pl.last_igemm_plan()
        assert plan["kgroup"] >= (2 if kg else 1), plan
        assert np.array_equal(got, want)
