"""Pins the oracle (CPU, no GPU needed): KATs from the reference's own tests,
the reference's naive conv_basic/basic_gemm compiled into oracle/_ref, and the
calib goldens written as plain C in the reference's calib tests."""
import json
import os

import numpy as np
import pytest

import orc

HERE = os.path.dirname(os.path.abspath(__file__))
KAT = json.load(open(os.path.join(HERE, "golden", "kat_cuda_conv_int8.json")))


def test_kat1_all_ones_fp32_out():
    # lite/kernels/cuda/conv_compute_test.cc:232-298
    k = KAT["kat1"]
    x = np.full(k["x_shape"], 1, np.int8)
    w = np.full(k["w_shape"], 1, np.int8)
    sf, bf, _ = orc.fold_scales(k["weight_scale"], 4, 1.0, 1.0, orc.OUT_FP32)
    y = orc.conv2d(x, w, sf, None, out=orc.OUT_FP32, rnd=orc.ROUND_CUDA)
    assert y.reshape(-1).tolist() == k["expect"]


def test_kat2_mt19937_int8_out():
    # lite/kernels/cuda/conv_compute_test.cc:300-376
    k = KAT["kat2"]
    x = np.array(k["x"], np.int8).reshape(k["x_shape"])
    w = np.array(k["w"], np.int8).reshape(k["w_shape"])
    acc = orc.conv2d(x, w, None, None, out=orc.OUT_INT32)
    assert acc.reshape(-1).tolist() == k["acc"]
    sf, bf, _ = orc.fold_scales(k["weight_scale"], 4, k["input_scale"],
                                k["output_scale"], orc.OUT_INT8, k["bias"])
    for rnd in (orc.ROUND_CUDA, orc.ROUND_ARM):
        y = orc.conv2d(x, w, sf, bf, act=orc.ACT_RELU, out=orc.OUT_INT8, rnd=rnd)
        assert y.reshape(-1).tolist() == k["expect"]


needs_ref = pytest.mark.skipif(orc.ref() is None,
                               reason="oracle/_ref not built (needs /root/reference)")

CONV_CASES = [
    # N,H,W,C,K,R,S,stride,pad,dil,groups
    (1, 8, 8, 16, 16, 3, 3, 1, 1, 1, 1),
    (2, 9, 7, 8, 12, 3, 3, 2, 1, 1, 1),
    (1, 10, 10, 6, 4, 1, 1, 1, 0, 1, 1),
    (1, 11, 11, 4, 8, 5, 5, 2, 2, 1, 1),
    (1, 12, 12, 8, 8, 3, 3, 1, 2, 2, 1),
    (2, 9, 9, 16, 16, 3, 3, 1, 1, 1, 16),   # depthwise
    (1, 9, 9, 12, 12, 5, 5, 2, 2, 1, 12),   # depthwise 5x5 s2
    (1, 7, 7, 8, 16, 3, 3, 1, 1, 1, 2),     # grouped
    (1, 15, 15, 3, 32, 3, 3, 2, 1, 1, 1),   # first layer C=3
    (1, 16, 16, 3, 8, 7, 7, 2, 3, 1, 1),    # resnet stem shape
]


@needs_ref
@pytest.mark.parametrize("case", CONV_CASES)
def test_conv_acc_vs_reference_conv_basic(case):
    N, H, W, Cc, K, R, S, st, pd, dl, g = case
    rng = np.random.default_rng(hash(case) & 0xFFFF)
    x = rng.integers(-127, 128, (N, H, W, Cc), dtype=np.int8)
    w = rng.integers(-127, 128, (K, R, S, Cc // g), dtype=np.int8)
    mine = orc.conv2d(x, w, None, None, (st, st), (pd, pd, pd, pd), (dl, dl), g,
                      out=orc.OUT_INT32)
    theirs = orc.ref_conv_acc_nchw(orc.nhwc_to_nchw(x),
                                   np.ascontiguousarray(w.transpose(0, 3, 1, 2)),
                                   (st, st), (pd, pd), (dl, dl), g)
    assert np.array_equal(mine, orc.nchw_to_nhwc(theirs))


@needs_ref
@pytest.mark.parametrize("mnk", [(1, 8, 16), (5, 7, 33), (16, 32, 64), (3, 1000, 128)])
@pytest.mark.parametrize("b_is_nk", [False, True])
def test_gemm_acc_vs_reference_basic_gemm(mnk, b_is_nk):
    M, N, K = mnk
    rng = np.random.default_rng(M * 131 + N * 7 + K)
    A = rng.integers(-127, 128, (M, K), dtype=np.int8)
    B = rng.integers(-127, 128, (N, K) if b_is_nk else (K, N), dtype=np.int8)
    mine = orc.gemm(A, B, None, None, b_is_nk=b_is_nk, out=orc.OUT_INT32)
    theirs = orc.ref_gemm_acc(A, B, trans_b=b_is_nk)
    assert np.array_equal(mine, theirs)
    assert np.array_equal(mine, A.astype(np.int32) @ (B.T if b_is_nk else B).astype(np.int32))


def test_layout_roundtrip_and_numpy():
    rng = np.random.default_rng(3)
    x = rng.integers(-128, 128, (2, 5, 4, 7), dtype=np.int8)  # NCHW
    y = orc.nchw_to_nhwc(x)
    assert np.array_equal(y, x.transpose(0, 2, 3, 1))
    assert np.array_equal(orc.nhwc_to_nchw(y), x)
    xf = rng.standard_normal((1, 3, 6, 5)).astype(np.float32)
    assert np.array_equal(orc.nchw_to_nhwc(xf), xf.transpose(0, 2, 3, 1))


def test_calib_goldens():
    rng = np.random.default_rng(11)
    x = (rng.standard_normal(4096) * 3).astype(np.float32)
    scale = np.float32(0.0371)
    # ARM golden: lite/kernels/arm/calib_compute_test.cc:54-69
    inv = np.float32(1.0) / scale
    t = (x * inv).astype(np.float32)
    r = np.where(t >= 0, np.floor(t + np.float32(0.5)), np.ceil(t - np.float32(0.5)))
    # roundf == half away; recompute exactly to dodge float add error
    r = np.sign(t) * np.floor(np.abs(t).astype(np.float64) + 0.5)
    arm = np.clip(r, -127, 127).astype(np.int8)
    assert np.array_equal(orc.quantize(x, scale, orc.ROUND_ARM), arm)
    # CUDA kernel: lite/kernels/cuda/calib_compute.cu:27-36 (x/scale, clamp, rint)
    t = (x / scale).astype(np.float32)
    cu = np.rint(np.clip(t, -128, 127)).astype(np.int8)
    assert np.array_equal(orc.quantize(x, scale, orc.ROUND_CUDA), cu)
    q = rng.integers(-128, 128, 1000, dtype=np.int8)
    assert np.array_equal(orc.dequantize(q, scale), q.astype(np.float32) * scale)


def test_rounding_conventions_on_ties():
    L = orc.lib()
    f = lambda v, m: int(L.orc_round_sat_i8(orc._f(v), m))
    assert [f(v, orc.ROUND_ARM) for v in (0.5, 1.5, 2.5, -0.5, -2.5, 0.49999997, 127.6, -127.6, -500)] == \
        [1, 2, 3, -1, -3, 0, 127, -127, -127]
    assert [f(v, orc.ROUND_CUDA) for v in (0.5, 1.5, 2.5, -0.5, -2.5, 127.6, -127.6, -500)] == \
        [0, 2, 2, 0, -2, 127, -128, -128]


def test_epilogue_matches_cvt_kernel_text():
    # lite/backends/arm/math/conv_block_utils.h:3185-3226 restated in numpy with
    # an exact fma (float64 product+sum of float32 values rounds once).
    rng = np.random.default_rng(5)
    acc = rng.integers(-70_000_000, 70_000_000, (64, 16), dtype=np.int32)
    acc[0, :4] = [0, 1, -1, 2**24 + 1]
    sc = (rng.random(16, dtype=np.float32) * 1e-6 + 1e-7).astype(np.float32)
    bs = (rng.standard_normal(16) * 4).astype(np.float32)
    accf = acc.astype(np.float32)
    v = (accf.astype(np.float64) * sc.astype(np.float64) + bs.astype(np.float64)).astype(np.float32)
    for act, fn in ((orc.ACT_NONE, lambda t: t),
                    (orc.ACT_RELU, lambda t: np.maximum(t, 0)),
                    (orc.ACT_RELU6, lambda t: np.minimum(np.maximum(t, 0), np.float32(6.0))),
                    (orc.ACT_LEAKY, lambda t: np.where(t > 0, t, t * np.float32(0.1)))):
        want = fn(v).astype(np.float32)
        y = np.empty_like(want)
        arg = 6.0 if act == orc.ACT_RELU6 else 0.1
        orc.lib().orc_epilogue(orc._p(acc), orc.C.c_int64(64), 16, orc._p(sc), orc._p(bs), act,
                               orc._f(arg), orc.OUT_FP32, orc.ROUND_ARM, orc._p(y))
        assert np.array_equal(y, want)
        q = np.empty(want.shape, np.int8)
        orc.lib().orc_epilogue(orc._p(acc), orc.C.c_int64(64), 16, orc._p(sc), orc._p(bs), act,
                               orc._f(arg), orc.OUT_INT8, orc.ROUND_ARM, orc._p(q))
        r = np.sign(want) * np.floor(np.abs(want).astype(np.float64) + 0.5)
        assert np.array_equal(q, np.clip(r, -127, 127).astype(np.int8))


def test_fold_scales_rules():
    ws = np.array([0.5, 0.25, 2.0], np.float32)
    b = np.array([1.0, -2.0, 3.0], np.float32)
    sf, bf, six = orc.fold_scales(ws, 3, 0.1, 0.4, orc.OUT_INT8, b, orc.ACT_RELU6, 6.0)
    assert np.array_equal(sf, ws * np.float32(0.1) / np.float32(0.4))
    assert np.array_equal(bf, b / np.float32(0.4))
    assert six == np.float32(6.0) / np.float32(0.4)
    sf, bf, six = orc.fold_scales(ws[:1], 3, 0.1, 0.4, orc.OUT_FP32, None, orc.ACT_RELU6, 6.0)
    assert np.array_equal(sf, np.full(3, ws[0] * np.float32(0.1), np.float32))
    assert np.array_equal(bf, np.zeros(3, np.float32)) and six == 6.0


# ---- channel movement: the oracle's NCHW loops pinned on the baselines the
# reference's own tests use (lite/tests/kernels/shuffle_channel_compute_test.cc:38-66 is the
# memcpy permutation below; concat_compute_test.cc / split are plain axis-1 concatenation)
@pytest.mark.parametrize("shape,group", [((10, 16, 4, 4), 4), ((2, 72, 3, 5), 2), ((1, 12, 2, 2), 3),
                                         ((1, 6, 1, 1), 1), ((1, 6, 1, 1), 6)])
def test_shuffle_channel_oracle_vs_reference_baseline(shape, group):
    rng = np.random.default_rng(shape[1] + group)
    x = rng.integers(-128, 128, shape, dtype=np.int8)
    N, Cc, H, W = shape
    want = np.empty_like(x)
    for n in range(N):  # the reference test's RunBaseline, element for element
        for i in range(group):
            for j in range(Cc // group):
                want[n, j * group + i] = x[n, i * (Cc // group) + j]
    got = orc.shuffle_channel_nchw(x, group)
    assert np.array_equal(got, want)
    assert np.array_equal(got, x.reshape(N, group, Cc // group, H, W).transpose(0, 2, 1, 3, 4).reshape(shape))
    xf = x.astype(np.float32)  # the reference kernel is fp32: same movement on 4-byte elements
    assert np.array_equal(orc.shuffle_channel_nchw(xf, group), want.astype(np.float32))


@pytest.mark.parametrize("cs", [(2, 3), (64, 64), (9, 18, 3), (5,)])
def test_concat_split_oracle_vs_numpy(cs):
    rng = np.random.default_rng(sum(cs))
    xs = [rng.integers(-128, 128, (3, c, 4, 5), dtype=np.int8) for c in cs]
    y = orc.concat_nchw(xs)
    assert np.array_equal(y, np.concatenate(xs, axis=1))
    for p, x in zip(orc.split_nchw(y, list(cs)), xs):
        assert np.array_equal(p, x)


def test_relu_i8_oracle():
    x = np.arange(-128, 128, dtype=np.int8)
    assert np.array_equal(orc.relu_i8(x), np.maximum(x, 0))
    assert np.array_equal(orc.relu_i8(x, 47), np.clip(x, 0, 47))


# ---- round 2: pins against more of the reference's own object code (oracle/_ref) -------------
needs_ref = pytest.mark.skipif(orc.ref() is None or not hasattr(orc.ref(), "ref_saturate_cast_i8_f32_n"),
                               reason="oracle/_ref not built (needs /root/reference)")
needs_ref_extra = pytest.mark.skipif(orc.ref_extra() is None, reason="oracle/_ref/libref_extra.so not built")


@needs_ref
def test_round_saturate_is_the_references_saturate_cast():
    """orc_round_sat_i8 (ARM convention) == max(saturate_cast<int8_t>(v), -127) with the reference's own
    saturate.h:144-147 compiled into oracle/_ref: every tie k + 0.5, both float neighbours of every tie,
    the clamp edges, and random values."""
    ties = np.arange(-300, 300, dtype=np.float32) + np.float32(0.5)
    vals = np.concatenate([ties, np.nextafter(ties, np.float32(np.inf)), np.nextafter(ties, np.float32(-np.inf)),
                           np.arange(-300, 301, dtype=np.float32),
                           np.array([0.0, -0.0, 126.5, 127.49999, 127.5, -126.5, -127.5, -127.49999, -128.5, 1e6, -1e6,
                                     0.49999997, -0.49999997], np.float32),
                           np.random.default_rng(3).uniform(-140, 140, 200000).astype(np.float32)])
    want = np.maximum(orc.ref_saturate_i8(vals).astype(np.int32), -127).astype(np.int8)
    L = orc.lib()
    got = np.array([L.orc_round_sat_i8(orc._f(v), orc.ROUND_ARM) for v in vals[:3000]], np.int8)
    assert np.array_equal(got, want[:3000])
    # the vector path the kernels are compared with: quantize with scale 1 is round + saturate
    assert np.array_equal(orc.quantize(vals, 1.0, orc.ROUND_ARM), want)


@needs_ref_extra
@pytest.mark.parametrize("k,s,p", [((3, 3), (2, 2), (1, 1, 1, 1)), ((2, 2), (2, 2), (0, 0, 0, 0)),
                                   ((3, 3), (1, 1), (1, 1, 1, 1)), ((5, 3), (2, 1), (2, 2, 1, 1)),
                                   ((7, 7), (1, 1), (0, 0, 0, 0))])
@pytest.mark.parametrize("avg,exclusive", [(False, True), (True, True), (True, False)])
def test_pool_oracle_vs_reference_pool_compute_ref(k, s, p, avg, exclusive):
    """int8 pooling = the reference's fp32 golden on the integers: fp32 output identical, int8 output its
    requantisation (lite/kernels/cuda/pool_compute_test.cc:150-249 compiled where it lies)."""
    rng = np.random.default_rng(11)
    x = rng.integers(-128, 128, (2, 9, 11, 5), dtype=np.int8)  # NHWC
    ref = orc.ref_pool_nchw(x.transpose(0, 3, 1, 2).astype(np.float32), k, s, p, avg, exclusive)
    ref = ref.transpose(0, 2, 3, 1)
    typ = orc.POOL_AVG if avg else orc.POOL_MAX
    got = orc.pool2d(x, k, s, p, typ, exclusive, 1.0, orc.OUT_FP32)
    assert np.array_equal(got, ref)
    for rescale in (1.0, 0.37):
        q = orc.pool2d(x, k, s, p, typ, exclusive, rescale, orc.OUT_INT8)
        want = orc.quantize(ref * np.float32(rescale), 1.0)
        assert np.array_equal(q, want)


@needs_ref_extra
def test_pool_oracle_global_vs_reference():
    rng = np.random.default_rng(12)
    x = rng.integers(-128, 128, (3, 7, 7, 6), dtype=np.int8)
    for avg in (False, True):
        ref = orc.ref_pool_nchw(x.transpose(0, 3, 1, 2).astype(np.float32), (7, 7), (1, 1), (0, 0, 0, 0), avg,
                                global_pooling=True).transpose(0, 2, 3, 1)
        got = orc.pool2d(x, (7, 7), (1, 1), (0, 0, 0, 0), orc.POOL_AVG if avg else orc.POOL_MAX, True, 1.0,
                         orc.OUT_FP32, global_pooling=True)
        assert np.array_equal(got, ref)


@needs_ref_extra
@pytest.mark.parametrize("tx,ty", [(False, False), (False, True), (True, False), (True, True)])
def test_matmul_acc_vs_reference_matrix_mul(tx, ty):
    """int32 accumulators of the int8 mul / matmul contract == the kernels-test baseline (matrix_mul +
    transpose, lite/tests/kernels/matmul_compute_test.cc:23-112) for every transpose combination; small
    K keeps the fp32 baseline exact."""
    rng = np.random.default_rng(21)
    M, K, N = 13, 40, 9
    X = rng.integers(-127, 128, (K, M) if tx else (M, K), dtype=np.int8)
    Y = rng.integers(-127, 128, (N, K) if ty else (K, N), dtype=np.int8)
    ref = orc.ref_matmul_f32(X.astype(np.float32), Y.astype(np.float32), tx, ty, 1.0)
    acc = orc.gemm(X, Y, None, None, trans_a=tx, b_is_nk=ty, out=orc.OUT_INT32)
    assert np.array_equal(acc.astype(np.float32), ref)
    # alpha folds into the per-column scale: fp32 out within 1e-5 relative of the baseline with alpha
    alpha = 0.25
    sc = np.full(N, alpha, np.float32)
    got = orc.gemm(X, Y, sc, None, trans_a=tx, b_is_nk=ty, out=orc.OUT_FP32)
    ref_a = orc.ref_matmul_f32(X.astype(np.float32), Y.astype(np.float32), tx, ty, alpha)
    assert np.allclose(got, ref_a, rtol=1e-5, atol=1e-6)


def test_matmul_kat_all_ones():
    """lite/kernels/cuda/matmul_compute_test.cc:33-95: x [4,1,2] ones, y [4,1,2] ones, transpose_Y -> 2."""
    x = np.ones((4, 1, 2), np.int8)
    y = np.ones((4, 1, 2), np.int8)
    for b in range(4):
        acc = orc.gemm(x[b], y[b], None, None, b_is_nk=True, out=orc.OUT_INT32)
        assert acc.shape == (1, 1) and int(acc[0, 0]) == 2


def test_adaptive_pool_index_rule():
    """floor / ceil window rule of lite/kernels/cuda/pool_compute.cu:89-94 restated in numpy (double arithmetic as
    the reference computes it); P == H reproduces the identity, P == 1 the global pool."""
    rng = np.random.default_rng(17)
    x = rng.integers(-127, 128, (2, 11, 7, 5), dtype=np.int8)  # -128 would hit the ARM -127 floor
    for P, Q in ((3, 2), (11, 7), (1, 1), (4, 5)):
        for typ in (orc.POOL_MAX, orc.POOL_AVG):
            got = orc.adaptive_pool2d(x, (P, Q), typ, 1.0, orc.OUT_FP32)
            want = np.empty_like(got)
            for p_ in range(P):
                h0, h1 = int(np.floor(p_ * 11 / P)), int(np.ceil((p_ + 1) * 11 / P))
                for q_ in range(Q):
                    w0, w1 = int(np.floor(q_ * 7 / Q)), int(np.ceil((q_ + 1) * 7 / Q))
                    win = x[:, h0:h1, w0:w1, :].astype(np.float32)
                    want[:, p_, q_, :] = win.max((1, 2)) if typ == orc.POOL_MAX else \
                        (win.sum((1, 2)) / np.float32((h1 - h0) * (w1 - w0))).astype(np.float32)
            assert np.array_equal(got, want)
    assert np.array_equal(orc.adaptive_pool2d(x, (11, 7), orc.POOL_MAX, 1.0, orc.OUT_INT8), x)
