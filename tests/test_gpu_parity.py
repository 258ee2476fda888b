"""GPU parity tests (-m gpu): the sm_100a kernels, called through the C-ABI
(include/pl_int8.h), against the CPU oracle on the same seeded inputs.

Bar: int32 accumulators and int8 outputs bit-exact; fp32 outputs within 1e-5
relative (north_star) -- both sides use one explicit fmaf, so they are in fact
compared for equality and the tolerance is only the documented bound.
Shapes follow the reference's own sweeps: lite/tests/math/conv_int8_compute_test.cc
:481-731 (dw 3x3/5x5, 1x1, 3x3 s1/s2, random) and gemm_int8_compute_test.cc.
"""
import json
import os

import numpy as np
import pytest
import torch

import orc

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
DEV = "cuda:0"
FP32_RTOL = 1e-5


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).to(DEV)


def host(t):
    torch.cuda.synchronize()
    return t.cpu().numpy()


def pad_c(x, ld):
    """NHWC int8 -> channel pitch ld (zero pad)."""
    if x.shape[-1] == ld:
        return x
    out = np.zeros(x.shape[:-1] + (ld,), x.dtype)
    out[..., : x.shape[-1]] = x
    return out


def rup(v, m):
    return (v + m - 1) // m * m


def conv_inputs(case, seed):
    N, H, W, Cc, K, R, S, st, pads, dl = case
    rng = np.random.default_rng(seed)
    x = rng.integers(-127, 128, (N, H, W, Cc), dtype=np.int8)
    w = rng.integers(-127, 128, (K, R, S, Cc), dtype=np.int8)
    ws = ((1 + np.arange(K) % 7) / (7 * 127)).astype(np.float32)
    bias = rng.uniform(-1, 1, K).astype(np.float32)
    in_scale = np.float32(1 / 127)
    out_scale = np.float32(Cc * R * S / 127 / 4)
    return x, w, ws, bias, in_scale, out_scale


def run_conv(pl, case, x, w, sf, bf, act, act_arg, out, rnd):
    N, H, W, Cc, K, R, S, st, pads, dl = case
    ldx = rup(Cc, 16)
    xd = dev(pad_c(x, ldx))
    wp = pl.pack_conv_weight(dev(w))
    y = pl.conv2d(xd, wp, K, R, S, None if sf is None else dev(sf), None if bf is None else dev(bf),
                  (st, st), pads, (dl, dl), act, act_arg, out, rnd, Cc=Cc)
    return host(y)


CONV_CASES = [
    # N, H, W, C, K, R, S, stride, (pt,pb,pl,pr), dil
    (1, 56, 56, 64, 64, 3, 3, 1, (1, 1, 1, 1), 1),      # BASELINE config 1
    (2, 14, 14, 64, 64, 3, 3, 1, (1, 1, 1, 1), 1),
    (3, 9, 7, 32, 48, 3, 3, 2, (1, 1, 1, 1), 1),        # tiles straddle images, K tail
    (1, 28, 28, 128, 256, 1, 1, 1, (0, 0, 0, 0), 1),    # pointwise -> plain GEMM path
    (2, 28, 28, 256, 128, 1, 1, 2, (0, 0, 0, 0), 1),    # 1x1 stride 2 (ResNet projection)
    (1, 20, 20, 16, 32, 5, 5, 2, (2, 2, 2, 2), 1),
    (1, 17, 17, 32, 32, 3, 3, 1, (2, 2, 2, 2), 2),      # dilation 2
    (1, 12, 15, 96, 80, 3, 3, 1, (0, 1, 1, 0), 1),      # asymmetric pads, C=96 -> BK=128 OOB
    (1, 14, 14, 256, 512, 3, 3, 1, (1, 1, 1, 1), 1),    # 2 channel blocks, 2 n-tiles
    (2, 32, 32, 3, 32, 3, 3, 2, (1, 1, 1, 1), 1),       # MobileNet conv0: C=3 (pitch 16)
    (1, 38, 38, 3, 64, 7, 7, 2, (3, 3, 3, 3), 1),       # ResNet stem shape
    (1, 8, 8, 72, 72, 1, 1, 1, (0, 0, 0, 0), 1),        # humanseg widths (C,K not /16)
    (1, 7, 7, 512, 1000, 1, 1, 1, (0, 0, 0, 0), 1),
]


@pytest.mark.parametrize("case", CONV_CASES)
def test_conv_int32_accumulators_exact(pl, case):
    x, w, *_ = conv_inputs(case, 11)
    N, H, W, Cc, K, R, S, st, pads, dl = case
    want = orc.conv2d(x, w, None, None, (st, st), pads, (dl, dl), 1, out=orc.OUT_INT32)
    got = run_conv(pl, case, x, w, None, None, pl.ACT_NONE, 0.0, pl.OUT_INT32, pl.ROUND_ARM)
    assert got.shape == want.shape
    assert np.array_equal(got, want)


@pytest.mark.parametrize("case", CONV_CASES)
@pytest.mark.parametrize("act", ["relu", "none", "relu6", "leaky"])
def test_conv_int8_out_exact(pl, case, act):
    x, w, ws, bias, si, so = conv_inputs(case, 23)
    N, H, W, Cc, K, R, S, st, pads, dl = case
    a = {"none": orc.ACT_NONE, "relu": orc.ACT_RELU, "relu6": orc.ACT_RELU6, "leaky": orc.ACT_LEAKY}[act]
    arg = {"none": 0.0, "relu": 0.0, "relu6": 6.0, "leaky": 0.1}[act]
    if act == "relu6":
        so = np.float32(6.0 / 127)
    sf, bf, argf = orc.fold_scales(ws, K, si, so, orc.OUT_INT8, bias, a, arg)
    want = orc.conv2d(x, w, sf, bf, (st, st), pads, (dl, dl), 1, a, argf, orc.OUT_INT8)
    got = run_conv(pl, case, x, w, sf, bf, a, argf, pl.OUT_INT8, pl.ROUND_ARM)
    assert np.array_equal(got, want)
    assert want.min() < 0 or act != "none"  # the scale recipe really exercises the range
    assert len(np.unique(want)) > 8


@pytest.mark.parametrize("case", CONV_CASES[:6])
@pytest.mark.parametrize("rnd", ["arm", "cuda"])
def test_conv_fp32_out(pl, case, rnd):
    x, w, ws, bias, si, so = conv_inputs(case, 31)
    N, H, W, Cc, K, R, S, st, pads, dl = case
    r = orc.ROUND_ARM if rnd == "arm" else orc.ROUND_CUDA
    sf, bf, _ = orc.fold_scales(ws, K, si, so, orc.OUT_FP32, bias, orc.ACT_LEAKY, 0.2)
    want = orc.conv2d(x, w, sf, bf, (st, st), pads, (dl, dl), 1, orc.ACT_LEAKY, 0.2, orc.OUT_FP32, r)
    got = run_conv(pl, case, x, w, sf, bf, pl.ACT_LEAKY, 0.2, pl.OUT_FP32, r)
    np.testing.assert_allclose(got, want, rtol=FP32_RTOL, atol=0)
    assert np.array_equal(got, want)


def test_conv_cuda_legacy_rounding_int8(pl):
    case = CONV_CASES[1]
    x, w, ws, bias, si, so = conv_inputs(case, 5)
    N, H, W, Cc, K, R, S, st, pads, dl = case
    sf, bf, _ = orc.fold_scales(ws, K, si, so, orc.OUT_INT8, bias)
    want = orc.conv2d(x, w, sf, bf, (st, st), pads, (dl, dl), 1, orc.ACT_RELU, 0, orc.OUT_INT8, orc.ROUND_CUDA)
    got = run_conv(pl, case, x, w, sf, bf, pl.ACT_RELU, 0.0, pl.OUT_INT8, pl.ROUND_CUDA)
    assert np.array_equal(got, want)


def test_reference_kats_on_gpu(pl):
    """lite/kernels/cuda/conv_compute_test.cc:232-376 (the reference's own kCUDA int8 vectors)."""
    kat = json.load(open(os.path.join(HERE, "golden", "kat_cuda_conv_int8.json")))
    k = kat["kat1"]
    x = np.full(k["x_shape"], 1, np.int8)
    w = np.full(k["w_shape"], 1, np.int8)
    sf, bf, _ = orc.fold_scales(k["weight_scale"], 4, 1.0, 1.0, orc.OUT_FP32)
    case = (1, 3, 3, 4, 4, 3, 3, 1, (0, 0, 0, 0), 1)
    got = run_conv(pl, case, x, w, sf, None, pl.ACT_NONE, 0.0, pl.OUT_FP32, pl.ROUND_CUDA)
    assert got.reshape(-1).tolist() == k["expect"]
    k = kat["kat2"]
    x = np.array(k["x"], np.int8).reshape(k["x_shape"])
    w = np.array(k["w"], np.int8).reshape(k["w_shape"])
    sf, bf, _ = orc.fold_scales(k["weight_scale"], 4, k["input_scale"], k["output_scale"], orc.OUT_INT8, k["bias"])
    case = (1, 3, 3, 3, 4, 3, 3, 1, (0, 0, 0, 0), 1)
    acc = run_conv(pl, case, x, w, None, None, pl.ACT_NONE, 0.0, pl.OUT_INT32, pl.ROUND_CUDA)
    assert acc.reshape(-1).tolist() == k["acc"]
    for rnd in (pl.ROUND_CUDA, pl.ROUND_ARM):
        got = run_conv(pl, case, x, w, sf, bf, pl.ACT_RELU, 0.0, pl.OUT_INT8, rnd)
        assert got.reshape(-1).tolist() == k["expect"]


def test_conv_full_size_slices_match_oracle(pl):
    """ResNet-50 res2 3x3 at batch 64: images are independent, so the first and last
    image of the big launch must equal the oracle run on those images alone."""
    case = (64, 56, 56, 64, 64, 3, 3, 1, (1, 1, 1, 1), 1)
    x, w, ws, bias, si, so = conv_inputs(case, 77)
    sf, bf, _ = orc.fold_scales(ws, 64, si, so, orc.OUT_INT8, bias)
    got = run_conv(pl, case, x, w, sf, bf, pl.ACT_RELU, 0.0, pl.OUT_INT8, pl.ROUND_ARM)
    for n in (0, 31, 63):
        want = orc.conv2d(x[n:n + 1], w, sf, bf, (1, 1), (1, 1, 1, 1), act=orc.ACT_RELU)
        assert np.array_equal(got[n:n + 1], want)
    # linearity of the accumulator in the filter: conv(x, w1) + conv(x, w2) == conv(x, w1 + w2)
    rng = np.random.default_rng(3)
    w1 = rng.integers(-60, 61, w.shape, dtype=np.int8)
    w2 = rng.integers(-60, 61, w.shape, dtype=np.int8)
    a1 = run_conv(pl, case, x, w1, None, None, 0, 0.0, pl.OUT_INT32, 0)
    a2 = run_conv(pl, case, x, w2, None, None, 0, 0.0, pl.OUT_INT32, 0)
    a12 = run_conv(pl, case, x, (w1 + w2).astype(np.int8), None, None, 0, 0.0, pl.OUT_INT32, 0)
    assert np.array_equal(a1 + a2, a12)


@pytest.mark.parametrize("case", [CONV_CASES[1], CONV_CASES[3], CONV_CASES[11], CONV_CASES[12]])
@pytest.mark.parametrize("out", ["int8", "fp32"])
def test_conv_fused_residual(pl, case, out):
    """conv + int8 shortcut + relu in one epilogue (ConvParam::residualData)."""
    x, w, ws, bias, si, so = conv_inputs(case, 41)
    N, H, W, Cc, K, R, S, st, pads, dl = case
    o = orc.OUT_INT8 if out == "int8" else orc.OUT_FP32
    sf, bf, _ = orc.fold_scales(ws, K, si, so, o, bias)
    P = orc.conv_out_size(H, R, dl, pads[0], pads[1], st)
    Q = orc.conv_out_size(W, S, dl, pads[2], pads[3], st)
    res = np.random.default_rng(8).integers(-127, 128, (N, P, Q, K), dtype=np.int8)
    rs = 0.37 if out == "int8" else 0.011
    want = orc.conv2d(x, w, sf, bf, (st, st), pads, (dl, dl), 1, orc.ACT_RELU, 0.0, o, residual=res, res_scale=rs)
    ldx = rup(Cc, 16)
    wp = pl.pack_conv_weight(dev(w))
    got = host(pl.conv2d(dev(pad_c(x, ldx)), wp, K, R, S, dev(sf), dev(bf), (st, st), pads, (dl, dl),
                         pl.ACT_RELU, 0.0, o, Cc=Cc, residual=dev(res), res_scale=rs))
    assert np.array_equal(got, want)


HALO_CASES = [
    # stride-1 undilated windows that take the halo-reuse path (mode 3): WP = Q + S - 1 <= 128
    (2, 28, 28, 128, 128, 3, 3, 1, (1, 1, 1, 1), 1),    # WP=30, 4 rows per band, 7 full bands
    (3, 30, 30, 32, 48, 3, 3, 1, (1, 1, 1, 1), 1),      # WP=32: last band of every image is a 2-row tail
    (1, 9, 9, 48, 24, 3, 3, 1, (1, 1, 1, 1), 1),        # whole image in one band, C=48 < BK, K tail
    (2, 20, 20, 16, 32, 5, 5, 1, (2, 2, 2, 2), 1),      # 5x5 window
    (1, 56, 56, 64, 320, 3, 3, 1, (1, 1, 1, 1), 1),     # two n-tiles (256 + 64) over the same halo tiles
    (1, 13, 40, 32, 32, 3, 1, 1, (1, 1, 0, 0), 1),      # 3x1 window: WP == Q
    (2, 10, 10, 192, 64, 2, 2, 1, (0, 1, 0, 1), 1),     # even window, asymmetric pads, 2 channel blocks
    (1, 7, 7, 512, 512, 3, 3, 1, (1, 1, 1, 1), 1),      # res5 3x3 shape
]


@pytest.mark.parametrize("case", [
    # stride-1 windows whose halo buffers + filter ring do not fit the shared memory: the launch falls back to the
    # im2col path instead of failing (round-1 advisor finding)
    (2, 96, 96, 128, 256, 5, 5, 1, (2, 2, 2, 2), 1),
    (1, 40, 110, 128, 256, 7, 7, 1, (3, 3, 3, 3), 1),
    (1, 30, 120, 192, 64, 5, 5, 1, (2, 2, 2, 2), 1),
])
@pytest.mark.parametrize("out", ["int8", "fp32"])
def test_conv_halo_falls_back_when_smem_is_short(pl, case, out):
    x, w, ws, bias, si, so = conv_inputs(case, 71)
    N, H, W, Cc, K, R, S, st, pads, dl = case
    o = orc.OUT_INT8 if out == "int8" else orc.OUT_FP32
    sf, bf, _ = orc.fold_scales(ws, K, si, so, o, bias)
    want = orc.conv2d(x, w, sf, bf, (st, st), pads, (dl, dl), 1, orc.ACT_RELU, 0.0, o)
    got = host(pl.conv2d(dev(pad_c(x, rup(Cc, 16))), pl.pack_conv_weight(dev(w)), K, R, S, dev(sf), dev(bf),
                         (st, st), pads, (dl, dl), orc.ACT_RELU, 0.0, o, Cc=Cc))
    if o == orc.OUT_FP32:
        assert np.allclose(got, want, rtol=1e-5, atol=1e-6)
    else:
        assert np.array_equal(got, want)


@pytest.mark.parametrize("case", [
    # N, H, W, C, K, R, stride, pad, groups  (1 < groups < C: lite/kernels/arm/conv_gemmlike.cc:84-264)
    (2, 14, 14, 64, 64, 3, 1, 1, 4),
    (1, 28, 28, 128, 96, 1, 1, 0, 2),
    (1, 15, 17, 96, 48, 3, 2, 1, 3),
    (3, 8, 8, 256, 512, 3, 1, 1, 8),
])
@pytest.mark.parametrize("out", ["int8", "fp32", "int32"])
def test_grouped_dense_conv(pl, case, out):
    N, H, W, Cc, K, R, st, pd, G = case
    rng = np.random.default_rng(hash(case) & 0xFFFF)
    x = rng.integers(-127, 128, (N, H, W, Cc), dtype=np.int8)
    w = rng.integers(-127, 128, (K, R, R, Cc // G), dtype=np.int8)
    ws = ((1 + np.arange(K) % 7) / (7 * 127)).astype(np.float32)
    bias = rng.uniform(-1, 1, K).astype(np.float32)
    o = {"int8": orc.OUT_INT8, "fp32": orc.OUT_FP32, "int32": orc.OUT_INT32}[out]
    sf, bf, _ = orc.fold_scales(ws, K, 1 / 127, (Cc // G) * R * R / 127 / 4, o if o != orc.OUT_INT32 else orc.OUT_FP32, bias)
    want = orc.conv2d(x, w, sf, bf, (st, st), (pd,) * 4, (1, 1), G, orc.ACT_RELU, 0.0, o)
    got = host(pl.conv2d(dev(x), pl.pack_conv_weight(dev(w)), K, R, R, dev(sf), dev(bf), (st, st), (pd,) * 4, (1, 1),
                         orc.ACT_RELU, 0.0, o, Cc=Cc, groups=G))
    if o == orc.OUT_FP32:
        assert np.allclose(got, want, rtol=1e-5, atol=1e-6)
    else:
        assert np.array_equal(got, want)


@pytest.mark.parametrize("case", HALO_CASES)
@pytest.mark.parametrize("mode", ["int32", "int8_relu", "fp32_res", "int8_res"])
def test_conv_halo_reuse_path(pl, case, mode, monkeypatch):
    """Halo-reuse tiles vs the oracle, and vs the im2col path (PL_IGEMM_NO_HALO=1) on the same input."""
    x, w, ws, bias, si, so = conv_inputs(case, 53)
    N, H, W, Cc, K, R, S, st, pads, dl = case
    P = orc.conv_out_size(H, R, dl, pads[0], pads[1], st)
    Q = orc.conv_out_size(W, S, dl, pads[2], pads[3], st)
    res = rs = None
    if mode == "int32":
        o, sf, bf, act = orc.OUT_INT32, None, None, orc.ACT_NONE
    else:
        o = orc.OUT_FP32 if mode.startswith("fp32") else orc.OUT_INT8
        sf, bf, _ = orc.fold_scales(ws, K, si, so, o, bias)
        act = orc.ACT_RELU
        if mode.endswith("res"):
            res = np.random.default_rng(9).integers(-127, 128, (N, P, Q, K), dtype=np.int8)
            rs = 0.41 if o == orc.OUT_INT8 else 0.013
    kw = {} if res is None else {"residual": res, "res_scale": rs}
    want = orc.conv2d(x, w, sf, bf, (st, st), pads, (dl, dl), 1, act, 0.0, o, **kw)
    xd = dev(pad_c(x, rup(Cc, 16)))
    wp = pl.pack_conv_weight(dev(w))

    def go():
        kw2 = {} if res is None else {"residual": dev(res), "res_scale": rs}
        return host(pl.conv2d(xd, wp, K, R, S, None if sf is None else dev(sf), None if bf is None else dev(bf),
                              (st, st), pads, (dl, dl), act, 0.0, o, Cc=Cc, **kw2))

    got = go()
    assert got.shape == want.shape
    assert np.array_equal(got, want)
    monkeypatch.setenv("PL_IGEMM_NO_HALO", "1")
    pl.lib().pl_refresh_knobs()  # knobs are cached per process
    try:
        assert np.array_equal(go(), want)
    finally:
        monkeypatch.delenv("PL_IGEMM_NO_HALO")
        pl.lib().pl_refresh_knobs()


def test_conv_writes_into_channel_slice(pl):
    """ldy > K: the conv writes a channel slice of a wider buffer (free concat)."""
    case = (1, 10, 10, 32, 32, 3, 3, 1, (1, 1, 1, 1), 1)
    x, w, ws, bias, si, so = conv_inputs(case, 9)
    sf, bf, _ = orc.fold_scales(ws, 32, si, so, orc.OUT_INT8, bias)
    want = orc.conv2d(x, w, sf, bf, (1, 1), (1, 1, 1, 1))
    buf = torch.full((1, 10, 10, 64), 99, dtype=torch.int8, device=DEV)
    wp = pl.pack_conv_weight(dev(w))
    view = buf[..., 32:]  # pointer offset 32, pitch 64
    d = pl.conv_desc(1, 10, 10, 32, 32, 3, 3, (1, 1), (1, 1, 1, 1), (1, 1), 1, 32, 64)
    sfd, bfd, xd = dev(sf), dev(bf), dev(x)
    e = pl.make_epilogue(sfd, bfd)
    import ctypes as C
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    # NULL arguments are rejected, not dereferenced
    assert pl.lib().pl_conv2d_i8_nhwc(None, None, None, None, None, st) != 0
    rc = pl.lib().pl_conv2d_i8_nhwc(C.c_void_p(xd.data_ptr()), C.c_void_p(wp.data_ptr()),
                                    C.c_void_p(view.data_ptr()), C.byref(d), C.byref(e), st)
    assert rc == 0
    out = host(buf)
    assert np.array_equal(out[..., 32:], want)
    assert np.all(out[..., :32] == 99)


def test_conv_error_codes(pl):
    x = torch.zeros((1, 8, 8, 24), dtype=torch.int8, device=DEV)  # pitch 24: not /16
    wp = torch.zeros(4096, dtype=torch.int8, device=DEV)
    s = torch.ones(16, device=DEV)
    with pytest.raises(pl.PlError):
        pl.conv2d(x, wp, 16, 3, 3, s, None, (1, 1), (1, 1, 1, 1))


GEMM_CASES = [(128, 64, 64), (256, 256, 128), (300, 1000, 1024), (1, 1000, 1024), (64, 1000, 2048),
              (1024, 512, 2304), (130, 48, 96), (17, 24, 48), (2048, 2048, 512)]


@pytest.mark.parametrize("mnk", GEMM_CASES)
def test_gemm_parity(pl, mnk):
    M, N, K = mnk
    rng = np.random.default_rng(M + 3 * N + 7 * K)
    A = rng.integers(-127, 128, (M, K), dtype=np.int8)
    Bt = rng.integers(-127, 128, (N, K), dtype=np.int8)
    s = ((rng.random(N, dtype=np.float32) + 0.5) / 127 / 127 * (64.0 / K) * 8).astype(np.float32)
    b = rng.standard_normal(N).astype(np.float32)
    Ad, Bd, sd, bd = dev(A), dev(Bt), dev(s), dev(b)
    acc = host(pl.gemm(Ad, Bd, None, None, out=pl.OUT_INT32))
    assert np.array_equal(acc, A.astype(np.int32) @ Bt.astype(np.int32).T)
    want = orc.gemm(A, Bt, s, b, b_is_nk=True, act=orc.ACT_RELU, out=orc.OUT_INT8)
    got = host(pl.gemm(Ad, Bd, sd, bd, act=pl.ACT_RELU, out=pl.OUT_INT8))
    assert np.array_equal(got, want)
    want = orc.gemm(A, Bt, s, b, b_is_nk=True, out=orc.OUT_FP32)
    got = host(pl.gemm(Ad, Bd, sd, bd, out=pl.OUT_FP32))
    np.testing.assert_allclose(got, want, rtol=FP32_RTOL, atol=0)
    assert np.array_equal(got, want)


@pytest.mark.parametrize("mnk", [(640, 512, 2080), (1000, 300, 4096), (256, 256, 64)])
def test_gemm_tile_pair_mode(pl, mnk, monkeypatch):
    """Long-K GEMMs run as 256 x 256 tile pairs (one B stage feeds two accumulators); the
    heuristic only picks it for large problems, so force it here on ragged small ones."""
    monkeypatch.setenv("PL_MPAIR_FORCE", "1")
    M, N, K = mnk
    rng = np.random.default_rng(M + N + K)
    A = rng.integers(-127, 128, (M, K), dtype=np.int8)
    Bt = rng.integers(-127, 128, (N, K), dtype=np.int8)
    s = ((rng.random(N, dtype=np.float32) + 0.5) / 127 / 127 * (64.0 / K) * 8).astype(np.float32)
    b = rng.standard_normal(N).astype(np.float32)
    for act, oact in ((pl.ACT_RELU, orc.ACT_RELU), (pl.ACT_NONE, orc.ACT_NONE)):
        want = orc.gemm(A, Bt, s, b, b_is_nk=True, act=oact, out=orc.OUT_INT8)
        got = host(pl.gemm(dev(A), dev(Bt), dev(s), dev(b), act=act, out=pl.OUT_INT8))
        assert np.array_equal(got, want)
    acc = host(pl.gemm(dev(A), dev(Bt), None, None, out=pl.OUT_INT32))
    assert np.array_equal(acc, A.astype(np.int32) @ Bt.astype(np.int32).T)
    # fp32 out also runs as pairs (stored straight from registers), act none and relu
    for act, oact in ((pl.ACT_RELU, orc.ACT_RELU), (pl.ACT_NONE, orc.ACT_NONE)):
        want = orc.gemm(A, Bt, s, b, b_is_nk=True, act=oact, out=orc.OUT_FP32)
        got = host(pl.gemm(dev(A), dev(Bt), dev(s), dev(b), act=act, out=pl.OUT_FP32))
        assert np.array_equal(got, want)


@pytest.mark.parametrize("pairs", [False, True])
def test_gemm_grouped_tile_order(pl, pairs, monkeypatch):
    """Long-K GEMMs walk their tiles in bands of group_m rows (L2 locality); 20 tile rows in bands of 8 leave
    a ragged last band of 4 (2 with pairs: bands of 4), the last row itself is ragged (M % 128 != 0)."""
    monkeypatch.setenv("PL_GROUP_M", "4" if pairs else "8")
    if pairs:
        monkeypatch.setenv("PL_MPAIR_FORCE", "1")
    else:
        monkeypatch.setenv("PL_NO_MPAIR", "1")
    M, N, K = 2482, 2048, 2048
    rng = np.random.default_rng(77)
    A = rng.integers(-127, 128, (M, K), dtype=np.int8)
    Bt = rng.integers(-127, 128, (N, K), dtype=np.int8)
    s = ((rng.random(N, dtype=np.float32) + 0.5) / 127 / 127 * (64.0 / K) * 8).astype(np.float32)
    b = rng.standard_normal(N).astype(np.float32)
    want = orc.gemm(A, Bt, s, b, b_is_nk=True, act=orc.ACT_RELU, out=orc.OUT_INT8)
    got = host(pl.gemm(dev(A), dev(Bt), dev(s), dev(b), act=pl.ACT_RELU, out=pl.OUT_INT8))
    assert np.array_equal(got, want)
    monkeypatch.setenv("PL_GROUP_M", "1")  # and the row-major order gives the same bytes
    assert np.array_equal(host(pl.gemm(dev(A), dev(Bt), dev(s), dev(b), act=pl.ACT_RELU, out=pl.OUT_INT8)), want)


def test_gemm_fc_weight_layout(pl):
    """fc keeps W as [K,N] (lite/kernels/cuda/fc_compute.cu:199-353); it is transposed
    once (PrepareForRun) with pl_transpose_i8 and then used K-major."""
    rng = np.random.default_rng(2)
    M, N, K = 32, 1000, 1024
    A = rng.integers(-127, 128, (M, K), dtype=np.int8)
    Wkn = rng.integers(-127, 128, (K, N), dtype=np.int8)
    s = np.full(N, 1e-4, np.float32)
    Bt = pl.transpose_i8(dev(Wkn))
    assert np.array_equal(host(Bt), Wkn.T)
    want = orc.gemm(A, Wkn, s, None, b_is_nk=False, out=orc.OUT_INT8)
    got = host(pl.gemm(dev(A), Bt, dev(s), None, out=pl.OUT_INT8))
    assert np.array_equal(got, want)


DW_CASES = [
    # N, H, W, C, R, stride, pad, dil
    (2, 28, 28, 64, 3, 1, 1, 1),
    (2, 28, 28, 64, 3, 2, 1, 1),
    (1, 112, 112, 32, 3, 1, 1, 1),   # MobileNetV1 dw1
    (1, 15, 13, 128, 3, 2, 1, 1),    # odd sizes
    (1, 14, 14, 512, 3, 1, 1, 1),
    (1, 12, 12, 72, 3, 1, 1, 1),     # humanseg widths
    (1, 24, 24, 36, 3, 2, 1, 1),
    (1, 9, 9, 9, 3, 1, 1, 1),        # C=9 in a pitch-16 buffer
    (1, 16, 16, 32, 5, 1, 2, 1),     # 5x5 (generic path)
    (1, 16, 16, 32, 5, 2, 2, 1),
    (1, 16, 16, 16, 3, 1, 2, 2),     # dilation (generic path)
    (1, 7, 7, 1024, 3, 1, 1, 1),
    # shared-memory staged kernel: several tiles per CTA (ring reuse), column tiles (Q > 128), channel
    # chunks with a tail, bands with a ragged last row block, pad 0, asymmetric sizes
    (9, 56, 56, 128, 3, 1, 1, 1),
    (9, 56, 56, 128, 3, 2, 1, 1),
    (2, 20, 150, 48, 3, 1, 1, 1),
    (2, 21, 290, 24, 3, 2, 1, 1),
    (3, 37, 41, 336, 3, 1, 1, 1),
    (3, 37, 41, 336, 3, 2, 1, 1),
    (2, 30, 30, 40, 3, 1, 0, 1),
    (2, 31, 29, 40, 3, 2, 0, 1),
    (40, 14, 14, 512, 3, 1, 1, 1),
    (40, 14, 14, 512, 3, 2, 1, 1),
]


@pytest.mark.parametrize("case", DW_CASES)
@pytest.mark.parametrize("out", ["int8", "fp32", "int32"])
def test_depthwise_parity(pl, case, out):
    N, H, W, Cc, R, st, pd, dl = case
    rng = np.random.default_rng(hash(case) & 0xFFFF)
    x = rng.integers(-127, 128, (N, H, W, Cc), dtype=np.int8)
    w = rng.integers(-127, 128, (Cc, R, R), dtype=np.int8)
    ws = ((1 + np.arange(Cc) % 5) / (5 * 127)).astype(np.float32)
    bias = rng.uniform(-1, 1, Cc).astype(np.float32)
    o = {"int8": orc.OUT_INT8, "fp32": orc.OUT_FP32, "int32": orc.OUT_INT32}[out]
    sf, bf, _ = orc.fold_scales(ws, Cc, 1 / 127, R * R / 127 / 2, o if o != orc.OUT_INT32 else orc.OUT_FP32, bias)
    want = orc.conv2d(x, w.reshape(Cc, R, R, 1), sf, bf, (st, st), (pd, pd, pd, pd), (dl, dl), Cc,
                      orc.ACT_RELU, 0.0, o)
    ld = rup(Cc, 16)
    got = host(pl.dwconv2d(dev(pad_c(x, ld)), dev(w), dev(sf), dev(bf), R, R, (st, st), (pd, pd, pd, pd),
                           (dl, dl), pl.ACT_RELU, 0.0, o, Cc=Cc))
    assert np.array_equal(got[..., :Cc], want)
    assert not got[..., Cc:].any()


@pytest.mark.parametrize("n", [0, 1, 15, 16, 17, 4096, 100003])
@pytest.mark.parametrize("rnd", ["arm", "cuda"])
def test_calib_quantize_dequantize(pl, n, rnd):
    rng = np.random.default_rng(n + 1)
    x = (rng.standard_normal(n) * 3).astype(np.float32)
    if n >= 16:
        x[:8] = np.array([0.5, 1.5, 2.5, -0.5, -2.5, 0.49999997, 1000.0, -1000.0], np.float32) * np.float32(0.0371)
    r = orc.ROUND_ARM if rnd == "arm" else orc.ROUND_CUDA
    scale = np.float32(0.0371)
    got = host(pl.quantize(dev(x), scale, r)) if n else np.zeros(0, np.int8)
    assert np.array_equal(got, orc.quantize(x, scale, r))
    q = rng.integers(-128, 128, n, dtype=np.int8)
    got = host(pl.dequantize(dev(q), scale)) if n else np.zeros(0, np.float32)
    assert np.array_equal(got, orc.dequantize(q, scale))


def test_rounding_every_tie_and_edge(pl):
    """Every k+0.5 tie, the values next to them, and the clamp edges, both conventions."""
    base = np.arange(-130, 131, dtype=np.float32)
    vals = np.concatenate([base, base + np.float32(0.5), np.nextafter(base + np.float32(0.5), np.float32(1e9)),
                           np.nextafter(base + np.float32(0.5), np.float32(-1e9)),
                           np.array([1e9, -1e9, 127.49999, -127.5, 126.5, -126.5], np.float32)]).astype(np.float32)
    for r in (orc.ROUND_ARM, orc.ROUND_CUDA):
        got = host(pl.quantize(dev(vals), 1.0, r))
        assert np.array_equal(got, orc.quantize(vals, 1.0, r))


@pytest.mark.parametrize("shape", [(2, 3, 16, 18), (3, 4, 8, 8), (2, 1, 6, 6), (1, 2, 10, 14), (2, 3, 224, 224)])
def test_layout_thin_int8(pl, shape):
    """image feeds (C <= 4, HW % 4 == 0) take the thin int8 layout kernel: dense pitch and one-word pitch"""
    x8 = np.random.default_rng(sum(shape)).integers(-128, 128, shape, dtype=np.int8)
    assert np.array_equal(host(pl.nchw_to_nhwc(dev(x8))), orc.nchw_to_nhwc(x8))
    y4 = host(pl.nchw_to_nhwc(dev(x8), 4))
    assert np.array_equal(y4[..., : shape[1]], orc.nchw_to_nhwc(x8)) and not y4[..., shape[1]:].any()


@pytest.mark.parametrize("shape", [(2, 3, 17, 19), (1, 64, 56, 56), (3, 72, 12, 12), (1, 1000, 1, 1)])
def test_layout_transforms(pl, shape):
    rng = np.random.default_rng(sum(shape))
    x8 = rng.integers(-128, 128, shape, dtype=np.int8)
    xf = rng.standard_normal(shape).astype(np.float32)
    for x in (x8, xf):
        y = host(pl.nchw_to_nhwc(dev(x)))
        assert np.array_equal(y, orc.nchw_to_nhwc(x))
        assert np.array_equal(host(pl.nhwc_to_nchw(dev(y))), x)
    ld = rup(shape[1], 16)
    y = host(pl.nchw_to_nhwc(dev(x8), ld))
    assert np.array_equal(y[..., : shape[1]], orc.nchw_to_nhwc(x8)) and not y[..., shape[1]:].any()
    assert np.array_equal(host(pl.nhwc_to_nchw(dev(y), shape[1])), x8)
    # fused calib + layout + channel pad (entry of the int8 region)
    scale = np.float32(0.02)
    if 64 * ld <= 48 * 1024:
        y = host(pl.quantize_nchw_to_nhwc(dev(xf), scale, ld))
        want = orc.nchw_to_nhwc(orc.quantize(xf, scale))
        assert np.array_equal(y[..., : shape[1]], want) and not y[..., shape[1]:].any()


POOL_CASES = [
    # N,H,W,C, k, s, pad, type, exclusive, ceil, global
    (2, 112, 112, 64, 3, 2, 1, "max", True, False, False),   # ResNet-50 maxpool
    (4, 7, 7, 2048, 7, 1, 0, "avg", True, False, True),      # global average
    (2, 7, 7, 1024, 7, 1, 0, "avg", True, False, True),
    (1, 96, 96, 36, 3, 2, 1, "max", True, False, False),     # humanseg
    (1, 13, 13, 24, 3, 2, 1, "avg", True, True, False),
    (1, 13, 13, 24, 3, 2, 1, "avg", False, False, False),
    (1, 10, 10, 9, 2, 2, 0, "max", True, False, False),
]


@pytest.mark.parametrize("case", POOL_CASES)
@pytest.mark.parametrize("out", ["int8", "fp32"])
def test_pool_parity(pl, case, out):
    N, H, W, Cc, k, s, pd, typ, excl, ceil, glob = case
    rng = np.random.default_rng(hash(case) & 0xFFFF)
    x = rng.integers(-127, 128, (N, H, W, Cc), dtype=np.int8)
    t = orc.POOL_MAX if typ == "max" else orc.POOL_AVG
    o = orc.OUT_INT8 if out == "int8" else orc.OUT_FP32
    resc = 0.7 if out == "int8" else 0.013
    want = orc.pool2d(x, (k, k), (s, s), (pd, pd, pd, pd), t, excl, resc, o, ceil_mode=ceil, global_pooling=glob)
    ld = rup(Cc, 16)
    got = host(pl.pool2d(dev(pad_c(x, ld)), (k, k), (s, s), (pd, pd, pd, pd), t, excl, resc, o,
                         ceil_mode=ceil, global_pooling=glob, Cc=Cc))
    assert np.array_equal(got[..., :Cc], want)


@pytest.mark.parametrize("shape", [(2, 112, 112, 64), (1, 96, 96, 36), (2, 9, 9, 24), (1, 10, 10, 9)])
def test_max_pool_same_scale(pl, shape):
    """rescale == 1 (the scale-preserving max pool of every net here) takes the packed path: the result is
    the window maximum itself, clamped at -127 like the ARM requantiser (inputs include -128)."""
    N, H, W, Cc = shape
    rng = np.random.default_rng(Cc)
    x = rng.integers(-128, 128, shape, dtype=np.int8)
    x[:, ::3, ::3, :] = -128  # whole windows of -128 at the borders
    want = orc.pool2d(x, (3, 3), (2, 2), (1, 1, 1, 1), orc.POOL_MAX, True, 1.0, orc.OUT_INT8)
    ld = rup(Cc, 16)
    got = host(pl.pool2d(dev(pad_c(x, ld)), (3, 3), (2, 2), (1, 1, 1, 1), orc.POOL_MAX, True, 1.0, orc.OUT_INT8, Cc=Cc))
    assert np.array_equal(got[..., :Cc], want)
    assert want.min() >= -127


@pytest.mark.parametrize("n", [1, 16, 1000, 256 * 56 * 56 * 4 + 5])
def test_residual_add_parity(pl, n):
    rng = np.random.default_rng(n)
    a = rng.integers(-127, 128, n, dtype=np.int8)
    b = rng.integers(-127, 128, n, dtype=np.int8)
    for act in (orc.ACT_RELU, orc.ACT_NONE):
        want = orc.add(a, b, 0.61, 0.47, act)
        assert np.array_equal(host(pl.add(dev(a), dev(b), 0.61, 0.47, act)), want)
    want = orc.add(a, b, 0.011, 0.02, orc.ACT_RELU, out=orc.OUT_FP32)
    assert np.array_equal(host(pl.add(dev(a), dev(b), 0.011, 0.02, pl.ACT_RELU, out=pl.OUT_FP32)), want)


def test_launch_counter_and_peak_probe(pl):
    n0 = pl.launch_count()
    out = pl.i8_mma_peak(4, 8, 256)
    torch.cuda.synchronize()
    assert pl.launch_count() == n0 + 1
    # operands are all ones: every accumulator of the last-written stage is a multiple of 128
    assert int(out[0].item()) % 128 == 0 and int(out[0].item()) > 0


@pytest.mark.parametrize("case", [(2, 20, 22, 3, 32, 3, 2, 1), (1, 38, 38, 3, 64, 7, 2, 3), (1, 9, 9, 4, 16, 5, 1, 2)])
def test_thin_first_layer_im2col_gemm(pl, case):
    """C <= 4 first layers: pack transform (explicit im2col, one word per pixel) + GEMM
    equals the oracle's direct convolution."""
    N, H, W, Cc, K, k, st, pd = case
    rng = np.random.default_rng(sum(case))
    x = rng.integers(-127, 128, (N, H, W, Cc), dtype=np.int8)
    w = rng.integers(-127, 128, (K, k, k, Cc), dtype=np.int8)
    ws = ((1 + np.arange(K) % 7) / (7 * 127)).astype(np.float32)
    bias = rng.uniform(-1, 1, K).astype(np.float32)
    sf, bf, _ = orc.fold_scales(ws, K, 1 / 127, Cc * k * k / 127 / 4, orc.OUT_INT8, bias)
    want = orc.conv2d(x, w, sf, bf, (st, st), (pd,) * 4, act=orc.ACT_RELU)
    A = pl.im2col(dev(pad_c(x, 4)), k, k, (st, st), (pd,) * 4)
    P = orc.conv_out_size(H, k, 1, pd, pd, st)
    Q = orc.conv_out_size(W, k, 1, pd, pd, st)
    # the packed rows are the (r, s, c4) patches
    a = host(A).reshape(N, P, Q, -1)
    xp = np.zeros((N, H + 2 * pd, W + 2 * pd, 4), np.int8)
    xp[:, pd:pd + H, pd:pd + W, :Cc] = x
    for (n, p_, q_) in ((0, 0, 0), (N - 1, P - 1, Q - 1), (0, P // 2, Q // 3)):
        patch = xp[n, p_ * st:p_ * st + k, q_ * st:q_ * st + k, :].reshape(-1)
        assert np.array_equal(a[n, p_, q_, : k * k * 4], patch)
        assert not a[n, p_, q_, k * k * 4:].any()
    w4 = np.zeros((K, k, k, 4), np.int8)
    w4[..., :Cc] = w
    bt = np.zeros((K, A.shape[1]), np.int8)
    bt[:, : k * k * 4] = w4.reshape(K, -1)
    got = host(pl.gemm(A, dev(bt), dev(sf), dev(bf), act=pl.ACT_RELU)).reshape(N, P, Q, K)
    assert np.array_equal(got, want)


# ---- thin first layers, direct (on-chip window gather, no im2col matrix) -------------
THIN_DIRECT = [
    # N, H, W, C, K, R, S, stride, pads(t,b,l,r)
    (2, 32, 32, 3, 64, 7, 7, 2, (3, 3, 3, 3)),     # ResNet stem shape (scaled down)
    (3, 30, 34, 3, 32, 3, 3, 2, (1, 1, 1, 1)),     # MobileNet stem shape
    (1, 19, 23, 4, 24, 5, 5, 1, (2, 2, 2, 2)),
    (2, 17, 16, 1, 16, 3, 3, 1, (1, 1, 1, 1)),
    (1, 21, 20, 2, 40, 3, 7, 2, (1, 0, 3, 2)),     # 3 x 7 window, asymmetric pads
    (1, 9, 9, 3, 8, 7, 5, 1, (3, 3, 2, 2)),
    (5, 13, 11, 3, 130, 3, 3, 1, (0, 0, 0, 0)),    # N tail + two N tiles
]


@pytest.mark.parametrize("case", THIN_DIRECT)
@pytest.mark.parametrize("mode", ["int8_relu", "int8_none", "fp32", "int32"])
def test_thin_direct_conv(pl, case, mode):
    N, H, W, Cc, K, R, S, st, pads = case
    rng = np.random.default_rng(sum(case[:8]) + len(mode))
    x = rng.integers(-128, 128, (N, H, W, Cc), dtype=np.int8)
    w = rng.integers(-127, 128, (K, R, S, Cc), dtype=np.int8)
    assert pl.thin_conv_supported(R, S, Cc)
    ws = ((1 + np.arange(K) % 7) / (7 * 127)).astype(np.float32)
    bias = rng.uniform(-1, 1, K).astype(np.float32)
    out = {"int8_relu": orc.OUT_INT8, "int8_none": orc.OUT_INT8, "fp32": orc.OUT_FP32, "int32": orc.OUT_INT32}[mode]
    act = orc.ACT_RELU if mode == "int8_relu" else orc.ACT_NONE
    sf, bf, _ = orc.fold_scales(ws, K, 1 / 127, Cc * R * S / 127 / 4, orc.OUT_INT8 if out == orc.OUT_INT8 else orc.OUT_FP32, bias)
    want = orc.conv2d(x, w, sf, bf, (st, st), pads, act=act, out=out)
    x4 = np.full((N, H, W, 4), 77, np.int8)  # junk in the pad channels: the packed filter zeroes them
    x4[..., :Cc] = x
    wt = pl.pack_thin_conv_weight(dev(w))
    got = host(pl.thin_conv2d(dev(x4), wt, K, R, S, Cc, dev(sf), dev(bf), (st, st), pads, act=act, out=out))
    if out == orc.OUT_FP32:
        np.testing.assert_allclose(got, want, rtol=1e-5, atol=1e-6)
    else:
        assert np.array_equal(got, want)


def test_thin_direct_full_size_stem(pl):
    """ResNet-50 stem at full spatial size, batch 8: direct kernel == explicit im2col + GEMM."""
    rng = np.random.default_rng(5)
    N, H, W, Cc, K = 8, 224, 224, 3, 64
    x4 = np.zeros((N, H, W, 4), np.int8)
    x4[..., :Cc] = rng.integers(-127, 128, (N, H, W, Cc), dtype=np.int8)
    w = rng.integers(-127, 128, (K, 7, 7, Cc), dtype=np.int8)
    sf = np.full(K, 1e-3, np.float32)
    bf = rng.uniform(-1, 1, K).astype(np.float32)
    xd = dev(x4)
    got = pl.thin_conv2d(xd, pl.pack_thin_conv_weight(dev(w)), K, 7, 7, Cc, dev(sf), dev(bf), (2, 2), (3, 3, 3, 3),
                         act=pl.ACT_RELU)
    A = pl.im2col(xd, 7, 7, (2, 2), (3, 3, 3, 3))
    w4 = np.zeros((K, 7, 7, 4), np.int8)
    w4[..., :Cc] = w
    bt = np.zeros((K, A.shape[1]), np.int8)
    bt[:, : 49 * 4] = w4.reshape(K, -1)
    ref = pl.gemm(A, dev(bt), dev(sf), dev(bf), act=pl.ACT_RELU).reshape(got.shape)
    assert torch.equal(got, ref)
    # and a slice against the oracle
    want = orc.conv2d(x4[:1, :, :, :Cc], w, sf, bf, (2, 2), (3, 3, 3, 3), act=orc.ACT_RELU)
    assert np.array_equal(host(got[:1]), want)


def test_thin_direct_rejects_unsupported(pl):
    assert not pl.thin_conv_supported(3, 3, 8)
    assert not pl.thin_conv_supported(9, 9, 3)
    assert not pl.thin_conv_supported(3, 3, 3, (2, 2))


# ---- stride-2 thin stems in space-to-depth form (bulk-copied spans, overlapping descriptors) ----
S2D_CASES = [
    # N, H, W, C, K, R, stride, pads(t,b,l,r)
    (2, 32, 32, 3, 64, 7, 2, (3, 3, 3, 3)),      # ResNet stem (scaled down)
    (3, 30, 34, 3, 32, 3, 2, (1, 1, 1, 1)),      # MobileNet stem
    (1, 64, 224, 3, 64, 7, 2, (3, 3, 3, 3)),     # full-width rows (Q = 112)
    (2, 20, 256, 4, 16, 3, 2, (1, 1, 1, 1)),     # Q = 128: every accumulator row is an output
    (1, 18, 22, 1, 24, 5, 2, (2, 2, 2, 2)),      # even pad: shift 0, 3 tap rows padded to 4 taps per row
    (4, 12, 16, 2, 130, 3, 2, (1, 0, 1, 0)),     # no bottom / right pad, K not a multiple of 16
    (1, 16, 16, 3, 8, 3, 2, (0, 0, 0, 0)),       # no padding at all
]


@pytest.mark.parametrize("case", S2D_CASES)
@pytest.mark.parametrize("mode", ["int8_relu", "int8_none", "fp32", "int32"])
def test_s2d_stem_conv(pl, case, mode):
    N, H, W, Cc, K, R, st, pads = case
    rng = np.random.default_rng(sum(case[:7]) + len(mode))
    g = pl.s2d_geometry(N, H, W, Cc, R, R, (st, st), pads)
    assert g.supported
    xf = rng.uniform(-1, 1, (N, Cc, H, W)).astype(np.float32)
    in_scale = 1 / 127
    x = orc.nchw_to_nhwc(orc.quantize(xf, in_scale))  # what the s2d quantiser must hold, in NHWC
    w = rng.integers(-127, 128, (K, R, R, Cc), dtype=np.int8)
    ws = ((1 + np.arange(K) % 7) / (7 * 127)).astype(np.float32)
    bias = rng.uniform(-1, 1, K).astype(np.float32)
    out = {"int8_relu": orc.OUT_INT8, "int8_none": orc.OUT_INT8, "fp32": orc.OUT_FP32, "int32": orc.OUT_INT32}[mode]
    act = orc.ACT_RELU if mode == "int8_relu" else orc.ACT_NONE
    sf, bf, _ = orc.fold_scales(ws, K, in_scale, Cc * R * R / 127 / 4, orc.OUT_INT8 if out == orc.OUT_INT8 else orc.OUT_FP32, bias)
    want = orc.conv2d(x, w, sf, bf, (st, st), pads, act=act, out=out)
    xs = pl.quantize_nchw_to_s2d(dev(xf), in_scale, g)
    wp = pl.pack_s2d_conv_weight(dev(w), g)
    got = host(pl.s2d_conv2d(xs, wp, N, H, W, Cc, K, R, R, dev(sf), dev(bf), (st, st), pads, act=act, out=out))
    if out == orc.OUT_FP32:
        np.testing.assert_allclose(got, want, rtol=1e-5, atol=1e-6)
    else:
        assert np.array_equal(got, want)


def test_s2d_quantiser_layout(pl):
    """The 16-byte pixel is (dy, dx, c) of a 2x2 block; borders stay zero."""
    rng = np.random.default_rng(11)
    N, Cc, H, W = 2, 3, 8, 12
    g = pl.s2d_geometry(N, H, W, Cc, 7, 7, (2, 2), (3, 3, 3, 3))
    xf = rng.uniform(-1, 1, (N, Cc, H, W)).astype(np.float32)
    q = orc.quantize(xf, 1 / 127)
    buf = host(pl.quantize_nchw_to_s2d(dev(xf), 1 / 127, g))
    img = buf[: N * g.HP * g.WS * 16].reshape(N, g.HP, g.WS, 2, 2, 4)
    want = np.zeros_like(img)
    for dy in range(2):
        for dx in range(2):
            want[:, g.pad_s:g.pad_s + H // 2, g.pad_s:g.pad_s + W // 2, dy, dx, :Cc] = \
                q[:, :, dy::2, dx::2].transpose(0, 2, 3, 1)
    assert np.array_equal(img, want)
    assert not buf[N * g.HP * g.WS * 16:].any()


@pytest.mark.parametrize("W", [12, 10, 224])
@pytest.mark.parametrize("feed", ["f16", "u8"])
def test_s2d_quantiser_narrow_feeds(pl, feed, W):
    """fp16 / uint8 feeds of the s2d entry kernel == the oracle's calib on the same fp32 values (u8: after the
    image normalisation (u8 - mean) * norm of lite/utils/cv/image2tensor.cc, evaluated in fp32)."""
    rng = np.random.default_rng(13)
    N, Cc, H = 2, 3, 8
    g = pl.s2d_geometry(N, H, W, Cc, 7, 7, (2, 2), (3, 3, 3, 3))
    if feed == "f16":
        xh = rng.uniform(-1, 1, (N, Cc, H, W)).astype(np.float16)
        xf = xh.astype(np.float32)
        scale = 1 / 127
        buf = host(pl.quantize_nchw_to_s2d(dev(xh), scale, g))
    else:
        xu = rng.integers(0, 256, (N, Cc, H, W), dtype=np.uint8)
        mean = np.array([123.675, 116.28, 103.53], np.float32)
        norm = (1.0 / np.array([58.395, 57.12, 57.375], np.float32)).astype(np.float32)
        xf = (xu.astype(np.float32) - mean.reshape(1, 3, 1, 1)) * norm.reshape(1, 3, 1, 1)
        scale = float(np.abs(xf).max()) / 127
        buf = host(pl.quantize_nchw_to_s2d(dev(xu), scale, g, mean=mean, norm=norm))
    q = orc.quantize(xf, scale)
    img = buf[: N * g.HP * g.WS * 16].reshape(N, g.HP, g.WS, 2, 2, 4)
    want = np.zeros_like(img)
    for dy in range(2):
        for dx in range(2):
            want[:, g.pad_s:g.pad_s + H // 2, g.pad_s:g.pad_s + W // 2, dy, dx, :Cc] = \
                q[:, :, dy::2, dx::2].transpose(0, 2, 3, 1)
    assert np.array_equal(img, want)


def test_s2d_geometry_rejects(pl):
    assert not pl.s2d_geometry(1, 32, 32, 3, 7, 7, (1, 1), (3, 3, 3, 3)).supported      # stride 1
    assert not pl.s2d_geometry(1, 31, 32, 3, 7, 7, (2, 2), (3, 3, 3, 3)).supported      # odd height
    assert not pl.s2d_geometry(1, 32, 32, 8, 3, 3, (2, 2), (1, 1, 1, 1)).supported      # not thin
    assert not pl.s2d_geometry(1, 32, 600, 3, 3, 3, (2, 2), (1, 1, 1, 1)).supported     # Q > 128
    assert not pl.s2d_geometry(1, 32, 32, 3, 9, 9, (2, 2), (4, 4, 4, 4)).supported      # 5 taps per axis


def test_humanseg_model_layers(pl):
    """BASELINE config 3: every distinct conv2d / depthwise_conv2d shape of the shipped
    humanseg_shfv75_0119_quant `__model__` (channels 3/9/18/36/72/144 -> padded pitches, s2d stem, 3x3 s1/s2
    depthwise), bit-exact against the oracle.  Shapes: tests/golden/humanseg_layers.json."""
    import importlib.util
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    spec = importlib.util.spec_from_file_location("humanseg_layers", os.path.join(root, "tools", "humanseg_layers.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    assert mod.check(batch=2) == 20


@pytest.mark.parametrize("shape,out_hw", [((2, 14, 14, 64), (3, 3)), ((1, 7, 9, 20), (2, 4)), ((3, 10, 10, 33), (1, 1)),
                                          ((1, 5, 5, 16), (5, 5)), ((2, 13, 6, 48), (4, 6))])
@pytest.mark.parametrize("typ", ["max", "avg"])
@pytest.mark.parametrize("out", ["int8", "fp32"])
def test_adaptive_pool(pl, shape, out_hw, typ, out):
    """PoolParam::adaptive (lite/kernels/cuda/pool_compute.cu:69-187 index rules) on int8 NHWC."""
    rng = np.random.default_rng(31)
    x = rng.integers(-128, 128, shape, dtype=np.int8)
    t = orc.POOL_MAX if typ == "max" else orc.POOL_AVG
    o = orc.OUT_INT8 if out == "int8" else orc.OUT_FP32
    want = orc.adaptive_pool2d(x, out_hw, t, 0.61, o)
    got = host(pl.adaptive_pool2d(dev(x), out_hw, t, 0.61, o))
    assert np.array_equal(got, want)


@pytest.mark.parametrize("M,K,N", [(1, 4096, 4096), (1, 1024, 1000), (2, 2048, 1000), (3, 100, 33), (4, 777, 130),
                                   (5, 1600, 64), (8, 4111, 257), (1, 16, 4), (1, 65600, 8)])
@pytest.mark.parametrize("out", ["int8", "fp32", "int32"])
def test_gemv_skinny_m(pl, M, K, N, out):
    """M <= 8 takes the weight-streaming kernel (gemv_i8.cu): K tails with garbage in A's pad bytes, N tails,
    k-splits across the warps of a CTA.  Same contract as the tensor-core GEMM."""
    rng = np.random.default_rng(M * 1000 + N)
    ldk = rup(K, 16)
    A = rng.integers(-127, 128, (M, ldk), dtype=np.int8)       # pad bytes are NOT zero
    Bt = np.zeros((N, ldk), np.int8)
    Bt[:, :K] = rng.integers(-127, 128, (N, K), dtype=np.int8)
    scale = (rng.random(N) * 2e-3 / np.sqrt(K) + 1e-5).astype(np.float32)
    bias = rng.standard_normal(N).astype(np.float32)
    o = {"int8": orc.OUT_INT8, "fp32": orc.OUT_FP32, "int32": orc.OUT_INT32}[out]
    want = orc.gemm(A[:, :K], Bt[:, :K], scale, bias, b_is_nk=True, act=orc.ACT_RELU, out=o)
    got = host(pl.gemm(dev(A), dev(Bt), dev(scale), dev(bias), act=pl.ACT_RELU, out=o, K=K))
    if o == orc.OUT_FP32:
        assert np.allclose(got, want, rtol=1e-5, atol=1e-6)
    else:
        assert np.array_equal(got, want)


@pytest.mark.parametrize("case", [(2, 32, 32, 3, 7, 3), (3, 20, 28, 3, 3, 1), (1, 16, 16, 4, 7, 3), (2, 12, 12, 1, 3, 1)])
def test_nhwc_to_s2d_matches_fused_quantiser(pl, case):
    """the int8 entry of the space-to-depth stems (plugin chain calib -> layout -> conv) writes the very bytes the
    fused fp32 quantiser writes"""
    N, H, W, Cc, k, pad = case
    rng = np.random.default_rng(3)
    xf = rng.uniform(-1, 1, (N, Cc, H, W)).astype(np.float32)
    g = pl.s2d_geometry(N, H, W, Cc, k, k, (2, 2), (pad,) * 4)
    assert g.supported
    want = host(pl.quantize_nchw_to_s2d(dev(xf), 1 / 127, g))
    q = orc.nchw_to_nhwc(orc.quantize(xf, 1 / 127))  # int8 NHWC, dense C channels
    got = host(pl.nhwc_to_s2d(dev(q), g))
    assert np.array_equal(got, want)


@pytest.mark.parametrize("C,ld", [(3, 4), (9, 16), (18, 32), (72, 80), (16, 16), (5, 8)])
def test_repitch(pl, C, ld):
    rng = np.random.default_rng(C)
    x = rng.integers(-128, 128, (3, 5, 7, C), dtype=np.int8)
    got = host(pl.repitch(dev(x), ld))
    assert got.shape == (3, 5, 7, ld)
    assert np.array_equal(got[..., :C], x) and not got[..., C:].any()
