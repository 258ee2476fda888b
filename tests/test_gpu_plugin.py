"""GPU parity of the kernel CLASSES (REGISTER_LITE_KERNEL surface) against the oracle:
each test creates the kernel through the registry, SetParam()s the reference's param
struct, Launch()es it and compares with orc on the same seeded inputs.  Modelled on
lite/kernels/cuda/conv_compute_test.cc:277-420 and lite/tests/math/conv_int8_compute_test.cc."""
import numpy as np
import pytest

import orc

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def lh():
    import __graft_entry__ as ge
    return ge.load_package().lite_harness


def _rand_i8(rng, shape, lo=-127, hi=127):
    return rng.integers(lo, hi + 1, size=shape, dtype=np.int64).astype(np.int8)


def _conv_case(lh, rng, N, H, W, Cin, K, R, S, stride, pads, dil, groups, act, act_coef, out_int8, bias=True,
               per_channel=True, residual=False, launches=1, op=None):
    cg = Cin // groups
    x = _rand_i8(rng, (N, H, W, Cin))
    w = _rand_i8(rng, (K, R, S, cg), -63, 63)
    b = rng.standard_normal(K).astype(np.float32) if bias else None
    in_scale = 0.023
    ws = (rng.random(K).astype(np.float32) * 0.01 + 0.001) if per_channel else np.array([0.004], np.float32)
    out_scale = 0.11
    out = orc.OUT_INT8 if out_int8 else orc.OUT_FP32
    sf, bf, aa = orc.fold_scales(ws, K, in_scale, out_scale, out, b, act, act_coef)
    P = orc.conv_out_size(H, R, dil[0], pads[0], pads[1], stride[0])
    Q = orc.conv_out_size(W, S, dil[1], pads[2], pads[3], stride[1])
    res = _rand_i8(rng, (N, P, Q, K)) if residual else None
    rs = 0.05
    want = orc.conv2d(x, w, sf, bf if bias else None, stride, pads, dil, groups, act, aa, out,
                      residual=res, res_scale=(rs / out_scale if out_int8 else rs))
    op = op or ("depthwise_conv2d" if groups == Cin and groups == K and groups > 1 else "conv2d")
    got, bias_after = lh.run_conv(op, x, w, b, in_scale, ws, out_scale, stride, pads, dil, groups, act, act_coef,
                                  out_int8, res, rs, launches)
    assert got.shape == want.shape
    if bias:
        np.testing.assert_array_equal(bias_after, b)  # the shared bias tensor is NOT scaled in place
    if out_int8:
        np.testing.assert_array_equal(got, want)
    else:
        np.testing.assert_allclose(got, want, rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("out_int8", [True, False])
@pytest.mark.parametrize("act,coef", [(0, 0.0), (1, 0.0), (2, 6.0), (4, 0.1)])
def test_conv2d_dense_3x3(lh, out_int8, act, coef):
    _conv_case(lh, np.random.default_rng(1), 2, 14, 15, 32, 48, 3, 3, (1, 1), (1, 1, 1, 1), (1, 1), 1, act, coef, out_int8)


@pytest.mark.parametrize("Cin,K", [(3, 32), (4, 16), (20, 24), (72, 40), (64, 256), (130, 8)])
def test_conv2d_channel_shapes(lh, Cin, K):
    # C <= 4 takes the im2col+GEMM path, C % 16 != 0 the padded-copy path
    _conv_case(lh, np.random.default_rng(Cin), 2, 12, 13, Cin, K, 3, 3, (2, 2), (1, 1, 1, 1), (1, 1), 1, 1, 0.0, True)


@pytest.mark.parametrize("R,S,stride,pads,dil", [
    (1, 1, (1, 1), (0, 0, 0, 0), (1, 1)),
    (1, 1, (2, 2), (0, 0, 0, 0), (1, 1)),
    (5, 5, (1, 1), (2, 2, 2, 2), (1, 1)),
    (7, 7, (2, 2), (3, 3, 3, 3), (1, 1)),
    (3, 3, (1, 1), (2, 2, 2, 2), (2, 2)),
    (3, 2, (2, 1), (1, 0, 0, 1), (1, 1)),
])
def test_conv2d_filter_geometry(lh, R, S, stride, pads, dil):
    _conv_case(lh, np.random.default_rng(R * 10 + S), 1, 17, 19, 32, 32, R, S, stride, pads, dil, 1, 0, 0.0, True,
               per_channel=False)


def test_conv2d_launch_twice_is_idempotent(lh):
    # cudnn_conv.cc:361-364 rescales the shared bias on every init(); ours must not
    _conv_case(lh, np.random.default_rng(5), 1, 9, 9, 16, 16, 3, 3, (1, 1), (1, 1, 1, 1), (1, 1), 1, 1, 0.0, True,
               launches=3)


@pytest.mark.parametrize("out_int8", [True, False])
def test_conv2d_fused_residual(lh, out_int8):
    _conv_case(lh, np.random.default_rng(6), 2, 8, 8, 64, 128, 1, 1, (1, 1), (0, 0, 0, 0), (1, 1), 1, 1, 0.0,
               out_int8, residual=True)


@pytest.mark.parametrize("C,R,stride,pad", [(32, 3, 1, 1), (48, 3, 2, 1), (64, 5, 1, 2), (16, 5, 2, 2), (24, 3, 1, 0)])
@pytest.mark.parametrize("out_int8", [True, False])
def test_depthwise_conv2d(lh, C, R, stride, pad, out_int8):
    _conv_case(lh, np.random.default_rng(C + R), 2, 13, 11, C, C, R, R, (stride, stride), (pad,) * 4, (1, 1), C, 2,
               6.0, out_int8)


def test_depthwise_registered_under_conv2d_too(lh):
    # the op type stays "conv2d" with groups == C in many models (conv_op.h picks by param)
    _conv_case(lh, np.random.default_rng(9), 1, 10, 10, 32, 32, 3, 3, (1, 1), (1, 1, 1, 1), (1, 1), 32, 1, 0.0, True,
               op="conv2d")


@pytest.mark.parametrize("out_int8", [True, False])
def test_grouped_conv_through_the_registry(lh, out_int8):
    # 1 < groups < C through ConvComputeInt8 (the reference's int8 conv takes any groups, conv_gemmlike.cc:84-264)
    _conv_case(lh, np.random.default_rng(12), 2, 12, 12, 64, 96, 3, 3, (1, 1), (1, 1, 1, 1), (1, 1), 2, 1, 0.0, out_int8)
    _conv_case(lh, np.random.default_rng(13), 1, 9, 9, 128, 64, 1, 1, (1, 1), (0, 0, 0, 0), (1, 1), 4, 0, 0.0, out_int8)


def test_conv2d_errors(lh):
    rng = np.random.default_rng(3)
    x, w = _rand_i8(rng, (1, 8, 8, 32)), _rand_i8(rng, (32, 3, 3, 8))
    with pytest.raises(lh.LiteError, match="group"):  # C / groups = 8: slices not 16-byte aligned
        lh.run_conv("conv2d", x, w, None, 0.1, [0.1], 0.1, groups=4)
    w = _rand_i8(rng, (32, 3, 3, 32))
    with pytest.raises(lh.LiteError, match="weight_scale"):
        lh.run_conv("conv2d", x, w, None, 0.1, [0.1, 0.2], 0.1)


@pytest.mark.parametrize("M,K,N", [(1, 1024, 1000), (7, 100, 33), (64, 512, 512), (3, 17, 5)])
@pytest.mark.parametrize("out_int8", [True, False])
def test_fc(lh, M, K, N, out_int8):
    rng = np.random.default_rng(M + K + N)
    x, w = _rand_i8(rng, (M, K)), _rand_i8(rng, (K, N), -63, 63)
    b = rng.standard_normal(N).astype(np.float32)
    ws = rng.random(N).astype(np.float32) * 0.01 + 0.001
    out = orc.OUT_INT8 if out_int8 else orc.OUT_FP32
    sf, bf, _ = orc.fold_scales(ws, N, 0.02, 0.3, out, b)
    want = orc.gemm(x, w, sf, bf, act=orc.ACT_RELU, out=out)
    got = lh.run_fc(x, w, b, 0.02, ws, 0.3, 1, "relu", out_int8, launches=2)
    if out_int8:
        np.testing.assert_array_equal(got, want)
    else:
        np.testing.assert_allclose(got, want, rtol=1e-5, atol=1e-6)


def test_fc_flattens_leading_dims(lh):
    rng = np.random.default_rng(11)
    x, w = _rand_i8(rng, (2, 3, 4, 8)), _rand_i8(rng, (32, 10))
    sf, bf, _ = orc.fold_scales([0.01], 10, 0.02, 0.3, orc.OUT_INT8, None)
    want = orc.gemm(x.reshape(6, 32), w, sf, None)
    got = lh.run_fc(x, w, None, 0.02, [0.01], 0.3, in_num_col_dims=2)
    assert got.shape == (2, 3, 10)
    np.testing.assert_array_equal(got.reshape(6, 10), want)
    with pytest.raises(lh.LiteError):
        lh.run_fc(x, w, None, 0.02, [0.01], 0.3, in_num_col_dims=1, act="tanh")


@pytest.mark.parametrize("out_int8", [True, False])
def test_mul(lh, out_int8):
    rng = np.random.default_rng(12)
    x, y = _rand_i8(rng, (4, 5, 6, 8)), _rand_i8(rng, (6, 8, 7, 3))
    out = orc.OUT_INT8 if out_int8 else orc.OUT_FP32
    sf, _, _ = orc.fold_scales([0.003], 21, 0.02, 0.7, out, None)
    want = orc.gemm(x.reshape(20, 48), y.reshape(48, 21), sf, None, out=out)
    got = lh.run_mul(x, y, 0.02, [0.003], 0.7, 2, 2, out_int8)
    assert got.shape == (4, 5, 7, 3)
    if out_int8:
        np.testing.assert_array_equal(got.reshape(20, 21), want)
    else:
        np.testing.assert_allclose(got.reshape(20, 21), want, rtol=1e-5)


@pytest.mark.parametrize("tx,ty", [(False, False), (True, False), (False, True), (True, True)])
@pytest.mark.parametrize("batched_y", [False, True])
def test_matmul(lh, tx, ty, batched_y):
    rng = np.random.default_rng(13 + tx * 2 + ty)
    B, M, K, N = 3, 9, 40, 21
    x = _rand_i8(rng, (B, K, M) if tx else (B, M, K))
    yshape = (N, K) if ty else (K, N)
    y = _rand_i8(rng, ((B,) + yshape) if batched_y else yshape)
    alpha = 0.5
    sf, _, _ = orc.fold_scales([0.004 * alpha], N, 0.02, 0.9, orc.OUT_INT8, None)
    got = lh.run_matmul(x, y, 0.02, [0.004], 0.9, tx, ty, alpha, True)
    assert got.shape == (B, M, N)
    for b in range(B):
        yb = y[b] if batched_y else y
        want = orc.gemm(x[b], yb, sf, None, trans_a=tx, b_is_nk=ty)
        np.testing.assert_array_equal(got[b], want)


def test_calib_both_directions(lh):
    rng = np.random.default_rng(14)
    x = (rng.standard_normal((3, 5, 7, 9)) * 3).astype(np.float32)
    x.flat[:6] = [0.5 * 0.05, -0.5 * 0.05, 1.5 * 0.05, 200.0, -200.0, 0.0]
    q = lh.run_calib(x, 0.05, True)
    np.testing.assert_array_equal(q, orc.quantize(x, 0.05))
    f = lh.run_calib(q, 0.05, False)
    np.testing.assert_array_equal(f, orc.dequantize(q, 0.05))


@pytest.mark.parametrize("dtype", [np.int8, np.float32])
def test_layout_both_directions(lh, dtype):
    rng = np.random.default_rng(15)
    x = (rng.standard_normal((2, 5, 6, 7)) * 50).astype(dtype)
    y = lh.run_layout(x, True)
    np.testing.assert_array_equal(y, x.transpose(0, 2, 3, 1))
    np.testing.assert_array_equal(lh.run_layout(y, False), x)


def test_io_copy(lh):
    x = np.arange(1000, dtype=np.float32).reshape(10, 100)
    np.testing.assert_array_equal(lh.run_io_copy_roundtrip(x), x)


@pytest.mark.parametrize("ptype,k,s,p,glob,excl,ceil", [
    ("max", 3, 2, 1, False, True, False),
    ("max", 3, 2, 0, False, True, True),
    ("avg", 2, 2, 0, False, True, False),
    ("avg", 3, 1, 1, False, False, False),
    ("avg", 7, 1, 0, True, True, False),
])
@pytest.mark.parametrize("out_int8", [True, False])
def test_pool2d(lh, ptype, k, s, p, glob, excl, ceil, out_int8):
    rng = np.random.default_rng(16)
    x = _rand_i8(rng, (2, 14, 14, 48))
    out = orc.OUT_INT8 if out_int8 else orc.OUT_FP32
    resc = np.float32(0.03) / np.float32(0.05) if out_int8 else np.float32(0.03)
    want = orc.pool2d(x, (k, k), (s, s), (p,) * 4, 0 if ptype == "max" else 1, excl, resc, out, ceil_mode=ceil,
                      global_pooling=glob)
    got = lh.run_pool(x, ptype, (k, k), (s, s), (p,) * 4, 0.03, 0.05, glob, excl, ceil, out_int8)
    assert got.shape == want.shape
    if out_int8:
        np.testing.assert_array_equal(got, want)
    else:
        np.testing.assert_allclose(got, want, rtol=1e-6)


@pytest.mark.parametrize("act", ["", "relu"])
@pytest.mark.parametrize("out_int8", [True, False])
def test_elementwise_add(lh, act, out_int8):
    rng = np.random.default_rng(17)
    a, b = _rand_i8(rng, (2, 7, 7, 64)), _rand_i8(rng, (2, 7, 7, 64))
    out = orc.OUT_INT8 if out_int8 else orc.OUT_FP32
    so = np.float32(0.08)
    sa = np.float32(0.03) / so if out_int8 else np.float32(0.03)
    sb = np.float32(0.05) / so if out_int8 else np.float32(0.05)
    want = orc.add(a, b, sa, sb, orc.ACT_RELU if act else orc.ACT_NONE, 0.0, out)
    got = lh.run_add(a, b, 0.03, 0.05, 0.08, act, out_int8)
    if out_int8:
        np.testing.assert_array_equal(got, want)
    else:
        np.testing.assert_allclose(got, want, rtol=1e-6)


# ---- channel movement kernels, through the registry ----------------
def _nchw(x):
    return np.ascontiguousarray(np.moveaxis(x, -1, 1))


def _nhwc(x):
    return np.ascontiguousarray(np.moveaxis(x, 1, -1))


@pytest.mark.parametrize("cs", [(36, 36), (64, 64), (9, 18, 3)])
def test_concat_and_split_kernels(lh, cs):
    rng = np.random.default_rng(sum(cs))
    xs = [_rand_i8(rng, (2, 12, 12, c), -128, 127) for c in cs]
    want = _nhwc(orc.concat_nchw([_nchw(x) for x in xs]))
    got = lh.run_concat(xs, axis=1)
    np.testing.assert_array_equal(got, want)
    parts = lh.run_split(want, axis=1, sections=cs)
    for p, x in zip(parts, xs):
        np.testing.assert_array_equal(p, x)
    if len(set(cs)) == 1:  # `num` equal parts
        for p, x in zip(lh.run_split(want, axis=1, num=len(cs)), xs):
            np.testing.assert_array_equal(p, x)
    with pytest.raises(lh.LiteError):
        lh.run_concat(xs, axis=2)  # only the channel axis keeps NHWC tensors resident


@pytest.mark.parametrize("C,group", [(72, 2), (16, 4), (18, 2), (144, 2)])
def test_shuffle_channel_kernel(lh, C, group):
    rng = np.random.default_rng(C + group)
    x = _rand_i8(rng, (2, 12, 12, C), -128, 127)
    want = _nhwc(orc.shuffle_channel_nchw(_nchw(x), group))
    np.testing.assert_array_equal(lh.run_shuffle_channel(x, group), want)
    with pytest.raises(lh.LiteError):
        lh.run_shuffle_channel(x, 5)


def test_relu_kernels(lh):
    rng = np.random.default_rng(23)
    x = _rand_i8(rng, (2, 9, 9, 40), -128, 127)
    np.testing.assert_array_equal(lh.run_relu(x), orc.relu_i8(x))
    # relu6 on a tensor of scale 0.1: the clip sits at round(6 / 0.1) = 60
    np.testing.assert_array_equal(lh.run_relu(x, relu6=True, clip=6.0, input_scale=0.1), orc.relu_i8(x, 60))


@pytest.mark.parametrize("typ", ["max", "avg"])
def test_adaptive_pool_through_the_registry(lh, typ):
    """PoolParam::adaptive (ksize = output size, lite/kernels/cuda/pool_compute.cu:256-350)."""
    rng = np.random.default_rng(41)
    x = _rand_i8(rng, (2, 13, 9, 24))
    want = orc.adaptive_pool2d(x, (4, 3), orc.POOL_MAX if typ == "max" else orc.POOL_AVG, 0.02 / 0.03, orc.OUT_INT8)
    got = lh.run_pool(x, typ, [4, 3], [1, 1], [0, 0, 0, 0], 0.02, 0.03, adaptive=True)
    np.testing.assert_array_equal(got, want)
