"""ctypes window onto oracle/liboracle.so (+ oracle/_ref/libref_naive.so).

Checker only: imported by tests/, __graft_entry__.smoke() and bench.py's CPU
legs.  Never imported by the product package.
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ODIR = os.path.join(ROOT, "oracle")

ACT_NONE, ACT_RELU, ACT_RELU6, ACT_LEAKY = 0, 1, 2, 4
OUT_INT8, OUT_FP32, OUT_INT32 = 0, 1, 2
ROUND_ARM, ROUND_CUDA = 0, 1
POOL_MAX, POOL_AVG = 0, 1

_NP_OUT = {OUT_INT8: np.int8, OUT_FP32: np.float32, OUT_INT32: np.int32}


def build(ref=True):
    subprocess.check_call(["make", "-s", "-C", ODIR, "all"])
    if ref and os.path.isdir("/root/reference"):
        subprocess.check_call(["make", "-s", "-C", ODIR, "ref"])


def _load(path):
    return C.CDLL(path) if os.path.exists(path) else None


_lib = None
_ref = None


def lib():
    global _lib
    if _lib is None:
        p = os.path.join(ODIR, "liboracle.so")
        if not os.path.exists(p):
            build(ref=False)
        _lib = C.CDLL(p)
        _lib.orc_epilogue_f32.restype = C.c_float
        _lib.orc_round_sat_i8.restype = C.c_int8
    return _lib


def ref():
    """The reference's own naive math (None when oracle/_ref was never built)."""
    global _ref
    if _ref is None:
        _ref = _load(os.path.join(ODIR, "_ref", "libref_naive.so"))
    return _ref


_ref_extra = None


def ref_extra():
    """Golden functions that live inside reference test files (pool_compute_ref, matrix_mul)."""
    global _ref_extra
    if _ref_extra is None:
        _ref_extra = _load(os.path.join(ODIR, "_ref", "libref_extra.so"))
    return _ref_extra


def set_num_threads(n):
    lib().orc_set_num_threads(int(n))
    if ref() is not None and hasattr(ref(), "ref_set_num_threads"):
        ref().ref_set_num_threads(int(n))


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _f(v):
    return C.c_float(float(v))


def conv_out_size(i, k, d, p0, p1, s):
    return (i + p0 + p1 - (d * (k - 1) + 1)) // s + 1


def fold_scales(w_scale, n_out, in_scale, out_scale, out_dtype, bias=None,
                act=ACT_NONE, act_arg=0.0):
    w_scale = np.ascontiguousarray(w_scale, np.float32)
    sf = np.empty(n_out, np.float32)
    bf = np.empty(n_out, np.float32)
    b = None if bias is None else np.ascontiguousarray(bias, np.float32)
    aa = C.c_float(0)
    rc = lib().orc_fold_scales(_p(w_scale), len(w_scale), n_out, _f(in_scale),
                               _f(out_scale), out_dtype, _p(b), _p(sf), _p(bf),
                               act, _f(act_arg), C.byref(aa))
    assert rc == 0
    return sf, bf, aa.value


def conv2d(x, w, scale_f, bias_f, strides=(1, 1), pads=(0, 0, 0, 0),
           dil=(1, 1), groups=1, act=ACT_NONE, act_arg=0.0, out=OUT_INT8,
           rnd=ROUND_ARM, residual=None, res_scale=0.0):
    """x NHWC int8, w OHWI int8 -> y NHWC.  residual: int8 NHWC of the output
    shape, fused as v = fmaf(res, res_scale, v) before the activation."""
    if residual is not None:
        acc = conv2d(x, w, None, None, strides, pads, dil, groups, out=OUT_INT32)
        N_, P_, Q_, K_ = acc.shape
        y = np.empty(acc.shape, _NP_OUT[out])
        res = np.ascontiguousarray(residual, np.int8)
        assert res.shape == acc.shape
        sf = np.ascontiguousarray(scale_f, np.float32)
        bf = None if bias_f is None else np.ascontiguousarray(bias_f, np.float32)
        rc = lib().orc_epilogue_res(_p(acc), C.c_int64(N_ * P_ * Q_), K_, _p(sf), _p(bf), _p(res), K_,
                                    _f(res_scale), act, _f(act_arg), out, rnd, _p(y))
        assert rc == 0
        return y
    N, H, W, Cc = x.shape
    K, R, S, _ = w.shape
    P = conv_out_size(H, R, dil[0], pads[0], pads[1], strides[0])
    Q = conv_out_size(W, S, dil[1], pads[2], pads[3], strides[1])
    y = np.empty((N, P, Q, K), _NP_OUT[out])
    x = np.ascontiguousarray(x)
    w = np.ascontiguousarray(w)
    sf = np.ascontiguousarray(scale_f, np.float32) if out != OUT_INT32 else None
    bf = None if bias_f is None else np.ascontiguousarray(bias_f, np.float32)
    rc = lib().orc_conv2d_i8_nhwc(_p(x), _p(w), _p(bf), _p(sf), _p(y), N, H, W,
                                  Cc, K, R, S, strides[0], strides[1], pads[0],
                                  pads[1], pads[2], pads[3], dil[0], dil[1],
                                  groups, act, _f(act_arg), out, rnd)
    assert rc == 0, rc
    return y


def gemm(A, B, scale_f, bias_f, trans_a=False, b_is_nk=False, act=ACT_NONE,
         act_arg=0.0, out=OUT_INT8, rnd=ROUND_ARM):
    A = np.ascontiguousarray(A)
    B = np.ascontiguousarray(B)
    M, K = (A.shape[1], A.shape[0]) if trans_a else A.shape
    N = B.shape[0] if b_is_nk else B.shape[1]
    y = np.empty((M, N), _NP_OUT[out])
    sf = np.ascontiguousarray(scale_f, np.float32) if out != OUT_INT32 else None
    bf = None if bias_f is None else np.ascontiguousarray(bias_f, np.float32)
    rc = lib().orc_gemm_i8(_p(A), _p(B), _p(bf), _p(sf), _p(y), M, N, K,
                           int(trans_a), int(b_is_nk), act, _f(act_arg), out,
                           rnd)
    assert rc == 0
    return y


def quantize(x, scale, rnd=ROUND_ARM):
    x = np.ascontiguousarray(x, np.float32)
    y = np.empty(x.shape, np.int8)
    lib().orc_quantize_f32_i8(_p(x), _p(y), C.c_int64(x.size), _f(scale), rnd)
    return y


def dequantize(x, scale):
    x = np.ascontiguousarray(x, np.int8)
    y = np.empty(x.shape, np.float32)
    lib().orc_dequantize_i8_f32(_p(x), _p(y), C.c_int64(x.size), _f(scale))
    return y


def nchw_to_nhwc(x):
    x = np.ascontiguousarray(x)
    N, Cc, H, W = x.shape
    y = np.empty((N, H, W, Cc), x.dtype)
    lib().orc_nchw_to_nhwc(_p(x), _p(y), N, Cc, H, W, x.itemsize)
    return y


def nhwc_to_nchw(x):
    x = np.ascontiguousarray(x)
    N, H, W, Cc = x.shape
    y = np.empty((N, Cc, H, W), x.dtype)
    lib().orc_nhwc_to_nchw(_p(x), _p(y), N, Cc, H, W, x.itemsize)
    return y


def pool2d(x, ksize, strides, pads, pool_type, exclusive=True, rescale=1.0,
           out=OUT_INT8, rnd=ROUND_ARM, ceil_mode=False, global_pooling=False):
    x = np.ascontiguousarray(x, np.int8)
    N, H, W, Cc = x.shape
    kh, kw = ksize
    pt, pb, pl, pr = pads
    if global_pooling:
        kh, kw, pt, pb, pl, pr = H, W, 0, 0, 0, 0
    if ceil_mode:
        P = -(-(H + pt + pb - kh) // strides[0]) + 1
        Q = -(-(W + pl + pr - kw) // strides[1]) + 1
    else:
        P = (H + pt + pb - kh) // strides[0] + 1
        Q = (W + pl + pr - kw) // strides[1] + 1
    y = np.empty((N, P, Q, Cc), _NP_OUT[out])
    lib().orc_pool2d_i8_nhwc(_p(x), _p(y), N, H, W, Cc, kh, kw, strides[0],
                             strides[1], pt, pl, P, Q, pool_type,
                             int(exclusive), _f(rescale), out, rnd)
    return y


def adaptive_pool2d(x, out_hw, pool_type, rescale=1.0, out=OUT_INT8, rnd=ROUND_ARM):
    x = np.ascontiguousarray(x, np.int8)
    N, H, W, Cc = x.shape
    P, Q = out_hw
    y = np.empty((N, P, Q, Cc), _NP_OUT[out])
    lib().orc_adaptive_pool2d_i8_nhwc(_p(x), _p(y), N, H, W, Cc, P, Q, pool_type, _f(rescale), out, rnd)
    return y


def conv2d_transpose_f32(x_nhwc, w_iohw, bias, strides, pads_tl, relu=False, out_scale=0.0):
    """x [N,H,W,Cin] fp32, w [Cin,Cout,kh,kw]; pads_tl = (top, left); symmetric pads assumed for the output size."""
    x = np.ascontiguousarray(x_nhwc, np.float32)
    w = np.ascontiguousarray(w_iohw, np.float32)
    N, H, W, Ci = x.shape
    _, Co, kh, kw = w.shape
    P = (H - 1) * strides[0] - 2 * pads_tl[0] + kh
    Q = (W - 1) * strides[1] - 2 * pads_tl[1] + kw
    y = np.empty((N, P, Q, Co), np.int8 if out_scale > 0 else np.float32)
    b = None if bias is None else np.ascontiguousarray(bias, np.float32)
    lib().orc_conv2d_transpose_f32_nhwc(_p(x), _p(w), _p(b), _p(y), N, H, W, Ci, Co, kh, kw, strides[0], strides[1],
                                        pads_tl[0], pads_tl[1], P, Q, int(relu), _f(out_scale))
    return y


def add(a, b, sa, sb, act=ACT_NONE, act_arg=0.0, out=OUT_INT8, rnd=ROUND_ARM):
    a = np.ascontiguousarray(a, np.int8)
    b = np.ascontiguousarray(b, np.int8)
    y = np.empty(a.shape, _NP_OUT[out])
    lib().orc_add_i8(_p(a), _p(b), _p(y), C.c_int64(a.size), _f(sa), _f(sb),
                     act, _f(act_arg), out, rnd)
    return y


def _ptr_array(arrs):
    return (C.c_void_p * len(arrs))(*[a.ctypes.data for a in arrs])


def concat_nchw(xs):
    """concat along axis 1 of NCHW arrays (any dtype)"""
    xs = [np.ascontiguousarray(x) for x in xs]
    N = xs[0].shape[0]
    inner = int(np.prod(xs[0].shape[2:])) * xs[0].itemsize
    cs = (C.c_int * len(xs))(*[x.shape[1] for x in xs])
    y = np.empty((N, sum(x.shape[1] for x in xs)) + xs[0].shape[2:], xs[0].dtype)
    lib().orc_concat_nchw(len(xs), _ptr_array(xs), cs, _p(y), N, C.c_int64(inner))
    return y


def split_nchw(x, sections):
    x = np.ascontiguousarray(x)
    N = x.shape[0]
    inner = int(np.prod(x.shape[2:])) * x.itemsize
    ys = [np.empty((N, c) + x.shape[2:], x.dtype) for c in sections]
    cs = (C.c_int * len(ys))(*sections)
    lib().orc_split_nchw(_p(x), len(ys), _ptr_array(ys), cs, N, C.c_int64(inner))
    return ys


def shuffle_channel_nchw(x, group):
    x = np.ascontiguousarray(x)
    N, Cc = x.shape[:2]
    y = np.empty_like(x)
    rc = lib().orc_shuffle_channel_nchw(_p(x), _p(y), group, N, Cc,
                                        C.c_int64(int(np.prod(x.shape[2:])) * x.itemsize))
    assert rc == 0
    return y


def relu_i8(x, hi=127):
    x = np.ascontiguousarray(x, np.int8)
    y = np.empty_like(x)
    lib().orc_relu_i8(_p(x), _p(y), C.c_int64(x.size), hi)
    return y


# ---- the reference's own naive math (oracle/_ref) ---------------------------
def ref_conv_acc_nchw(x_nchw, w_oihw, strides, pads_hw, dil, groups):
    r = ref()
    assert r is not None
    x = np.ascontiguousarray(x_nchw, np.int8)
    w = np.ascontiguousarray(w_oihw, np.int8)
    N, Cc, H, W = x.shape
    K, _, R, S = w.shape
    P = conv_out_size(H, R, dil[0], pads_hw[0], pads_hw[0], strides[0])
    Q = conv_out_size(W, S, dil[1], pads_hw[1], pads_hw[1], strides[1])
    y = np.zeros((N, K, P, Q), np.int32)
    r.ref_conv_basic_i8(_p(x), _p(y), N, K, P, Q, Cc, H, W, _p(w), groups, S, R,
                        strides[1], strides[0], dil[1], dil[0], pads_hw[1],
                        pads_hw[0])
    return y


def ref_gemm_acc(A, B, trans_a=False, trans_b=False):
    r = ref()
    assert r is not None
    A = np.ascontiguousarray(A, np.int8)
    B = np.ascontiguousarray(B, np.int8)
    M, K = (A.shape[1], A.shape[0]) if trans_a else A.shape
    N = B.shape[0] if trans_b else B.shape[1]
    c = np.zeros((M, N), np.int32)
    r.ref_basic_gemm_i8(int(trans_a), int(trans_b), M, N, K, _p(A), A.shape[1],
                        _p(B), B.shape[1], _p(c), N)
    return c


def ref_saturate_i8(v):
    """saturate_cast<int8_t>(float) of lite/backends/arm/math/saturate.h:144-147, elementwise."""
    r = ref()
    assert r is not None
    v = np.ascontiguousarray(v, np.float32)
    out = np.empty(v.shape, np.int8)
    r.ref_saturate_cast_i8_f32_n(_p(v), _p(out), C.c_longlong(v.size))
    return out


def ref_pool_nchw(x_nchw_f32, ksize, strides, pads, is_avg, exclusive=True, global_pooling=False, out_hw=None):
    """pool_compute_ref of lite/kernels/cuda/pool_compute_test.cc:150-249 (fp32 NCHW)."""
    r = ref_extra()
    assert r is not None
    x = np.ascontiguousarray(x_nchw_f32, np.float32)
    N, Cc, H, W = x.shape
    pt, pb, pl_, pr = pads
    if global_pooling:
        oh, ow = 1, 1
    elif out_hw:
        oh, ow = out_hw
    else:
        oh = (H + pt + pb - ksize[0]) // strides[0] + 1
        ow = (W + pl_ + pr - ksize[1]) // strides[1] + 1
    y = np.zeros((N, Cc, oh, ow), np.float32)
    r.ref_pool_compute(_p(x), _p(y), N, Cc, H, W, oh, ow, ksize[0], ksize[1], strides[0], strides[1], pt, pb,
                       pl_, pr, int(is_avg), int(exclusive), int(global_pooling))
    return y


def ref_matmul_f32(x, y, trans_x=False, trans_y=False, alpha=1.0):
    """mul_low_efficiency's recipe (lite/tests/kernels/matmul_compute_test.cc:48-112) on the reference's
    own matrix_mul / transpose: 2-D x, y."""
    r = ref_extra()
    assert r is not None
    x = np.ascontiguousarray(x, np.float32)
    y = np.ascontiguousarray(y, np.float32)
    if trans_x:
        xt = np.empty((x.shape[1], x.shape[0]), np.float32)
        r.ref_transpose(x.shape[0], x.shape[1], _p(x), _p(xt))
        x = xt
    if trans_y:
        yt = np.empty((y.shape[1], y.shape[0]), np.float32)
        r.ref_transpose(y.shape[0], y.shape[1], _p(y), _p(yt))
        y = yt
    m, k = x.shape
    n = y.shape[1]
    out = np.empty((m, n), np.float32)
    r.ref_matrix_mul(m, k, n, _f(alpha), _p(x), _p(y), _p(out))
    return out


def ref_deconv_nchw(x, w, bias, strides, pads, dil=(1, 1), group=1, relu=False):
    """deconv_basic<float,float> of naive_math_impl.h:540-642.  x [N,Cin,H,W], w [Cin, Cout/g, kh, kw];
    pads = (top, bottom, left, right)."""
    r = ref()
    assert r is not None
    x = np.ascontiguousarray(x, np.float32)
    w = np.ascontiguousarray(w, np.float32)
    N, Cin, H, W = x.shape
    _, cog, kh, kw = w.shape
    Cout = cog * group
    oh = (H - 1) * strides[0] - pads[0] - pads[1] + dil[0] * (kh - 1) + 1
    ow = (W - 1) * strides[1] - pads[2] - pads[3] + dil[1] * (kw - 1) + 1
    y = np.zeros((N, Cout, oh, ow), np.float32)
    b = None if bias is None else np.ascontiguousarray(bias, np.float32)
    r.ref_deconv_basic_f32(_p(x), _p(y), N, Cout, oh, ow, Cin, H, W, _p(w), _p(b), group, kw, kh, strides[1],
                           strides[0], dil[1], dil[0], pads[2], pads[3], pads[0], pads[1], int(b is not None),
                           int(relu))
    return y
