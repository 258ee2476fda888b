"""ctypes window onto libpl_int8_b200.so (the C-ABI in include/pl_int8.h).

torch is used for device memory and streams only; every compute call goes
through the C-ABI with raw device pointers.  Missing library -> ImportError
(no fallback of any kind).
"""
import ctypes as C
import os
import subprocess

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
# PL_INT8_LIB: developer knob, points the loader at an experimental build of the same C-ABI
# (tools/build_variant.sh); never a different implementation
LIB_PATH = os.environ.get("PL_INT8_LIB") or os.path.join(HERE, "libpl_int8_b200.so")

ACT_NONE, ACT_RELU, ACT_RELU6, ACT_LEAKY = 0, 1, 2, 4
OUT_INT8, OUT_FP32, OUT_INT32 = 0, 1, 2
ROUND_ARM, ROUND_CUDA = 0, 1
POOL_MAX, POOL_AVG = 0, 1
_TORCH_OUT = {OUT_INT8: torch.int8, OUT_FP32: torch.float32, OUT_INT32: torch.int32}


class Epilogue(C.Structure):
    _fields_ = [("scale", C.c_void_p), ("bias", C.c_void_p), ("act", C.c_int),
                ("act_arg", C.c_float), ("out_dtype", C.c_int), ("round_mode", C.c_int),
                ("residual", C.c_void_p), ("ld_res", C.c_int), ("res_scale", C.c_float)]


class ConvDesc(C.Structure):
    _fields_ = [(n, C.c_int) for n in (
        "N", "H", "W", "C", "ldx", "K", "ldy", "R", "S", "stride_h", "stride_w",
        "pad_t", "pad_b", "pad_l", "pad_r", "dil_h", "dil_w", "groups")]


class S2dGeometry(C.Structure):
    _fields_ = [("supported", C.c_int), ("HP", C.c_int), ("WS", C.c_int), ("pad_s", C.c_int), ("shift", C.c_int),
                ("TA", C.c_int), ("TB", C.c_int), ("P", C.c_int), ("Q", C.c_int), ("bytes", C.c_size_t)]


def build(verbose=False):
    """Compile the CUDA library in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    out = subprocess.run(["make", "-C", os.path.join(HERE, "csrc"), "-j8"],
                         capture_output=True, text=True)
    if out.returncode != 0:
        raise RuntimeError("nvcc build failed:\n" + out.stdout + out.stderr)
    if verbose:
        print(out.stdout)


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: run __graft_entry__.build() (there is no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        L.pl_launch_count.restype = C.c_int64
        L.pl_version.restype = C.c_char_p
        L.pl_conv_packed_weight_bytes.restype = C.c_size_t
        L.pl_s2d_conv_geometry.restype = S2dGeometry
        L.pl_s2d_conv_packed_weight_bytes.restype = C.c_size_t
        _lib = L
    return _lib


class PlError(RuntimeError):
    pass


def _check(rc, what):
    if rc != 0:
        raise PlError(f"{what} failed with code {rc}")


def _ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _f(v):
    return C.c_float(float(v))


def launch_count():
    return int(lib().pl_launch_count())


PLAN_KEYS = ("mode", "BN", "BK", "stages", "cta_pairs", "tile_pairs", "band_rows", "grid", "filter_resident", "epi")


def last_igemm_plan():
    """Tile plan of this thread's most recent conv / gemm launch (pl_last_igemm_plan)."""
    buf = (C.c_int * len(PLAN_KEYS))()
    n = lib().pl_last_igemm_plan(buf, len(PLAN_KEYS))
    d = dict(zip(PLAN_KEYS[:n], list(buf)[:n]))
    if "stages" in d:  # ring slots per barrier group ride in the second byte
        d["kgroup"] = (d["stages"] >> 8) & 0xff
        d["stages"] &= 0xff
    return d


def refresh_knobs():
    lib().pl_refresh_knobs()


def conv_out_size(i, k, d, p0, p1, s):
    return (i + p0 + p1 - (d * (k - 1) + 1)) // s + 1


def fold_scales(w_scale, n_out, in_scale, out_scale, out_dtype, bias=None, act=ACT_NONE, act_arg=0.0):
    """Host-side scale folding (pl_fold_scales); returns numpy (scale, bias, act_arg)."""
    ws = np.ascontiguousarray(w_scale, np.float32)
    sf = np.empty(n_out, np.float32)
    bf = np.empty(n_out, np.float32)
    b = None if bias is None else np.ascontiguousarray(bias, np.float32)
    aa = C.c_float(0)
    rc = lib().pl_fold_scales(ws.ctypes.data_as(C.c_void_p), len(ws), n_out, _f(in_scale), _f(out_scale),
                              out_dtype, None if b is None else b.ctypes.data_as(C.c_void_p),
                              sf.ctypes.data_as(C.c_void_p), bf.ctypes.data_as(C.c_void_p), act,
                              _f(act_arg), C.byref(aa))
    _check(rc, "pl_fold_scales")
    return sf, bf, aa.value


def make_epilogue(scale, bias, act=ACT_NONE, act_arg=0.0, out=OUT_INT8, rnd=ROUND_ARM,
                  residual=None, ld_res=0, res_scale=0.0):
    e = Epilogue()
    e.scale = None if scale is None else scale.data_ptr()
    e.bias = None if bias is None else bias.data_ptr()
    e.act, e.act_arg, e.out_dtype, e.round_mode = act, float(act_arg), out, rnd
    e.residual = None if residual is None else residual.data_ptr()
    e.ld_res, e.res_scale = ld_res, float(res_scale)
    return e


def pack_conv_weight(w_ohwi):
    """OHWI int8 (device) -> packed filter for pl_conv2d_i8_nhwc."""
    K, R, S, Cc = w_ohwi.shape
    nbytes = lib().pl_conv_packed_weight_bytes(K, R, S, Cc)
    out = torch.empty(nbytes, dtype=torch.int8, device=w_ohwi.device)
    _check(lib().pl_pack_conv_weight_i8(_ptr(w_ohwi.contiguous()), _ptr(out), K, R, S, Cc, _stream()),
           "pl_pack_conv_weight_i8")
    return out


def conv_desc(N, H, W, Cc, K, R, S, strides=(1, 1), pads=(0, 0, 0, 0), dil=(1, 1), groups=1,
              ldx=None, ldy=None):
    d = ConvDesc()
    d.N, d.H, d.W, d.C, d.K, d.R, d.S = N, H, W, Cc, K, R, S
    d.ldx = Cc if ldx is None else ldx
    d.ldy = K if ldy is None else ldy
    d.stride_h, d.stride_w = strides
    d.pad_t, d.pad_b, d.pad_l, d.pad_r = pads
    d.dil_h, d.dil_w = dil
    d.groups = groups
    return d


def conv2d(x, w_packed, K, R, S, scale, bias, strides=(1, 1), pads=(0, 0, 0, 0), dil=(1, 1),
           act=ACT_NONE, act_arg=0.0, out=OUT_INT8, rnd=ROUND_ARM, y=None, Cc=None, ldy=None,
           residual=None, res_scale=0.0, groups=1):
    """x: [N,H,W,ldx] int8 device tensor (logical channels Cc <= ldx).  groups > 1: w_packed was packed from
    OHWI [K, R, S, Cc / groups]."""
    N, H, W, Cx = x.shape
    ldx = x.stride(-2)  # a channel slice of a wider buffer keeps the wide pitch
    Cc = Cx if Cc is None else Cc
    P = conv_out_size(H, R, dil[0], pads[0], pads[1], strides[0])
    Q = conv_out_size(W, S, dil[1], pads[2], pads[3], strides[1])
    if ldy is None:
        ldy = K if y is None else y.stride(-2)
    if y is None:
        y = torch.empty((N, P, Q, ldy), dtype=_TORCH_OUT[out], device=x.device)
        if ldy != K:
            y.zero_()
    d = conv_desc(N, H, W, Cc, K, R, S, strides, pads, dil, groups, ldx, ldy)
    e = make_epilogue(scale, bias, act, act_arg, out, rnd, residual,
                      0 if residual is None else residual.stride(-2), res_scale)
    _check(lib().pl_conv2d_i8_nhwc(_ptr(x), _ptr(w_packed), _ptr(y), C.byref(d), C.byref(e), _stream()),
           "pl_conv2d_i8_nhwc")
    return y


def dwconv2d(x, w_crs, scale, bias, R, S, strides=(1, 1), pads=(0, 0, 0, 0), dil=(1, 1), act=ACT_NONE,
             act_arg=0.0, out=OUT_INT8, rnd=ROUND_ARM, y=None, Cc=None):
    N, H, W, Cx = x.shape
    ldx = x.stride(-2)
    Cc = Cx if Cc is None else Cc
    P = conv_out_size(H, R, dil[0], pads[0], pads[1], strides[0])
    Q = conv_out_size(W, S, dil[1], pads[2], pads[3], strides[1])
    if y is None:
        y = torch.zeros((N, P, Q, Cx), dtype=_TORCH_OUT[out], device=x.device)
    d = conv_desc(N, H, W, Cc, Cc, R, S, strides, pads, dil, Cc, ldx, y.stride(-2))
    e = make_epilogue(scale, bias, act, act_arg, out, rnd)
    _check(lib().pl_dwconv2d_i8_nhwc(_ptr(x), _ptr(w_crs), _ptr(y), C.byref(d), C.byref(e), _stream()),
           "pl_dwconv2d_i8_nhwc")
    return y


def gemm(A, Bt, scale, bias, act=ACT_NONE, act_arg=0.0, out=OUT_INT8, rnd=ROUND_ARM, y=None,
         M=None, N=None, K=None):
    """C[M,N] = epi(A[M,K] . Bt[N,K]^T); A, Bt int8 row-major device tensors."""
    M = A.shape[0] if M is None else M
    K = A.shape[1] if K is None else K
    N = Bt.shape[0] if N is None else N
    if y is None:
        y = torch.empty((M, N), dtype=_TORCH_OUT[out], device=A.device)
    e = make_epilogue(scale, bias, act, act_arg, out, rnd)
    _check(lib().pl_gemm_i8(_ptr(A), A.stride(0), _ptr(Bt), Bt.stride(0), _ptr(y), y.stride(0), M, N, K,
                            C.byref(e), _stream()), "pl_gemm_i8")
    return y


def gemm_peers(A, Bt, scale, bias, y, peer_ptrs, maps_dev, act=ACT_NONE, act_arg=0.0, rnd=ROUND_ARM, N=None, K=None):
    """pl_gemm_i8_peers: int8-out GEMM whose tiles are stored to `y` (a column block view of the local
    full result) AND to the raw device addresses `peer_ptrs` (the same block in the peers' results)."""
    M = A.shape[0]
    K = A.shape[1] if K is None else K
    N = Bt.shape[0] if N is None else N
    e = make_epilogue(scale, bias, act, act_arg, OUT_INT8, rnd)
    if isinstance(peer_ptrs, int):  # maps already uploaded by gemm_peers_prepare: launch only
        arr, n_peers = None, peer_ptrs
    else:
        arr = (C.c_void_p * max(len(peer_ptrs), 1))(*[C.c_void_p(int(p_)) for p_ in peer_ptrs])
        n_peers = len(peer_ptrs)
    _check(lib().pl_gemm_i8_peers(_ptr(A), A.stride(0), _ptr(Bt), Bt.stride(0), _ptr(y), arr, n_peers,
                                  _ptr(maps_dev), y.stride(0), M, N, K, C.byref(e), _stream()), "pl_gemm_i8_peers")
    return y


def gemm_mcast(A, Bt, scale, bias, y, mc_ptr, act=ACT_NONE, act_arg=0.0, rnd=ROUND_ARM, N=None, K=None, out=OUT_INT8):
    """pl_gemm_i8_mcast: GEMM (int8 or fp32 out) whose tiles leave through the NVSwitch multicast address `mc_ptr`
    (an int: the multicast VA of the column block `y` views in this rank's copy of the full result)."""
    M = A.shape[0]
    K = A.shape[1] if K is None else K
    N = Bt.shape[0] if N is None else N
    e = make_epilogue(scale, bias, act, act_arg, out, rnd)
    _check(lib().pl_gemm_i8_mcast(_ptr(A), A.stride(0), _ptr(Bt), Bt.stride(0), _ptr(y), C.c_void_p(int(mc_ptr)),
                                  y.stride(0), M, N, K, C.byref(e), _stream()), "pl_gemm_i8_mcast")
    return y


def mcast_copy_2d(src, mc_ptr, ld_dst_bytes):
    """pl_mcast_copy_2d: broadcast the 2-D device tensor `src` (a strided view is fine) to the multicast address
    `mc_ptr` (int) with row pitch `ld_dst_bytes`."""
    rows, cols = src.shape
    esz = src.element_size()
    _check(lib().pl_mcast_copy_2d(_ptr(src), C.c_int64(src.stride(0) * esz), C.c_void_p(int(mc_ptr)),
                                  C.c_int64(ld_dst_bytes), C.c_int64(rows), C.c_int64(cols * esz), _stream()),
           "pl_mcast_copy_2d")


def gemm_peers_prepare(y, peer_ptrs, maps_dev, scale, bias, M, N, K, act=ACT_NONE, act_arg=0.0, rnd=ROUND_ARM):
    """pl_gemm_i8_peers_prepare: one-off synchronous upload of the peer tensor maps; later gemm_peers calls pass
    peer_ptrs = len(peer_ptrs) (an int) and are a single kernel launch."""
    e = make_epilogue(scale, bias, act, act_arg, OUT_INT8, rnd)
    arr = (C.c_void_p * len(peer_ptrs))(*[C.c_void_p(int(p_)) for p_ in peer_ptrs])
    _check(lib().pl_gemm_i8_peers_prepare(_ptr(y), arr, len(peer_ptrs), _ptr(maps_dev), y.stride(0), M, N, K,
                                          C.byref(e)), "pl_gemm_i8_peers_prepare")


def gemm_peers_static(A, Bt, scale, bias, y, peer_ptrs, act=ACT_NONE, act_arg=0.0, rnd=ROUND_ARM, N=None, K=None):
    """pl_gemm_i8_peers_static: pl_gemm_i8_peers without caller scratch (the peer tensor maps live in a
    file-scope __device__ table of PL_STATIC_PEERS_MAX entries inside the library); allocates nothing."""
    M = A.shape[0]
    K = A.shape[1] if K is None else K
    N = Bt.shape[0] if N is None else N
    e = make_epilogue(scale, bias, act, act_arg, OUT_INT8, rnd)
    arr = (C.c_void_p * max(len(peer_ptrs), 1))(*[C.c_void_p(int(p_)) for p_ in peer_ptrs])
    _check(lib().pl_gemm_i8_peers_static(_ptr(A), A.stride(0), _ptr(Bt), Bt.stride(0), _ptr(y), arr, len(peer_ptrs),
                                         y.stride(0), M, N, K, C.byref(e), _stream()), "pl_gemm_i8_peers_static")
    return y


def transpose_i8(src):
    rows, cols = src.shape
    dst = torch.empty((cols, rows), dtype=torch.int8, device=src.device)
    _check(lib().pl_transpose_i8(_ptr(src), src.stride(0), _ptr(dst), rows, rows, cols, _stream()),
           "pl_transpose_i8")
    return dst


def quantize(x, scale, rnd=ROUND_ARM):
    x = x.contiguous()
    y = torch.empty(x.shape, dtype=torch.int8, device=x.device)
    _check(lib().pl_quantize_f32_i8(_ptr(x), _ptr(y), C.c_int64(x.numel()), _f(scale), rnd, _stream()),
           "pl_quantize_f32_i8")
    return y


def dequantize(x, scale):
    x = x.contiguous()
    y = torch.empty(x.shape, dtype=torch.float32, device=x.device)
    _check(lib().pl_dequantize_i8_f32(_ptr(x), _ptr(y), C.c_int64(x.numel()), _f(scale), _stream()),
           "pl_dequantize_i8_f32")
    return y


def quantize_nchw_to_nhwc(x, scale, ld=None, rnd=ROUND_ARM, y=None):
    N, Cc, H, W = x.shape
    ld = Cc if ld is None else ld
    if y is None:
        y = torch.empty((N, H, W, ld), dtype=torch.int8, device=x.device)
    _check(lib().pl_quantize_nchw_f32_to_nhwc_i8(_ptr(x.contiguous()), _ptr(y), N, Cc, H, W, ld, _f(scale),
                                                 rnd, _stream()), "pl_quantize_nchw_f32_to_nhwc_i8")
    return y


def im2col(x, R, S, strides=(1, 1), pads=(0, 0, 0, 0), dil=(1, 1), A=None, cbytes=None):
    """Explicit im2col of a thin-channel NHWC int8 tensor [N,H,W,ld] (ld % 4 == 0) ->
    A [N*P*Q, ldA], ldA = roundup(R*S*cbytes, 16)."""
    N, H, W, ld = x.shape
    cbytes = ld if cbytes is None else cbytes
    P = conv_out_size(H, R, dil[0], pads[0], pads[1], strides[0])
    Q = conv_out_size(W, S, dil[1], pads[2], pads[3], strides[1])
    ldA = (R * S * cbytes + 15) // 16 * 16
    if A is None:
        A = torch.empty((N * P * Q, ldA), dtype=torch.int8, device=x.device)
    _check(lib().pl_im2col_i8_nhwc(_ptr(x), ld, _ptr(A), A.stride(0), N, H, W, cbytes, R, S, strides[0],
                                   strides[1], pads[0], pads[2], dil[0], dil[1], P, Q, _stream()),
           "pl_im2col_i8_nhwc")
    return A


def thin_conv_supported(R, S, Cc, dil=(1, 1)):
    return bool(lib().pl_thin_conv_supported(R, S, Cc, dil[0], dil[1]))


def pack_thin_conv_weight(w_ohwi):
    """OHWI int8 (device, C <= 4) -> packed filter for thin_conv2d."""
    K, R, S, Cc = w_ohwi.shape
    nbytes = lib().pl_thin_conv_packed_weight_bytes(K, R, S)
    out = torch.empty(nbytes, dtype=torch.int8, device=w_ohwi.device)
    _check(lib().pl_pack_thin_conv_weight_i8(_ptr(w_ohwi.contiguous()), _ptr(out), K, R, S, Cc, _stream()),
           "pl_pack_thin_conv_weight_i8")
    return out


def thin_conv2d(x, w_thin, K, R, S, Cc, scale, bias, strides=(1, 1), pads=(0, 0, 0, 0), act=ACT_NONE,
                act_arg=0.0, out=OUT_INT8, rnd=ROUND_ARM, y=None, ldy=None):
    """x: [N,H,W,4] int8 (one word per pixel, Cc <= 4 logical channels): direct conv, A tiles
    gathered on chip."""
    N, H, W, ldx = x.shape
    assert ldx == 4
    P = conv_out_size(H, R, 1, pads[0], pads[1], strides[0])
    Q = conv_out_size(W, S, 1, pads[2], pads[3], strides[1])
    ldy = K if ldy is None else ldy
    if y is None:
        y = torch.empty((N, P, Q, ldy), dtype=_TORCH_OUT[out], device=x.device)
        if ldy != K:
            y.zero_()
    d = conv_desc(N, H, W, Cc, K, R, S, strides, pads, (1, 1), 1, 4, ldy)
    e = make_epilogue(scale, bias, act, act_arg, out, rnd)
    _check(lib().pl_thin_conv2d_i8_nhwc(_ptr(x), _ptr(w_thin), _ptr(y), C.byref(d), C.byref(e), _stream()),
           "pl_thin_conv2d_i8_nhwc")
    return y


def s2d_geometry(N, H, W, Cc, R, S, strides, pads, dil=(1, 1)):
    """Space-to-depth form of a stride-2 thin stem (include/pl_int8.h: pl_s2d_geometry); .supported == 0
    when the conv does not qualify."""
    return lib().pl_s2d_conv_geometry(N, H, W, Cc, R, S, strides[0], strides[1], pads[0], pads[1], pads[2],
                                      pads[3], dil[0], dil[1])


def s2d_input_buffer(g, device):
    """Zero-filled buffer for the s2d input (the borders must stay zero)."""
    return torch.zeros(g.bytes, dtype=torch.int8, device=device)


def quantize_nchw_to_s2d(x, scale, g, y=None, rnd=ROUND_ARM, mean=None, norm=None):
    """fp32 / fp16 / uint8 NCHW (by dtype of `x`) -> int8 space-to-depth input.  uint8 feeds take the image
    normalisation x = (u8 - mean[c]) * norm[c] (host sequences of C floats) in front of the calib step."""
    N, Cc, H, W = x.shape
    if y is None:
        y = s2d_input_buffer(g, x.device)
    x = x.contiguous()
    if x.dtype == torch.float32:
        _check(lib().pl_quantize_nchw_f32_to_s2d_i8(_ptr(x), _ptr(y), N, Cc, H, W, C.byref(g), _f(scale),
                                                    rnd, _stream()), "pl_quantize_nchw_f32_to_s2d_i8")
    elif x.dtype == torch.float16:
        _check(lib().pl_quantize_nchw_f16_to_s2d_i8(_ptr(x), _ptr(y), N, Cc, H, W, C.byref(g), _f(scale),
                                                    rnd, _stream()), "pl_quantize_nchw_f16_to_s2d_i8")
    elif x.dtype == torch.uint8:
        m = (C.c_float * Cc)(*[float(v) for v in (mean if mean is not None else [0.0] * Cc)])
        s_ = (C.c_float * Cc)(*[float(v) for v in (norm if norm is not None else [1.0] * Cc)])
        _check(lib().pl_quantize_nchw_u8_to_s2d_i8(_ptr(x), _ptr(y), N, Cc, H, W, C.byref(g), m, s_, _f(scale),
                                                   rnd, _stream()), "pl_quantize_nchw_u8_to_s2d_i8")
    else:
        raise TypeError(f"unsupported feed dtype {x.dtype}")
    return y


def nhwc_to_s2d(x, g, Cc=None, y=None):
    """int8 NHWC [N,H,W,ld] (Cc <= 4 logical channels) -> space-to-depth input buffer (pl_nhwc_i8_to_s2d_i8)"""
    N, H, W, ld = x.shape
    Cc = ld if Cc is None else Cc
    if y is None:
        y = s2d_input_buffer(g, x.device)
    _check(lib().pl_nhwc_i8_to_s2d_i8(_ptr(x), x.stride(-2), _ptr(y), N, Cc, H, W, C.byref(g), _stream()),
           "pl_nhwc_i8_to_s2d_i8")
    return y


def repitch(x, ld_dst, Cc=None):
    """int8 [..., ld_src] -> [..., ld_dst] with zero pad channels (pl_repitch_i8_nhwc)"""
    ld_src = x.shape[-1]
    Cc = ld_src if Cc is None else Cc
    pixels = x.numel() // ld_src
    y = torch.empty(tuple(x.shape[:-1]) + (ld_dst,), dtype=torch.int8, device=x.device)
    _check(lib().pl_repitch_i8_nhwc(_ptr(x), ld_src, _ptr(y), ld_dst, C.c_int64(pixels), Cc, _stream()),
           "pl_repitch_i8_nhwc")
    return y


def pack_s2d_conv_weight(w_ohwi, g):
    K, R, S, Cc = w_ohwi.shape
    out = torch.empty(lib().pl_s2d_conv_packed_weight_bytes(K, C.byref(g)), dtype=torch.int8, device=w_ohwi.device)
    _check(lib().pl_pack_s2d_conv_weight_i8(_ptr(w_ohwi.contiguous()), _ptr(out), K, R, S, Cc, C.byref(g),
                                            _stream()), "pl_pack_s2d_conv_weight_i8")
    return out


def s2d_conv2d(x_s2d, w_s2d, N, H, W, Cc, K, R, S, scale, bias, strides=(2, 2), pads=(0, 0, 0, 0), act=ACT_NONE,
               act_arg=0.0, out=OUT_INT8, rnd=ROUND_ARM, y=None, ldy=None):
    """Stride-2 thin stem on the space-to-depth input written by quantize_nchw_to_s2d."""
    P = conv_out_size(H, R, 1, pads[0], pads[1], strides[0])
    Q = conv_out_size(W, S, 1, pads[2], pads[3], strides[1])
    ldy = K if ldy is None else ldy
    if y is None:
        y = torch.empty((N, P, Q, ldy), dtype=_TORCH_OUT[out], device=x_s2d.device)
        if ldy != K:
            y.zero_()
    d = conv_desc(N, H, W, Cc, K, R, S, strides, pads, (1, 1), 1, 4, ldy)
    e = make_epilogue(scale, bias, act, act_arg, out, rnd)
    _check(lib().pl_s2d_conv2d_i8(_ptr(x_s2d), _ptr(w_s2d), _ptr(y), C.byref(d), C.byref(e), _stream()),
           "pl_s2d_conv2d_i8")
    return y


def nchw_to_nhwc(x, ld=None):
    N, Cc, H, W = x.shape
    ld = Cc if ld is None else ld
    y = torch.empty((N, H, W, ld), dtype=x.dtype, device=x.device)
    _check(lib().pl_nchw_to_nhwc(_ptr(x.contiguous()), _ptr(y), N, Cc, H, W, ld, x.element_size(),
                                 _stream()), "pl_nchw_to_nhwc")
    return y


def nhwc_to_nchw(x, Cc=None):
    N, H, W, ld = x.shape
    Cc = ld if Cc is None else Cc
    y = torch.empty((N, Cc, H, W), dtype=x.dtype, device=x.device)
    _check(lib().pl_nhwc_to_nchw(_ptr(x.contiguous()), _ptr(y), N, Cc, H, W, ld, x.element_size(),
                                 _stream()), "pl_nhwc_to_nchw")
    return y


def set_tile_order(reverse):
    """Scheduling hint for the next launches of this thread (pl_set_tile_order): walk the output tiles last-to-first"""
    lib().pl_set_tile_order(int(bool(reverse)))


def pool2d(x, ksize, strides, pads, pool_type, exclusive=True, rescale=1.0, out=OUT_INT8,
           rnd=ROUND_ARM, ceil_mode=False, global_pooling=False, Cc=None, y=None):
    N, H, W, ldx = x.shape
    Cc = ldx if Cc is None else Cc
    kh, kw = ksize
    pt, pb, pl_, pr = pads
    if global_pooling:
        kh, kw, pt, pb, pl_, pr = H, W, 0, 0, 0, 0
    if ceil_mode:
        P = -(-(H + pt + pb - kh) // strides[0]) + 1
        Q = -(-(W + pl_ + pr - kw) // strides[1]) + 1
    else:
        P = (H + pt + pb - kh) // strides[0] + 1
        Q = (W + pl_ + pr - kw) // strides[1] + 1
    if y is None:
        y = torch.zeros((N, P, Q, ldx), dtype=_TORCH_OUT[out], device=x.device)
    _check(lib().pl_pool2d_i8_nhwc(_ptr(x), ldx, _ptr(y), y.shape[-1], N, H, W, Cc, kh, kw, strides[0],
                                   strides[1], pt, pl_, P, Q, pool_type, int(exclusive), _f(rescale), out,
                                   rnd, _stream()), "pl_pool2d_i8_nhwc")
    return y


def adaptive_pool2d(x, out_hw, pool_type, rescale=1.0, out=OUT_INT8, rnd=ROUND_ARM, Cc=None):
    N, H, W, Cx = x.shape
    Cc = Cx if Cc is None else Cc
    P, Q = out_hw
    y = torch.zeros((N, P, Q, Cx), dtype=_TORCH_OUT[out], device=x.device)
    _check(lib().pl_adaptive_pool2d_i8_nhwc(_ptr(x), x.stride(-2), _ptr(y), y.stride(-2), N, H, W, Cc, P, Q, pool_type,
                                            _f(rescale), out, rnd, _stream()), "pl_adaptive_pool2d_i8_nhwc")
    return y


def add(a, b, sa, sb, act=ACT_NONE, act_arg=0.0, out=OUT_INT8, rnd=ROUND_ARM, y=None):
    if y is None:
        y = torch.empty(a.shape, dtype=_TORCH_OUT[out], device=a.device)
    _check(lib().pl_add_i8(_ptr(a), _ptr(b), _ptr(y), C.c_int64(a.numel()), _f(sa), _f(sb), act, _f(act_arg),
                           out, rnd, _stream()), "pl_add_i8")
    return y


def _view(t, Cc):
    """(pointer, channels, pixel pitch) of an NHWC tensor or a channel slice of one (a
    slice t[..., a:b] is a pointer offset + the wide pitch)"""
    assert t.stride(-1) == 1 and t.dim() >= 2
    ld = t.stride(-2)
    for d in range(t.dim() - 2):  # the pixels must be evenly spaced
        assert t.stride(d) == t.stride(d + 1) * t.shape[d + 1], "pixel dims must be dense"
    return t.data_ptr(), (t.shape[-1] if Cc is None else Cc), ld


def _pixels(t):
    n = 1
    for d in t.shape[:-1]:
        n *= d
    return n


def concat(xs, y=None):
    """int8 NHWC concat along channels (lite/backends/arm/math/concat.h:29-60, axis = C)"""
    views = [_view(x, None) for x in xs]
    ctot = sum(v[1] for v in views)
    if y is None:
        y = torch.empty(tuple(xs[0].shape[:-1]) + (ctot,), dtype=torch.int8, device=xs[0].device)
    yp, _, ldy = _view(y, None)
    n = len(xs)
    ptrs = (C.c_void_p * n)(*[v[0] for v in views])
    cs = (C.c_int * n)(*[v[1] for v in views])
    lds = (C.c_int * n)(*[v[2] for v in views])
    _check(lib().pl_concat_i8_nhwc(n, ptrs, cs, lds, C.c_void_p(yp), ldy, C.c_int64(_pixels(xs[0])), _stream()),
           "pl_concat_i8_nhwc")
    return y


def split(x, sections, ys=None):
    """int8 NHWC split along channels (lite/backends/arm/math/split.cc:54-83)"""
    xp, _, ldx = _view(x, None)
    if ys is None:
        ys = [torch.empty(tuple(x.shape[:-1]) + (c,), dtype=torch.int8, device=x.device) for c in sections]
    views = [_view(t, None) for t in ys]
    n = len(ys)
    ptrs = (C.c_void_p * n)(*[v[0] for v in views])
    cs = (C.c_int * n)(*sections)
    lds = (C.c_int * n)(*[v[2] for v in views])
    _check(lib().pl_split_i8_nhwc(C.c_void_p(xp), ldx, n, ptrs, cs, lds, C.c_int64(_pixels(x)), _stream()),
           "pl_split_i8_nhwc")
    return ys


def shuffle_channel(x, group, y=None):
    """int8 NHWC shuffle_channel (lite/backends/arm/math/shuffle_channel.cc:24-76)"""
    xp, Cc, ldx = _view(x, None)
    if y is None:
        y = torch.empty(tuple(x.shape), dtype=torch.int8, device=x.device)
    yp, _, ldy = _view(y, None)
    _check(lib().pl_shuffle_channel_i8_nhwc(C.c_void_p(xp), ldx, C.c_void_p(yp), ldy, Cc, group,
                                            C.c_int64(_pixels(x)), _stream()), "pl_shuffle_channel_i8_nhwc")
    return y


def concat_shuffle2(a, b, y=None):
    """concat(a, b) along channels + shuffle_channel(group 2) fused: y[..., 2j] = a[..., j], y[..., 2j+1] = b[..., j]"""
    ap, ca, lda = _view(a, None)
    bp, cb, ldb = _view(b, None)
    assert ca == cb
    if y is None:
        y = torch.empty(tuple(a.shape[:-1]) + (2 * ca,), dtype=torch.int8, device=a.device)
    yp, _, ldy = _view(y, None)
    _check(lib().pl_concat_shuffle2_i8_nhwc(C.c_void_p(ap), lda, C.c_void_p(bp), ldb, C.c_void_p(yp), ldy, ca,
                                            C.c_int64(_pixels(a)), _stream()), "pl_concat_shuffle2_i8_nhwc")
    return y


def relu_i8(x, hi=127, y=None):
    """relu / quantised relu6 on a dense int8 tensor"""
    assert x.is_contiguous()
    if y is None:
        y = torch.empty_like(x)
    _check(lib().pl_relu_i8(_ptr(x), _ptr(y), C.c_int64(x.numel()), hi, _stream()), "pl_relu_i8")
    return y


def i8_mma_peak(ctas, iters, bn=256):
    out = torch.zeros(4, dtype=torch.int32, device="cuda")
    _check(lib().pl_i8_mma_peak(ctas, iters, bn, _ptr(out), _stream()), "pl_i8_mma_peak")
    return out


def measure_i8_mma_peak(iters=20000, bn=256):
    """attainable tcgen05 kind::i8 rate in TOPS: back-to-back 128 x bn x 32 MMAs from
    resident smem on one CTA per SM, timed with CUDA events."""
    ctas = torch.cuda.get_device_properties(torch.cuda.current_device()).multi_processor_count
    i8_mma_peak(ctas, 1000, bn)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    i8_mma_peak(ctas, iters, bn)
    b.record()
    torch.cuda.synchronize()
    return ctas * iters * 8 * 2.0 * 128 * bn * 32 / (a.elapsed_time(b) * 1e-3) / 1e12  # 8 K=32 MMAs per iteration


# ---- fp32 ops between / after the int8 convs of a quantised model (fp32_tail.cu) -------------------------
def channel_gather(srcs, table_dev, C_dst, y, out=OUT_FP32, scale=0.0, rnd=ROUND_ARM):
    """srcs: fp32 NHWC device tensors of equal N, H, W; table_dev: int32 [C_dst, 2] (source index, channel);
    y: fp32 [N,H,W,C_dst] or int8 [N,H,W,ld] written in place."""
    n = len(srcs)
    ptrs = (C.c_void_p * n)(*[C.c_void_p(s.data_ptr()) for s in srcs])
    lds = (C.c_int * n)(*[int(s.stride(-2)) for s in srcs])
    pixels = srcs[0].shape[0] * srcs[0].shape[1] * srcs[0].shape[2]
    _check(lib().pl_channel_gather_f32_nhwc(n, ptrs, lds, _ptr(table_dev), _ptr(y), int(y.stride(-2)), C_dst,
                                            C.c_int64(pixels), out, _f(scale), rnd, _stream()),
           "pl_channel_gather_f32_nhwc")
    return y


def conv2d_transpose_f32(x, w_iohw, bias, stride, pad, y, Cout, act=ACT_NONE, out=OUT_FP32, out_scale=0.0):
    N, H, W, Ci = x.shape
    kh, kw = w_iohw.shape[2], w_iohw.shape[3]
    P, Q = y.shape[1], y.shape[2]
    _check(lib().pl_conv2d_transpose_f32_nhwc(_ptr(x), _ptr(w_iohw), _ptr(bias), _ptr(y), N, H, W, Ci, Cout, kh, kw,
                                              stride, stride, pad, pad, P, Q, int(y.stride(-2)), act, _f(0.0), out,
                                              _f(out_scale), ROUND_ARM, _stream()), "pl_conv2d_transpose_f32_nhwc")
    return y


def add_f32(a, b, y, act=ACT_NONE):
    _check(lib().pl_add_f32(_ptr(a), _ptr(b), _ptr(y), C.c_int64(a.numel()), act, _stream()), "pl_add_f32")
    return y


def softmax_f32(x, y):
    pixels = x.numel() // x.shape[-1]
    _check(lib().pl_softmax_f32_nhwc(_ptr(x), _ptr(y), C.c_int64(pixels), x.shape[-1], int(x.stride(-2)),
                                     int(y.stride(-2)), _stream()), "pl_softmax_f32_nhwc")
    return y


def scale_f32(x, y, scale, bias, bias_after_scale=True):
    _check(lib().pl_scale_f32(_ptr(x), _ptr(y), C.c_int64(x.numel()), _f(scale), _f(bias), int(bias_after_scale),
                              _stream()), "pl_scale_f32")
    return y
