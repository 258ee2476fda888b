"""ctypes driver for libpl_lite_kernels.so (the C++ mirror of Paddle-Lite's kernel
plugin surface): tensors + `run_*` helpers that create a kernel through the registry,
SetParam / SetContext / Launch() it and read the result back.  Used by tests/ and
INTEGRATION.md examples; numpy in, numpy out."""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libpl_lite_kernels.so")

K_FLOAT, K_INT8, K_INT32 = 1, 2, 3
T_HOST, T_CUDA = 1, 3
ACT_RELU, ACT_RELU6, ACT_LEAKY = 1, 2, 4
_NP2P = {np.dtype(np.float32): K_FLOAT, np.dtype(np.int8): K_INT8, np.dtype(np.int32): K_INT32}
_P2NP = {K_FLOAT: np.float32, K_INT8: np.int8, K_INT32: np.int32}

_lib = None


class LiteError(RuntimeError):
    pass


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} missing: run __graft_entry__.build()")
        C.CDLL(os.path.join(HERE, "libpl_int8_b200.so"), mode=C.RTLD_GLOBAL)
        L = C.CDLL(LIB_PATH)
        L.plk_tensor_create.restype = C.c_void_p
        L.plk_last_error.restype = C.c_char_p
        L.plk_tensor_bytes.restype = C.c_int64
        _lib = L
    return _lib


def _ck(rc):
    if rc != 0:
        raise LiteError(lib().plk_last_error().decode())


class Tensor:
    def __init__(self, array=None, target=T_CUDA):
        self.h = C.c_void_p(lib().plk_tensor_create())
        if array is not None:
            self.assign(array, target)

    def assign(self, array, target=T_CUDA):
        a = np.ascontiguousarray(array)
        if a.dtype not in _NP2P:
            raise LiteError(f"unsupported precision {a.dtype} (kFloat / kInt8 / kInt32 only)")
        dims = (C.c_int64 * a.ndim)(*a.shape)
        _ck(lib().plk_tensor_assign(self.h, _NP2P[a.dtype], target, dims, a.ndim, a.ctypes.data_as(C.c_void_p)))
        return self

    @property
    def shape(self):
        n = lib().plk_tensor_ndim(self.h)
        d = (C.c_int64 * max(n, 1))()
        lib().plk_tensor_dims(self.h, d)
        return tuple(d[i] for i in range(n))

    @property
    def precision(self):
        return lib().plk_tensor_precision(self.h)

    @property
    def target(self):
        return lib().plk_tensor_target(self.h)

    def numpy(self):
        out = np.empty(self.shape, _P2NP[self.precision])
        _ck(lib().plk_tensor_to_host(self.h, out.ctypes.data_as(C.c_void_p), C.c_int64(out.nbytes)))
        return out

    def __del__(self):
        try:
            lib().plk_tensor_destroy(self.h)
        except Exception:
            pass


def _h(t):
    return None if t is None else t.h


def _fa(v):
    a = np.ascontiguousarray(v, np.float32).reshape(-1)
    return a, a.ctypes.data_as(C.c_void_p), len(a)


def _ia(v):
    return (C.c_int * len(v))(*[int(x) for x in v])


def list_kernels():
    buf = C.create_string_buffer(1 << 16)
    n = lib().plk_list_kernels(buf, len(buf))
    assert n >= 0
    return [k for k in buf.value.decode().split("\n") if k]


def pick(op, enable_int8, consumers_int8, consumer_scales=None):
    """static_kernel_pick for one op -> ('op/alias/target/precision/layout', output_scale)."""
    n = len(consumers_int8)
    ci = _ia([1 if c else 0 for c in consumers_int8]) if n else None
    cs = (C.c_float * n)(*(consumer_scales or [1.0] * n)) if n else None
    buf = C.create_string_buffer(256)
    os_ = C.c_float(0)
    _ck(lib().plk_pick(op.encode(), int(enable_int8), n, ci, cs, buf, len(buf), C.byref(os_)))
    return buf.value.decode(), os_.value


def score_kernels(op, places, enable_int8, in_types=None):
    """KernelGrade of every candidate of `op`, best first: [(score, 'op/alias/target/precision/layout')].
    places: [(target, precision, layout)] ints of lite/api/paddle_place.h; in_types: {argument: precision}."""
    flat = _ia([v for p in places for v in p])
    in_types = in_types or {}
    names = (C.c_char_p * max(len(in_types), 1))(*[k.encode() for k in in_types])
    precs = _ia(list(in_types.values())) if in_types else None
    buf = C.create_string_buffer(1 << 14)
    _ck(lib().plk_score_kernels(op.encode(), len(places), flat, int(enable_int8), len(in_types), names, precs, buf, len(buf)))
    out = []
    for line in buf.value.decode().splitlines():
        sc, key = line.split(" ", 1)
        out.append((int(sc), key))
    return out


def run_conv(op, x, w, bias, input_scale, weight_scale, output_scale, strides=(1, 1), pads=(0, 0, 0, 0),
             dil=(1, 1), groups=1, act=0, act_coef=0.0, out_int8=True, residual=None, residual_scale=1.0,
             launches=1):
    """x NHWC int8, w OHWI int8 ([K,R,S,C/groups]), bias fp32 -> NHWC int8/fp32."""
    tx, tw = Tensor(x), Tensor(w)
    tb = None if bias is None else Tensor(np.asarray(bias, np.float32))
    tr = None if residual is None else Tensor(residual)
    out = Tensor()
    ws, wsp, nws = _fa(weight_scale)
    _ck(lib().plk_run_conv(op.encode(), tx.h, tw.h, _h(tb), _h(tr), out.h, _ia(strides), _ia(pads), _ia(dil), groups,
                           act, C.c_float(act_coef), C.c_float(input_scale), wsp, nws, C.c_float(output_scale),
                           C.c_float(residual_scale), int(out_int8), launches))
    res = out.numpy()
    return (res, tb.numpy()) if bias is not None else (res, None)


def run_fc(x, w, bias, input_scale, weight_scale, output_scale, in_num_col_dims=1, act="", out_int8=True, launches=1):
    tx, tw = Tensor(x), Tensor(w)
    tb = None if bias is None else Tensor(np.asarray(bias, np.float32))
    out = Tensor()
    ws, wsp, nws = _fa(weight_scale)
    _ck(lib().plk_run_fc(tx.h, tw.h, _h(tb), out.h, in_num_col_dims, act.encode(), C.c_float(input_scale), wsp, nws,
                         C.c_float(output_scale), int(out_int8), launches))
    return out.numpy()


def run_mul(x, y, input_scale, weight_scale, output_scale, x_num_col_dims=1, y_num_col_dims=1, out_int8=True):
    tx, ty, out = Tensor(x), Tensor(y), Tensor()
    ws, wsp, nws = _fa(weight_scale)
    _ck(lib().plk_run_mul(tx.h, ty.h, out.h, x_num_col_dims, y_num_col_dims, C.c_float(input_scale), wsp, nws,
                          C.c_float(output_scale), int(out_int8)))
    return out.numpy()


def run_matmul(x, y, input_scale, weight_scale, output_scale, transpose_x=False, transpose_y=False, alpha=1.0,
               out_int8=True):
    tx, ty, out = Tensor(x), Tensor(y), Tensor()
    ws, wsp, nws = _fa(weight_scale)
    _ck(lib().plk_run_matmul(tx.h, ty.h, out.h, int(transpose_x), int(transpose_y), C.c_float(alpha),
                             C.c_float(input_scale), wsp, nws, C.c_float(output_scale), int(out_int8)))
    return out.numpy()


def run_calib(x, scale, to_int8):
    tx, out = Tensor(x), Tensor()
    _ck(lib().plk_run_calib(tx.h, out.h, C.c_float(scale), int(to_int8)))
    return out.numpy()


def run_layout(x, to_nhwc):
    tx, out = Tensor(x), Tensor()
    _ck(lib().plk_run_layout(tx.h, out.h, int(to_nhwc)))
    return out.numpy()


def run_io_copy_roundtrip(x):
    th = Tensor(x, T_HOST)
    td, back = Tensor(), Tensor()
    _ck(lib().plk_run_io_copy(th.h, td.h, 1))
    assert td.target == T_CUDA
    _ck(lib().plk_run_io_copy(td.h, back.h, 0))
    assert back.target == T_HOST
    return back.numpy()


def run_pool(x, pooling_type, ksize, strides, pads, input_scale, output_scale, global_pooling=False, exclusive=True,
             ceil_mode=False, out_int8=True, adaptive=False):
    tx, out = Tensor(x), Tensor()
    _ck(lib().plk_run_pool(tx.h, out.h, pooling_type.encode(), _ia(ksize), _ia(strides), _ia(pads),
                           int(global_pooling), int(exclusive), int(ceil_mode) | (2 if adaptive else 0), C.c_float(input_scale),
                           C.c_float(output_scale), int(out_int8)))
    return out.numpy()


def run_add(x, y, x_scale, y_scale, out_scale, act="", out_int8=True):
    tx, ty, out = Tensor(x), Tensor(y), Tensor()
    _ck(lib().plk_run_add(tx.h, ty.h, out.h, C.c_float(x_scale), C.c_float(y_scale), C.c_float(out_scale),
                          act.encode(), int(out_int8)))
    return out.numpy()


def _ha(tensors):
    return (C.c_void_p * len(tensors))(*[t.h for t in tensors])


def run_concat(xs, axis=1):
    """xs: int8 NHWC arrays; axis is the model's (NCHW) axis, 1 = channels"""
    ts, out = [Tensor(x) for x in xs], Tensor()
    _ck(lib().plk_run_concat(_ha(ts), len(ts), out.h, axis))
    return out.numpy()


def run_split(x, axis=1, num=0, sections=()):
    n = len(sections) if sections else num
    tx, outs = Tensor(x), [Tensor() for _ in range(n)]
    _ck(lib().plk_run_split(tx.h, _ha(outs), n, axis, num, _ia(list(sections) or [0]), len(sections)))
    return [o.numpy() for o in outs]


def run_shuffle_channel(x, group):
    tx, out = Tensor(x), Tensor()
    _ck(lib().plk_run_shuffle_channel(tx.h, out.h, group))
    return out.numpy()


def run_relu(x, relu6=False, clip=6.0, input_scale=1.0):
    tx, out = Tensor(x), Tensor()
    _ck(lib().plk_run_relu(tx.h, out.h, int(relu6), C.c_float(clip), C.c_float(input_scale)))
    return out.numpy()


def plan_types(ops, extension=True, reference_decls=True):
    """ops: [(type, [inputs], [outputs], enable_int8)] in topological order, "feed" / "fetch" at the
    boundary.  Returns {"picked": [...], "casts": [(kind, var, consumer, from, to, bytes)], "inside": n,
    "bytes_per_elem": total} from the miniature type passes (lite/core/mir/type_cast_plan.h)."""
    text = "\n".join(f"{t};{','.join(i)};{','.join(o)};{int(q)}" for t, i, o, q in ops)
    buf = C.create_string_buffer(1 << 16)
    _ck(lib().plk_plan_types(text.encode(), int(extension), int(reference_decls), buf, len(buf)))
    res = {"picked": [], "casts": [], "inside": 0, "bytes_per_elem": 0}
    for ln in buf.value.decode().splitlines():
        f = ln.split()
        if f[0] == "op":
            res["picked"].append(f[3])
        elif f[0] == "cast":
            res["casts"].append((f[1], f[2], f[4], f[5], f[7], int(f[9])))
        elif f[0] == "casts":
            res["inside"], res["bytes_per_elem"] = int(f[3]), int(f[5])
    return res
