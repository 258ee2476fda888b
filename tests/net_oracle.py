"""Walks a paddle-lite_b200 `nets` program on the CPU with the oracle (checker only).

`engine="oracle"` uses oracle/liboracle.so (NHWC restatement); `engine="reference"`
runs every contraction through the reference's own naive conv_basic<int8_t,int> /
basic_gemm<int8_t,int> (oracle/_ref/libref_naive.so, NCHW/OIHW as the reference's CPU
path stores them) and the oracle epilogue -- the timed CPU baseline of bench.py.
"""
import time

import numpy as np

import orc

_ACT = {"none": orc.ACT_NONE, "relu": orc.ACT_RELU, "relu6": orc.ACT_RELU6, "leaky": orc.ACT_LEAKY}


def _conv_reference(x, op, sf, bf, act, arg, out):
    """x NHWC int8 -> conv via the reference's conv_basic (NCHW) + oracle epilogue."""
    g = op.K if op.kind == "dwconv" else 1
    x_nchw = np.ascontiguousarray(x.transpose(0, 3, 1, 2))
    w_oihw = np.ascontiguousarray(op.w.transpose(0, 3, 1, 2))
    acc = orc.ref_conv_acc_nchw(x_nchw, w_oihw, (op.stride, op.stride), (op.pad, op.pad), (1, 1), g)
    acc = np.ascontiguousarray(acc.transpose(0, 2, 3, 1))
    N, P, Q, K = acc.shape
    y = np.empty(acc.shape, orc._NP_OUT[out])
    orc.lib().orc_epilogue(orc._p(acc), orc.C.c_int64(N * P * Q), K, orc._p(sf), orc._p(bf), act,
                           orc._f(arg), out, orc.ROUND_ARM, orc._p(y))
    return y


def run_program(prog, x_nchw, engine="oracle", keep=False, timings=None):
    """x_nchw: fp32 [N,3,H,W].  Returns fp32 logits [N,1000] (and all tensors if keep)."""
    class _Vals(dict):  # channel slices (prog.views) read their parent
        def __missing__(self, k):
            parent, c0, c1 = prog.views[k]
            return np.ascontiguousarray(self[parent][..., c0:c1])

    vals = _Vals()
    for op in prog.ops:
        t0 = time.perf_counter()
        if op.kind == "quant_in":
            vals[op.dst] = orc.nchw_to_nhwc(orc.quantize(x_nchw, op.out_scale))
        elif op.kind in ("conv", "dwconv"):
            out = orc.OUT_FP32 if op.out_fp32 else orc.OUT_INT8
            act = _ACT[op.act]
            arg = 6.0 if op.act == "relu6" else 0.1
            sf, bf, argf = orc.fold_scales(op.w_scale, op.K, op.in_scale, 1.0 if op.out_fp32 else op.out_scale,
                                           out, op.bias, act, arg)
            x = vals[op.src]
            if op.get("res"):
                rs = op.res_scale if op.out_fp32 else float(np.float32(op.res_scale) / np.float32(op.out_scale))
                g = op.K if op.kind == "dwconv" else 1
                vals[op.dst] = orc.conv2d(x, op.w, sf, bf, (op.stride, op.stride), (op.pad,) * 4, (1, 1), g,
                                          act, argf, out, residual=vals[op.res], res_scale=rs)
            elif engine == "reference":
                vals[op.dst] = _conv_reference(x, op, sf, bf, act, argf, out)
            else:
                g = op.K if op.kind == "dwconv" else 1
                vals[op.dst] = orc.conv2d(x, op.w, sf, bf, (op.stride, op.stride), (op.pad,) * 4, (1, 1), g,
                                          act, argf, out)
        elif op.kind == "pool":
            typ = orc.POOL_MAX if op.typ == "max" else orc.POOL_AVG
            vals[op.dst] = orc.pool2d(vals[op.src], (op.k, op.k), (op.stride, op.stride), (op.pad,) * 4, typ,
                                      True, op.in_scale / op.out_scale)
        elif op.kind == "add":
            vals[op.dst] = orc.add(vals[op.src], vals[op.src2], op.a_scale / op.out_scale,
                                   op.b_scale / op.out_scale, _ACT[op.act])
        elif op.kind == "cat_shuffle":
            # the reference's two ops, on NCHW: concat (axis 1) then shuffle_channel (group 2)
            cat = orc.concat_nchw([np.moveaxis(vals[op.src], -1, 1), np.moveaxis(vals[op.src2], -1, 1)])
            vals[op.dst] = np.ascontiguousarray(np.moveaxis(orc.shuffle_channel_nchw(cat, 2), 1, -1))
        elif op.kind == "fc":
            x = vals[op.src]
            A = np.ascontiguousarray(x.reshape(x.shape[0], -1))
            sf, bf, _ = orc.fold_scales(op.w_scale, op.N, op.in_scale, 1.0, orc.OUT_FP32, op.bias)
            if engine == "reference":
                acc = orc.ref_gemm_acc(A, op.w)
                y = np.empty(acc.shape, np.float32)
                orc.lib().orc_epilogue(orc._p(acc), orc.C.c_int64(acc.shape[0]), op.N, orc._p(sf), orc._p(bf),
                                       orc.ACT_NONE, orc._f(0), orc.OUT_FP32, orc.ROUND_ARM, orc._p(y))
            else:
                y = orc.gemm(A, op.w, sf, bf, out=orc.OUT_FP32)
            vals[op.dst] = y
            last = y
        if timings is not None:
            timings.append((op.kind, time.perf_counter() - t0))
    return (last, vals) if keep else last
