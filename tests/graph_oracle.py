"""CPU walker of a fused model program (paddle-lite_b200/model.py: Program) on the oracle -- TEST INFRASTRUCTURE.
Every op is evaluated with the oracle's restatement of the reference arithmetic (tests/orc.py): int8 convs through
orc.conv2d (fold + fused epilogue), gathers as numpy indexing + the ARM calib, conv2d_transpose through the
pinned-order C loop, softmax in float64.  Returns every tensor of the program, NHWC, logical channels only."""
import numpy as np

import orc

_ACT = {"none": orc.ACT_NONE, "relu": orc.ACT_RELU, "relu6": orc.ACT_RELU6, "leaky": orc.ACT_LEAKY}


def run_graph(prog, x_nchw):
    x_nchw = np.ascontiguousarray(x_nchw, np.float32)
    vals = {"input": x_nchw}
    for op in prog.ops:
        k = op.kind
        if k == "quant_in":
            vals[op.dst] = orc.quantize(x_nchw.transpose(0, 2, 3, 1), op.out_scale)
        elif k in ("conv", "dwconv"):
            out = orc.OUT_FP32 if op.out_fp32 else orc.OUT_INT8
            sf, bf, arg = orc.fold_scales(op.w_scale, op.K, op.in_scale, 1.0 if op.out_fp32 else op.out_scale, out,
                                          op.bias, _ACT[op.act], 6.0)
            groups = op.K if k == "dwconv" else op.get("groups", 1)
            vals[op.dst] = orc.conv2d(vals[op.src], op.w, sf, bf, (op.stride, op.stride), (op.pad,) * 4, (1, 1), groups,
                                      _ACT[op.act], arg, out)
        elif k == "pool":
            typ = orc.POOL_MAX if op.typ == "max" else orc.POOL_AVG
            vals[op.dst] = orc.pool2d(vals[op.src], (op.k, op.k), (op.stride, op.stride), (op.pad,) * 4, typ, True,
                                      op.in_scale / op.out_scale, orc.OUT_INT8)
        elif k == "gather":
            srcs = [vals[s] for s in op.srcs]
            tab = np.asarray(op.table)
            y = np.stack([srcs[si][..., ch] for si, ch in tab], axis=-1).astype(np.float32)
            vals[op.dst] = y if op.out_scale is None else orc.quantize(y, op.out_scale)
        elif k == "deconv":
            vals[op.dst] = orc.conv2d_transpose_f32(vals[op.src], op.w, op.bias, (op.stride, op.stride), (op.pad, op.pad),
                                                    relu=op.act == "relu", out_scale=op.out_scale or 0.0)
        elif k == "addf":
            vals[op.dst] = (vals[op.src] + vals[op.src2]).astype(np.float32)
        elif k == "softmax":
            v = vals[op.src].astype(np.float64)
            e = np.exp(v - v.max(-1, keepdims=True))
            vals[op.dst] = (e / e.sum(-1, keepdims=True)).astype(np.float32)
        elif k == "scalef":
            v = vals[op.src]
            vals[op.dst] = (v * np.float32(op.scale) + np.float32(op.bias)).astype(np.float32) if op.bias_after_scale \
                else ((v + np.float32(op.bias)) * np.float32(op.scale)).astype(np.float32)
        elif k == "to_nhwc":
            vals[op.dst] = np.ascontiguousarray(x_nchw.transpose(0, 2, 3, 1))
        elif k == "to_nchw":
            vals[op.dst] = np.ascontiguousarray(vals[op.src].transpose(0, 3, 1, 2))
        else:
            raise NotImplementedError(k)
    return vals
