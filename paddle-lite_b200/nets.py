"""Synthetic int8 programs (MobileNetV1, ResNet-50) and their device executor.

A *program* is a plain list of op records (numpy weights, scales, shapes) -- the
state Paddle-Lite's optimiser leaves behind after quant_dequant_fuse / conv_bn_fuse
/ conv_activation_fuse / static_kernel_pick: every conv carries
`input_scale`, per-channel `weight_scale`, `output_scale`, folded-BN bias and a
fused activation; consecutive int8 ops keep their tensors int8 + NHWC + on device.
`DeviceNet` binds a program to pre-allocated device buffers and enqueues the whole
network through the C-ABI on the current stream (optionally as one CUDA graph).
tests/ walk the same program with the CPU oracle.

Model files for MobileNetV1/ResNet-50 int8 are not in the reference repo
(lite/api/mobilenetv1_int8_test.cc needs --model_dir), so weights are random-init
with seeded generators, as BASELINE.json says.
"""
import os

import numpy as np
import torch

from . import capi as pl


def rup(v, m):
    return (v + m - 1) // m * m


class Op(dict):
    __getattr__ = dict.__getitem__
    __setattr__ = dict.__setitem__


class Builder:
    """Builds a program; activation scales are filled in later by calibration."""

    def __init__(self, name, seed, in_shape):
        self.name = name
        self.rng = np.random.default_rng(seed)
        self.ops = []
        self.shapes = {"input": in_shape}  # name -> (H, W, C)   (per image)
        self.views = {}  # name -> (parent, c0, c1): a channel slice (pointer offset + the parent's pitch)
        self.n_tensors = 0

    def _new(self, shape):
        self.n_tensors += 1
        name = f"t{self.n_tensors}"
        self.shapes[name] = shape
        return name

    def quant_in(self):
        H, W, C = self.shapes["input"]
        out = self._new((H, W, C))
        self.ops.append(Op(kind="quant_in", src="input", dst=out, out_scale=None))
        return out

    def slice(self, src, c0, c1):
        """Channel slice [c0, c1) of a tensor: free (split is a view in NHWC with an explicit pitch)."""
        H, W, C = self.shapes[src]
        assert 0 <= c0 < c1 <= C and src not in self.views
        self.n_tensors += 1
        name = f"t{self.n_tensors}"
        self.shapes[name] = (H, W, c1 - c0)
        self.views[name] = (src, c0, c1)
        return name

    def cat_shuffle(self, a, b):
        """concat(a, b) along channels + shuffle_channel(group 2): the tail of a ShuffleNet-v2 unit.
        Pure byte movement, so both inputs must carry the same scale (see conv(out_scale_like=))."""
        H, W, C = self.shapes[a]
        assert self.shapes[b] == (H, W, C)
        out = self._new((H, W, 2 * C))
        self.ops.append(Op(kind="cat_shuffle", src=a, src2=b, dst=out, in_scale=None, out_scale=None))
        return out

    def conv(self, src, K, k, stride, pad, act="relu", depthwise=False, out_fp32=False, res=None,
             out_scale_like=None):
        H, W, C = self.shapes[src]
        P = (H + 2 * pad - k) // stride + 1
        Q = (W + 2 * pad - k) // stride + 1
        cg = 1 if depthwise else C
        w = self.rng.integers(-127, 128, (K, k, k, cg), dtype=np.int8)
        w_scale = (self.rng.uniform(0.5, 2.0, K) / 127.0 / np.sqrt(k * k * cg)).astype(np.float32)
        bias = self.rng.uniform(-0.5, 0.5, K).astype(np.float32)
        out = self._new((P, Q, K))
        self.ops.append(Op(kind="dwconv" if depthwise else "conv", src=src, dst=out, w=w, w_scale=w_scale,
                           bias=bias, k=k, stride=stride, pad=pad, act=act, K=K, C=C,
                           in_scale=None, out_scale=None, out_fp32=out_fp32, res=res, res_scale=None,
                           out_scale_like=out_scale_like))
        return out

    def pool(self, src, k, stride, pad, typ, global_pooling=False):
        H, W, C = self.shapes[src]
        if global_pooling:
            k, stride, pad = H, 1, 0
        P = (H + 2 * pad - k) // stride + 1
        Q = (W + 2 * pad - k) // stride + 1
        out = self._new((P, Q, C))
        self.ops.append(Op(kind="pool", src=src, dst=out, k=k, stride=stride, pad=pad, typ=typ,
                           in_scale=None, out_scale=None))
        return out

    def add(self, a, b, act="relu"):
        out = self._new(self.shapes[a])
        self.ops.append(Op(kind="add", src=a, src2=b, dst=out, act=act, a_scale=None, b_scale=None,
                           out_scale=None))
        return out

    def fc(self, src, N):
        H, W, C = self.shapes[src]
        Kd = H * W * C
        w = self.rng.integers(-127, 128, (Kd, N), dtype=np.int8)  # fc W is [K, N]
        w_scale = (self.rng.uniform(0.5, 2.0, N) / 127.0 / np.sqrt(Kd)).astype(np.float32)
        bias = self.rng.uniform(-0.5, 0.5, N).astype(np.float32)
        out = self._new((1, 1, N))
        self.ops.append(Op(kind="fc", src=src, dst=out, w=w, w_scale=w_scale, bias=bias, K=Kd, N=N,
                           in_scale=None, out_fp32=True))
        return out


def mobilenet_v1(seed=1234, act="relu", res=224):
    """MobileNetV1 1.0: conv0 3x3 s2 + 13 x (dw3x3 + pw1x1) + gap + fc."""
    b = Builder("mobilenet_v1", seed, (res, res, 3))
    t = b.quant_in()
    t = b.conv(t, 32, 3, 2, 1, act)
    cfg = [(64, 1), (128, 2), (128, 1), (256, 2), (256, 1), (512, 2), (512, 1), (512, 1), (512, 1),
           (512, 1), (512, 1), (1024, 2), (1024, 1)]
    for K, s in cfg:
        C = b.shapes[t][2]
        t = b.conv(t, C, 3, s, 1, act, depthwise=True)
        t = b.conv(t, K, 1, 1, 0, act)
    t = b.pool(t, 7, 1, 0, "avg", global_pooling=True)
    b.fc(t, 1000)
    return b


def shufflenet_v2_like(seed=1234, res=224, width=(128, 256, 512), repeats=(4, 8, 4)):
    """ShuffleNet-v2 topology (the backbone family of humanseg_shfv75_0119_quant) with channel counts that keep
    every half a multiple of 16 bytes: 3x3 s2 stem + max pool, three stages of one stride-2 unit + stride-1
    units, 1x1 head, gap, fc.  Stride-1 unit: split (free views) -> [x1 | x2 -> 1x1 -> dw3x3 -> 1x1] ->
    concat + shuffle_channel(2) (one fused pass); stride-2 unit: both branches see the whole input."""
    b = Builder("shufflenet_v2_like", seed, (res, res, 3))
    t = b.quant_in()
    t = b.conv(t, 32, 3, 2, 1, "relu")
    t = b.pool(t, 3, 2, 1, "max")
    for C, rep in zip(width, repeats):
        half = C // 2
        Cin = b.shapes[t][2]
        # stride-2 unit
        l = b.conv(t, Cin, 3, 2, 1, "none", depthwise=True)
        l = b.conv(l, half, 1, 1, 0, "relu")
        r = b.conv(t, half, 1, 1, 0, "relu")
        r = b.conv(r, half, 3, 2, 1, "none", depthwise=True)
        r = b.conv(r, half, 1, 1, 0, "relu", out_scale_like=l)
        t = b.cat_shuffle(l, r)
        for _ in range(rep - 1):
            x1, x2 = b.slice(t, 0, half), b.slice(t, half, C)
            r = b.conv(x2, half, 1, 1, 0, "relu")
            r = b.conv(r, half, 3, 1, 1, "none", depthwise=True)
            r = b.conv(r, half, 1, 1, 0, "relu", out_scale_like=x1)
            t = b.cat_shuffle(x1, r)
    t = b.conv(t, 1024, 1, 1, 0, "relu")
    t = b.pool(t, 7, 1, 0, "avg", global_pooling=True)
    b.fc(t, 1000)
    return b


def resnet50(seed=1234, res=224, fuse_residual=True):
    """ResNet-50 v1: 7x7 s2 stem, 3x3 s2 max pool, [3,4,6,3] bottlenecks (stride on the 3x3), gap, fc.
    fuse_residual: the shortcut add + relu runs in the epilogue of the block's last 1x1 conv
    (ConvParam::residualData; the reference's conv_elementwise fuse pass does this for fp32)
    instead of a separate int8 elementwise_add."""
    b = Builder("resnet50", seed, (res, res, 3))
    t = b.quant_in()
    t = b.conv(t, 64, 7, 2, 3, "relu")
    t = b.pool(t, 3, 2, 1, "max")
    for stage, (mid, blocks) in enumerate([(64, 3), (128, 4), (256, 6), (512, 3)]):
        for blk in range(blocks):
            stride = 2 if (blk == 0 and stage > 0) else 1
            short = t
            if blk == 0:
                short = b.conv(t, mid * 4, 1, stride, 0, "none")
            u = b.conv(t, mid, 1, 1, 0, "relu")
            u = b.conv(u, mid, 3, stride, 1, "relu")
            if fuse_residual:
                t = b.conv(u, mid * 4, 1, 1, 0, "relu", res=short)
            else:
                u = b.conv(u, mid * 4, 1, 1, 0, "none")
                t = b.add(u, short, "relu")
    t = b.pool(t, 7, 1, 0, "avg", global_pooling=True)
    b.fc(t, 1000)
    return b


def program_macs(prog):
    """MACs per image by op kind (the reference's OpCharacter formula, lite/operators/conv_op.h:43-62)."""
    tot = {"conv": 0, "dwconv": 0, "fc": 0}
    for op in prog.ops:
        if op.kind in ("conv", "dwconv"):
            P, Q, K = prog.shapes[op.dst]
            cg = 1 if op.kind == "dwconv" else op.C
            tot[op.kind] += P * Q * K * op.k * op.k * cg
        elif op.kind == "fc":
            tot["fc"] += op.K * op.N
    return tot


def program_bytes(prog):
    """Algorithmic bytes per image: every activation read/written once at its storage width,
    weights + scale/bias once per batch (returned separately)."""
    act = 0
    wts = 0
    for op in prog.ops:
        H, W, C = prog.shapes[op.src]
        P, Q, K = prog.shapes[op.dst]
        if op.kind == "quant_in":
            act += H * W * C * 4 + P * Q * K
        elif op.kind in ("conv", "dwconv"):
            act += H * W * C + P * Q * K * (4 if op.out_fp32 else 1) + (P * Q * K if op.get("res") else 0)
            wts += op.w.size + 8 * K
        elif op.kind == "pool":
            act += H * W * C + P * Q * K
        elif op.kind == "add":
            act += 3 * H * W * C
        elif op.kind == "cat_shuffle":
            act += 2 * P * Q * K
        elif op.kind == "fc":
            act += op.K + op.N * 4
            wts += op.w.size + 8 * op.N
    return act, wts


_FORCE_IM2COL = os.environ.get("PL_THIN_IM2COL", "0") == "1"  # A/B switch: old explicit-im2col first layer
_ACT = {"none": pl.ACT_NONE, "relu": pl.ACT_RELU, "relu6": pl.ACT_RELU6, "leaky": pl.ACT_LEAKY}


class DeviceNet:
    """A program bound to device buffers for a fixed batch.  `calibrate()` fixes the
    activation scales (abs-max / 127 on a calibration batch, computed with the fp32-out
    kernels); `run()` enqueues the int8 network on the current stream."""

    FEEDS = {"f32": torch.float32, "f16": torch.float16, "u8": torch.uint8}

    def __init__(self, prog, batch, device="cuda", feed="f32", mean=None, norm=None):
        """feed: element type of the NCHW input tensor.  "u8" is a raw image: the entry kernel applies
        x = (u8 - mean[c]) * norm[c] (the reference's host-side image2tensor step) before the calib."""
        self.prog = prog
        self.batch = batch
        self.dev = device
        self.bufs = {}
        self.layers = []
        self.graph = None
        H, W, C = prog.shapes["input"]
        self.feed = feed
        self.mean = list(mean) if mean is not None else [0.0] * C
        self.norm = list(norm) if norm is not None else [1.0] * C
        self.input = torch.zeros((batch, C, H, W), dtype=self.FEEDS[feed], device=device)  # NCHW
        self.calibrated = all(op.get("out_scale", 1.0) is not None or op.kind == "fc" for op in prog.ops)
        self._weights_to_device()

    # -- one-off (PrepareForRun) ------------------------------------------------
    def _s2d_geometry(self, op, n):
        H, W, _ = self.prog.shapes[op.src]
        return pl.s2d_geometry(n, H, W, op.C, op.k, op.k, (op.stride, op.stride), (op.pad,) * 4)

    def _weights_to_device(self):
        self.s2d_src = {}  # tensor name -> stem conv that reads it in space-to-depth form
        quant_dsts = {op.dst for op in self.prog.ops if op.kind == "quant_in"}
        for op in self.prog.ops:
            readers = [o for o in self.prog.ops if op.get("src") in (o.get("src"), o.get("src2"), o.get("res"))]
            if (op.kind == "conv" and op.C <= 4 and op.src in quant_dsts and len(readers) == 1 and not _FORCE_IM2COL
                    and not os.environ.get("PL_NO_S2D") and self._s2d_geometry(op, self.batch).supported):
                # stride-2 thin stem: space-to-depth input, spans bulk-copied and read in place by the MMA
                op["_wp"] = pl.pack_s2d_conv_weight(torch.from_numpy(op.w).to(self.dev), self._s2d_geometry(op, self.batch))
                op["_s2d"] = True
                self.s2d_src[op.src] = op
            elif op.kind == "conv" and op.C <= 4 and pl.thin_conv_supported(op.k, op.k, op.C) and not _FORCE_IM2COL:
                # thin first layer, direct: window words gathered on chip by the conv kernel
                op["_wp"] = pl.pack_thin_conv_weight(torch.from_numpy(op.w).to(self.dev))
                op["_thin"] = True
            elif op.kind == "conv" and op.C <= 4:
                # thin first layer, fallback: explicit im2col (one word per pixel) + tensor-core GEMM
                K, R, S, C = op.w.shape
                w4 = np.zeros((K, R, S, 4), np.int8)
                w4[..., :C] = op.w
                ldA = rup(R * S * 4, 16)
                bt = np.zeros((K, ldA), np.int8)
                bt[:, : R * S * 4] = w4.reshape(K, -1)
                op["_wp"] = torch.from_numpy(bt).to(self.dev)
                op["_ldA"] = ldA
            elif op.kind == "conv":
                op["_wp"] = pl.pack_conv_weight(torch.from_numpy(op.w).to(self.dev))
            elif op.kind == "dwconv":
                op["_wp"] = torch.from_numpy(np.ascontiguousarray(op.w.reshape(op.K, op.k, op.k))).to(self.dev)
            elif op.kind == "fc":
                op["_wp"] = pl.transpose_i8(torch.from_numpy(op.w).to(self.dev))  # [K,N] -> [N,K]
        torch.cuda.synchronize()

    def _buf(self, name, dtype=torch.int8):
        if name in self.prog.views:  # a channel slice of its parent's buffer
            parent, c0, c1 = self.prog.views[name]
            return self._buf(parent, dtype)[..., c0:c1]
        key = (name, dtype)
        if key not in self.bufs:
            H, W, C = self.prog.shapes[name]
            ld = (4 if C <= 4 else rup(C, 16)) if dtype == torch.int8 else C
            self.bufs[key] = torch.zeros((self.batch, H, W, ld), dtype=dtype, device=self.dev)
        return self.bufs[key]

    def _fold(self, op, n_out, out_fp32):
        out_dtype = pl.OUT_FP32 if out_fp32 else pl.OUT_INT8
        arg = 6.0 if op.get("act") == "relu6" else 0.1
        sf, bf, argf = pl.fold_scales(op.w_scale, n_out, op.in_scale, 1.0 if out_fp32 else op.out_scale,
                                      out_dtype, op.bias, _ACT[op.get("act", "none")], arg)
        return torch.from_numpy(sf).to(self.dev), torch.from_numpy(bf).to(self.dev), argf

    # -- scale calibration ------------------------------------------------------
    def calibrate(self, x_nchw):
        """Propagates scales layer by layer: out_scale = absmax(fp32 output)/127."""
        assert x_nchw.shape[0] <= self.batch
        n = x_nchw.shape[0]
        class _Views(dict):  # a channel slice reads its parent's value / scale
            def __init__(self, views, sl):
                super().__init__()
                self.views, self.sl = views, sl

            def __missing__(self, k):
                parent, c0, c1 = self.views[k]
                return self[parent][..., c0:c1] if self.sl else self[parent]

        scales = _Views(self.prog.views, False)
        vals = _Views(self.prog.views, True)
        for op in self.prog.ops:
            if op.kind == "quant_in":
                xf = x_nchw.float()
                if self.feed == "u8":
                    C_ = xf.shape[1]
                    xf = (xf - torch.tensor(self.mean, device=xf.device).view(1, C_, 1, 1)) * \
                        torch.tensor(self.norm, device=xf.device).view(1, C_, 1, 1)
                op.out_scale = float(xf.abs().max().item()) / 127.0
                C_in = x_nchw.shape[1]
                if op.dst in self.s2d_src:
                    vals[op.dst] = pl.quantize_nchw_to_s2d(x_nchw.contiguous(), op.out_scale,
                                                           self._s2d_geometry(self.s2d_src[op.dst], n),
                                                           mean=self.mean, norm=self.norm)
                    vals[op.dst]._pl_n = n
                else:
                    assert self.feed == "f32", "narrow feeds exist for the space-to-depth stems only"
                    vals[op.dst] = pl.quantize_nchw_to_nhwc(x_nchw.contiguous(), op.out_scale,
                                                            4 if C_in <= 4 else rup(C_in, 16))
                scales[op.dst] = op.out_scale
            elif op.kind in ("conv", "dwconv"):
                op.in_scale = scales[op.src]
                res = vals[op.res] if op.get("res") else None
                if res is not None:
                    op.res_scale = scales[op.res]
                sf, bf, argf = self._fold(op, op.K, True)
                y = self._exec_conv(op, vals[op.src], sf, bf, argf, pl.OUT_FP32, None, res, op.res_scale)
                op.out_scale = max(float(y.abs().max().item()), 1e-6) / 127.0
                if op.get("out_scale_like"):  # shares a concat with that tensor: requantise to its scale
                    op.out_scale = scales[op.out_scale_like]
                sf, bf, argf = self._fold(op, op.K, op.out_fp32)
                vals[op.dst] = self._exec_conv(op, vals[op.src], sf, bf, argf,
                                               pl.OUT_FP32 if op.out_fp32 else pl.OUT_INT8, None, res,
                                               self._res_scale(op))
                scales[op.dst] = op.out_scale
            elif op.kind == "pool":
                op.in_scale = scales[op.src]
                op.out_scale = op.in_scale  # pooling keeps the scale (max exact; avg requantised 1:1)
                vals[op.dst] = self._exec_pool(op, vals[op.src], None)
                scales[op.dst] = op.out_scale
            elif op.kind == "add":
                op.a_scale, op.b_scale = scales[op.src], scales[op.src2]
                y = pl.add(vals[op.src], vals[op.src2], op.a_scale, op.b_scale, _ACT[op.act], out=pl.OUT_FP32)
                op.out_scale = max(float(y.abs().max().item()), 1e-6) / 127.0
                vals[op.dst] = pl.add(vals[op.src], vals[op.src2], op.a_scale / op.out_scale,
                                      op.b_scale / op.out_scale, _ACT[op.act])
                scales[op.dst] = op.out_scale
            elif op.kind == "cat_shuffle":
                assert scales[op.src] == scales[op.src2], "concat inputs must share a scale"
                op.in_scale = op.out_scale = scales[op.src]
                vals[op.dst] = pl.concat_shuffle2(vals[op.src], vals[op.src2])
                scales[op.dst] = op.out_scale
            elif op.kind == "fc":
                op.in_scale = scales[op.src]
                vals[op.dst] = None
        torch.cuda.synchronize()
        self.calibrated = True
        del vals
        return self

    # -- per-op launchers -------------------------------------------------------
    def _res_scale(self, op):
        if not op.get("res"):
            return 0.0
        # single-precision division, as the kernel class computes it from its float attributes
        # (conv_compute_int8.cc: scale_in_eltwise / output_scale): both paths then feed identical epilogue inputs
        return op.res_scale if op.out_fp32 else float(np.float32(op.res_scale) / np.float32(op.out_scale))

    def _exec_conv(self, op, x, sf, bf, argf, out, y, res=None, res_scale=0.0):
        C = op.C
        if op.kind == "dwconv":
            return pl.dwconv2d(x, op._wp, sf, bf, op.k, op.k, (op.stride, op.stride), (op.pad,) * 4,
                               (1, 1), _ACT[op.act], argf, out, y=y, Cc=C)
        K = op.K
        if op.get("_s2d"):
            H, W, _ = self.prog.shapes[op.src]
            return pl.s2d_conv2d(x, op._wp, x._pl_n, H, W, C, K, op.k, op.k, sf, bf, (op.stride, op.stride),
                                 (op.pad,) * 4, _ACT[op.act], argf, out, y=y,
                                 ldy=(rup(K, 16) if out == pl.OUT_INT8 else K))
        if C <= 4 and op.get("_thin"):
            return pl.thin_conv2d(x, op._wp, K, op.k, op.k, C, sf, bf, (op.stride, op.stride), (op.pad,) * 4,
                                  _ACT[op.act], argf, out, y=y, ldy=(rup(K, 16) if out == pl.OUT_INT8 else K))
        if C <= 4:
            N_, H_, W_, _ = x.shape
            P = (H_ + 2 * op.pad - op.k) // op.stride + 1
            Q = (W_ + 2 * op.pad - op.k) // op.stride + 1
            A = self._im2col_buf(op, N_ * P * Q)
            pl.im2col(x, op.k, op.k, (op.stride, op.stride), (op.pad,) * 4, A=A)
            ldy = rup(K, 16) if out == pl.OUT_INT8 else K
            if y is None:
                y = torch.zeros((N_, P, Q, ldy), dtype=pl._TORCH_OUT[out], device=x.device)
            pl.gemm(A, op._wp, sf, bf, _ACT[op.act], argf, out, y=y.view(N_ * P * Q, ldy), M=N_ * P * Q, N=K,
                    K=op._ldA)
            return y
        return pl.conv2d(x, op._wp, K, op.k, op.k, sf, bf, (op.stride, op.stride), (op.pad,) * 4, (1, 1),
                         _ACT[op.act], argf, out, y=y, Cc=C, ldy=(rup(K, 16) if out == pl.OUT_INT8 else K),
                         residual=res, res_scale=res_scale or 0.0)

    def _im2col_buf(self, op, rows):
        key = ("im2col", id(op))
        buf = self.bufs.get(key)
        if buf is None or buf.shape[0] < rows:
            buf = torch.empty((rows, op._ldA), dtype=torch.int8, device=self.dev)
            self.bufs[key] = buf
        return buf[:rows]

    def _exec_pool(self, op, x, y):
        H, W, C = self.prog.shapes[op.src]
        typ = pl.POOL_MAX if op.typ == "max" else pl.POOL_AVG
        return pl.pool2d(x, (op.k, op.k), (op.stride, op.stride), (op.pad,) * 4, typ, True,
                         op.in_scale / op.out_scale, pl.OUT_INT8, Cc=C, y=y)

    # -- bind + run ---------------------------------------------------------------
    def bind(self, schedule=None):
        """Pre-computes folded scales and output buffers; builds the launch list.  Every entry runs its op on
        the images [i0, i1) of the batch (images are independent units on this path).

        schedule: [(first_op, last_op, chunks), ...] -- the ops of a segment run chunk by chunk (all ops on
        images [0, B/chunks), then on the next slice, ...) so that the activations between them stay in the
        126 MB L2 instead of making a round trip through HBM at full batch.  Default: PL_SUBBATCH
        ("first-last:chunks,..."), else every op once on the whole batch."""
        assert self.calibrated, "calibrate() or load scales first"
        self.layers = []
        self.s2d_bufs = {}
        self.schedule = self._parse_schedule(schedule)
        B = self.batch
        for op in self.prog.ops:
            if op.kind == "quant_in" and op.dst in self.s2d_src:
                g = self._s2d_geometry(self.s2d_src[op.dst], self.batch)
                y = pl.s2d_input_buffer(g, self.dev)  # zero borders, written once here
                y._pl_n = self.batch
                self.s2d_bufs[op.dst] = (y, g)
                img = g.HP * g.WS * 16
                self.layers.append((op, lambda i0=0, i1=B, op=op, y=y, g=g, img=img: pl.quantize_nchw_to_s2d(
                    self.input[i0:i1], op.out_scale, g, y=y[i0 * img:], mean=self.mean, norm=self.norm)))
            elif op.kind == "quant_in":
                assert self.feed == "f32", "narrow feeds exist for the space-to-depth stems only"
                y = self._buf(op.dst)
                self.layers.append((op, lambda i0=0, i1=B, op=op, y=y: pl.quantize_nchw_to_nhwc(
                    self.input[i0:i1], op.out_scale, y.shape[-1], y=y[i0:i1])))
            elif op.kind in ("conv", "dwconv"):
                sf, bf, argf = self._fold(op, op.K, op.out_fp32)
                y = self._buf(op.dst, torch.float32 if op.out_fp32 else torch.int8)
                out = pl.OUT_FP32 if op.out_fp32 else pl.OUT_INT8
                res = self._buf(op.res) if op.get("res") else None
                rs = self._res_scale(op)
                if op.src in self.s2d_bufs:
                    xs, g = self.s2d_bufs[op.src]
                    img = g.HP * g.WS * 16

                    def fn(i0=0, i1=B, op=op, xs=xs, img=img, y=y, sf=sf, bf=bf, argf=argf, out=out):
                        x = xs[i0 * img:]
                        x._pl_n = i1 - i0
                        return self._exec_conv(op, x, sf, bf, argf, out, y[i0:i1], None, 0.0)
                    self.layers.append((op, fn))
                else:
                    x = self._buf(op.src)
                    self.layers.append((op, lambda i0=0, i1=B, op=op, x=x, y=y, sf=sf, bf=bf, argf=argf, out=out,
                                        res=res, rs=rs:
                                        self._exec_conv(op, x[i0:i1], sf, bf, argf, out, y[i0:i1],
                                                        None if res is None else res[i0:i1], rs)))
            elif op.kind == "pool":
                x, y = self._buf(op.src), self._buf(op.dst)
                self.layers.append((op, lambda i0=0, i1=B, op=op, x=x, y=y: self._exec_pool(op, x[i0:i1], y[i0:i1])))
            elif op.kind == "add":
                a, b2, y = self._buf(op.src), self._buf(op.src2), self._buf(op.dst)
                sa, sb = op.a_scale / op.out_scale, op.b_scale / op.out_scale
                self.layers.append((op, lambda i0=0, i1=B, op=op, a=a, b2=b2, y=y, sa=sa, sb=sb:
                                    pl.add(a[i0:i1], b2[i0:i1], sa, sb, _ACT[op.act], y=y[i0:i1])))
            elif op.kind == "cat_shuffle":
                a, b2, y = self._buf(op.src), self._buf(op.src2), self._buf(op.dst)
                self.layers.append((op, lambda i0=0, i1=B, a=a, b2=b2, y=y:
                                    pl.concat_shuffle2(a[i0:i1], b2[i0:i1], y=y[i0:i1])))
            elif op.kind == "fc":
                sf, bf, _ = pl.fold_scales(op.w_scale, op.N, op.in_scale, 1.0, pl.OUT_FP32, op.bias)
                sf, bf = torch.from_numpy(sf).to(self.dev), torch.from_numpy(bf).to(self.dev)
                x = self._buf(op.src)
                A = x.view(self.batch, -1)
                y = self._buf(op.dst, torch.float32).view(self.batch, op.N)
                self.output = y
                self.layers.append((op, lambda i0=0, i1=B, op=op, A=A, y=y, sf=sf, bf=bf:
                                    pl.gemm(A[i0:i1], op._wp, sf, bf, out=pl.OUT_FP32, y=y[i0:i1], K=op.K)))
        # Tile order (pl_set_tile_order): an op walks its output tiles in the opposite direction of the op that wrote
        # its input, so it starts on the bytes that were written last and are still in the 126 MB L2 (results do not
        # depend on the order; PL_NET_ZIGZAG=0 keeps every op first-to-last)
        if os.environ.get("PL_NET_ZIGZAG", "1") != "0":
            wrote_reversed = {}
            for k, (op, fn) in enumerate(self.layers):
                r = (not wrote_reversed.get(op.get("src"), True)) if op.kind != "quant_in" else False
                wrote_reversed[op.dst] = r

                def ordered(i0=0, i1=B, fn=fn, r=r):
                    pl.set_tile_order(r)
                    try:
                        return fn(i0, i1)
                    finally:
                        pl.set_tile_order(False)
                self.layers[k] = (op, ordered)
        return self

    def _parse_schedule(self, schedule):
        n = len(self.prog.ops)
        if schedule is None:
            env = os.environ.get("PL_SUBBATCH", "")
            schedule = []
            for part in filter(None, env.split(",")):
                rng, chunks = part.split(":")
                lo, hi = rng.split("-")
                schedule.append((int(lo), int(hi), int(chunks)))
        segs, at = [], 0
        for lo, hi, chunks in sorted(schedule):
            hi = min(hi, n - 1)
            chunks = max(1, min(int(chunks), self.batch))
            if lo < at or lo > hi:
                continue
            if lo > at:
                segs.append((at, lo - 1, 1))
            segs.append((lo, hi, chunks))
            at = hi + 1
        if at < n:
            segs.append((at, n - 1, 1))
        return segs

    def s2d_as_nhwc(self, name):
        """Host copy of a space-to-depth input buffer, unfolded to int8 NHWC [N,H,W,4] (for checks)."""
        y, g = self.s2d_bufs[name]
        H, W, _ = self.prog.shapes[name]
        img = y[: self.batch * g.HP * g.WS * 16].view(self.batch, g.HP, g.WS, 2, 2, 4).cpu().numpy()
        img = img[:, g.pad_s:g.pad_s + H // 2, g.pad_s:g.pad_s + W // 2]
        return img.transpose(0, 1, 3, 2, 4, 5).reshape(self.batch, H, W, 4)

    def launch_order(self):
        """Layer index of every launch of one step, in execution order (chunked segments repeat their ops)."""
        order = []
        for lo, hi, chunks in self.schedule:
            for _ in range(chunks):
                order.extend(range(lo, hi + 1))
        return order

    def run(self):
        B = self.batch
        for lo, hi, chunks in self.schedule:
            for c in range(chunks):
                i0, i1 = c * B // chunks, (c + 1) * B // chunks
                for _, fn in self.layers[lo:hi + 1]:
                    fn(i0, i1)
        return self.output

    def capture(self):
        """One CUDA graph for the whole network (launch-bound at small batch)."""
        self.run()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            self.run()
        self.graph = g
        return self

    def capture_inputs(self, n=2):
        """One CUDA graph per fp32 input buffer (`self.inputs[i]`): a feeder can copy batch i+1 from pinned
        host memory straight into the other buffer while graph i runs -- no staging copy on the device.
        The graphs differ only in the pointer their first kernel reads; every activation buffer is shared."""
        self.inputs = [self.input] + [torch.zeros_like(self.input) for _ in range(n - 1)]
        self.graphs = []
        first = self.input
        for buf in self.inputs:
            self.input = buf
            self.capture()
            self.graphs.append(self.graph)
        self.input, self.graph = first, self.graphs[0]
        return self

    def replay(self, i=None):
        (self.graph if i is None else self.graphs[i]).replay()
        return self.output

    def forward(self, x_nchw):
        """Public call: fp32 NCHW device tensor in, fp32 logits out (device)."""
        self.input.copy_(x_nchw, non_blocking=True)
        return self.replay() if self.graph is not None else self.run()
