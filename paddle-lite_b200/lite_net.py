"""A whole network through the reference-facing plugin surface: every op is a KernelLite instance created through
KernelRegistry (REGISTER_LITE_KERNEL keys), given its param struct and a CUDAContext, and run by Launch() from a
runtime program (lite/api/harness.cc: Program -- the RuntimeProgram::Run loop of lite/core/program.cc:265-315).
The op chain is the one the reference's passes produce for a quantised conv net on kCUDA:

    feed (fp32 NCHW, device) -> calib fp32_to_int8 -> layout int8 nchw2nhwc -> conv2d / depthwise_conv2d (int8_out,
    fused bias + act + residual) -> pool2d -> ... -> fc (fp32_out)

`PluginNet` has the surface bench.py's end-to-end loop drives (inputs / replay / output / batch), so `e2e_plugin` is
measured by exactly the loop that measures `e2e`: the difference is the path -- libpl_lite_kernels.so's kernel classes
instead of the ctypes C-ABI calls of nets.DeviceNet.  Feed and fetch tensors are torch allocations shared into
lite::Tensor (ShareExternalMemory); everything in between is owned by the lite tensors."""
import ctypes as C

import numpy as np
import torch

from . import lite_harness as lh

_ACT = {"none": 0, "relu": lh.ACT_RELU, "relu6": lh.ACT_RELU6, "leaky": lh.ACT_LEAKY}


def _share(t):
    """lite::Tensor over a torch device tensor"""
    prec = {torch.float32: lh.K_FLOAT, torch.int8: lh.K_INT8, torch.int32: lh.K_INT32}[t.dtype]
    out = lh.Tensor()
    dims = (C.c_int64 * t.dim())(*t.shape)
    lh._ck(lh.lib().plk_tensor_share_external(out.h, prec, dims, t.dim(), C.c_void_p(t.data_ptr()),
                                              C.c_int64(t.numel() * t.element_size())))
    out._keep = t
    return out


class PluginNet:
    def __init__(self, prog, batch, device="cuda", n_inputs=2):
        """prog: a calibrated nets.Program (scales set, e.g. by nets.DeviceNet.calibrate on the same object)"""
        L = lh.lib()
        L.plk_program_begin.restype = C.c_void_p
        self.prog, self.batch, self.dev = prog, batch, device
        H, W, Cc = prog.shapes["input"]
        self.inputs = [torch.zeros((batch, Cc, H, W), dtype=torch.float32, device=device) for _ in range(n_inputs)]
        self.input = self.inputs[0]
        self.keep = []
        self.t = {}
        self._stream = torch.cuda.current_stream().cuda_stream
        stream = C.c_void_p(self._stream)
        ops = prog.ops
        assert ops[0].kind == "quant_in"
        q = ops[0]
        # one head program per input buffer (calib + layout), all writing the same NHWC int8 tensor
        self.heads = []
        nhwc = lh.Tensor()
        self.t[q.dst] = nhwc
        for buf in self.inputs:
            feed, qt = _share(buf), lh.Tensor()
            self.keep += [feed, qt]
            head = C.c_void_p(L.plk_program_begin(stream))
            lh._ck(L.plk_run_calib(feed.h, qt.h, C.c_float(q.out_scale), 1))
            L.plk_tensor_set_precision(qt.h, lh.K_INT8)  # what the calib declares: the layout kernel is picked from it
            lh._ck(L.plk_run_layout(qt.h, nhwc.h, 1))
            L.plk_program_end()
            self.heads.append(head)
        self.body = C.c_void_p(L.plk_program_begin(stream))
        try:
            for op in ops[1:]:
                self._add(op)
        finally:
            L.plk_program_end()
        self.n_instructions = L.plk_program_size(self.body) + L.plk_program_size(self.heads[0])
        self.graphs = None

    def _tensor(self, name):
        if name not in self.t:
            self.t[name] = lh.Tensor()
        return self.t[name]

    def _const(self, a):
        t = lh.Tensor(np.ascontiguousarray(a))
        self.keep.append(t)
        return t

    def _add(self, op):
        L = lh.lib()
        k = op.kind
        if k in ("conv", "dwconv"):
            w = op.w if k == "conv" else np.ascontiguousarray(op.w).reshape(op.K, op.k, op.k, 1)
            x, y = self._tensor(op.src), self._tensor(op.dst)
            res = self._tensor(op.res) if op.get("res") else None
            ws, wsp, nws = lh._fa(op.w_scale)
            if op.out_fp32:  # fetch tensor of a conv-only net
                self._set_output(op, y)
            lh._ck(L.plk_run_conv(b"depthwise_conv2d" if k == "dwconv" else b"conv2d", x.h, self._const(w).h,
                                  self._const(np.asarray(op.bias, np.float32)).h if op.get("bias") is not None else None,
                                  lh._h(res), y.h, lh._ia((op.stride,) * 2), lh._ia((op.pad,) * 4), lh._ia((1, 1)),
                                  op.C if k == "dwconv" else op.get("groups", 1), _ACT[op.get("act", "none")],
                                  C.c_float(6.0 if op.get("act") == "relu6" else 0.1), C.c_float(op.in_scale), wsp, nws,
                                  C.c_float(op.out_scale if not op.out_fp32 else 1.0),
                                  C.c_float(op.get("res_scale") or 0.0), int(not op.out_fp32), 1))
        elif k == "pool":
            x, y = self._tensor(op.src), self._tensor(op.dst)
            lh._ck(L.plk_run_pool(x.h, y.h, op.typ.encode(), lh._ia((op.k,) * 2), lh._ia((op.stride,) * 2),
                                  lh._ia((op.pad,) * 4), 0, 1, 0, C.c_float(op.in_scale), C.c_float(op.out_scale), 1))
        elif k == "add":
            a, b, y = self._tensor(op.src), self._tensor(op.src2), self._tensor(op.dst)
            lh._ck(L.plk_run_add(a.h, b.h, y.h, C.c_float(op.a_scale), C.c_float(op.b_scale), C.c_float(op.out_scale),
                                 (op.act if op.act != "none" else "").encode(), 1))
        elif k == "fc":
            x, y = self._tensor(op.src), self._tensor(op.dst)
            self._set_output(op, y, (self.batch, op.N))
            ws, wsp, nws = lh._fa(op.w_scale)
            lh._ck(L.plk_run_fc(x.h, self._const(op.w).h, self._const(np.asarray(op.bias, np.float32)).h, y.h, 1, b"",
                                C.c_float(op.in_scale), wsp, nws, C.c_float(1.0), 0, 1))
        else:
            raise NotImplementedError(f"{k}: no KernelLite instruction for this op record")

    def _set_output(self, op, y, shape=None):
        if shape is None:
            P, Q, K = self.prog.shapes[op.dst]
            shape = (self.batch, P, Q, K)
        self.output = torch.zeros(shape, dtype=torch.float32, device=self.dev)
        self._out_t = y
        dims = (C.c_int64 * len(shape))(*shape)
        lh._ck(lh.lib().plk_tensor_share_external(y.h, lh.K_FLOAT, dims, len(shape), C.c_void_p(self.output.data_ptr()),
                                                  C.c_int64(self.output.numel() * 4)))

    def run(self, i=0):
        """Instruction by instruction (the eager RuntimeProgram::Run)"""
        L = lh.lib()
        cur = torch.cuda.current_stream().cuda_stream
        if cur != self._stream:  # e.g. the side stream of a graph capture
            for p_ in self.heads + [self.body]:
                lh._ck(L.plk_program_set_stream(p_, C.c_void_p(cur)))
            self._stream = cur
        lh._ck(L.plk_program_run(self.heads[i], 0))
        lh._ck(L.plk_program_run(self.body, 0))
        return self.output

    def capture(self):
        """One CUDA graph per input buffer around the same Launch() loop (after a warm-up run: PrepareForRun packs
        weights and the first Run sizes every tensor)."""
        for i in range(len(self.inputs)):
            self.run(i)
        torch.cuda.synchronize()
        out_ptr = self.output.data_ptr()
        self.graphs = []
        for i in range(len(self.inputs)):
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                self.run(i)
            self.graphs.append(g)
        raw = lh.lib().plk_tensor_raw
        raw.restype = C.c_void_p
        assert self.output.data_ptr() == out_ptr and raw(self._out_t.h) == out_ptr, "the fetch tensor was re-allocated"
        return self

    def replay(self, i=0):
        if self.graphs is None:
            return self.run(i or 0)
        self.graphs[i or 0].replay()
        return self.output

    def tensor_numpy(self, name):
        torch.cuda.synchronize()
        return self.t[name].numpy()

    def __del__(self):
        try:
            L = lh.lib()
            for h in self.heads:
                L.plk_program_destroy(h)
            L.plk_program_destroy(self.body)
        except Exception:
            pass
