"""Predictor-level API of the hot path: the reference's public C++ surface, same names and
argument meaning, over this repo's front end (model.py) and device executor.

  CxxConfig / Place / CreatePaddlePredictor      lite/api/paddle_api.h:138-362, lite/api/paddle_place.h:45-100
  PaddlePredictor (GetInput / GetOutput / Run / GetInputNames / GetOutputNames / GetInputByName / GetTensor /
                   Clone / SaveOptimizedModel / GetVersion)   lite/api/paddle_api.h:90-131
  Tensor (Resize / shape / mutable_data / data / CopyFromCpu / CopyToCpu / target / precision)
                                                 lite/api/paddle_api.h:47-86
  Build: load ProgramDesc + params, run the fusion passes, pick kernels for the valid places
                                                 lite/api/cxx_api.cc:248-355
  Run: one instruction per op record, optional NVTX range per instruction
                                                 lite/core/program.cc:265-315 (NVTX: :272-293)
  Profiler: per-instruction timers and the Summary table, MACs by the reference's formula
                                                 lite/core/profile/profiler.cc:120-260, lite/operators/conv_op.h:43-62

Differences that follow from the target: the only valid places are kCUDA ones (there is no CPU fallback -- a
config without a kCUDA place raises, like the reference's "no kernel found" LOG(FATAL)); the batch is fixed by
`GetInput(0)->Resize()` before the first Run (buffers are bound once, the network is one CUDA graph); Run()
enqueues on the predictor's stream and returns, reading an output synchronises.
"""
import os
import time

import numpy as np

from . import model as _model

VERSION = "paddle-lite_b200 r2 (sm_100a int8)"


# ---- lite/api/paddle_place.h ------------------------------------------------------------------------
class TargetType:
    kUnk, kHost, kX86, kCUDA, kARM, kOpenCL, kAny = range(7)


class PrecisionType:
    kUnk, kFloat, kInt8, kInt32, kAny, kFP16, kBool, kInt64, kInt16, kUInt8, kFP64 = range(11)


class DataLayoutType:
    kUnk, kNCHW, kAny, kNHWC = range(4)


_TARGET_STR = {TargetType.kHost: "host", TargetType.kX86: "x86", TargetType.kCUDA: "cuda", TargetType.kARM: "arm",
               TargetType.kAny: "any"}
_PREC_STR = {PrecisionType.kFloat: "float", PrecisionType.kInt8: "int8_t", PrecisionType.kInt32: "int32_t",
             PrecisionType.kAny: "any"}
_LAYOUT_STR = {DataLayoutType.kNCHW: "NCHW", DataLayoutType.kNHWC: "NHWC", DataLayoutType.kAny: "any"}


class Place:
    def __init__(self, target=TargetType.kUnk, precision=PrecisionType.kFloat, layout=DataLayoutType.kNCHW, device=0):
        self.target, self.precision, self.layout, self.device = target, precision, layout, device

    def __repr__(self):
        return f"{_TARGET_STR.get(self.target, '?')}/{_PREC_STR.get(self.precision, '?')}/{_LAYOUT_STR.get(self.layout, '?')}"

    def __eq__(self, o):
        return (self.target, self.precision, self.layout) == (o.target, o.precision, o.layout)


class LiteModelType:
    kProtobuf, kNaiveBuffer = range(2)


# ---- lite/api/paddle_api.h: ConfigBase / CxxConfig --------------------------------------------------
class CxxConfig:
    def __init__(self):
        self._model_dir = ""
        self._model_file = ""
        self._param_file = ""
        self._valid_places = []
        self._device_id = 0
        self._threads = 1
        self._quant_model = False
        self._cuda_graph = True
        self._profile = False
        self._nvtx = os.environ.get("PL_NVTX", "0") == "1"

    def set_model_dir(self, x):
        self._model_dir = x

    def model_dir(self):
        return self._model_dir

    def set_model_file(self, path):
        self._model_file = path

    def model_file(self):
        return self._model_file

    def set_param_file(self, path):
        self._param_file = path

    def param_file(self):
        return self._param_file

    def set_valid_places(self, places):
        self._valid_places = list(places)

    def valid_places(self):
        return list(self._valid_places)

    def set_device_id(self, device_id):
        self._device_id = int(device_id)

    def device_id(self):
        return self._device_id

    def set_threads(self, threads):  # host threads: nothing on this path runs on the host
        self._threads = int(threads)

    def threads(self):
        return self._threads

    def set_power_mode(self, mode):  # ARM big.LITTLE binding: no meaning on this target
        pass

    def set_quant_model(self, quant_model):
        self._quant_model = bool(quant_model)

    def quant_model(self):
        return self._quant_model

    # -- knobs of this target (no counterpart in the reference) --
    def set_cuda_graph(self, on):
        """Run() replays one CUDA graph of the whole program (default) or launches instruction by instruction."""
        self._cuda_graph = bool(on)

    def set_profile(self, on):
        """LITE_WITH_PROFILE is a build flag in the reference; here a config switch: Run() times every instruction."""
        self._profile = bool(on)

    def set_nvtx(self, on):
        """LITE_WITH_NVTX: an NVTX range per instruction and one per Run (eager mode)."""
        self._nvtx = bool(on)


# ---- lite/api/paddle_api.h: Tensor ------------------------------------------------------------------
class Tensor:
    """Handle on a feed or fetch tensor of a predictor.  A feed lives in pinned host memory until Run() copies
    it to the device (mutable_data / CopyFromCpu write there); a fetch reads the device result."""

    def __init__(self, pred, name, is_input):
        self._p, self._name, self._in = pred, name, is_input

    def Resize(self, shape):
        assert self._in, "only feeds are resized by the caller"
        self._p._resize_input(tuple(int(v) for v in shape))

    def shape(self):
        return list(self._p._in_shape if self._in else self._p._fetch_shape(self._name))

    def mutable_data(self, target=TargetType.kHost):
        assert self._in, "outputs are read-only"
        self._p._host_dirty = True
        if target == TargetType.kCUDA:  # the device-side feed: written by the caller's own kernels
            self._p._host_dirty = False
            return self._p._net_ready().input
        return self._p._host_input().numpy()

    def CopyFromCpu(self, data):
        a = np.ascontiguousarray(data, np.float32).reshape(self._p._in_shape)
        self._p._host_input().numpy()[...] = a
        self._p._host_dirty = True

    def data(self):
        """host copy of the tensor (synchronises)"""
        if self._in:
            return self._p._host_input().numpy()
        return self._p._fetch(self._name)

    def CopyToCpu(self, out=None):
        a = self.data()
        if out is None:
            return a
        np.copyto(np.asarray(out).reshape(a.shape), a)
        return out

    def target(self):
        return TargetType.kHost if self._in else TargetType.kCUDA

    def precision(self):
        return PrecisionType.kFloat

    def IsInitialized(self):
        return self._p._net is not None or self._in


# ---- lite/core/profile/profiler.cc ------------------------------------------------------------------
class OpCharacter:
    def __init__(self, op, prog, batch, plan=None):
        k = op["kind"]
        self.op_type = {"conv": "conv2d", "dwconv": "depthwise_conv2d", "deconv": "conv2d_transpose", "fc": "fc",
                        "pool": "pool2d", "quant_in": "calib", "gather": "concat/shuffle_channel/split",
                        "addf": "elementwise_add", "add": "elementwise_add", "softmax": "softmax", "scalef": "scale",
                        "to_nchw": "layout", "to_nhwc": "layout", "cat_shuffle": "concat+shuffle_channel"}.get(k, k)
        int8 = k in ("conv", "dwconv", "pool", "quant_in", "add", "cat_shuffle", "fc")
        self.kernel_attr = "cuda/" + ("int8_t" if int8 else "float") + "/NHWC"
        self.kernel_func_name = {"conv": "igemm_i8", "dwconv": "dwconv_i8", "fc": "igemm_i8", "pool": "pool2d_i8",
                                 "quant_in": "quantize_nchw", "gather": "channel_gather", "deconv": "conv2d_transpose_f32",
                                 "addf": "add_f32", "softmax": "softmax_f32", "scalef": "scale_f32"}.get(k, k)
        if plan and k in ("conv", "fc"):
            self.kernel_func_name += f"<m{plan['mode']},n{plan['BN']}" + (",pairs" if plan.get("cta_pairs") else "") + ">"
        src = op.get("src") or (op.get("srcs") or ["input"])[0]
        H, W, C = prog.shapes.get(src, prog.shapes["input"])
        P, Q, K = prog.shapes.get(op.get("dst"), (1, 1, 1))
        self.input_shape = f"{batch}x{C}x{H}x{W}"
        self.output_shape = f"{batch}x{K}x{P}x{Q}"
        self.filter_shape = ""
        self.remark = ""
        self.macs = 0.0
        if k in ("conv", "dwconv", "deconv") and "k" in op:
            g = op.get("groups", 1) if k != "dwconv" else C
            cin = C
            self.filter_shape = f"{K}x{cin // g}x{op['k']}x{op['k']}"
            self.remark = (f"{op['k']}x{op['k']}p{op.get('pad', 0)}s{op.get('stride', 1)}g{g}d1"
                           + ("Bias" if op.get("bias") is not None else "") + str(op.get("act", "")))
            # MACs = 2.f * kw * kh * batchsize * out_c * out_h * out_w * in_c / group   (conv_op.h:57-62)
            if k == "deconv":  # every input pixel feeds k*k*out_c MACs (conv_transpose_op.h, same 2x convention)
                self.macs = 2.0 * op["k"] * op["k"] * batch * H * W * cin * K / g
            else:
                self.macs = 2.0 * op["k"] * op["k"] * (batch * K * P * Q) * cin / g
        elif k == "fc":
            self.filter_shape = f"{op['K']}x{op['N']}"
            self.macs = 2.0 * batch * op["K"] * op["N"]

    def key(self):
        return (self.op_type, self.kernel_attr, self.kernel_func_name)


class Profiler:
    """Per-instruction timers (CUDA events around every instruction of an eager Run) and the reference's two
    summary tables: concise (one row per op type / kernel) and detailed (one row per instruction)."""

    def __init__(self, name, characters):
        self.name = name
        self.ch = characters
        self.laps = [[] for _ in characters]  # ms per Run

    def record(self, ms_list):
        for lap, ms in zip(self.laps, ms_list):
            lap.append(ms)

    @staticmethod
    def _stat(v, w):
        v = v[w:] if len(v) > w else v
        return (sum(v) / len(v), min(v), max(v), v[-1]) if v else (0.0, 0.0, 0.0, 0.0)

    def Summary(self, concise=True, warm_up=10):
        n = len(self.laps[0]) if self.laps else 0
        w = warm_up if n > warm_up else 0
        rows = [self._stat(v, w) for v in self.laps]
        total = sum(r[0] for r in rows) or 1.0
        out = []
        if concise:
            out.append(f"Timing cycle = {n}")
            out.append(f"===== Concise Dispatch Profiler Summary: {self.name}, Exclude {w} warm-ups =====")
            out.append(f"{'OperatorType':<28} {'KerneAttr(Place)':<22} {'KernelFuncName':<28} {'Avg(ms)':>9} {'Min(ms)':>9} "
                       f"{'Max(ms)':>9} {'Avg(%)':>7} {'GOPs':>9} {'CalledTimes':>11}")
            agg = {}
            for c, r in zip(self.ch, rows):
                a = agg.setdefault(c.key(), [0.0, 0.0, 0.0, 0.0, 0])
                a[0] += r[0]
                a[1] += r[1]
                a[2] += r[2]
                a[3] += c.macs * 1e-9
                a[4] += 1
            for key, a in sorted(agg.items(), key=lambda kv: -kv[1][0]):
                out.append(f"{key[0]:<28} {key[1]:<22} {key[2]:<28} {a[0]:9.4f} {a[1]:9.4f} {a[2]:9.4f} "
                           f"{100 * a[0] / total:6.1f}% {a[3]:9.3f} {a[4]:11d}")
        else:
            out.append(f"===== Detailed Dispatch Profiler Summary: {self.name}, Exclude {w} warm-ups =====")
            out.append(f"{'OperatorType':<28} {'KerneAttr(Place)':<18} {'KernelFuncName':<26} {'Remark':<22} {'InDim':<16} "
                       f"{'FilterDim':<14} {'OutDim':<16} {'Avg(ms)':>8} {'Min(ms)':>8} {'Max(ms)':>8} {'Last(ms)':>8} "
                       f"{'Avg(%)':>7} {'GOPs':>8} {'GOPS':>9}")
            for c, r in zip(self.ch, rows):
                gops = c.macs * 1e-9
                out.append(f"{c.op_type:<28} {c.kernel_attr:<18} {c.kernel_func_name:<26} {c.remark:<22} {c.input_shape:<16} "
                           f"{c.filter_shape:<14} {c.output_shape:<16} {r[0]:8.4f} {r[1]:8.4f} {r[2]:8.4f} {r[3]:8.4f} "
                           f"{100 * r[0] / total:6.1f}% {gops:8.3f} {(gops / (r[0] * 1e-3) if r[0] > 0 else 0):9.1f}")
        out.append(f"total: {total:.4f} ms per Run, "
                   f"{sum(c.macs for c in self.ch) * 1e-9 / (total * 1e-3) / 1e3:.2f} TOP/s")
        return "\n".join(out)


# ---- lite/api/cxx_api.cc: Predictor -----------------------------------------------------------------
class PaddlePredictor:
    def __init__(self, config, _prog=None):
        self._cfg = config
        places = config.valid_places()
        if not any(p.target == TargetType.kCUDA for p in places):
            raise RuntimeError("valid_places holds no kCUDA place: this library has kCUDA kernels only (no CPU "
                               "fallback); the reference LOG(FATAL)s the same way when no kernel matches a place")
        self._int8_ok = any(p.target == TargetType.kCUDA and p.precision == PrecisionType.kInt8 for p in places)
        if _prog is not None:
            self._prog = _prog
        else:
            self._prog, self._fuse_report = self._build(config)
        quantised = any(op["kind"] in ("conv", "dwconv", "quant_in") for op in self._prog.ops)
        if quantised and not self._int8_ok:
            raise RuntimeError("the model is quantised (enable_int8 ops) but no Place(kCUDA, kInt8) is valid")
        H, W, C = self._prog.shapes["input"]
        self._in_shape = (1, C, H, W)
        self._net = None
        self._host = None
        self._host_dirty = False
        self._profiler = None

    # Build (cxx_api.cc:248-355): LoadModelPb -> optimizer passes -> program
    @staticmethod
    def _build(config):
        d = config.model_dir()
        if config.model_file() and config.param_file():
            raise NotImplementedError("combined-parameter models: the shipped quant models use one file per param")
        if d and os.path.exists(os.path.join(d, "__model__")):
            g = _model.Graph.load(d)
            report = _model.optimize(g)
            return _model.build_program(g, os.path.basename(d.rstrip("/"))), report
        if d and os.path.exists(d + ".json"):  # a program saved by SaveOptimizedModel
            return _model.load_saved_program(d), {}
        raise FileNotFoundError(f"no __model__ under {d!r} and no optimized program {d!r}.json")

    # -- feeds / fetches --
    def GetInputNames(self):
        return ["input"]

    def GetOutputNames(self):
        return [self._prog.output or "output"]

    def GetInput(self, i):
        assert i == 0
        return Tensor(self, "input", True)

    def GetInputByName(self, name):
        assert name in self.GetInputNames()
        return Tensor(self, name, True)

    def GetOutput(self, i):
        assert i == 0
        return Tensor(self, self.GetOutputNames()[0], False)

    def GetTensor(self, name):
        """any named tensor of the program (activations are NHWC on the device; int8 ones come back as int8)"""
        return Tensor(self, name, name in self.GetInputNames())

    def GetVersion(self):
        return VERSION

    def GetParamNames(self):
        return [f"{i}.{k}" for i, op in enumerate(self._prog.ops) for k in ("w", "bias", "w_scale") if k in op]

    def Clone(self):
        """a second predictor over the same fused program (own buffers, own graph), cxx_api.cc Clone()"""
        return PaddlePredictor(self._cfg, _prog=self._prog)

    def SaveOptimizedModel(self, model_dir, model_type=LiteModelType.kNaiveBuffer, record_info=False):
        """the fused program (op records + int8 weights + scales): `model_dir`.json / .npz; loads back through
        CxxConfig.set_model_dir(model_dir)"""
        _model.save_program(self._prog, model_dir)

    # -- run --
    def _resize_input(self, shape):
        H, W, C = self._prog.shapes["input"]
        if tuple(shape[1:]) != (C, H, W):
            raise ValueError(f"the program was built for {C}x{H}x{W} inputs, got {shape}")
        if shape != self._in_shape:
            self._in_shape, self._net, self._host = shape, None, None

    def _host_input(self):
        import torch
        if self._host is None:
            self._host = torch.zeros(self._in_shape, dtype=torch.float32).pin_memory()
        return self._host

    def _net_ready(self):
        if self._net is None:
            import torch
            torch.cuda.set_device(self._cfg.device_id())
            net = _model.GraphNet(self._prog, self._in_shape[0], device=f"cuda:{self._cfg.device_id()}")
            net.bind()
            if self._cfg._cuda_graph and not self._cfg._profile:
                net.capture()
            self._net = net
            if self._cfg._profile:
                from . import capi as pl
                chars = []
                for op, fn in net.layers:  # one eager pass records which tile plan every conv takes
                    fn()
                    chars.append(OpCharacter(op, self._prog, self._in_shape[0],
                                             pl.last_igemm_plan() if op["kind"] in ("conv", "fc") else None))
                torch.cuda.synchronize()
                self._profiler = Profiler(self._prog.name, chars)
        return self._net

    def Run(self):
        import torch
        net = self._net_ready()
        if self._host_dirty:
            net.input.copy_(self._host, non_blocking=True)
            self._host_dirty = False
        if self._cfg._profile:
            evs = []
            for op, fn in net.layers:
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                fn()
                b.record()
                evs.append((a, b))
            torch.cuda.synchronize()
            self._profiler.record([a.elapsed_time(b) for a, b in evs])
        elif net.graph is not None:
            net.replay()
        elif self._cfg._nvtx:
            nv = torch.cuda.nvtx
            nv.range_push(self._prog.name)  # one range per Run, one per instruction (program.cc:272-293)
            for i, (op, fn) in enumerate(net.layers):
                nv.range_push(f"{i}:{op['kind']}:{op.get('dst', '')}")
                fn()
                nv.range_pop()
            nv.range_pop()
        else:
            net.run()

    def _fetch_shape(self, name):
        if name == self.GetOutputNames()[0] and self._net is not None:
            return tuple(self._net.output.shape)
        H, W, C = self._prog.shapes[name]
        return (self._in_shape[0], H, W, C)

    def _fetch(self, name):
        import torch
        net = self._net_ready()
        torch.cuda.synchronize()
        if name == self.GetOutputNames()[0]:
            return net.output.cpu().numpy()
        t = net.tensor(name)
        C = self._prog.shapes[name][2]
        return t[..., :C].cpu().numpy()

    def profiler_summary(self, concise=True, warm_up=10):
        if self._profiler is None:
            raise RuntimeError("set_profile(True) on the config first (the reference needs a LITE_WITH_PROFILE build)")
        return self._profiler.Summary(concise, warm_up)


def CreatePaddlePredictor(config):
    return PaddlePredictor(config)


# ---- lite/api/benchmark.cc --------------------------------------------------------------------------
def benchmark(model_dir, input_shape, warmup=10, repeats=100, result_filename=None, show_output=False,
              input_img_path="", profile=False, cuda_graph=True, device_id=0):
    """The reference's benchmark_bin flow: build a predictor, fill the input with 1.0 (or the numbers of
    --input_img_path), `warmup` untimed Runs, `repeats` timed ones (wall clock around Run + synchronise),
    min / max / average in ms appended to result_filename.  Returns the record."""
    import torch
    cfg = CxxConfig()
    cfg.set_model_dir(model_dir)
    cfg.set_device_id(device_id)
    cfg.set_valid_places([Place(TargetType.kCUDA, PrecisionType.kInt8, DataLayoutType.kNHWC),
                          Place(TargetType.kCUDA, PrecisionType.kFloat, DataLayoutType.kNHWC),
                          Place(TargetType.kCUDA, PrecisionType.kFloat, DataLayoutType.kNCHW)])
    cfg.set_cuda_graph(cuda_graph)
    cfg.set_profile(profile)
    pred = CreatePaddlePredictor(cfg)
    x = pred.GetInput(0)
    x.Resize(input_shape)
    data = x.mutable_data()
    if input_img_path:
        data[...] = np.loadtxt(input_img_path, dtype=np.float32).reshape(data.shape)
    else:
        data[...] = 1.0
    for _ in range(warmup):
        pred.Run()
    torch.cuda.synchronize()
    perf = []
    for _ in range(repeats):
        t0 = time.perf_counter()
        pred.Run()
        torch.cuda.synchronize()
        perf.append((time.perf_counter() - t0) * 1e3)
    name = os.path.basename(model_dir.rstrip("/"))
    rec = {"model": name, "input_shape": list(input_shape), "warmup": warmup, "repeats": repeats,
           "min": min(perf), "max": max(perf), "average": sum(perf) / len(perf),
           "img_s": input_shape[0] / (sum(perf) / len(perf) * 1e-3)}
    if result_filename:
        with open(result_filename, "a") as f:
            f.write(f"{name:<30}min = {rec['min']:<12.5f}max = {rec['max']:<12.5f}average = {rec['average']:<12.5f}\n")
    if show_output:
        out = pred.GetOutput(0).data().ravel()
        rec["max_value"], rec["max_index"] = float(out.max()), int(out.argmax())
        rec["output_head"] = [float(v) for v in out[:10]]
    if profile:
        rec["summary"] = pred.profiler_summary(concise=True, warm_up=0)
        rec["summary_detailed"] = pred.profiler_summary(concise=False, warm_up=0)
    return rec
