"""Quantised-model front end: reads a Paddle `__model__` + its parameter files, applies
the reference's fusion passes as host code, and binds the fused graph to the C-ABI kernels.

  reader   ProgramDesc protobuf through a minimal wire-format parser (lite/core/framework.proto; no protoc in
           this image) and the LoDTensor files (lite/model_parser/model_parser.cc:44-147 LoadLoDTensor /
           LoadParam, tensor_io.cc:65-105: u32 version, u64 lod levels, u32 version, i32 desc size, TensorDesc,
           raw data)
  fusers   delete fake_quantize_* -> input_scale                quant_dequant_op_fuser.cc:59-92
           fake_dequantize_max_abs -> enable_int8, weight_scale, int8 weights   quant_dequant_op_fuser.cc:132-202
           conv / conv2d_transpose + batch_norm (int8: weight_scale[i] *= |alpha_i|, sign flip; fp32: w *= alpha)
                                                                conv_bn_fuser.cc:100-250, conv_bn_fuser.h:36-50
           conv + relu                                          conv_activation_fuser.cc
           reshape2 + transpose2 + reshape2 -> shuffle_channel  shuffle_channel_fuser.cc:1-109
           output_scale = input_scale of the adjacent quantised op   quantized_op_attributes_inference_pass.cc:29-90
  program  `build_program`: the fused graph as the op records `GraphNet` executes (a conv writes int8 when its
           only consumer is a quantised op, fp32 otherwise -- static_kernel_pick's int8_out / fp32_out choice;
           concat / shuffle_channel / split chains become channel maps of ONE gather kernel per materialised
           tensor, quantising on the way out when the consumer is int8)
  GraphNet the device executor (same surface as nets.DeviceNet: input / inputs / capture_inputs / replay / output)
  Predictor  the CxxConfig / Predictor-shaped public API (lite/api/paddle_api.h, cxx_api.cc:248-355) and the
           profiler summary (lite/core/profile/profiler.cc, MACs formula lite/operators/conv_op.h:43-62)

torch is used for device memory, streams and CUDA graphs only; every compute call goes through the C-ABI.
"""
import json
import os
import struct

import numpy as np

# ----------------------------------------------------------------------------------------------------
# protobuf wire format (ProgramDesc, lite/core/framework.proto)
# ----------------------------------------------------------------------------------------------------


def _varint(b, i):
    v = s = 0
    while True:
        c = b[i]
        i += 1
        v |= (c & 0x7F) << s
        s += 7
        if not c & 0x80:
            return v, i


def _fields(b):
    i = 0
    while i < len(b):
        key, i = _varint(b, i)
        fn, wt = key >> 3, key & 7
        if wt == 0:
            v, i = _varint(b, i)
        elif wt == 1:
            v = b[i:i + 8]
            i += 8
        elif wt == 2:
            n, i = _varint(b, i)
            v = b[i:i + n]
            i += n
        elif wt == 5:
            v = b[i:i + 4]
            i += 4
        else:
            raise ValueError(f"unsupported protobuf wire type {wt}")
        yield fn, wt, v


def _sint(v, bits=64):
    return v - (1 << bits) if v >= (1 << (bits - 1)) else v


def _ints(wt, v):
    if wt == 0:
        return [_sint(v)]
    out, i = [], 0
    while i < len(v):
        x, i = _varint(v, i)
        out.append(_sint(x))
    return out


def _parse_attr(b):
    """OpDesc.Attr: name=1 type=2 i=3 f=4 s=5 ints=6 floats=7 strings=8 b=10 bools=11 block_idx=12 l=13"""
    name, val, typ = None, None, None
    ints, floats, strings = [], [], []
    for fn, wt, v in _fields(b):
        if fn == 1:
            name = v.decode()
        elif fn == 2:
            typ = v
        elif fn == 3:
            val = _sint(v & 0xFFFFFFFF, 32) if v < (1 << 32) else _sint(v)
        elif fn == 4:
            val = struct.unpack("<f", v)[0]
        elif fn == 5:
            val = v.decode(errors="replace")
        elif fn == 6:
            ints += [_sint(x & 0xFFFFFFFF, 32) if 0 <= x < (1 << 32) else x for x in _ints(wt, v)]
        elif fn == 7:
            floats += [struct.unpack("<f", v)[0]] if wt == 5 else list(struct.unpack(f"<{len(v) // 4}f", v))
        elif fn == 8:
            strings.append(v.decode(errors="replace"))
        elif fn == 10:
            val = bool(v)
        elif fn == 13:
            val = _sint(v)
    # AttrType: INT=0 FLOAT=1 STRING=2 INTS=3 FLOATS=4 STRINGS=5 BOOLEAN=6 BOOLEANS=7 BLOCK=8 LONG=9
    if typ == 3:
        val = ints
    elif typ == 4:
        val = floats
    elif typ == 5:
        val = strings
    return name, val


def _parse_opvar(b):
    name, args = None, []
    for fn, _, v in _fields(b):
        if fn == 1:
            name = v.decode()
        elif fn == 2:
            args.append(v.decode())
    return name, args


def _parse_op(b):
    op = {"type": None, "inputs": {}, "outputs": {}, "attrs": {}}
    for fn, _, v in _fields(b):
        if fn == 3:
            op["type"] = v.decode()
        elif fn == 1:
            k, a = _parse_opvar(v)
            op["inputs"][k] = a
        elif fn == 2:
            k, a = _parse_opvar(v)
            op["outputs"][k] = a
        elif fn == 4:
            k, a = _parse_attr(v)
            op["attrs"][k] = a
    return op


def _parse_var(b):
    name, dims, persistable = None, None, False
    for fn, _, v in _fields(b):
        if fn == 1:
            name = v.decode()
        elif fn == 3:
            persistable = bool(v)
        elif fn == 2:  # VarType
            for f2, _, v2 in _fields(v):
                if f2 == 3:  # LoDTensorDesc
                    for f3, _, v3 in _fields(v2):
                        if f3 == 1:  # TensorDesc: data_type=1 dims=2
                            dims = []
                            for f4, w4, v4 in _fields(v3):
                                if f4 == 2:
                                    dims += _ints(w4, v4)
    return name, dims, persistable


def load_program(model_dir):
    """-> (ops, var_dims, persistable names) of block 0 of `model_dir/__model__` (LoadModelPb, model_parser.cc)."""
    b = open(os.path.join(model_dir, "__model__"), "rb").read()
    ops, var_dims, persist = [], {}, set()
    for fn, _, v in _fields(b):
        if fn != 1:  # ProgramDesc.blocks
            continue
        for f2, _, v2 in _fields(v):  # BlockDesc: idx=1 parent=2 vars=3 ops=4
            if f2 == 3:
                n, d, p = _parse_var(v2)
                var_dims[n] = d
                if p:
                    persist.add(n)
            elif f2 == 4:
                ops.append(_parse_op(v2))
        break  # block 0 only (no control flow in the models on this path)
    return ops, var_dims, persist


_PB_DTYPE = {0: np.bool_, 1: np.int16, 2: np.int32, 3: np.int64, 5: np.float32, 6: np.float64, 20: np.uint8,
             21: np.int8}


def load_lod_tensor(path):
    """One parameter file (LoadParam -> LoDTensorDeserializer::LoadWithForwardReader, tensor_io.cc:65-105)."""
    b = open(path, "rb").read()
    (version,) = struct.unpack_from("<I", b, 0)
    if version != 0:
        raise ValueError(f"{path}: tensor version {version} is not supported")
    (lod_level,) = struct.unpack_from("<Q", b, 4)
    i = 12
    for _ in range(lod_level):
        (size,) = struct.unpack_from("<Q", b, i)
        i += 8 + size
    (inner,) = struct.unpack_from("<I", b, i)
    if inner != 0:
        raise ValueError(f"{path}: tensor inner version {inner}")
    (desc_size,) = struct.unpack_from("<i", b, i + 4)
    i += 8
    dtype, dims = None, []
    for fn, wt, v in _fields(b[i:i + desc_size]):
        if fn == 1:
            dtype = _PB_DTYPE[v]
        elif fn == 2:
            dims += _ints(wt, v)
    i += desc_size
    n = int(np.prod(dims)) if dims else 1
    return np.frombuffer(b, dtype=dtype, count=n, offset=i).reshape(dims).copy()


# ----------------------------------------------------------------------------------------------------
# graph + fusion passes (host code; the arithmetic follows the reference's fusers line by line)
# ----------------------------------------------------------------------------------------------------
_CONV_TYPES = ("conv2d", "depthwise_conv2d", "conv2d_transpose")
_QUANT_OPS = ("fake_quantize_range_abs_max", "fake_quantize_moving_average_abs_max")


class Graph:
    """Ops of block 0 in program order + the parameter tensors (the reference's SSAGraph + Scope, flattened)."""

    def __init__(self, ops, var_dims, params):
        self.ops = [op for op in ops if op["type"] not in ("feed", "fetch")]
        self.feeds = [op["outputs"]["Out"][0] for op in ops if op["type"] == "feed"]
        self.fetches = [op["inputs"]["X"][0] for op in ops if op["type"] == "fetch"]
        self.var_dims = dict(var_dims)
        self.params = params

    @classmethod
    def load(cls, model_dir):
        ops, var_dims, persist = load_program(model_dir)
        params = {}
        for n in sorted(persist):
            p = os.path.join(model_dir, n)
            if n not in ("feed", "fetch") and os.path.exists(p):
                params[n] = load_lod_tensor(p)
        return cls(ops, var_dims, params)

    def consumers(self, var):
        return [op for op in self.ops if any(var in a for a in op["inputs"].values())]

    def producer(self, var):
        for op in self.ops:
            if any(var in a for a in op["outputs"].values()):
                return op
        return None

    def remove(self, *dead):
        ids = {id(o) for o in dead}
        self.ops = [op for op in self.ops if id(op) not in ids]

    @staticmethod
    def rename_input(op, old, new):
        for k, a in op["inputs"].items():
            op["inputs"][k] = [new if v == old else v for v in a]

    def counts(self):
        c = {}
        for op in self.ops:
            c[op["type"]] = c.get(op["type"], 0) + 1
        return c


def fuse_delete_quant(g):
    """DeleteQuantOpFuser (quant_dequant_op_fuser.cc:59-92): scale = OutScale[0] / (2^(bits-1) - 1) becomes the
    consumers' input scale of that argument; the consumers read the un-quantised variable."""
    n = 0
    for q in [op for op in g.ops if op["type"] in _QUANT_OPS]:
        bits = q["attrs"].get("bit_length", 8)
        rng = (1 << (bits - 1)) - 1
        scale_var = q["outputs"]["OutScale"][0]
        scale = np.float32(g.params[scale_var].reshape(-1)[0]) / np.float32(rng)
        x, out = q["inputs"]["X"][0], q["outputs"]["Out"][0]
        for c in g.consumers(out):
            c.setdefault("input_scale", {})[x] = float(scale)
            c["attrs"]["bit_length"] = bits
            g.rename_input(c, out, x)
        g.remove(q)
        n += 1
    return n


def fuse_dequant(g):
    """DequantOpFuser (quant_dequant_op_fuser.cc:132-202): weight_scale = range * range / max_range / range for
    every output channel, enable_int8, weights float -> int8 by truncating cast; the quantised op writes the
    dequant op's output."""
    n = 0
    for dq in [op for op in g.ops if op["type"] == "fake_dequantize_max_abs"]:
        x, out = dq["inputs"]["X"][0], dq["outputs"]["Out"][0]
        qop = g.producer(x)
        if qop is None or qop["type"] not in _CONV_TYPES + ("mul", "matmul"):
            continue
        bits = qop["attrs"].get("bit_length", 8)
        rng = (1 << (bits - 1)) - 1
        max_range = np.float32(dq["attrs"]["max_range"])
        whole = np.float32(rng * rng) / max_range / np.float32(rng)
        wname = qop["inputs"]["Filter" if qop["type"] in _CONV_TYPES else "Y"][0]
        w = g.params[wname]
        n_out = w.shape[0] if qop["type"] in _CONV_TYPES else w.shape[1]
        qop["weight_scale"] = np.full(n_out, whole, np.float32)
        qop["enable_int8"] = True
        if w.dtype != np.int8:
            g.params[wname] = w.astype(np.float32).astype(np.int8)  # static_cast<int8_t>(float): truncation
        okey = "Output" if qop["type"] in _CONV_TYPES else "Out"
        qop["outputs"][okey] = [out]
        g.remove(dq)
        n += 1
    return n


def fuse_conv_bn(g):
    """ConvBNFuser (conv_bn_fuser.cc:100-250): alpha = scale / sqrt(var + eps), beta = -mean * alpha;
    int8: weight_scale[i] *= |alpha_i| and the filter row changes sign when alpha_i < 0; fp32: w *= alpha;
    bias = bn_bias + beta (+ alpha * conv_bias)."""
    n = 0
    for conv in [op for op in g.ops if op["type"] in _CONV_TYPES]:
        out = conv["outputs"]["Output"][0]
        cons = g.consumers(out)
        if len(cons) != 1 or cons[0]["type"] != "batch_norm":
            continue
        bn = cons[0]
        sc, bi = g.params[bn["inputs"]["Scale"][0]], g.params[bn["inputs"]["Bias"][0]]
        mean, var = g.params[bn["inputs"]["Mean"][0]], g.params[bn["inputs"]["Variance"][0]]
        eps = np.float32(bn["attrs"].get("epsilon", 1e-5))
        alpha = (sc / np.sqrt(var + eps)).astype(np.float32)
        beta = (-mean * alpha).astype(np.float32)
        wname = conv["inputs"]["Filter"][0]
        w = g.params[wname]
        transpose = conv["type"] == "conv2d_transpose"
        groups = conv["attrs"].get("groups", 1)
        if conv.get("enable_int8"):
            ws = conv["weight_scale"].copy()
            w = w.copy()
            if transpose:  # [Cin, Cout/g, kh, kw]: output channel = dim 1 (the reference loops k over Cin/g rows)
                for i in range(w.shape[1] * groups):
                    ws[i] *= abs(alpha[i])
                    if alpha[i] < 0:
                        w[:, i % w.shape[1]] *= -1
            else:
                ws *= np.abs(alpha)
                w[alpha < 0] *= -1
            conv["weight_scale"] = ws.astype(np.float32)
        else:
            w = w.astype(np.float32).copy()
            if transpose:
                co = w.shape[1]
                for gi in range(groups):
                    cin_g = w.shape[0] // groups
                    w[gi * cin_g:(gi + 1) * cin_g] *= alpha[gi * co:(gi + 1) * co].reshape(1, co, 1, 1)
            else:
                w *= alpha.reshape(-1, 1, 1, 1)
        new_w = wname + ".bn_fused"
        g.params[new_w] = w
        conv["inputs"]["Filter"] = [new_w]
        bias = bi.astype(np.float32).copy()
        if conv["inputs"].get("Bias"):
            bias += alpha * g.params[conv["inputs"]["Bias"][0]]
        bias += beta
        bname = bn["inputs"]["Bias"][0] + ".bn_fused"
        g.params[bname] = bias.astype(np.float32)
        conv["inputs"]["Bias"] = [bname]
        conv["outputs"]["Output"] = [bn["outputs"]["Y"][0]]
        g.remove(bn)
        n += 1
    return n


def fuse_conv_elementwise_bias(g):
    """conv + elementwise_add(persistable per-channel Y, axis 1) -> conv bias (conv_elementwise_fuser.cc)."""
    n = 0
    for conv in [op for op in g.ops if op["type"] in _CONV_TYPES]:
        out = conv["outputs"]["Output"][0]
        cons = g.consumers(out)
        if len(cons) != 1 or cons[0]["type"] != "elementwise_add":
            continue
        add = cons[0]
        y = add["inputs"]["Y"][0]
        if add["inputs"]["X"][0] != out or y not in g.params or add["attrs"].get("axis", -1) != 1:
            continue
        b = g.params[y].astype(np.float32).reshape(-1)
        if conv["inputs"].get("Bias"):
            b = b + g.params[conv["inputs"]["Bias"][0]]
        bname = y + ".as_conv_bias"
        g.params[bname] = b
        conv["inputs"]["Bias"] = [bname]
        conv["outputs"]["Output"] = [add["outputs"]["Out"][0]]
        g.remove(add)
        n += 1
    return n


def fuse_conv_activation(g):
    """conv + relu -> ActivationParam{has_active, kRelu} on the conv (conv_activation_fuser.cc)."""
    n = 0
    for conv in [op for op in g.ops if op["type"] in _CONV_TYPES]:
        out = conv["outputs"]["Output"][0]
        cons = g.consumers(out)
        if len(cons) != 1 or cons[0]["type"] not in ("relu", "relu6", "leaky_relu"):
            continue
        act = cons[0]
        conv["act"] = {"relu": "relu", "relu6": "relu6", "leaky_relu": "leaky"}[act["type"]]
        conv["act_arg"] = act["attrs"].get("threshold", 6.0) if act["type"] == "relu6" else act["attrs"].get("alpha", 0.0)
        conv["outputs"]["Output"] = [act["outputs"]["Out"][0]]
        g.remove(act)
        n += 1
    return n


def fuse_shuffle_channel(g):
    """reshape2 [-1, group, C/group, H, W] + transpose2 [0, 2, 1, 3, 4] + reshape2 [-1, C, H, W] -> shuffle_channel
    (shuffle_channel_fuser.cc:1-109)."""
    n = 0
    for r1 in [op for op in g.ops if op["type"] == "reshape2"]:
        shp = r1["attrs"].get("shape") or []
        if len(shp) != 5:
            continue
        c1 = g.consumers(r1["outputs"]["Out"][0])
        if len(c1) != 1 or c1[0]["type"] != "transpose2" or list(c1[0]["attrs"].get("axis", [])) != [0, 2, 1, 3, 4]:
            continue
        c2 = g.consumers(c1[0]["outputs"]["Out"][0])
        if len(c2) != 1 or c2[0]["type"] != "reshape2" or len(c2[0]["attrs"].get("shape") or []) != 4:
            continue
        op = {"type": "shuffle_channel", "inputs": {"X": r1["inputs"]["X"]}, "outputs": {"Out": c2[0]["outputs"]["Out"]},
              "attrs": {"group": int(shp[1])}}
        g.ops[g.ops.index(r1)] = op
        g.remove(c1[0], c2[0])
        n += 1
    return n


def infer_output_scales(g):
    """QuantizedOpAttributesInferencePass (quantized_op_attributes_inference_pass.cc:29-90): the output scale of an
    op is the input scale its adjacent quantised consumer holds for that variable."""
    for op in g.ops:
        for outs in op["outputs"].values():
            for v in outs:
                for c in g.consumers(v):
                    s = c.get("input_scale", {}).get(v)
                    if s is not None:
                        op.setdefault("output_scale", {})[v] = s
    return g


def optimize(g):
    """The fusion passes in the reference's order (lite/core/optimizer.h:77-135, the ones on this path)."""
    report = {"delete_quant": fuse_delete_quant(g), "dequant": fuse_dequant(g),
              "conv_elementwise_bias": fuse_conv_elementwise_bias(g), "conv_bn": fuse_conv_bn(g),
              "conv_activation": fuse_conv_activation(g), "shuffle_channel": fuse_shuffle_channel(g)}
    infer_output_scales(g)
    return report


# ----------------------------------------------------------------------------------------------------
# fused graph -> executable program
# ----------------------------------------------------------------------------------------------------
class Op(dict):
    __getattr__ = dict.__getitem__
    __setattr__ = dict.__setitem__


class Program:
    """Executable op records (same conventions as nets.Builder: shapes are per image (H, W, C), activations NHWC)."""

    def __init__(self, name, in_shape, in_scale):
        self.name = name
        self.ops = []
        self.shapes = {"input": in_shape}
        self.views = {}
        self.in_scale = in_scale
        self.output = None  # name of the fetched tensor (fp32 NCHW)


_MOVE = ("concat", "split", "shuffle_channel")


def build_program(g, name="model"):
    """Precision / materialisation decisions over the fused graph:
    * a quantised conv reads int8 at its input_scale; it writes int8 at output_scale when every consumer is a
      quantised op (int8_out registration), fp32 otherwise (fp32_out) -- static_kernel_pick_pass.cc:144-166;
    * max pool2d between quantised ops runs on the int8 tensor (max commutes with the monotone quantiser);
    * concat / shuffle_channel / split never run: they compose into channel maps, and ONE gather launch
      materialises a mapped tensor where a compute op needs it -- in fp32, or quantised at the consumer's
      input_scale (the calib node type_precision_cast_pass.cc:171-260 would insert there);
    * conv2d_transpose, elementwise_add, softmax, scale are fp32 (the model leaves them un-quantised)."""
    dims = g.var_dims
    feed = g.feeds[0]
    _, C0, H0, W0 = dims[feed]
    prog = Program(name, (H0, W0, C0), None)

    def shape_of(v):
        d = dims[v]
        return (int(d[2]), int(d[3]), int(d[1]))

    # --- channel maps of movement ops: var -> list of (materialised source var, channel) ---------------
    cmap = {}

    def channels(v):
        if v in cmap:
            return cmap[v]
        return [(v, c) for c in range(shape_of(v)[2])]

    for op in g.ops:
        if op["type"] == "concat":
            assert op["attrs"].get("axis", 1) == 1, "concat along channels only"
            cmap[op["outputs"]["Out"][0]] = [ch for x in op["inputs"]["X"] for ch in channels(x)]
        elif op["type"] == "shuffle_channel":
            src = channels(op["inputs"]["X"][0])
            grp = op["attrs"]["group"]
            per = len(src) // grp
            # out[j * group + i] = in[i * (C / group) + j]   (lite/backends/arm/math/shuffle_channel.cc:24-76)
            cmap[op["outputs"]["Out"][0]] = [src[(o % grp) * per + o // grp] for o in range(len(src))]
        elif op["type"] == "split":
            assert op["attrs"].get("axis", 1) == 1
            src = channels(op["inputs"]["X"][0])
            outs = op["outputs"]["Out"]
            secs = op["attrs"].get("sections") or [len(src) // len(outs)] * len(outs)
            c0 = 0
            for o, s in zip(outs, secs):
                cmap[o] = src[c0:c0 + s]
                c0 += s

    compute = [op for op in g.ops if op["type"] not in _MOVE]

    def need(op, v):
        if op.get("enable_int8"):
            return ("i8", op["input_scale"][v])
        if op["type"] == "pool2d" and op["attrs"].get("pooling_type") == "max" and op.get("output_scale"):
            return ("i8", list(op["output_scale"].values())[0])
        return ("f32", None)

    def consumers_needs(v):
        """forms in which the materialised var v is read, through any chain of movement ops"""
        needs, seen, stack = set(), set(), [v]
        while stack:
            x = stack.pop()
            if x in seen:
                continue
            seen.add(x)
            for c in g.consumers(x):
                if c["type"] in _MOVE:
                    needs.add(("f32", None))  # gathers read fp32 sources
                    stack += [o for outs in c["outputs"].values() for o in outs]
                else:
                    needs.add(need(c, x))
            if x in g.fetches:
                needs.add(("f32", None))
        return needs

    avail = {}  # (var, form, scale) -> tensor name in the program

    def tensor_for(v, form, scale):
        """name of a materialised tensor holding var v in `form`; emits a gather when it does not exist yet"""
        key = (v, form, scale)
        if key in avail:
            return avail[key]
        if v == feed and form == "f32":
            channels_of_feed = [(feed, c) for c in range(C0)]
            assert channels(v) == channels_of_feed
        srcs = channels(v)
        names, table = [], []
        for sv, ch in srcs:
            t = avail.get((sv, "f32", None))
            if t is None and sv == feed:  # the fp32 NCHW feed read by an un-quantised op: one layout op
                t = f"{feed}@f32"
                prog.shapes[t] = prog.shapes["input"]
                prog.ops.append(Op(kind="to_nhwc", src="input", dst=t, bytes=8 * H0 * W0 * C0))
                avail[(feed, "f32", None)] = t
            assert t is not None, f"{v}: source {sv} has no fp32 form"
            if t not in names:
                names.append(t)
            table.append((names.index(t), ch))
        H, W, _ = shape_of(srcs[0][0])
        out = f"{v}@{form}" + ("" if scale is None else f"{scale:.6g}")
        prog.shapes[out] = (H, W, len(table))
        prog.ops.append(Op(kind="gather", src=names[0], srcs=names, dst=out, table=np.array(table, np.int32),
                           out_scale=scale, bytes=H * W * len(table) * (4 + (1 if form == "i8" else 4))))
        avail[key] = out
        return out

    for op in compute:
        t = op["type"]
        if t in ("conv2d", "depthwise_conv2d") and op.get("enable_int8"):
            xv = op["inputs"]["Input"][0]
            yv = op["outputs"]["Output"][0]
            s_in = op["input_scale"][xv]
            if xv == feed:
                prog.in_scale = s_in
                prog.shapes["input.q"] = prog.shapes["input"]
                prog.ops.append(Op(kind="quant_in", src="input", dst="input.q", out_scale=s_in))
                avail[(xv, "i8", s_in)] = "input.q"
            x = tensor_for(xv, "i8", s_in)
            w = g.params[op["inputs"]["Filter"][0]]
            K, Cg, R, S = w.shape
            a = op["attrs"]
            st, pd = a["strides"], a["paddings"]
            assert st[0] == st[1] and R == S and list(a.get("dilations", [1, 1])) == [1, 1]
            pads = pd if len(pd) == 4 else [pd[0], pd[0], pd[1], pd[1]]
            assert len(set(pads)) == 1
            needs = consumers_needs(yv)
            i8_needs = {n for n in needs if n[0] == "i8"}
            out_fp32 = ("f32", None) in needs or len(i8_needs) != 1
            s_out = None if out_fp32 else list(i8_needs)[0][1]
            depthwise = t == "depthwise_conv2d" or (a.get("groups", 1) == K and Cg == 1 and K > 1)
            bias = g.params[op["inputs"]["Bias"][0]] if op["inputs"].get("Bias") else np.zeros(K, np.float32)
            out = f"{yv}@" + ("f32" if out_fp32 else f"i8{s_out:.6g}")
            prog.shapes[out] = shape_of(yv)
            Cin = shape_of(xv)[2]
            prog.ops.append(Op(kind="dwconv" if depthwise else "conv", src=x, dst=out,
                               w=np.ascontiguousarray(w.transpose(0, 2, 3, 1)),  # OIHW -> OHWI
                               w_scale=op["weight_scale"].astype(np.float32), bias=bias.astype(np.float32), k=int(R),
                               stride=int(st[0]), pad=int(pads[0]), act=op.get("act", "none"), K=int(K), C=int(Cin),
                               groups=int(a.get("groups", 1)), in_scale=float(s_in), out_scale=s_out, out_fp32=out_fp32,
                               res=None, res_scale=None, var=yv))
            avail[(yv, "f32", None) if out_fp32 else (yv, "i8", s_out)] = out
        elif t == "pool2d":
            xv, yv = op["inputs"]["X"][0], op["outputs"]["Out"][0]
            a = op["attrs"]
            form, s = need(op, xv)
            assert form == "i8", "fp32 pooling is not on this model's path"
            x = tensor_for(xv, "i8", s)
            out = f"{yv}@i8{s:.6g}"
            prog.shapes[out] = shape_of(yv)
            prog.ops.append(Op(kind="pool", src=x, dst=out, k=int(a["ksize"][0]), stride=int(a["strides"][0]),
                               pad=int(a["paddings"][0]), typ=a["pooling_type"], in_scale=s, out_scale=s, var=yv))
            avail[(yv, "i8", s)] = out
            assert consumers_needs(yv) == {("i8", s)}, "pool output is read at another precision"
        elif t == "conv2d_transpose":
            xv, yv = op["inputs"]["Input"][0], op["outputs"]["Output"][0]
            x = tensor_for(xv, "f32", None)
            w = g.params[op["inputs"]["Filter"][0]].astype(np.float32)  # [Cin, Cout / g, kh, kw]
            a = op["attrs"]
            assert a.get("groups", 1) == 1 and list(a.get("dilations", [1, 1])) == [1, 1]
            needs = consumers_needs(yv)
            i8_needs = {n for n in needs if n[0] == "i8"}
            s_out = list(i8_needs)[0][1] if needs == i8_needs and len(i8_needs) == 1 else None
            bias = g.params[op["inputs"]["Bias"][0]] if op["inputs"].get("Bias") else np.zeros(w.shape[1], np.float32)
            out = f"{yv}@" + ("f32" if s_out is None else f"i8{s_out:.6g}")
            prog.shapes[out] = shape_of(yv)
            H, W, Ci = shape_of(xv)
            P, Q, Co = shape_of(yv)
            pd = a["paddings"]
            prog.ops.append(Op(kind="deconv", src=x, dst=out, w=np.ascontiguousarray(w), bias=bias.astype(np.float32),
                               k=int(w.shape[2]), stride=int(a["strides"][0]), pad=int(pd[0]), act=op.get("act", "none"),
                               out_scale=s_out, bytes=H * W * Ci * 4 + P * Q * Co * (4 if s_out is None else 1),
                               ops=2 * H * W * Ci * Co * int(w.shape[2]) * int(w.shape[3]), var=yv))
            avail[(yv, "f32", None) if s_out is None else (yv, "i8", s_out)] = out
        elif t == "elementwise_add":
            av, bv, yv = op["inputs"]["X"][0], op["inputs"]["Y"][0], op["outputs"]["Out"][0]
            a_, b_ = tensor_for(av, "f32", None), tensor_for(bv, "f32", None)
            assert shape_of(av) == shape_of(bv), "tensor + tensor of equal shape only"
            out = f"{yv}@f32"
            prog.shapes[out] = shape_of(yv)
            H, W, Cc = shape_of(yv)
            prog.ops.append(Op(kind="addf", src=a_, src2=b_, dst=out, bytes=12 * H * W * Cc, var=yv))
            avail[(yv, "f32", None)] = out
        elif t in ("softmax", "scale"):
            xv, yv = op["inputs"]["X"][0], op["outputs"]["Out"][0]
            x = tensor_for(xv, "f32", None)
            out = f"{yv}@f32"
            prog.shapes[out] = shape_of(yv)
            H, W, Cc = shape_of(yv)
            if t == "softmax":
                assert op["attrs"].get("axis", -1) in (1, -3), "softmax over the channel axis"
                prog.ops.append(Op(kind="softmax", src=x, dst=out, bytes=8 * H * W * Cc, var=yv))
            else:
                a = op["attrs"]
                prog.ops.append(Op(kind="scalef", src=x, dst=out, scale=float(a.get("scale", 1.0)),
                                   bias=float(a.get("bias", 0.0)), bias_after_scale=bool(a.get("bias_after_scale", True)),
                                   bytes=8 * H * W * Cc, var=yv))
            avail[(yv, "f32", None)] = out
        else:
            raise NotImplementedError(f"op {t} is not on the int8 conv path of this framework")
    fv = g.fetches[0]
    x = tensor_for(fv, "f32", None)
    H, W, Cc = shape_of(fv)
    prog.shapes["output"] = (H, W, Cc)
    prog.ops.append(Op(kind="to_nchw", src=x, dst="output", bytes=8 * H * W * Cc))
    prog.output = "output"
    return prog


# ----------------------------------------------------------------------------------------------------
# persistence of a fused program (tests/golden fixture: the GPU box has no reference tree)
# ----------------------------------------------------------------------------------------------------
_ARRAY_KEYS = ("w", "w_scale", "bias", "table")


def save_program(prog, prefix):
    """prefix.json (structure) + prefix.npz (arrays)"""
    arrays, ops = {}, []
    for i, op in enumerate(prog.ops):
        rec = {}
        for k, v in op.items():
            if k in _ARRAY_KEYS and isinstance(v, np.ndarray):
                arrays[f"{i}.{k}"] = v
            elif isinstance(v, (np.floating, np.integer)):
                rec[k] = v.item()
            else:
                rec[k] = v
        ops.append(rec)
    json.dump({"name": prog.name, "in_scale": prog.in_scale, "output": prog.output,
               "shapes": {k: list(v) for k, v in prog.shapes.items()}, "ops": ops}, open(prefix + ".json", "w"))
    np.savez_compressed(prefix + ".npz", **arrays)


def load_saved_program(prefix):
    meta = json.load(open(prefix + ".json"))
    arrays = np.load(prefix + ".npz")
    prog = Program(meta["name"], tuple(meta["shapes"]["input"]), meta["in_scale"])
    prog.shapes = {k: tuple(v) for k, v in meta["shapes"].items()}
    prog.output = meta["output"]
    for i, rec in enumerate(meta["ops"]):
        op = Op(rec)
        for k in _ARRAY_KEYS:
            if f"{i}.{k}" in arrays:
                op[k] = arrays[f"{i}.{k}"]
        prog.ops.append(op)
    return prog


_HERE = os.path.dirname(os.path.abspath(__file__))
HUMANSEG_DIR = "/root/reference/humanseg_shfv75_0119_quant"
HUMANSEG_FIXTURE = os.path.join(os.path.dirname(_HERE), "tests", "golden", "humanseg_fused")


def humanseg_program(model_dir=None):
    """BASELINE config 3.  From the model files when they are present (authoring container), else from the committed
    fixture of the FUSED program (tools/make_humanseg_fixture.py wrote it with this very front end)."""
    model_dir = model_dir or HUMANSEG_DIR
    if os.path.exists(os.path.join(model_dir, "__model__")):
        g = Graph.load(model_dir)
        optimize(g)
        return build_program(g, "humanseg")
    if os.path.exists(HUMANSEG_FIXTURE + ".json"):
        return load_saved_program(HUMANSEG_FIXTURE)
    raise FileNotFoundError("neither the humanseg model directory nor its fused-program fixture is available")


# ----------------------------------------------------------------------------------------------------
# device executor
# ----------------------------------------------------------------------------------------------------
def _rup(v, m):
    return (v + m - 1) // m * m


class GraphNet:
    """A fused Program bound to device buffers for a fixed batch (the RuntimeProgram of this framework: one
    instruction per op record, lite/core/program.cc:265-315).  Same surface as nets.DeviceNet."""

    calibrated = True  # scales come from the model file

    def __init__(self, prog, batch, device="cuda", feed="f32", mean=None, norm=None):
        import torch

        from . import capi as pl
        assert feed == "f32", "the model's feed is its fp32 NCHW image"
        self.torch, self.pl = torch, pl
        self.prog, self.batch, self.dev = prog, batch, device
        H, W, C = prog.shapes["input"]
        self.input = torch.zeros((batch, C, H, W), dtype=torch.float32, device=device)
        self.bufs, self.layers, self.graph = {}, [], None
        self.keep = []  # device constants (weights, scales, tables)

    def _buf(self, name, form):
        if name not in self.bufs:
            H, W, C = self.prog.shapes[name]
            t = self.torch
            if form == "i8":
                self.bufs[name] = t.zeros((self.batch, H, W, _rup(C, 16)), dtype=t.int8, device=self.dev)
            else:
                self.bufs[name] = t.zeros((self.batch, H, W, C), dtype=t.float32, device=self.dev)
        return self.bufs[name]

    def _dev(self, a):
        x = self.torch.from_numpy(np.ascontiguousarray(a)).to(self.dev)
        self.keep.append(x)
        return x

    def calibrate(self, x):
        return self

    def bind(self):
        pl, t = self.pl, self.torch
        act_code = {"none": pl.ACT_NONE, "relu": pl.ACT_RELU, "relu6": pl.ACT_RELU6, "leaky": pl.ACT_LEAKY}
        forms = {}  # tensor name -> "i8" / "f32"
        for op in self.prog.ops:
            k = op.kind
            if k == "quant_in":
                forms[op.dst] = "i8"
            elif k in ("conv", "dwconv"):
                forms[op.dst] = "f32" if op.out_fp32 else "i8"
            elif k == "pool":
                forms[op.dst] = "i8"
            elif k in ("gather", "deconv"):
                forms[op.dst] = "f32" if op.out_scale is None else "i8"
            else:
                forms[op.dst] = "f32"
        self.layers = []
        s2d = {}
        for idx, op in enumerate(self.prog.ops):
            k = op.kind
            if k == "quant_in":
                # the consumer decides the layout of the quantised input (stride-2 thin stems read space-to-depth pixels)
                cons = [o for o in self.prog.ops if o.get("src") == op.dst]
                c0 = cons[0]
                H, W, C = self.prog.shapes["input"]
                g = pl.s2d_geometry(self.batch, H, W, C, c0.k, c0.k, (c0.stride, c0.stride), (c0.pad,) * 4) \
                    if (len(cons) == 1 and c0.kind == "conv" and C <= 4) else None
                if g is not None and g.supported:
                    y = pl.s2d_input_buffer(g, self.dev)
                    y._pl_n = self.batch
                    s2d[op.dst] = (y, g)
                    self.layers.append((op, lambda op=op, y=y, g=g: pl.quantize_nchw_to_s2d(self.input, op.out_scale, g, y=y)))
                else:
                    ld = 4 if C <= 4 else _rup(C, 16)
                    y = t.zeros((self.batch, H, W, ld), dtype=t.int8, device=self.dev)
                    self.bufs[op.dst] = y
                    self.layers.append((op, lambda op=op, y=y, ld=ld: pl.quantize_nchw_to_nhwc(self.input, op.out_scale, ld, y=y)))
            elif k in ("conv", "dwconv"):
                out = pl.OUT_FP32 if op.out_fp32 else pl.OUT_INT8
                sf, bf, argf = pl.fold_scales(op.w_scale, op.K, op.in_scale, 1.0 if op.out_fp32 else op.out_scale, out,
                                              op.bias, act_code[op.act], 6.0)
                sf, bf = self._dev(sf), self._dev(bf)
                y = self._buf(op.dst, forms[op.dst])
                st, pd = (op.stride, op.stride), (op.pad,) * 4
                if k == "dwconv":
                    w = self._dev(op.w.reshape(op.K, op.k, op.k))
                    x = self._buf(op.src, "i8")
                    self.layers.append((op, lambda op=op, x=x, w=w, sf=sf, bf=bf, argf=argf, out=out, y=y, st=st, pd=pd:
                                        pl.dwconv2d(x, w, sf, bf, op.k, op.k, st, pd, (1, 1), act_code[op.act], argf, out,
                                                    y=y, Cc=op.C)))
                elif op.src in s2d:
                    xs, g = s2d[op.src]
                    wp = pl.pack_s2d_conv_weight(self._dev(op.w), g)
                    self.keep.append(wp)
                    H, W, C = self.prog.shapes["input"]
                    self.layers.append((op, lambda op=op, xs=xs, wp=wp, sf=sf, bf=bf, argf=argf, out=out, y=y, st=st, pd=pd,
                                        H=H, W=W: pl.s2d_conv2d(xs, wp, self.batch, H, W, op.C, op.K, op.k, op.k, sf, bf, st,
                                                                pd, act_code[op.act], argf, out, y=y, ldy=y.shape[-1])))
                elif op.C <= 4:
                    assert pl.thin_conv_supported(op.k, op.k, op.C), "thin first layer outside the direct kernel's windows"
                    wp = pl.pack_thin_conv_weight(self._dev(op.w))
                    self.keep.append(wp)
                    x = self.bufs[op.src]
                    self.layers.append((op, lambda op=op, x=x, wp=wp, sf=sf, bf=bf, argf=argf, out=out, y=y, st=st, pd=pd:
                                        pl.thin_conv2d(x, wp, op.K, op.k, op.k, op.C, sf, bf, st, pd, act_code[op.act],
                                                       argf, out, y=y, ldy=y.shape[-1])))
                else:
                    wp = pl.pack_conv_weight(self._dev(op.w))
                    self.keep.append(wp)
                    x = self._buf(op.src, "i8")
                    self.layers.append((op, lambda op=op, x=x, wp=wp, sf=sf, bf=bf, argf=argf, out=out, y=y, st=st, pd=pd:
                                        pl.conv2d(x, wp, op.K, op.k, op.k, sf, bf, st, pd, (1, 1), act_code[op.act], argf,
                                                  out, y=y, Cc=op.C, ldy=y.shape[-1], groups=op.get("groups", 1))))
            elif k == "pool":
                x, y = self._buf(op.src, "i8"), self._buf(op.dst, "i8")
                C = self.prog.shapes[op.src][2]
                typ = pl.POOL_MAX if op.typ == "max" else pl.POOL_AVG
                self.layers.append((op, lambda op=op, x=x, y=y, C=C, typ=typ:
                                    pl.pool2d(x, (op.k, op.k), (op.stride, op.stride), (op.pad,) * 4, typ, True,
                                              op.in_scale / op.out_scale, pl.OUT_INT8, Cc=C, y=y)))
            elif k == "gather":
                srcs = [self._buf(s, "f32") for s in op.srcs]
                table = self._dev(np.asarray(op.table, np.int32))
                y = self._buf(op.dst, forms[op.dst])
                Cd = self.prog.shapes[op.dst][2]
                out = pl.OUT_FP32 if op.out_scale is None else pl.OUT_INT8
                self.layers.append((op, lambda srcs=srcs, table=table, y=y, Cd=Cd, out=out, sc=op.out_scale or 0.0:
                                    pl.channel_gather(srcs, table, Cd, y, out, sc)))
            elif k == "deconv":
                x = self._buf(op.src, "f32")
                y = self._buf(op.dst, forms[op.dst])
                w, b = self._dev(op.w), self._dev(op.bias)
                Co = self.prog.shapes[op.dst][2]
                out = pl.OUT_FP32 if op.out_scale is None else pl.OUT_INT8
                self.layers.append((op, lambda op=op, x=x, y=y, w=w, b=b, Co=Co, out=out:
                                    pl.conv2d_transpose_f32(x, w, b, op.stride, op.pad, y, Co, act_code[op.act], out,
                                                            op.out_scale or 0.0)))
            elif k == "addf":
                a, b, y = self._buf(op.src, "f32"), self._buf(op.src2, "f32"), self._buf(op.dst, "f32")
                self.layers.append((op, lambda a=a, b=b, y=y: pl.add_f32(a, b, y)))
            elif k == "softmax":
                x, y = self._buf(op.src, "f32"), self._buf(op.dst, "f32")
                self.layers.append((op, lambda x=x, y=y: pl.softmax_f32(x, y)))
            elif k == "scalef":
                x, y = self._buf(op.src, "f32"), self._buf(op.dst, "f32")
                self.layers.append((op, lambda op=op, x=x, y=y: pl.scale_f32(x, y, op.scale, op.bias, op.bias_after_scale)))
            elif k == "to_nhwc":
                y = self._buf(op.dst, "f32")
                N, H, W, C = y.shape
                lib = pl.lib()
                import ctypes as C_
                self.layers.append((op, lambda y=y, N=N, H=H, W=W, C=C: pl._check(lib.pl_nchw_to_nhwc(
                    C_.c_void_p(self.input.data_ptr()), C_.c_void_p(y.data_ptr()), N, C, H, W, C, 4, pl._stream()),
                    "pl_nchw_to_nhwc")))
            elif k == "to_nchw":
                x = self._buf(op.src, "f32")
                N, H, W, C = x.shape
                y = t.zeros((N, C, H, W), dtype=t.float32, device=self.dev)
                self.output = y
                lib = pl.lib()
                import ctypes as C_
                self.layers.append((op, lambda x=x, y=y, N=N, H=H, W=W, C=C: pl._check(lib.pl_nhwc_to_nchw(
                    C_.c_void_p(x.data_ptr()), C_.c_void_p(y.data_ptr()), N, C, H, W, C, 4, pl._stream()), "pl_nhwc_to_nchw")))
            else:
                raise NotImplementedError(k)
        t.cuda.synchronize()
        return self

    def tensor(self, name):
        """device tensor of an intermediate (tests)"""
        return self.bufs[name]

    def run(self):
        for _, fn in self.layers:
            fn()
        return self.output

    def capture(self):
        t = self.torch
        self.run()
        t.cuda.synchronize()
        g = t.cuda.CUDAGraph()
        with t.cuda.graph(g):
            self.run()
        self.graph = g
        return self

    def capture_inputs(self, n=2):
        t = self.torch
        self.inputs = [self.input] + [t.zeros_like(self.input) for _ in range(n - 1)]
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
        self.input.copy_(x_nchw, non_blocking=True)
        return self.replay() if self.graph is not None else self.run()
