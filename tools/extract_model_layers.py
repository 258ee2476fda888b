"""Reads a Paddle `__model__` (ProgramDesc protobuf, lite/core/framework.proto) with a minimal wire-format
parser and writes the list of conv2d / depthwise_conv2d layers (shapes, strides, pads, groups, whether the
input is fake-quantised) as a JSON fixture.  Runs in the authoring container only (the reference tree is not
on the GPU box); the fixture is committed.
  python tools/extract_model_layers.py /root/reference/humanseg_shfv75_0119_quant tests/golden/humanseg_layers.json"""
import json
import os
import struct
import sys


def varint(b, i):
    v = s = 0
    while True:
        c = b[i]
        i += 1
        v |= (c & 0x7F) << s
        s += 7
        if not c & 0x80:
            return v, i


def fields(b):
    i = 0
    while i < len(b):
        key, i = varint(b, i)
        fn, wt = key >> 3, key & 7
        if wt == 0:
            v, i = varint(b, i)
        elif wt == 1:
            v = b[i:i + 8]
            i += 8
        elif wt == 2:
            n, i = varint(b, i)
            v = b[i:i + n]
            i += n
        elif wt == 5:
            v = b[i:i + 4]
            i += 4
        else:
            raise ValueError(f"wire type {wt}")
        yield fn, wt, v


def sint(v):  # int64 varint -> signed
    return v - (1 << 64) if v >= (1 << 63) else v


def packed_or_single(vals):
    out = []
    for wt, v in vals:
        if wt == 0:
            out.append(sint(v))
        else:
            i = 0
            while i < len(v):
                x, i = varint(v, i)
                out.append(sint(x))
    return out


def parse_attr(b):
    a = {"ints": [], "floats": []}
    for fn, wt, v in fields(b):
        if fn == 1:
            a["name"] = v.decode()
        elif fn == 2:
            a["type"] = v
        elif fn == 3:
            a["i"] = sint(v) if v < (1 << 63) else sint(v)
            if a["i"] >= (1 << 31):
                a["i"] -= (1 << 32)
        elif fn == 4:
            a["f"] = struct.unpack("<f", v)[0]
        elif fn == 5:
            a["s"] = v.decode(errors="replace")
        elif fn == 6:
            a["ints"] += packed_or_single([(wt, v)])
        elif fn == 7:
            if wt == 5:
                a["floats"].append(struct.unpack("<f", v)[0])
            else:
                a["floats"] += list(struct.unpack(f"<{len(v) // 4}f", v))
        elif fn == 10:
            a["b"] = bool(v)
    return a


def parse_opvar(b):
    name, args = None, []
    for fn, wt, v in fields(b):
        if fn == 1:
            name = v.decode()
        elif fn == 2:
            args.append(v.decode())
    return name, args


def parse_op(b):
    op = {"inputs": {}, "outputs": {}, "attrs": {}}
    for fn, wt, v in fields(b):
        if fn == 3:
            op["type"] = v.decode()
        elif fn == 1:
            k, a = parse_opvar(v)
            op["inputs"][k] = a
        elif fn == 2:
            k, a = parse_opvar(v)
            op["outputs"][k] = a
        elif fn == 4:
            a = parse_attr(v)
            op["attrs"][a.get("name")] = a
    return op


def parse_var(b):
    name, dims = None, None
    for fn, wt, v in fields(b):
        if fn == 1:
            name = v.decode()
        elif fn == 2:  # VarType
            for f2, w2, v2 in fields(v):
                if f2 == 3:  # LoDTensorDesc
                    for f3, w3, v3 in fields(v2):
                        if f3 == 1:  # TensorDesc
                            dims = packed_or_single([(w4, v4) for f4, w4, v4 in fields(v3) if f4 == 2])
    return name, dims


def main():
    model_dir, out_path = sys.argv[1], sys.argv[2]
    b = open(os.path.join(model_dir, "__model__"), "rb").read()
    ops, var_dims = [], {}
    for fn, wt, v in fields(b):
        if fn != 1:
            continue
        for f2, w2, v2 in fields(v):  # BlockDesc
            if f2 == 3:
                n, d = parse_var(v2)
                var_dims[n] = d
            elif f2 == 4:
                ops.append(parse_op(v2))
    counts = {}
    for op in ops:
        counts[op["type"]] = counts.get(op["type"], 0) + 1
    quantised = set()
    for op in ops:
        if op["type"].startswith("fake_quantize") or op["type"].startswith("fake_dequantize"):
            quantised.update(op["outputs"].get("Out", []))
    layers = []
    for op in ops:
        if op["type"] not in ("conv2d", "depthwise_conv2d"):
            continue
        x, w, y = op["inputs"]["Input"][0], op["inputs"]["Filter"][0], op["outputs"]["Output"][0]
        a = op["attrs"]
        layers.append({
            "type": op["type"], "input": x, "in_dims": var_dims.get(x), "filter_dims": var_dims.get(w),
            "out_dims": var_dims.get(y), "strides": a["strides"]["ints"], "paddings": a["paddings"]["ints"],
            "dilations": a["dilations"]["ints"], "groups": a["groups"]["i"],
            "input_fake_quantised": x in quantised})
    feed = [op for op in ops if op["type"] == "feed"]
    json.dump({"source": os.path.basename(model_dir.rstrip("/")) + "/__model__", "op_counts": counts,
               "feed": [(op["outputs"]["Out"][0], var_dims.get(op["outputs"]["Out"][0])) for op in feed],
               "layers": layers}, open(out_path, "w"), indent=1)
    print(counts)
    print(len(layers), "conv layers;", sum(l["input_fake_quantised"] for l in layers), "with fake-quantised input")
    for l in layers[:6]:
        print(l)


if __name__ == "__main__":
    main()
