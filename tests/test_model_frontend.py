"""Quantised-model front end (paddle-lite_b200/model.py) on the CPU: wire-format reader, LoDTensor reader, the
fusion passes restated from the reference, the fused-program fixture, and the oracle walk of the shipped model."""
import importlib
import os

import numpy as np
import pytest

import graph_oracle
import orc


@pytest.fixture(scope="module")
def m():
    import __graft_entry__ as ge
    ge.load_package()
    return importlib.import_module("paddle_lite_b200.model")


def _has_model(m):
    return os.path.exists(os.path.join(m.HUMANSEG_DIR, "__model__"))


def test_reader_parses_the_shipped_model(m):
    if not _has_model(m):
        pytest.skip("reference tree not present")
    ops, dims, persist = m.load_program(m.HUMANSEG_DIR)
    counts = {}
    for op in ops:
        counts[op["type"]] = counts.get(op["type"], 0) + 1
    # SURVEY.md 8d config 3
    assert counts["conv2d"] == 44 and counts["depthwise_conv2d"] == 18 and counts["conv2d_transpose"] == 4
    assert counts["fake_quantize_range_abs_max"] == 58 and counts["fake_dequantize_max_abs"] == 62
    assert counts["batch_norm"] == 65 and counts["concat"] == 16 and counts["split"] == 14
    assert dims["image"] == [-1, 3, 192, 192]
    w = m.load_lod_tensor(os.path.join(m.HUMANSEG_DIR, "conv2d_0.w_0"))
    assert w.shape == (36, 3, 3, 3) and w.dtype == np.float32
    assert np.array_equal(w, np.round(w)) and np.abs(w).max() <= 127  # slim stores integer-valued weights
    s = m.load_lod_tensor(os.path.join(m.HUMANSEG_DIR, "image.scale"))
    assert s.shape == (1,) and s[0] > 0


def _hand_graph(m, rng, alpha_sign):
    """x -> fake_quant -> conv2d(3x3) -> fake_dequant -> batch_norm -> relu -> fetch"""
    K, C = 4, 3
    w = rng.integers(-127, 128, (K, C, 3, 3)).astype(np.float32)
    params = {"w": w, "x.scale": np.array([2.5], np.float32),
              "bn.s": (rng.uniform(0.5, 1.5, K) * alpha_sign).astype(np.float32), "bn.b": rng.standard_normal(K).astype(np.float32),
              "bn.m": rng.standard_normal(K).astype(np.float32), "bn.v": rng.uniform(0.5, 2, K).astype(np.float32)}
    ops = [
        {"type": "feed", "inputs": {"X": ["feed"]}, "outputs": {"Out": ["x"]}, "attrs": {}},
        {"type": "fake_quantize_range_abs_max", "inputs": {"X": ["x"], "InScale": ["x.scale"]},
         "outputs": {"Out": ["x.q"], "OutScale": ["x.scale"]}, "attrs": {"bit_length": 8}},
        {"type": "conv2d", "inputs": {"Input": ["x.q"], "Filter": ["w"]}, "outputs": {"Output": ["c"]},
         "attrs": {"strides": [1, 1], "paddings": [1, 1], "dilations": [1, 1], "groups": 1}},
        {"type": "fake_dequantize_max_abs", "inputs": {"X": ["c"], "Scale": ["x.scale"]}, "outputs": {"Out": ["c.dq"]},
         "attrs": {"max_range": 127.0 * 127.0 / 0.8}},
        {"type": "batch_norm", "inputs": {"X": ["c.dq"], "Scale": ["bn.s"], "Bias": ["bn.b"], "Mean": ["bn.m"], "Variance": ["bn.v"]},
         "outputs": {"Y": ["bn"]}, "attrs": {"epsilon": 1e-5}},
        {"type": "relu", "inputs": {"X": ["bn"]}, "outputs": {"Out": ["y"]}, "attrs": {}},
        {"type": "fetch", "inputs": {"X": ["y"]}, "outputs": {"Out": ["fetch"]}, "attrs": {}},
    ]
    dims = {"x": [-1, C, 6, 6], "x.q": [-1, C, 6, 6], "c": [-1, K, 6, 6], "c.dq": [-1, K, 6, 6], "bn": [-1, K, 6, 6],
            "y": [-1, K, 6, 6]}
    return m.Graph(ops, dims, params), w, params


@pytest.mark.parametrize("alpha_sign", [1.0, -1.0])
def test_fusers_reproduce_the_reference_rules(m, alpha_sign):
    """quant_dequant_op_fuser.cc:59-92 / :132-202 and conv_bn_fuser.cc:155-189 on a hand graph, then the fused int8
    conv against the un-fused float chain the model file describes."""
    rng = np.random.default_rng(3)
    g, w, P = _hand_graph(m, rng, alpha_sign)
    rep = m.optimize(g)
    assert rep["delete_quant"] == 1 and rep["dequant"] == 1 and rep["conv_bn"] == 1 and rep["conv_activation"] == 1
    assert [op["type"] for op in g.ops] == ["conv2d"]
    conv = g.ops[0]
    in_scale = np.float32(2.5) / np.float32(127)
    assert conv["input_scale"]["x"] == pytest.approx(float(in_scale), rel=0, abs=0)
    whole = np.float32(127 * 127) / np.float32(127.0 * 127.0 / 0.8) / np.float32(127)
    alpha = P["bn.s"] / np.sqrt(P["bn.v"] + np.float32(1e-5))
    assert np.array_equal(conv["weight_scale"], (np.full(4, whole, np.float32) * np.abs(alpha)).astype(np.float32))
    wq = g.params[conv["inputs"]["Filter"][0]]
    assert wq.dtype == np.int8 and np.array_equal(wq, (w * np.sign(alpha).reshape(-1, 1, 1, 1)).astype(np.int8))
    bias = g.params[conv["inputs"]["Bias"][0]]
    assert np.allclose(bias, P["bn.b"] - P["bn.m"] * alpha, rtol=1e-6)
    assert conv["act"] == "relu" and conv["outputs"]["Output"] == ["y"]
    # numerics: fused program (oracle walk) == the float chain of the original graph
    prog = m.build_program(g, "hand")
    x = rng.uniform(-2.5, 2.5, (2, 3, 6, 6)).astype(np.float32)
    got = graph_oracle.run_graph(prog, x)["output"]
    xq = np.clip(np.round(x.astype(np.float64) / 2.5 * 127), -127, 127)
    acc = orc.ref_conv_acc_nchw(xq.astype(np.int8), w.astype(np.int8), (1, 1), (1, 1), (1, 1), 1).astype(np.float64)
    deq = acc * 2.5 / (127.0 * 127.0 / 0.8)  # fake_dequantize_max_abs: x * scale / max_range
    bn = (deq - P["bn.m"].reshape(1, -1, 1, 1)) * alpha.reshape(1, -1, 1, 1).astype(np.float64) + P["bn.b"].reshape(1, -1, 1, 1)
    assert np.allclose(got, np.maximum(bn, 0), rtol=2e-5, atol=2e-5)


def test_shuffle_channel_fuser_and_channel_maps(m):
    """reshape2 + transpose2 + reshape2 -> shuffle_channel (shuffle_channel_fuser.cc); the composed channel map of
    concat -> shuffle -> split equals the reference loops (orc.* restate concat.h / shuffle_channel.cc / split.cc)."""
    ops = [
        {"type": "feed", "inputs": {"X": ["feed"]}, "outputs": {"Out": ["a"]}, "attrs": {}},
        {"type": "scale", "inputs": {"X": ["a"]}, "outputs": {"Out": ["b"]}, "attrs": {"scale": 2.0, "bias": 1.0}},
        {"type": "concat", "inputs": {"X": ["a", "b"]}, "outputs": {"Out": ["c"]}, "attrs": {"axis": 1}},
        {"type": "reshape2", "inputs": {"X": ["c"]}, "outputs": {"Out": ["r1"]}, "attrs": {"shape": [-1, 2, 6, 4, 4]}},
        {"type": "transpose2", "inputs": {"X": ["r1"]}, "outputs": {"Out": ["t"]}, "attrs": {"axis": [0, 2, 1, 3, 4]}},
        {"type": "reshape2", "inputs": {"X": ["t"]}, "outputs": {"Out": ["r2"]}, "attrs": {"shape": [-1, 12, 4, 4]}},
        {"type": "split", "inputs": {"X": ["r2"]}, "outputs": {"Out": ["s0", "s1"]}, "attrs": {"axis": 1, "num": 2}},
        {"type": "scale", "inputs": {"X": ["s1"]}, "outputs": {"Out": ["y"]}, "attrs": {"scale": 1.0, "bias": 0.0}},
        {"type": "fetch", "inputs": {"X": ["y"]}, "outputs": {"Out": ["fetch"]}, "attrs": {}},
    ]
    dims = {k: [-1, 6, 4, 4] for k in ("a", "b", "s0", "s1", "y")}
    dims.update({"c": [-1, 12, 4, 4], "r2": [-1, 12, 4, 4]})
    g = m.Graph(ops, dims, {})
    assert m.fuse_shuffle_channel(g) == 1 and "shuffle_channel" in g.counts()
    prog = m.build_program(g, "shuffle")
    x = np.random.default_rng(5).standard_normal((2, 6, 4, 4)).astype(np.float32)
    got = graph_oracle.run_graph(prog, x)["output"]
    cat = orc.concat_nchw([x, x * 2 + 1])
    want = orc.split_nchw(orc.shuffle_channel_nchw(cat, 2), [6, 6])[1]
    assert np.array_equal(got, want)


def test_fixture_is_the_front_end_output(m):
    if not _has_model(m):
        pytest.skip("reference tree not present")
    g = m.Graph.load(m.HUMANSEG_DIR)
    m.optimize(g)
    fresh = m.build_program(g, "humanseg")
    saved = m.load_saved_program(m.HUMANSEG_FIXTURE)
    assert len(fresh.ops) == len(saved.ops) and fresh.shapes == saved.shapes
    for a, b in zip(fresh.ops, saved.ops):
        assert a.kind == b.kind and a.src == b.src and a.dst == b.dst
        for k in ("w", "w_scale", "bias", "table"):
            if k in a:
                assert np.array_equal(np.asarray(a[k]), np.asarray(b[k])), (a.kind, a.dst, k)
        for k in ("in_scale", "out_scale", "stride", "pad", "act"):
            if k in a:
                assert a[k] == b[k]


def test_oracle_walks_the_shipped_model(m):
    prog = m.humanseg_program()
    kinds = {}
    for op in prog.ops:
        kinds[op.kind] = kinds.get(op.kind, 0) + 1
    assert kinds["conv"] == 44 and kinds["dwconv"] == 18 and kinds["deconv"] == 4
    x = np.random.default_rng(1234).uniform(0, 1, (1, 3, 192, 192)).astype(np.float32)
    vals = graph_oracle.run_graph(prog, x)
    y = vals["output"]
    assert y.shape == (1, 2, 192, 192) and np.allclose(y.sum(1), 1.0, atol=1e-5)
    n_i8 = sum(1 for v in vals.values() if v.dtype == np.int8)
    assert n_i8 >= 50
    for k, v in vals.items():  # activations use their range: the file's scales fit the data
        if v.dtype == np.int8 and k != "input.q":
            assert np.abs(v).max() > 8, k


@pytest.mark.skipif(orc.ref() is None or not hasattr(orc.ref(), "ref_deconv_basic_f32"), reason="oracle/_ref not built")
@pytest.mark.parametrize("case", [(5, 6, 7, 3, 2, 2, 0), (4, 4, 8, 5, 3, 2, 1), (6, 5, 4, 4, 4, 2, 1), (3, 3, 6, 2, 3, 1, 1)])
def test_deconv_oracle_vs_reference_deconv_basic(case):
    """the pinned-order conv2d_transpose == deconv_basic (naive_math_impl.h:540-642) up to summation order"""
    H, W, Ci, Co, k, s, p = case
    rng = np.random.default_rng(0)
    x = rng.standard_normal((2, H, W, Ci)).astype(np.float32)
    w = rng.standard_normal((Ci, Co, k, k)).astype(np.float32)
    b = rng.standard_normal(Co).astype(np.float32)
    got = orc.conv2d_transpose_f32(x, w, b, (s, s), (p, p), relu=True)
    ref = orc.ref_deconv_nchw(x.transpose(0, 3, 1, 2), w, b, (s, s), (p, p, p, p), relu=True).transpose(0, 2, 3, 1)
    assert np.allclose(got, ref, rtol=1e-5, atol=1e-5)
