"""Predictor-level API (paddle-lite_b200/predictor.py): the CPU-side contract -- config / place validation, feed
handling, the profiler's tables and the MACs formula of lite/operators/conv_op.h:57-62.  No GPU is touched:
a predictor binds device buffers only at its first Run."""
import importlib
import os

import numpy as np
import pytest


@pytest.fixture(scope="module")
def P():
    import __graft_entry__ as ge
    ge.load_package()
    return importlib.import_module("paddle_lite_b200.predictor")


@pytest.fixture(scope="module")
def humanseg_dir(P):
    m = importlib.import_module("paddle_lite_b200.model")
    if os.path.exists(os.path.join(m.HUMANSEG_DIR, "__model__")):
        return m.HUMANSEG_DIR
    return m.HUMANSEG_FIXTURE  # the fused program as SaveOptimizedModel writes it


def cuda_places(P):
    return [P.Place(P.TargetType.kCUDA, P.PrecisionType.kInt8, P.DataLayoutType.kNHWC),
            P.Place(P.TargetType.kCUDA, P.PrecisionType.kFloat, P.DataLayoutType.kNHWC)]


def test_enum_values_match_paddle_place_h(P):
    # lite/api/paddle_place.h:45-83
    assert (P.TargetType.kHost, P.TargetType.kX86, P.TargetType.kCUDA, P.TargetType.kARM) == (1, 2, 3, 4)
    assert (P.PrecisionType.kFloat, P.PrecisionType.kInt8, P.PrecisionType.kInt32, P.PrecisionType.kAny) == (1, 2, 3, 4)
    assert (P.DataLayoutType.kNCHW, P.DataLayoutType.kAny, P.DataLayoutType.kNHWC) == (1, 2, 3)


def test_config_round_trip(P):
    c = P.CxxConfig()
    c.set_model_dir("/x/y")
    c.set_threads(4)
    c.set_power_mode(0)
    c.set_device_id(1)
    c.set_valid_places(cuda_places(P))
    assert c.model_dir() == "/x/y" and c.threads() == 4 and c.device_id() == 1
    assert [repr(p) for p in c.valid_places()] == ["cuda/int8_t/NHWC", "cuda/float/NHWC"]


def test_no_cuda_place_is_fatal(P, humanseg_dir):
    c = P.CxxConfig()
    c.set_model_dir(humanseg_dir)
    c.set_valid_places([P.Place(P.TargetType.kARM, P.PrecisionType.kInt8), P.Place(P.TargetType.kX86, P.PrecisionType.kFloat)])
    with pytest.raises(RuntimeError, match="kCUDA"):
        P.CreatePaddlePredictor(c)
    c.set_valid_places([P.Place(P.TargetType.kCUDA, P.PrecisionType.kFloat)])
    with pytest.raises(RuntimeError, match="kInt8"):  # a quantised model needs the int8 place
        P.CreatePaddlePredictor(c)


def test_missing_model_raises(P):
    c = P.CxxConfig()
    c.set_model_dir("/nonexistent/model")
    c.set_valid_places(cuda_places(P))
    with pytest.raises(FileNotFoundError):
        P.CreatePaddlePredictor(c)


def test_feed_names_shapes_and_resize(P, humanseg_dir):
    c = P.CxxConfig()
    c.set_model_dir(humanseg_dir)
    c.set_valid_places(cuda_places(P))
    pred = P.CreatePaddlePredictor(c)
    assert pred.GetInputNames() == ["input"] and pred.GetOutputNames() == ["output"]
    x = pred.GetInput(0)
    assert x.shape() == [1, 3, 192, 192]
    x.Resize([4, 3, 192, 192])
    assert pred.GetInputByName("input").shape() == [4, 3, 192, 192]
    with pytest.raises(ValueError):
        x.Resize([4, 3, 224, 224])
    assert pred.GetVersion().startswith("paddle-lite_b200")
    assert len(pred.GetParamNames()) > 100
    assert x.target() == P.TargetType.kHost and x.precision() == P.PrecisionType.kFloat


def test_op_character_macs_formula(P, humanseg_dir):
    m = importlib.import_module("paddle_lite_b200.model")
    prog = m.load_saved_program(humanseg_dir) if not os.path.isdir(humanseg_dir) else m.humanseg_program(humanseg_dir)
    conv = next(op for op in prog.ops if op.kind == "conv" and op.k == 3)
    ch = P.OpCharacter(conv, prog, 8)
    H, W, C = prog.shapes[conv.src]
    Pq, Q, K = prog.shapes[conv.dst]
    # MACs = 2.f * kw * kh * batchsize * out_c * out_h * out_w * in_c / group   (conv_op.h:57)
    assert ch.macs == 2.0 * 3 * 3 * 8 * K * Pq * Q * C / conv.get("groups", 1)
    assert ch.op_type == "conv2d" and ch.kernel_attr == "cuda/int8_t/NHWC"
    assert ch.remark.startswith("3x3p") and ch.input_shape == f"8x{C}x{H}x{W}"
    dw = next(op for op in prog.ops if op.kind == "dwconv")
    cd = P.OpCharacter(dw, prog, 8)
    Hd, Wd, Cd = prog.shapes[dw.src]
    Pd, Qd, Kd = prog.shapes[dw.dst]
    assert cd.macs == 2.0 * dw.k * dw.k * 8 * Kd * Pd * Qd * Cd / Cd  # groups == channels


def test_profiler_summary_tables(P, humanseg_dir):
    m = importlib.import_module("paddle_lite_b200.model")
    prog = m.load_saved_program(humanseg_dir) if not os.path.isdir(humanseg_dir) else m.humanseg_program(humanseg_dir)
    chars = [P.OpCharacter(op, prog, 2) for op in prog.ops]
    prof = P.Profiler("humanseg", chars)
    rng = np.random.default_rng(0)
    for it in range(15):
        prof.record(list(rng.uniform(0.01, 0.02, len(chars)) * (3.0 if it < 10 else 1.0)))  # slow warm-ups
    s = prof.Summary(concise=True, warm_up=10)
    assert "Timing cycle = 15" in s and "Exclude 10 warm-ups" in s
    assert "OperatorType" in s and "KerneAttr(Place)" in s and "CalledTimes" in s  # the reference's header, typo included
    row = next(l for l in s.splitlines() if l.startswith("conv2d "))
    assert int(row.split()[-1]) == sum(1 for op in prog.ops if op.kind == "conv")
    total = float(s.splitlines()[-1].split()[1])
    assert 0.01 * len(chars) <= total <= 0.02 * len(chars)  # warm-ups excluded
    d = prof.Summary(concise=False, warm_up=10)
    assert len(d.splitlines()) == len(chars) + 3 and "Last(ms)" in d and "GOPS" in d
