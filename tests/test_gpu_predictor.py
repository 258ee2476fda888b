"""Predictor API on the GPU: the shipped quantised model through CxxConfig / CreatePaddlePredictor / Run against
the device executor it wraps (bit for bit) and the oracle walk (1e-5 relative, north_star's fp32 bound); Clone,
SaveOptimizedModel round trip, the profiler summary and the benchmark flow of lite/api/benchmark.cc."""
import importlib
import json
import os

import numpy as np
import pytest
import torch

import graph_oracle

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def P():
    import __graft_entry__ as ge
    ge.load_package()
    return importlib.import_module("paddle_lite_b200.predictor")


@pytest.fixture(scope="module")
def M():
    return importlib.import_module("paddle_lite_b200.model")


def model_path(M):
    return M.HUMANSEG_DIR if os.path.exists(os.path.join(M.HUMANSEG_DIR, "__model__")) else M.HUMANSEG_FIXTURE


def make(P, M, **kw):
    c = P.CxxConfig()
    c.set_model_dir(kw.pop("path", model_path(M)))
    c.set_valid_places([P.Place(P.TargetType.kCUDA, P.PrecisionType.kInt8, P.DataLayoutType.kNHWC),
                        P.Place(P.TargetType.kCUDA, P.PrecisionType.kFloat, P.DataLayoutType.kNHWC)])
    for k, v in kw.items():
        getattr(c, "set_" + k)(v)
    return P.CreatePaddlePredictor(c)


def test_predictor_matches_oracle_and_executor(P, M):
    N = 2
    x = np.random.default_rng(5).uniform(0, 1, (N, 3, 192, 192)).astype(np.float32)
    pred = make(P, M)
    t = pred.GetInput(0)
    t.Resize([N, 3, 192, 192])
    t.CopyFromCpu(x)
    pred.Run()
    y = pred.GetOutput(0).CopyToCpu()
    assert y.shape == (N, 2, 192, 192) and pred.GetOutput(0).shape() == [N, 2, 192, 192]
    want = graph_oracle.run_graph(M.humanseg_program(), x)
    assert np.allclose(y, want["output"], rtol=1e-5, atol=1e-6)
    net = M.GraphNet(M.humanseg_program(), N, device="cuda:0").bind()
    net.input.copy_(torch.from_numpy(x))
    assert np.array_equal(y, net.run().cpu().numpy())
    # an intermediate int8 tensor by name, logical channels only
    conv = next(op for op in pred._prog.ops if op.kind == "conv")
    assert np.array_equal(pred.GetTensor(conv.dst).data(), want[conv.dst])
    # a second Run on new data through mutable_data, then Clone and the eager (no graph) mode agree
    x2 = np.random.default_rng(6).uniform(0, 1, x.shape).astype(np.float32)
    t.mutable_data()[...] = x2
    pred.Run()
    y2 = pred.GetOutput(0).data()
    assert not np.array_equal(y2, y)
    for other in (pred.Clone(), make(P, M, cuda_graph=False), make(P, M, cuda_graph=False, nvtx=True)):
        o = other.GetInput(0)
        o.Resize([N, 3, 192, 192])
        o.CopyFromCpu(x2)
        other.Run()
        assert np.array_equal(other.GetOutput(0).data(), y2)


def test_save_optimized_model_round_trip(P, M, tmp_path):
    pred = make(P, M)
    prefix = str(tmp_path / "humanseg_opt")
    pred.SaveOptimizedModel(prefix)
    assert os.path.exists(prefix + ".json") and os.path.exists(prefix + ".npz")
    again = make(P, M, path=prefix)
    x = np.random.default_rng(7).uniform(0, 1, (1, 3, 192, 192)).astype(np.float32)
    outs = []
    for p_ in (pred, again):
        p_.GetInput(0).CopyFromCpu(x)
        p_.Run()
        outs.append(p_.GetOutput(0).data())
    assert np.array_equal(outs[0], outs[1])


def test_profiler_and_benchmark_flow(P, M, tmp_path):
    res = str(tmp_path / "result.txt")
    rec = P.benchmark(model_path(M), (4, 3, 192, 192), warmup=2, repeats=5, result_filename=res, show_output=True,
                      profile=True)
    assert rec["repeats"] == 5 and rec["min"] <= rec["average"] <= rec["max"]
    line = open(res).read()
    assert "min = " in line and "max = " in line and "average = " in line  # benchmark.cc:176-181
    s = rec["summary"]
    assert "Concise Dispatch Profiler Summary" in s and "conv2d" in s and "depthwise_conv2d" in s
    assert "igemm_i8<" in s  # the tile plan each conv took
    assert len(rec["summary_detailed"].splitlines()) == len(M.humanseg_program().ops) + 3
    fast = P.benchmark(model_path(M), (4, 3, 192, 192), warmup=3, repeats=20)
    assert fast["average"] < rec["average"]  # one graph replay beats 110 timed eager launches
    json.dumps(fast)
