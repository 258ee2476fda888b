"""BASELINE config 3 end to end: the shipped quantised model humanseg_shfv75_0119_quant (its fused program, from the
model files or the committed fixture) on the GPU executor against the oracle walk of the same program: every int8
tensor bit for bit with the FILE's scales, every fp32 tensor and the softmax output within 1e-5 relative."""
import importlib

import numpy as np
import pytest
import torch

import graph_oracle

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def m():
    import __graft_entry__ as ge
    ge.load_package()
    return importlib.import_module("paddle_lite_b200.model")


def test_humanseg_every_tensor_vs_oracle(m):
    prog = m.humanseg_program()
    N = 2
    x = np.random.default_rng(1234).uniform(0, 1, (N, 3, 192, 192)).astype(np.float32)
    want = graph_oracle.run_graph(prog, x)
    net = m.GraphNet(prog, N, device="cuda:0").bind()
    net.input.copy_(torch.from_numpy(x))
    out = net.run()
    torch.cuda.synchronize()
    n_i8 = n_f32 = 0
    for op in prog.ops:
        if op.dst in ("output",) or op.dst not in net.bufs:
            continue
        got = net.bufs[op.dst].cpu().numpy()
        ref = want[op.dst]
        C = ref.shape[-1]
        if ref.dtype == np.int8:
            assert np.array_equal(got[..., :C], ref), f"{op.kind} {op.dst}"
            assert not got[..., C:].any(), f"pad channels of {op.dst} are not zero"
            n_i8 += 1
        else:
            assert np.allclose(got[..., :C], ref, rtol=1e-5, atol=1e-6), f"{op.kind} {op.dst}"
            n_f32 += 1
    assert n_i8 >= 50 and n_f32 >= 30
    y = out.cpu().numpy()
    assert y.shape == (N, 2, 192, 192)
    assert np.allclose(y, want["output"], rtol=1e-5, atol=1e-6)


def test_humanseg_graph_replay_matches_eager(m):
    prog = m.humanseg_program()
    net = m.GraphNet(prog, 4, device="cuda:0").bind()
    x = torch.rand((4, 3, 192, 192), generator=torch.Generator().manual_seed(7))
    net.input.copy_(x)
    eager = net.run().clone()
    net.capture_inputs(2)
    net.inputs[1].copy_(x)
    for i in (0, 1):
        assert torch.equal(net.replay(i), eager)


@pytest.mark.parametrize("case", [(2, 5, 6, 7, 3, 2, 2, 0, False), (1, 24, 24, 18, 18, 2, 2, 0, True),
                                  (2, 4, 4, 8, 5, 3, 2, 1, False), (1, 6, 5, 4, 4, 4, 2, 1, True)])
def test_conv2d_transpose_f32(m, case, pl):
    import orc
    N, H, W, Ci, Co, k, s, p, quant = case
    rng = np.random.default_rng(11)
    x = rng.standard_normal((N, H, W, Ci)).astype(np.float32)
    w = rng.standard_normal((Ci, Co, k, k)).astype(np.float32)
    b = rng.standard_normal(Co).astype(np.float32)
    sc = 0.05 if quant else 0.0
    want = orc.conv2d_transpose_f32(x, w, b, (s, s), (p, p), relu=True, out_scale=sc)
    dev = "cuda:0"
    y = torch.zeros(want.shape, dtype=torch.int8 if quant else torch.float32, device=dev)
    pl.conv2d_transpose_f32(torch.from_numpy(x).to(dev), torch.from_numpy(w).to(dev), torch.from_numpy(b).to(dev), s, p, y,
                            Co, pl.ACT_RELU, pl.OUT_INT8 if quant else pl.OUT_FP32, sc)
    torch.cuda.synchronize()
    assert np.array_equal(y.cpu().numpy(), want)  # pinned accumulation order: bit-exact even in fp32
