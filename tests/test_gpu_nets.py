"""End-to-end parity: whole int8 programs (MobileNetV1, ResNet-50) on the GPU through
the C-ABI vs the CPU oracle walking the same program with the same scales."""
import numpy as np
import pytest
import torch

import net_oracle

pytestmark = pytest.mark.gpu


def _nets():
    import __graft_entry__ as ge
    ge.load_package()
    import importlib
    return importlib.import_module("paddle_lite_b200.nets")


def _check(builder, batch, res, graph):
    nets = _nets()
    prog = builder(seed=7, res=res)
    g = torch.Generator(device="cpu").manual_seed(5)
    x = torch.rand((batch, 3, res, res), generator=g, dtype=torch.float32)
    net = nets.DeviceNet(prog, batch).calibrate(x.cuda()).bind()
    if graph:
        net.capture()
    got = net.forward(x.cuda())
    torch.cuda.synchronize()
    got = got.cpu().numpy().copy()
    want, vals = net_oracle.run_program(prog, x.numpy(), keep=True)
    # every intermediate int8 tensor is bit-exact, not only the logits
    for key, buf in net.bufs.items():
        name, dtype = key
        if dtype == torch.int8 and name in vals:
            C = prog.shapes[name][2]
            assert np.array_equal(buf.cpu().numpy()[..., :C], vals[name]), f"tensor {name} differs"
    for name in net.s2d_bufs:  # stem input kept in space-to-depth form
        C = prog.shapes[name][2]
        assert np.array_equal(net.s2d_as_nhwc(name)[..., :C], vals[name]), f"tensor {name} (s2d) differs"
    np.testing.assert_allclose(got, want, rtol=1e-5, atol=0)
    assert np.array_equal(got, want)
    assert np.isfinite(got).all() and np.std(got) > 0


def test_mobilenet_v1_small(pl):
    _check(_nets().mobilenet_v1, 3, 64, graph=False)


def test_resnet50_small(pl):
    _check(_nets().resnet50, 2, 64, graph=True)


def test_resnet50_small_unfused_add(pl):
    import functools
    _check(functools.partial(_nets().resnet50, fuse_residual=False), 2, 64, graph=False)


def test_mobilenet_v1_224(pl):
    _check(_nets().mobilenet_v1, 2, 224, graph=True)


def test_resnet50_224(pl):
    _check(_nets().resnet50, 1, 224, graph=False)


def test_reference_engine_equals_oracle_engine(pl):
    """The timed CPU baseline (reference's conv_basic / basic_gemm) computes the same program."""
    import orc
    if orc.ref() is None:
        pytest.skip("oracle/_ref not built")
    nets = _nets()
    prog = nets.mobilenet_v1(seed=3, res=32)
    x = torch.rand((1, 3, 32, 32), generator=torch.Generator().manual_seed(1))
    nets.DeviceNet(prog, 1).calibrate(x.cuda())
    a = net_oracle.run_program(prog, x.numpy(), engine="oracle")
    b = net_oracle.run_program(prog, x.numpy(), engine="reference")
    assert np.array_equal(a, b)


def test_shufflenet_v2_like_small(pl):
    """ShuffleNet-v2 topology: split as channel-slice views, fused concat + shuffle_channel, everything int8
    NHWC resident between the stem calib and the fc."""
    _check(_nets().shufflenet_v2_like, 2, 64, graph=True)


def test_shufflenet_v2_like_224(pl):
    _check(_nets().shufflenet_v2_like, 1, 224, graph=False)
