"""Whole networks through the plugin surface (lite_net.PluginNet: KernelLite instances from the registry, run by
Launch() from a runtime program) against the C-ABI executor on the same calibrated program: the int8 tensors in
between and the logits must be identical bytes -- both paths end in the same kernels, the plugin path adds the
reference's op chain in front (calib -> layout -> thin-input conv instead of the fused space-to-depth entry)."""
import importlib

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def pkg():
    import __graft_entry__ as ge
    p = ge.load_package()
    return p, importlib.import_module("paddle_lite_b200.nets"), importlib.import_module("paddle_lite_b200.lite_net")


@pytest.mark.parametrize("name,batch", [("resnet50", 2), ("mobilenet_v1", 3)])
def test_plugin_net_matches_device_net(pkg, name, batch):
    _, nets, lite_net = pkg
    prog = getattr(nets, name)(seed=3)
    H, W, C = prog.shapes["input"]
    x = torch.rand((batch, C, H, W), generator=torch.Generator().manual_seed(5)).cuda() * 2 - 1
    ref = nets.DeviceNet(prog, batch).calibrate(x).bind()
    want = ref.forward(x).clone()
    torch.cuda.synchronize()
    net = lite_net.PluginNet(prog, batch)
    assert net.n_instructions == len(prog.ops) + 1  # calib + layout in place of the fused quantiser
    net.inputs[0].copy_(x)
    got = net.run(0)
    torch.cuda.synchronize()
    assert torch.equal(got, want)
    # an int8 tensor in the middle, then the graph replays of both input buffers
    mid = [op for op in prog.ops if op.kind in ("conv", "dwconv")][7]
    assert np.array_equal(net.tensor_numpy(mid.dst), ref._buf(mid.dst).cpu().numpy()[..., : prog.shapes[mid.dst][2]])
    net.capture()
    net.inputs[1].copy_(x)
    for i in (0, 1):
        assert torch.equal(net.replay(i), want)
