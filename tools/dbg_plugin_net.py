"""debug: PluginNet vs DeviceNet, tensor by tensor; then a large-batch eager run"""
import importlib
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

ge.load_package()
nets = importlib.import_module("paddle_lite_b200.nets")
lite_net = importlib.import_module("paddle_lite_b200.lite_net")
name = sys.argv[1] if len(sys.argv) > 1 else "resnet50"
batch = int(sys.argv[2]) if len(sys.argv) > 2 else 2
prog = getattr(nets, name)(seed=3)
H, W, C = prog.shapes["input"]
x = torch.rand((batch, C, H, W), generator=torch.Generator().manual_seed(5)).cuda() * 2 - 1
ref = nets.DeviceNet(prog, batch).calibrate(x).bind()
want = ref.forward(x).clone()
torch.cuda.synchronize()
net = lite_net.PluginNet(prog, batch)
net.inputs[0].copy_(x)
try:
    got = net.run(0)
    torch.cuda.synchronize()
except Exception as e:
    print("RUN FAILED:", e)
    sys.exit(1)
print("logits equal:", torch.equal(got, want), float((got - want).abs().max()))
if batch <= 8:
    for op in prog.ops:
        if op.kind == "fc":
            continue
        a = net.tensor_numpy(op.dst)
        if op.kind == "quant_in" and op.dst in ref.s2d_bufs:
            b = ref.s2d_as_nhwc(op.dst)[..., : a.shape[-1]]
        else:
            b = ref._buf(op.dst).cpu().numpy()[..., : a.shape[-1]]
        ok = np.array_equal(a, b)
        print(op.kind, op.dst, a.shape, "OK" if ok else f"MISMATCH {np.abs(a.astype(int) - b.astype(int)).max()} {np.mean(a != b):.4f}")
        if not ok:
            break
