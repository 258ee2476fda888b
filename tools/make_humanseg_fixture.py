"""Writes tests/golden/humanseg_fused.{json,npz}: the FUSED executable program of the shipped quantised model
/root/reference/humanseg_shfv75_0119_quant (BASELINE config 3), produced by this repo's own front end
(paddle-lite_b200/model.py: wire-format reader + LoDTensor reader + the fusion passes).  The GPU box has no reference
tree; the fixture lets its tests and bench.py run the real model.  Runs in the authoring container only.
  python tools/make_humanseg_fixture.py"""
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

ge.load_package()
m = importlib.import_module("paddle_lite_b200.model")
g = m.Graph.load(m.HUMANSEG_DIR)
print("fusions:", m.optimize(g))
prog = m.build_program(g, "humanseg")
m.save_program(prog, m.HUMANSEG_FIXTURE)
kinds = {}
for op in prog.ops:
    kinds[op.kind] = kinds.get(op.kind, 0) + 1
print(len(prog.ops), "ops:", kinds)
print({p: os.path.getsize(m.HUMANSEG_FIXTURE + p) for p in (".json", ".npz")})
