"""Experiment: does a tcgen05 smem descriptor whose start address is shifted by whole rows inside a
TMA-swizzled K-major tile read the rows the way TMA wrote them?  (needed for halo-reuse convs)
Run one (shift, baseoff) per process: PL_DBG_ASHIFT=s PL_DBG_BASEOFF=b python tools/dbg_rowshift.py K"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

pl = ge.load_package().capi
K = int(sys.argv[1]) if len(sys.argv) > 1 else 64
shift = int(os.environ.get("PL_DBG_ASHIFT", "0"))
M, N = 128, 64
rng = np.random.default_rng(0)
A = rng.integers(-127, 128, (M, K), dtype=np.int8)
B = rng.integers(-127, 128, (N, K), dtype=np.int8)
y = pl.gemm(torch.from_numpy(A).cuda(), torch.from_numpy(B).cuda(), None, None, out=pl.OUT_INT32)
torch.cuda.synchronize()
got = y.cpu().numpy()
ref = A.astype(np.int32) @ B.astype(np.int32).T
rows = M - shift
ok = np.array_equal(got[:rows], ref[shift:shift + rows])
bad = int((got[:rows] != ref[shift:shift + rows]).any(axis=1).sum())
print(f"K={K} shift={shift} baseoff={os.environ.get('PL_DBG_BASEOFF', '0')} -> rows_ok={ok} bad_rows={bad}/{rows}", flush=True)
