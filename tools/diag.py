"""Diagnostic runner for GPU trips: runs a ladder of small cases through the C-ABI,
each in this one process, and prints WHERE results differ from the oracle (not just
that they do).  Not a test; writes to stdout."""
import os
import sys
import traceback

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as ge  # noqa: E402
import orc  # noqa: E402

pl = ge.load_package().capi
DEV = "cuda:0"


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).to(DEV)


def report(name, got, want):
    if got.shape != want.shape:
        print(f"[{name}] SHAPE {got.shape} vs {want.shape}")
        return False
    bad = got != want
    nb = int(bad.sum())
    if nb == 0:
        print(f"[{name}] OK ({want.size} elems)")
        return True
    idx = np.argwhere(bad)
    print(f"[{name}] MISMATCH {nb}/{want.size}; first idx {idx[:6].tolist()}")
    flat = bad.reshape(-1, bad.shape[-1])
    rows = np.where(flat.any(1))[0]
    cols = np.where(flat.any(0))[0]
    print(f"   bad rows: n={len(rows)} first={rows[:12].tolist()} ; rows%8 hist={np.bincount(rows % 8, minlength=8).tolist()}")
    print(f"   bad cols: n={len(cols)} first={cols[:12].tolist()} ; cols%16 hist={np.bincount(cols % 16, minlength=16).tolist()}")
    g = got.reshape(-1, got.shape[-1])
    w = want.reshape(-1, want.shape[-1])
    r0 = rows[0]
    print(f"   row {r0}: got {g[r0, :8].tolist()} want {w[r0, :8].tolist()}")
    return False


def gemm_case(M, N, K, seed=0, ones=False):
    rng = np.random.default_rng(seed)
    if ones:
        A = np.ones((M, K), np.int8)
        Bt = np.ones((N, K), np.int8)
    else:
        A = rng.integers(-127, 128, (M, K), dtype=np.int8)
        Bt = rng.integers(-127, 128, (N, K), dtype=np.int8)
    got = pl.gemm(dev(A), dev(Bt), None, None, out=pl.OUT_INT32)
    torch.cuda.synchronize()
    return report(f"gemm {M}x{N}x{K}{' ones' if ones else ''}", got.cpu().numpy(),
                  A.astype(np.int32) @ Bt.astype(np.int32).T)


def conv_case(N, H, W, Cc, K, R, S, st, pads, dl, seed=0):
    rng = np.random.default_rng(seed)
    x = rng.integers(-127, 128, (N, H, W, Cc), dtype=np.int8)
    w = rng.integers(-127, 128, (K, R, S, Cc), dtype=np.int8)
    want = orc.conv2d(x, w, None, None, (st, st), pads, (dl, dl), 1, out=orc.OUT_INT32)
    ld = (Cc + 15) // 16 * 16
    xp = np.zeros((N, H, W, ld), np.int8)
    xp[..., :Cc] = x
    wp = pl.pack_conv_weight(dev(w))
    got = pl.conv2d(dev(xp), wp, K, R, S, None, None, (st, st), pads, (dl, dl), out=pl.OUT_INT32, Cc=Cc)
    torch.cuda.synchronize()
    return report(f"conv N{N} {H}x{W}x{Cc}->{K} k{R}x{S} s{st} p{pads} d{dl}", got.cpu().numpy(), want)


def main():
    print(torch.cuda.get_device_name(0), pl.lib().pl_version())
    steps = [
        lambda: gemm_case(128, 16, 32, ones=True),
        lambda: gemm_case(128, 16, 32),
        lambda: gemm_case(128, 64, 64),
        lambda: gemm_case(128, 64, 128),
        lambda: gemm_case(128, 256, 128),
        lambda: gemm_case(128, 64, 512),
        lambda: gemm_case(256, 64, 2048),
        lambda: gemm_case(100, 24, 48),
        lambda: gemm_case(5000, 512, 1024),
        lambda: gemm_case(40000, 64, 64),
        lambda: conv_case(1, 8, 16, 32, 16, 1, 1, 1, (0, 0, 0, 0), 1),
        lambda: conv_case(1, 8, 16, 32, 16, 1, 1, 2, (0, 0, 0, 0), 1),
        lambda: conv_case(1, 8, 16, 32, 16, 3, 3, 1, (0, 0, 0, 0), 1),
        lambda: conv_case(1, 8, 16, 32, 16, 3, 3, 1, (1, 1, 1, 1), 1),
        lambda: conv_case(2, 14, 14, 64, 64, 3, 3, 1, (1, 1, 1, 1), 1),
        lambda: conv_case(3, 9, 7, 32, 48, 3, 3, 2, (1, 1, 1, 1), 1),
        lambda: conv_case(1, 17, 17, 32, 32, 3, 3, 1, (2, 2, 2, 2), 2),
        lambda: conv_case(1, 14, 14, 256, 512, 3, 3, 1, (1, 1, 1, 1), 1),
        lambda: conv_case(2, 32, 32, 3, 32, 3, 3, 2, (1, 1, 1, 1), 1),
        lambda: conv_case(1, 56, 56, 64, 64, 3, 3, 1, (1, 1, 1, 1), 1),
    ]
    for i, s in enumerate(steps):
        try:
            s()
        except Exception as e:  # keep going unless the context died
            print(f"step {i} raised {type(e).__name__}: {e}")
            try:
                torch.cuda.synchronize()
            except Exception as e2:
                print("context is dead:", e2)
                traceback.print_exc()
                return 1
    return 0


if __name__ == "__main__":
    sys.exit(main())
