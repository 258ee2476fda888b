"""pl_gemm_i8_peers_static -- the allocation-free variant of pl_gemm_i8_peers (its peer tensor maps
live in a file-scope __device__ table inside the library, include/pl_int8.h) -- against the existing
entry point with caller scratch.  Both are called raw through the C-ABI (ctypes), on one GPU: the
"peers" are two more result buffers on the same device, which the TMA store warp reaches exactly like
NVLink-mapped ones.  Integer / byte work: the outputs must be identical bit for bit.

Collected last (file name) so that a failure here cannot hide the other GPU tests under `pytest -x`;
passes alone in a fresh process and after the rest of the suite.
"""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _case(seed, M, K, N):
    rng = np.random.default_rng(seed)
    ldk = (K + 15) // 16 * 16
    A = np.zeros((M, ldk), np.int8)
    A[:, :K] = rng.integers(-127, 128, (M, K))
    Bt = np.zeros((N, ldk), np.int8)
    Bt[:, :K] = rng.integers(-127, 128, (N, K))
    scale = (rng.random(N) * 2e-3 / np.sqrt(K) + 1e-5).astype(np.float32)
    bias = rng.standard_normal(N).astype(np.float32)
    return A, Bt, scale, bias


def _run(pl, entry, A, Bt, scale, bias, M, K, N, n_peers, maps=None):
    """One raw C-ABI call; returns [local result, peer results...] as numpy."""
    L = pl.lib()
    ldc = (N + 15) // 16 * 16
    dev = torch.device("cuda", 0)
    Ad, Btd = torch.from_numpy(A).to(dev), torch.from_numpy(Bt).to(dev)
    sd, bd = torch.from_numpy(scale).to(dev), torch.from_numpy(bias).to(dev)
    outs = [torch.full((M, ldc), 77, dtype=torch.int8, device=dev) for _ in range(1 + n_peers)]
    e = pl.make_epilogue(sd, bd, pl.ACT_RELU, 0.0, pl.OUT_INT8, pl.ROUND_ARM)
    peers = (C.c_void_p * max(n_peers, 1))(*[C.c_void_p(o.data_ptr()) for o in outs[1:]])
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    if entry == "static":
        rc = L.pl_gemm_i8_peers_static(C.c_void_p(Ad.data_ptr()), Ad.stride(0), C.c_void_p(Btd.data_ptr()),
                                       Btd.stride(0), C.c_void_p(outs[0].data_ptr()), peers, n_peers, ldc,
                                       M, N, K, C.byref(e), stream)
    else:
        rc = L.pl_gemm_i8_peers(C.c_void_p(Ad.data_ptr()), Ad.stride(0), C.c_void_p(Btd.data_ptr()),
                                Btd.stride(0), C.c_void_p(outs[0].data_ptr()), peers, n_peers,
                                C.c_void_p(maps.data_ptr()), ldc, M, N, K, C.byref(e), stream)
    assert rc == 0, f"{entry} entry point returned {rc}"
    torch.cuda.synchronize()
    return [o.cpu().numpy() for o in outs]


def test_static_peer_maps_match_the_scratch_entry_point_twice(pl):
    maps = torch.zeros(16 * 128, dtype=torch.uint8, device="cuda:0")  # caller scratch of the existing entry point
    # two calls in a row on different data and shapes: the shared table keeps the first call's maps,
    # the second call must overwrite every entry it uses
    for seed, (M, K, N, n_peers) in enumerate([(300, 520, 200, 2), (129, 96, 1000, 3)]):
        A, Bt, scale, bias = _case(seed, M, K, N)
        want = _run(pl, "scratch", A, Bt, scale, bias, M, K, N, n_peers, maps)
        got = _run(pl, "static", A, Bt, scale, bias, M, K, N, n_peers)
        assert len(got) == 1 + n_peers
        for w, g in zip(want, got):
            assert np.array_equal(w, g)  # bit-exact, pad columns (never written: still 77) included
        for g in got[1:]:
            assert np.array_equal(g[:, :N], got[0][:, :N])  # every peer holds the local tile data
        assert (got[0][:, :N] != 77).any()


def test_static_peer_maps_no_peers_equals_plain_gemm(pl):
    M, K, N = 64, 256, 48
    A, Bt, scale, bias = _case(5, M, K, N)
    got = _run(pl, "static", A, Bt, scale, bias, M, K, N, 0)[0]
    dev = torch.device("cuda", 0)
    want = pl.gemm(torch.from_numpy(A).to(dev), torch.from_numpy(Bt).to(dev), torch.from_numpy(scale).to(dev),
                   torch.from_numpy(bias).to(dev), act=pl.ACT_RELU, out=pl.OUT_INT8, N=N, K=K)
    torch.cuda.synchronize()
    assert np.array_equal(got[:, :N], want.cpu().numpy()[:, :N])


def test_static_peer_maps_limit_is_an_error(pl):
    L = pl.lib()
    M, K, N = 32, 64, 32
    A, Bt, scale, bias = _case(9, M, K, N)
    dev = torch.device("cuda", 0)
    Ad, Btd = torch.from_numpy(A).to(dev), torch.from_numpy(Bt).to(dev)
    sd, bd = torch.from_numpy(scale).to(dev), torch.from_numpy(bias).to(dev)
    y = torch.zeros((M, N), dtype=torch.int8, device=dev)
    e = pl.make_epilogue(sd, bd, pl.ACT_RELU, 0.0, pl.OUT_INT8, pl.ROUND_ARM)
    n_peers = 9  # PL_STATIC_PEERS_MAX is 8
    peers = (C.c_void_p * n_peers)(*[C.c_void_p(y.data_ptr())] * n_peers)
    rc = L.pl_gemm_i8_peers_static(C.c_void_p(Ad.data_ptr()), Ad.stride(0), C.c_void_p(Btd.data_ptr()),
                                   Btd.stride(0), C.c_void_p(y.data_ptr()), peers, n_peers, N, M, N, K,
                                   C.byref(e), None)
    assert rc == -6  # PL_ERR_LIMIT
    torch.cuda.synchronize()
    assert int(y.abs().sum()) == 0  # nothing was launched
