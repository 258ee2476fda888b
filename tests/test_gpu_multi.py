"""2-GPU NCCL tests (skipped on a 1-GPU box): N-split GEMM + all-gather of output
columns equals the single-GPU / oracle result bit for bit; batch-sharded conv equals
the unsharded one."""
import os
import socket
import sys

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, M, K, N, out_fp32, q):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        sys.path.insert(0, ROOT)
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import __graft_entry__ as ge
        import orc
        pkg = ge.load_package()
        pl, ns = pkg.capi, pkg.nsplit
        rng = np.random.default_rng(7)
        A = rng.integers(-127, 128, (M, K)).astype(np.int8)
        B = rng.integers(-127, 128, (K, N)).astype(np.int8)
        scale = (rng.random(N) * 1e-3 + 1e-4).astype(np.float32)
        bias = rng.standard_normal(N).astype(np.float32)
        out = orc.OUT_FP32 if out_fp32 else orc.OUT_INT8
        g = ns.NSplitGemm(pl, B, scale, bias, out, act=pl.ACT_RELU)
        ldk = g.ldk
        Ad = torch.zeros((M, ldk), dtype=torch.int8, device="cuda")
        Ad[:, :K] = torch.from_numpy(A).cuda()
        got = g(Ad).cpu().numpy()
        want = orc.gemm(A, B, scale, bias, act=orc.ACT_RELU, out=out)
        ok = np.array_equal(got, want) if not out_fp32 else np.allclose(got, want, rtol=1e-5, atol=1e-6)
        if not out_fp32:
            # fused GEMM + all-gather: tiles TMA-stored into the peers' results over NVLink
            for _ in range(2):
                fused = g.fused(Ad)
                torch.cuda.synchronize()
                ok = ok and np.array_equal(fused.cpu().numpy(), want)
                dist.barrier()
            # the same through the NVSwitch multicast address (multimem.st in the store warp), where the box has NVLS
            mc = "skipped"
            if g.mcast_supported():
                mc = True
                for _ in range(2):
                    fused = g.fused_mcast(Ad)
                    torch.cuda.synchronize()
                    mc = mc and np.array_equal(fused.cpu().numpy(), want)
                    dist.barrier()
                ok = ok and mc
            got = (got.shape, mc)
        elif N % 16 == 0:
            # fp32 results: 16-byte multimem.st straight from the epilogue registers
            g.fused_setup(M)
            if g.mcast_supported() and g.act in (pl.ACT_NONE, pl.ACT_RELU):
                for _ in range(2):
                    fused = g.fused_mcast(Ad)
                    torch.cuda.synchronize()
                    ok = ok and np.allclose(fused.cpu().numpy(), want, rtol=1e-5, atol=1e-6) and \
                        np.array_equal(fused.cpu().numpy(), got)
                    dist.barrier()
        q.put((rank, bool(ok), got if isinstance(got, tuple) else got.shape))
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
@pytest.mark.parametrize("M,K,N,out_fp32", [(64, 512, 1000, False), (256, 1024, 4096, True), (7, 100, 33, False),
                                            (1000, 2048, 1536, False)])
def test_nsplit_gemm_nccl(M, K, N, out_fp32):
    import torch.multiprocessing as mp
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, M, K, N, out_fp32, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(300)
        assert p.exitcode == 0
    res = [q.get(timeout=10) for _ in range(world)]
    assert all(ok for _, ok, _ in res), res
