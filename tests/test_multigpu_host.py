"""world_size-2 gloo tests (CPU) of the multi-GPU host logic: batch
sharding has no collective; the N-split GEMM gathers column blocks.  The local GEMM of
each rank is computed by the oracle here (checker only) -- on the GPU box the same
gather runs over NCCL with the tcgen05 GEMM (tests/test_gpu_multi.py)."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _nsplit():
    import importlib.util
    spec = importlib.util.spec_from_file_location("pl_nsplit", os.path.join(ROOT, "paddle-lite_b200", "nsplit.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def test_shard_batch_covers_everything_once():
    ns = _nsplit()
    for total in (0, 1, 7, 64, 256, 257):
        for world in (1, 2, 3, 4, 8):
            spans = [ns.shard_batch(total, world, r) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == total
            for a, b in zip(spans, spans[1:]):
                assert a[1] == b[0]
            sizes = [e - b for b, e in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        ns.shard_batch(10, 2, 2)


def test_shard_columns_aligned_and_complete():
    ns = _nsplit()
    for n in (1, 16, 1000, 1024, 4096, 16384, 33):
        for world in (1, 2, 4, 8):
            per = ns.block_width(n, world)
            assert per % 16 == 0
            spans = [ns.shard_columns(n, world, r) for r in range(world)]
            cover = []
            for r, (a, b) in enumerate(spans):
                assert a == min(r * per, n) and b - a <= per
                cover += list(range(a, b))
            assert cover == list(range(n))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, M, K, N, out_fp32, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import orc
        ns = _nsplit()
        rng = np.random.default_rng(7)  # same on every rank: A replicated, B sliced
        A = rng.integers(-127, 128, (M, K)).astype(np.int8)
        B = rng.integers(-127, 128, (K, N)).astype(np.int8)
        scale = (rng.random(N) * 1e-3 + 1e-4).astype(np.float32)
        bias = rng.standard_normal(N).astype(np.float32)
        out = orc.OUT_FP32 if out_fp32 else orc.OUT_INT8
        per = ns.block_width(N, world)
        n0, n1 = ns.shard_columns(N, world, rank)
        local = np.zeros((M, per), np.float32 if out_fp32 else np.int8)
        if n1 > n0:
            local[:, :n1 - n0] = orc.gemm(A, B[:, n0:n1], scale[n0:n1], bias[n0:n1], act=orc.ACT_RELU, out=out)
        full = ns.gather_columns(torch.from_numpy(local), N).numpy()
        want = orc.gemm(A, B, scale, bias, act=orc.ACT_RELU, out=out)
        q.put((rank, bool(np.array_equal(full, want)), full.shape))
        # batch sharding: per-rank results concatenate to the single-rank result, no collective
        b, e = ns.shard_batch(M, world, rank)
        part = orc.gemm(A[b:e], B, scale, bias, out=out) if e > b else np.zeros((0, N), want.dtype)
        whole = orc.gemm(A, B, scale, bias, out=out)
        q.put((rank, bool(np.array_equal(part, whole[b:e])), part.shape))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("M,K,N,out_fp32", [(5, 48, 40, False), (16, 64, 1000, True), (3, 32, 16, False)])
def test_nsplit_gather_world2_gloo(M, K, N, out_fp32):
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, M, K, N, out_fp32, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(180)
        assert p.exitcode == 0
    res = [q.get(timeout=10) for _ in range(2 * world)]
    assert all(ok for _, ok, _ in res), res


def test_bench_reference_arm_runs_every_net():
    """`bench.py --impl reference` (the CPU leg the driver times first) must fill every scale the
    program leaves to calibration -- including the fused-residual scale of ResNet-50."""
    import json
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for net in ("resnet50", "mobilenet_v1", "shufflenet_v2_like"):
        out = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--net", net,
                              "--steps", "1", "--warmup", "0"], capture_output=True, text=True, timeout=300)
        assert out.returncode == 0, out.stderr[-2000:]
        line = json.loads(out.stdout.strip().splitlines()[-1])
        assert line["impl"] == "reference" and line["value"] > 0
        assert line["cpu_baseline"]["kind"] in ("reference", "port")
        assert line["e2e"]["h2d_bytes_per_step"] == 0
