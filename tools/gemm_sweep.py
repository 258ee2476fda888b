"""BASELINE config 5: int8 fc/GEMM sweep.  Single GPU: python tools/gemm_sweep.py
N-split: torchrun --nproc-per-node G tools/gemm_sweep.py  (all-gather of output columns).
Prints one JSON line per shape: TOPS, GB/s, fraction of the measured int8 MMA peak."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402


def time_ms(fn, iters, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters


def main():
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    pkg = ge.load_package()
    pl, ns = pkg.capi, pkg.nsplit
    shapes = [(n, n, n) for n in (256, 512, 1024, 2048, 4096, 8192, 16384)]
    shapes += [(m, 4096, 4096) for m in (1, 16, 64)] + [(m, 16384, 16384) for m in (16, 64)]
    if len(sys.argv) > 1:
        shapes = [tuple(int(v) for v in a.split("x")) for a in sys.argv[1:]]
    peak = None
    if rank == 0:
        peak = round(pl.measure_i8_mma_peak(), 1)  # TOPS measured by the tcgen05 issue loop
    for out in (pl.OUT_INT8, pl.OUT_FP32):
        for M, N, K in shapes:
            g = torch.Generator().manual_seed(1234)
            B = torch.randint(-127, 128, (K, N), dtype=torch.int8, generator=g)
            scale = torch.full((N,), 1e-4)
            bias = torch.zeros(N)
            A = torch.randint(-127, 128, (M, K), dtype=torch.int8, generator=g).cuda()
            op = ns.NSplitGemm(pl, B.numpy(), scale.numpy(), bias.numpy(), out)
            iters = max(3, min(200, int(2e12 / (2.0 * M * N * K / world + 1))))
            ms_local = time_ms(lambda: op.local(A), iters)
            ms_full = time_ms(lambda: op(A), iters)
            ms_fused = 0.0
            if world > 1 and out == pl.OUT_INT8:
                op.fused_setup(M)
                ms_fused = time_ms(lambda: op.fused(A), iters)
            t = torch.tensor([ms_local, ms_full, ms_fused], device="cuda")
            if world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms_local, ms_full, ms_fused = t.tolist()
            if rank == 0:
                ops = 2.0 * M * N * K
                eb = 1 if out == pl.OUT_INT8 else 4
                byts = M * K + K * N + M * N * eb + 8 * N
                print(json.dumps({
                    "shape": [M, N, K], "out": "int8" if eb == 1 else "fp32", "n_gpus": world,
                    "ms_gemm": round(ms_local, 4), "ms_with_gather": round(ms_full, 4),
                    "tops_gemm": round(ops / ms_local / 1e9, 1), "tops_with_gather": round(ops / ms_full / 1e9, 1),
                    "ms_fused_peer_store": round(ms_fused, 4) if ms_fused else None,
                    "tops_fused_peer_store": round(ops / ms_fused / 1e9, 1) if ms_fused else None,
                    "gbs_gemm": round(byts / world / ms_local / 1e6, 1),
                    "frac_mma_peak": None if not peak else round(ops / world / ms_local / 1e9 / peak, 3),
                    "mma_peak_tops": peak}), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
