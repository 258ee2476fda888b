#!/usr/bin/env python
"""bench.py -- int8 images/s of the B200 hot path on BASELINE.json's headline config:
ResNet-50 int8 224x224, batch 256 per GPU, batch-sharded over N GPUs (no collective on
the per-image path), synthetic random-init weights and inputs.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
                  [--net resnet50|mobilenet_v1|shufflenet_v2_like] [--batch B]

A "step" is one pass of the whole int8 network (calib+layout -> convs/pool/add -> fc)
over one batch.  `value` is device-resident throughput (inputs in HBM), `e2e` goes
through the public call with pinned HOST fp32 NCHW input (H2D of every batch and D2H
of the logits inside the timed region, double-buffered).  `--impl reference` times the
reference's own CPU code for the path (lite/tests/utils/naive_math_impl.h conv_basic /
basic_gemm compiled into oracle/_ref, else the oracle port) on the host cores.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

METRIC = "int8 images/s"
UNIT = "img/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--net", default="resnet50", choices=["resnet50", "mobilenet_v1", "shufflenet_v2_like"])
    ap.add_argument("--batch", type=int, default=None, help="images per GPU (default 256; mobilenet 64)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-seconds", type=float, default=15.0)
    ap.add_argument("--breakdown", default=None, help="write a per-layer timing table to this file")
    return ap.parse_args()


def workload_name(net, batch):
    return f"{net} int8 224x224 batch {batch}/GPU (random-init weights, synthetic input)"


# ----------------------------------------------------------------------------
# clocks
# ----------------------------------------------------------------------------
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1]))
                mx = float(f[2])
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


# ----------------------------------------------------------------------------
# CPU legs (the only places that execute oracle/)
# ----------------------------------------------------------------------------
def cpu_reference_rate(net_name, seconds, steps=1, warmup=0):
    """Times the reference's CPU implementation of the path on a bounded sample: whole
    images of the same program at 224x224, one image per pass, all host threads."""
    import numpy as np
    import torch

    import __graft_entry__ as ge
    import net_oracle
    import orc

    ge.load_package()
    import importlib
    nets = importlib.import_module("paddle_lite_b200.nets")
    prog = getattr(nets, net_name)(seed=1234)
    # scales: fixed synthetic values (timing does not depend on them)
    s = 0.02
    for op in prog.ops:
        for k in ("in_scale", "out_scale", "a_scale", "b_scale", "res_scale"):
            if k in op and op[k] is None and (k != "res_scale" or op.get("res")):
                op[k] = s
    kind = "reference" if orc.ref() is not None else "port"
    engine = "reference" if kind == "reference" else "oracle"
    x = torch.rand((1, 3, 224, 224), generator=torch.Generator().manual_seed(1)).numpy()
    times = []
    for _ in range(warmup):
        net_oracle.run_program(prog, x, engine=engine)
    t_all = time.perf_counter()
    n = 0
    while n < steps or (steps == 0 and time.perf_counter() - t_all < seconds):
        t0 = time.perf_counter()
        net_oracle.run_program(prog, x, engine=engine)
        times.append(time.perf_counter() - t0)
        n += 1
        if steps == 0 and n >= 64:
            break
    mean = sum(times) / len(times)
    return {"value": 1.0 / mean, "unit": UNIT, "cores": orc.lib().orc_num_threads(), "kind": kind,
            "sample": f"{len(times)} pass(es) of 1 image through the whole {net_name} program "
                      f"(all convs/fc via {'conv_basic<int8_t,int>/basic_gemm from naive_math_impl.h' if kind == 'reference' else 'the oracle port'}"
                      f", OpenMP over all host threads), mean {mean * 1e3:.1f} ms/image"}, mean


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    batch = args.batch or (256 if args.net == "resnet50" else 64)
    cb, mean = cpu_reference_rate(args.net, args.cpu_seconds, steps=max(1, args.steps), warmup=max(0, min(args.warmup, 1)))
    line = {"metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": mean * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "int8", "data": "synthetic", "impl": "reference",
            "config": {"workload": workload_name(args.net, batch), "cpu_sample_batch": 1},
            "cpu_baseline": cb,
            "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))
    return 0


# ----------------------------------------------------------------------------
# B200 arm
# ----------------------------------------------------------------------------
def run_b200(args):
    import torch
    import torch.distributed as dist

    import __graft_entry__ as ge

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py --impl b200 needs a GPU (there is no CPU fallback)"
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    pkg = ge.load_package()
    pl = pkg.capi
    import importlib
    nets = importlib.import_module("paddle_lite_b200.nets")

    batch = args.batch or (256 if args.net == "resnet50" else 64)
    prog = getattr(nets, args.net)(seed=1234)
    gen = torch.Generator(device="cpu").manual_seed(1234 + rank)
    x_host = torch.rand((batch, 3, 224, 224), generator=gen, dtype=torch.float32).pin_memory()
    net = nets.DeviceNet(prog, batch, device=f"cuda:{local}")
    net.calibrate(x_host[:4].cuda())
    net.bind()
    net.input.copy_(x_host)
    net.capture()
    launches_per_step = len(net.layers)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident throughput ------------------------------------------------
    for _ in range(max(args.warmup, 3)):
        net.replay()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    n0 = pl.launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record()
    for _ in range(args.steps):
        net.replay()
    ev1.record()
    barrier()
    ms = ev0.elapsed_time(ev1)
    clocks = sampler.stop() if rank == 0 else None
    t = torch.tensor([ms], dtype=torch.float64, device=f"cuda:{local}")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_max = float(t.item())
    ms_per_step = ms_max / args.steps
    value = world * batch * args.steps / (ms_max * 1e-3)

    # ---- end to end: pinned host fp32 in, logits out, double-buffered --------------
    copy_stream = torch.cuda.Stream()
    staging = [torch.empty_like(net.input) for _ in range(2)]
    out_host = [torch.empty((batch, 1000), dtype=torch.float32).pin_memory() for _ in range(2)]
    ready = [torch.cuda.Event() for _ in range(2)]
    consumed = [torch.cuda.Event() for _ in range(2)]
    main = torch.cuda.current_stream()

    def e2e_steps(k):
        for i in range(k):
            b = i & 1
            with torch.cuda.stream(copy_stream):
                copy_stream.wait_event(consumed[b])
                staging[b].copy_(x_host, non_blocking=True)
                ready[b].record(copy_stream)
            main.wait_event(ready[b])
            net.input.copy_(staging[b], non_blocking=True)
            consumed[b].record(main)
            net.replay()
            out_host[b].copy_(net.output, non_blocking=True)

    for b in range(2):
        consumed[b].record(main)
    e2e_steps(max(3, min(args.warmup, 5)))
    barrier()
    t0 = time.perf_counter()
    ev0.record()
    e2e_steps(args.steps)
    ev1.record()
    barrier()
    e2e_ms = ev0.elapsed_time(ev1)
    t = torch.tensor([e2e_ms], dtype=torch.float64, device=f"cuda:{local}")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = world * batch * args.steps / (float(t.item()) * 1e-3)
    h2d = x_host.numel() * 4
    d2h = batch * 1000 * 4

    # ---- roofline of the dominant kernel (igemm_i8_kernel), timed per launch ----------
    roof, table = layer_roofline(net, prog, batch, nets, torch)
    if args.breakdown and rank == 0:
        with open(args.breakdown, "w") as f:
            json.dump(table, f, indent=1)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0
    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        cpu, _ = cpu_reference_rate(args.net, args.cpu_seconds, steps=0)
    macs = nets.program_macs(prog)
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "int8", "data": "synthetic",
            "config": {"workload": workload_name(args.net, batch), "global_batch": batch * world,
                       "parallelism": f"batch-sharded x{world}, no collective", "cuda_graph": True,
                       "l2": f"per-step working set ({(nets.program_bytes(prog)[0] * batch + nets.program_bytes(prog)[1]) / 1e6:.0f} MB of activations + weights) "
                             "exceeds the 126 MB L2; consecutive steps replay the whole network, no flush needed",
                       "gmac_per_image": sum(macs.values()) / 1e9},
            "tops": 2 * sum(macs.values()) * batch * world * args.steps / (ms_max * 1e-3) / 1e12,
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "input": "pinned host fp32 NCHW, copy overlapped with the previous step's compute"},
            "gpu_launches": launches_per_step * args.steps,
            "roofline": roof, "cpu_baseline": cpu}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


def layer_roofline(net, prog, batch, nets, torch):
    """Times every launch of one step with CUDA events (eager, same stream), then
    aggregates the tensor-core kernel: algorithmic ops and bytes / its summed time."""
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    hbm_peak, bf16_peak, src = 6650.0, 1590.0, "fallback"
    if os.path.exists(peaks_path):
        pk = json.load(open(peaks_path))
        hbm_peak, bf16_peak, src = pk["hbm_gbs"], pk.get("bf16_tflops_sustained", pk["bf16_tflops"]), "measured"
    reps = 5
    table = []
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in net.layers]
    acc = [0.0] * len(net.layers)
    for r in range(reps + 1):
        for i, (op, fn) in enumerate(net.layers):
            evs[i][0].record()
            fn()
            evs[i][1].record()
        torch.cuda.synchronize()
        if r > 0:
            for i in range(len(net.layers)):
                acc[i] += evs[i][0].elapsed_time(evs[i][1])
    tc_ms = tc_ops = tc_bytes = 0.0
    by_kind = {}
    for i, (op, _) in enumerate(net.layers):
        ms = acc[i] / reps
        H, W, C = prog.shapes[op.src]
        P, Q, K = prog.shapes[op.dst]
        ops = byts = 0
        if op.kind == "conv":
            ops = 2 * P * Q * K * op.k * op.k * op.C * batch
            byts = batch * (H * W * C + P * Q * K * (4 if op.out_fp32 else 1)) + op.w.size + 8 * K
            if op.get("res"):
                byts += batch * P * Q * K  # fused residual read
        elif op.kind == "dwconv":
            ops = 2 * P * Q * K * op.k * op.k * batch
            byts = batch * (H * W * C + P * Q * K) + op.w.size + 8 * K
        elif op.kind == "fc":
            ops = 2 * op.K * op.N * batch
            byts = batch * (op.K + op.N * 4) + op.w.size + 8 * op.N
        elif op.kind == "pool":
            byts = batch * (H * W * C + P * Q * K)
        elif op.kind == "add":
            byts = batch * 3 * H * W * C
        elif op.kind == "cat_shuffle":
            byts = batch * 2 * P * Q * K
        elif op.kind == "quant_in":
            byts = batch * (H * W * C * 4 + P * Q * 16)
        table.append({"i": i, "kind": op.kind, "in": [H, W, C], "out": [P, Q, K], "k": op.get("k"),
                      "stride": op.get("stride"), "ms": ms, "gops": ops / 1e9, "mbytes": byts / 1e6,
                      "tops": ops / ms / 1e9 if ms else None, "gbs": byts / ms / 1e6 if ms else None})
        d = by_kind.setdefault(op.kind, [0.0, 0.0, 0.0])
        d[0] += ms
        d[1] += ops
        d[2] += byts
        if op.kind in ("conv", "fc"):
            tc_ms += ms
            tc_ops += ops
            tc_bytes += byts
    total_ms = sum(acc) / reps
    i8_peak = 2.0 * bf16_peak  # dense int8 = 2x bf16 on the same tensor pipe
    t_tensor = tc_ops / (i8_peak * 1e12)
    t_hbm = tc_bytes / (hbm_peak * 1e9)
    n_tc = sum(1 for op, _ in net.layers if op.kind in ("conv", "fc"))
    if t_hbm >= t_tensor:
        achieved = tc_bytes / (tc_ms * 1e-3) / 1e9
        roof = {"bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s", "frac": achieved / hbm_peak}
    else:
        achieved = tc_ops / (tc_ms * 1e-3) / 1e12
        roof = {"bound": "tensor", "achieved": achieved, "peak": i8_peak, "unit": "TOP/s", "frac": achieved / i8_peak}
    # measured DRAM bytes per launch of this kernel: ncu dram__bytes_read.sum + dram__bytes_write.sum of
    # the same command (tools/gpu_profile_round.sh -> profiles/rNN_traffic.json), when captured
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        tj = json.load(open(tpath)).get(f"{prog.name}_b{batch}")
        if tj:
            traffic = tj["igemm_dram_bytes_per_launch"]
    try:
        probe = round(nets.pl.measure_i8_mma_peak(), 1)  # back-to-back tcgen05.mma.kind::i8, all SMs
    except Exception:
        probe = None
    roof.update({"tensor_peak_tops_probe": probe,
                 "tensor_frac_of_probe": (tc_ops / (tc_ms * 1e-3) / 1e12 / probe) if probe else None,
                 "traffic": traffic, "algorithmic_bytes_per_launch": tc_bytes / n_tc, "kernel": "igemm_i8_kernel", "launches_per_step": n_tc,
                 "avg_launch_ms": tc_ms / n_tc, "share_of_step": tc_ms / total_ms, "peak_source": src,
                 "tensor_tops": tc_ops / (tc_ms * 1e-3) / 1e12, "tensor_peak_tops_2x_bf16": i8_peak,
                 "hbm_gbs": tc_bytes / (tc_ms * 1e-3) / 1e9,
                 "by_kind_ms": {k: v[0] for k, v in by_kind.items()},
                 "by_kind_gbs": {k: (v[2] / v[0] / 1e6 if v[0] else None) for k, v in by_kind.items()}})
    return roof, table


def main():
    args = parse()
    if args.impl == "reference":
        return run_reference(args)
    return run_b200(args)


if __name__ == "__main__":
    sys.exit(main())
