#!/usr/bin/env python
"""bench.py -- int8 images/s of the B200 hot path on BASELINE.json's headline config:
ResNet-50 int8 224x224, batch 256 per GPU, batch-sharded over N GPUs (no collective on
the per-image path), synthetic random-init weights and inputs.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
                  [--net resnet50|mobilenet_v1|shufflenet_v2_like|humanseg] [--batch B] [--no-extras]

A "step" is one pass of the whole int8 network (calib+layout -> convs/pool/add -> fc)
over one batch.  `value` is device-resident throughput (inputs in HBM), `e2e` goes
through the public call with pinned HOST fp32 NCHW input (H2D of every batch and D2H
of the logits inside the timed region, double-buffered).  `--impl reference` times the
reference's own CPU code for the path (lite/tests/utils/naive_math_impl.h conv_basic /
basic_gemm compiled into oracle/_ref, else the oracle port) on the host cores.

The ONE JSON line also carries the other BASELINE.json configs as sub-records, measured in
the same process after the headline (each on every rank, max over ranks):
  nets.mobilenet_v1_b64 / nets.mobilenet_v1_b1   config 2
  nets.humanseg_b32                              config 3 (the shipped model, when its files travel)
  conv1                                          config 1 (latency + batched TOP/s)
  gemm                                           config 5 single-GPU sweep
  nsplit (WORLD_SIZE > 1)                        config 5 N-split: local / NCCL gather / fused peer store,
                                                 checked bit for bit against the single-GPU GEMM in-run
  sustained                                      >= 2 s of back-to-back replays of the headline, own clocks
  e2e_f16 / e2e_u8                               the same end-to-end loop with narrower host feeds
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
NETS = ["resnet50", "mobilenet_v1", "shufflenet_v2_like", "humanseg"]
IMAGENET_MEAN = [123.675, 116.28, 103.53]
IMAGENET_NORM = [1.0 / 58.395, 1.0 / 57.12, 1.0 / 57.375]


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--net", default="resnet50", choices=NETS)
    ap.add_argument("--batch", type=int, default=None, help="images per GPU (default 256; mobilenet 64; humanseg 32)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="headline only: skip the sub-records of the other configs")
    ap.add_argument("--cpu-seconds", type=float, default=15.0)
    ap.add_argument("--breakdown", default=None, help="write a per-layer timing table to this file")
    return ap.parse_args()


def default_batch(net):
    return {"resnet50": 256, "humanseg": 32}.get(net, 64)


def workload_name(net, batch, res=224):
    return f"{net} int8 {res}x{res} batch {batch}/GPU (random-init weights, synthetic input)"


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
        return self

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons, watts = [], None, set(), []
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1]))
                mx = float(f[2])
                watts.append(float(f[3]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm), "power_w_max": max(watts) if watts else None}


# ----------------------------------------------------------------------------
# CPU legs (the only places that execute oracle/)
# ----------------------------------------------------------------------------
def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def cpu_reference_rate(net_name, seconds, steps=1, warmup=0):
    """Times the reference's CPU implementation of the path on a bounded sample: whole
    images of the same program at 224x224, one image per pass, all host threads."""
    import torch

    import __graft_entry__ as ge
    import net_oracle
    import orc

    ge.load_package()
    import importlib
    nets = importlib.import_module("paddle_lite_b200.nets")
    # torchrun exports OMP_NUM_THREADS=1 to every rank: the CPU arm asks for every host core explicitly
    orc.set_num_threads(host_cores())
    if net_name == "humanseg":
        net_name = "shufflenet_v2_like"  # the CPU arm walks the synthetic program of the same family
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
    cores = orc.ref().ref_num_threads() if kind == "reference" else orc.lib().orc_num_threads()
    return {"value": 1.0 / mean, "unit": UNIT, "cores": cores, "kind": kind,
            "sample": f"{len(times)} pass(es) of 1 image through the whole {net_name} program "
                      f"(all convs/fc via {'conv_basic<int8_t,int>/basic_gemm from naive_math_impl.h' if kind == 'reference' else 'the oracle port'}"
                      f", OpenMP over all host threads), mean {mean * 1e3:.1f} ms/image"}, mean


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    batch = args.batch or default_batch(args.net)
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
def bind_to_gpu_numa_node(local_rank):
    """Pin this rank (and so its pinned host buffers, first touched after this call) to the CPUs of the NUMA node its GPU
    hangs off: with 8 ranks feeding 8 GPUs the H2D copies then cross no socket link.  Best effort: returns the node
    or None when the topology files are not there."""
    try:
        import torch
        bus = torch.cuda.get_device_properties(local_rank).pci_bus_id
        dom = torch.cuda.get_device_properties(local_rank).pci_domain_id
        dev = torch.cuda.get_device_properties(local_rank).pci_device_id
        path = f"/sys/bus/pci/devices/{dom:04x}:{bus:02x}:{dev:02x}.0/numa_node"
        node = int(open(path).read().strip())
        if node < 0:
            return None
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            return node
    except Exception:
        pass
    return None


class Ctx:
    """Process-wide plumbing: rank, device, barrier, max-over-ranks."""

    def __init__(self):
        import torch
        import torch.distributed as dist

        import __graft_entry__ as ge

        self.torch, self.dist = torch, dist
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        assert torch.cuda.is_available(), "bench.py --impl b200 needs a GPU (there is no CPU fallback)"
        torch.cuda.set_device(self.local)
        self.dev = f"cuda:{self.local}"
        self.numa = bind_to_gpu_numa_node(self.local) if self.world > 1 else None
        if self.world > 1:
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local))
        self.pkg = ge.load_package()
        self.pl = self.pkg.capi
        import importlib
        self.nets = importlib.import_module("paddle_lite_b200.nets")

    def barrier(self):
        self.torch.cuda.synchronize()
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, *vals):
        t = self.torch.tensor(list(vals), dtype=self.torch.float64, device=self.dev)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return [float(v) for v in t.tolist()]

    def min_over_ranks(self, *vals):
        t = self.torch.tensor(list(vals), dtype=self.torch.float64, device=self.dev)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MIN)
        return [float(v) for v in t.tolist()]

    def timed(self, fn, iters):
        """CUDA events on the current stream around `iters` calls, barrier + sync both sides, max over ranks."""
        torch = self.torch
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        self.barrier()
        ev0.record()
        for _ in range(iters):
            fn()
        ev1.record()
        self.barrier()
        return self.max_over_ranks(ev0.elapsed_time(ev1))[0]

    def close(self):
        if self.world > 1:
            self.dist.destroy_process_group()


def make_program(ctx, net_name):
    if net_name == "humanseg":
        model = importlib_model(ctx)
        return model.humanseg_program()
    return getattr(ctx.nets, net_name)(seed=1234)


def importlib_model(ctx):
    import importlib
    return importlib.import_module("paddle_lite_b200.model")


def host_input(ctx, prog, batch, feed):
    torch = ctx.torch
    H, W, C = prog.shapes["input"]
    gen = torch.Generator(device="cpu").manual_seed(1234 + ctx.rank)
    if feed == "u8":
        return torch.randint(0, 256, (batch, C, H, W), generator=gen, dtype=torch.uint8).pin_memory()
    x = torch.rand((batch, C, H, W), generator=gen, dtype=torch.float32)
    if feed == "f16":
        x = x.half()
    return x.pin_memory()


def build_net(ctx, net_name, batch, feed="f32"):
    prog = make_program(ctx, net_name)
    kw = {}
    if feed == "u8":
        kw = {"mean": IMAGENET_MEAN, "norm": IMAGENET_NORM}
    x_host = host_input(ctx, prog, batch, feed)
    if net_name == "humanseg":
        net = importlib_model(ctx).GraphNet(prog, batch, device=ctx.dev, feed=feed, **kw)
    else:
        net = ctx.nets.DeviceNet(prog, batch, device=ctx.dev, feed=feed, **kw)
    if not net.calibrated:
        net.calibrate(x_host[:min(4, batch)].to(ctx.dev))
    net.bind()
    net.input.copy_(x_host)
    net.capture_inputs(2)
    return prog, net, x_host


def e2e_loop(ctx, net, x_host, steps, warmup):
    """Public call with HOST buffers: every step copies its batch from pinned host memory straight into one of
    the network's two input buffers on a copy stream (overlapping the previous step's compute), replays the
    graph that reads that buffer, and copies the result back to pinned host memory."""
    torch = ctx.torch
    copy_stream = torch.cuda.Stream()
    out_host = [torch.empty(tuple(net.output.shape), dtype=net.output.dtype).pin_memory() for _ in range(2)]
    ready = [torch.cuda.Event() for _ in range(2)]
    consumed = [torch.cuda.Event() for _ in range(2)]
    main = torch.cuda.current_stream()

    def run(k):
        for i in range(k):
            b = i & 1
            with torch.cuda.stream(copy_stream):
                copy_stream.wait_event(consumed[b])
                net.inputs[b].copy_(x_host, non_blocking=True)
                ready[b].record(copy_stream)
            main.wait_event(ready[b])
            net.replay(b)
            consumed[b].record(main)
            out_host[b].copy_(net.output, non_blocking=True)

    for b in range(2):
        consumed[b].record(main)
    run(max(3, min(warmup, 5)))
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ctx.barrier()
    ev0.record()
    run(steps)
    ev1.record()
    ctx.barrier()
    ms = ctx.max_over_ranks(ev0.elapsed_time(ev1))[0]
    return {"value": ctx.world * net.batch * steps / (ms * 1e-3), "unit": UNIT,
            "h2d_bytes_per_step": x_host.numel() * x_host.element_size(),
            "d2h_bytes_per_step": net.output.numel() * net.output.element_size(),
            "input": f"pinned host {x_host.dtype} NCHW copied straight into the network's double-buffered input "
                     "(no staging copy), overlapped with the previous step's compute"}


def measure_net(ctx, net_name, batch, steps, warmup, clocks=False, e2e=True, profile=True, feed="f32"):
    """Device-resident rate (+ e2e, + per-kind breakdown from kernel records inside the graph replays)."""
    torch, nets = ctx.torch, ctx.nets
    prog, net, x_host = build_net(ctx, net_name, batch, feed)
    for _ in range(max(warmup, 3)):
        net.replay()
    ctx.barrier()
    sampler = ClockSampler(ctx.local).start() if (clocks and ctx.rank == 0) else None
    ms = ctx.timed(net.replay, steps)
    ck = sampler.stop() if sampler else None
    macs = nets.program_macs(prog)
    rec = {"ms_per_step": ms / steps, "img_s": ctx.world * batch * steps / (ms * 1e-3), "batch_per_gpu": batch,
           "launches_per_step": len(launch_order(net)),
           "tops": 2 * sum(macs.values()) * batch * ctx.world * steps / (ms * 1e-3) / 1e12}
    if ck is not None:
        rec["clocks"] = ck
    if e2e:
        rec["e2e"] = e2e_loop(ctx, net, x_host, steps, warmup)
    table = None
    if profile:
        table = layer_times(ctx, net, prog, batch, rec["ms_per_step"])
        rec["by_kind"] = by_kind(table)
    return rec, prog, net, x_host, table


# ---- per-launch times from inside the graph replay --------------------------------------------
def kernel_records(ctx, fn, reps):
    """CUPTI kernel activity records (torch.profiler) of `reps` calls of fn: [(name, start_us, end_us)] in
    start order.  None when the profiler is unavailable."""
    torch = ctx.torch
    try:
        from torch.profiler import ProfilerActivity, profile
        fn()
        torch.cuda.synchronize()
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            for _ in range(reps):
                fn()
            torch.cuda.synchronize()
        out = []
        for e in prof.events():
            if str(e.device_type).endswith("CUDA") and e.time_range is not None:
                nm = e.name
                if nm.startswith("Memcpy") or nm.startswith("Memset"):
                    continue
                out.append((nm, float(e.time_range.start), float(e.time_range.end)))
        out.sort(key=lambda r: r[1])
        return out or None
    except Exception as exc:  # pragma: no cover - depends on the box
        sys.stderr.write(f"[bench] kernel records unavailable: {exc!r}\n")
        return None


def layer_bytes_ops(op, prog, batch):
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
    elif "bytes" in op:
        byts, ops = op.bytes * batch, op.get("ops", 0) * batch
    return ops, byts


def launch_order(net):
    """layer index of every launch of one step (a sub-batched schedule launches a layer once per chunk)"""
    return net.launch_order() if hasattr(net, "launch_order") else list(range(len(net.layers)))


def layer_times(ctx, net, prog, batch, ms_per_step):
    """Per-launch time INSIDE the graph replay: exclusive time of every kernel record,
    end_i - max(start_i, end_{i-1}) (programmatic dependent launch lets kernel i+1 start while kernel i
    drains, so plain durations overlap), averaged over the replays and scaled so that the launches of one
    step sum to the CUDA-event step time.  Fallback without CUPTI: eager per-launch events, scaled the same way."""
    torch = ctx.torch
    L = len(net.layers)
    reps = 5
    per = None
    src = "cupti kernel records of graph replays (exclusive time), scaled to the CUDA-event step time"
    recs = kernel_records(ctx, net.replay, reps)
    order = launch_order(net)
    if recs is not None and len(recs) == len(order) * reps:
        per = [0.0] * L
        for r in range(reps):
            prev_end = None
            for j, i in enumerate(order):
                _, st, en = recs[r * len(order) + j]
                per[i] += (en - (st if prev_end is None else max(st, prev_end))) / 1e3
                prev_end = en if prev_end is None else max(prev_end, en)
        per = [v / reps for v in per]
    if per is None:
        src = "eager per-launch CUDA events, scaled to the CUDA-event step time"
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in net.layers]
        per = [0.0] * L
        for r in range(reps + 1):
            for i, (op, fn) in enumerate(net.layers):
                evs[i][0].record()
                fn()
                evs[i][1].record()
            torch.cuda.synchronize()
            if r > 0:
                for i in range(L):
                    per[i] += evs[i][0].elapsed_time(evs[i][1]) / reps
    raw_total = sum(per)
    k = ms_per_step / raw_total if raw_total > 0 else 1.0
    table = []
    for i, (op, _) in enumerate(net.layers):
        ops, byts = layer_bytes_ops(op, prog, batch)
        ms = per[i] * k
        H, W, C = prog.shapes[op.src]
        P, Q, K = prog.shapes[op.dst]
        table.append({"i": i, "kind": op.kind, "in": [H, W, C], "out": [P, Q, K], "k": op.get("k"),
                      "stride": op.get("stride"), "ms": ms, "gops": ops / 1e9, "mbytes": byts / 1e6,
                      "tops": ops / ms / 1e9 if ms else None, "gbs": byts / ms / 1e6 if ms else None})
    table.append({"timing": src, "raw_sum_ms": raw_total, "scale": k})
    return table


def peaks():
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    hbm_peak, bf16_peak, src = 6650.0, 1590.0, "fallback"
    if os.path.exists(peaks_path):
        pk = json.load(open(peaks_path))
        hbm_peak, bf16_peak, src = pk["hbm_gbs"], pk.get("bf16_tflops_sustained", pk["bf16_tflops"]), "measured"
    return hbm_peak, bf16_peak, src


def by_kind(table):
    hbm_peak, _, _ = peaks()
    agg = {}
    for row in table:
        if "kind" not in row:
            continue
        d = agg.setdefault(row["kind"], {"ms": 0.0, "gops": 0.0, "mbytes": 0.0, "launches": 0})
        d["ms"] += row["ms"]
        d["gops"] += row["gops"]
        d["mbytes"] += row["mbytes"]
        d["launches"] += 1
    for d in agg.values():
        d["gbs"] = d["mbytes"] / d["ms"] if d["ms"] else None  # MB / ms == GB/s
        d["hbm_frac"] = d["gbs"] / hbm_peak if d["gbs"] else None
        d["tops"] = d["gops"] / d["ms"] if d["ms"] else None  # GOP / ms == TOP/s
    return agg


def headline_roofline(ctx, table, prog, batch, ms_per_step):
    """Roofline of the dominant kernel (igemm_i8_kernel: every conv + the fc)."""
    hbm_peak, bf16_peak, src = peaks()
    rows = [r for r in table if r.get("kind") in ("conv", "fc")]
    tc_ms = sum(r["ms"] for r in rows)
    tc_ops = sum(r["gops"] for r in rows) * 1e9
    tc_bytes = sum(r["mbytes"] for r in rows) * 1e6
    total_ms = sum(r["ms"] for r in table if "kind" in r)
    n_tc = len(rows)
    i8_peak = 2.0 * bf16_peak  # dense int8 = 2x bf16 on the same tensor pipe
    t_tensor = tc_ops / (i8_peak * 1e12)
    t_hbm = tc_bytes / (hbm_peak * 1e9)
    if t_hbm >= t_tensor:
        achieved = tc_bytes / (tc_ms * 1e-3) / 1e9
        roof = {"bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s", "frac": achieved / hbm_peak}
    else:
        achieved = tc_ops / (tc_ms * 1e-3) / 1e12
        roof = {"bound": "tensor", "achieved": achieved, "peak": i8_peak, "unit": "TOP/s", "frac": achieved / i8_peak}
    traffic = None
    for name in ("traffic_r02.json", "traffic.json"):
        tpath = os.path.join(ROOT, "profiles", name)
        if os.path.exists(tpath):
            tj = json.load(open(tpath)).get(f"{prog.name}_b{batch}")
            if tj:
                traffic = tj["igemm_dram_bytes_per_launch"]
                break
    try:
        probe = round(ctx.pl.measure_i8_mma_peak(), 1)  # back-to-back tcgen05.mma.kind::i8, all SMs
    except Exception:
        probe = None
    roof.update({"tensor_peak_tops_probe": probe,
                 "tensor_frac_of_probe": (tc_ops / (tc_ms * 1e-3) / 1e12 / probe) if probe else None,
                 "traffic": traffic, "algorithmic_bytes_per_launch": tc_bytes / n_tc, "kernel": "igemm_i8_kernel",
                 "launches_per_step": n_tc, "avg_launch_ms": tc_ms / n_tc, "share_of_step": tc_ms / total_ms,
                 "sum_launch_ms": tc_ms, "ms_per_step": ms_per_step, "timing": table[-1].get("timing"),
                 "peak_source": src, "tensor_tops": tc_ops / (tc_ms * 1e-3) / 1e12,
                 "tensor_peak_tops_2x_bf16": i8_peak, "hbm_gbs": tc_bytes / (tc_ms * 1e-3) / 1e9,
                 "by_kind_ms": {k: v["ms"] for k, v in by_kind(table).items()},
                 "by_kind_gbs": {k: v["gbs"] for k, v in by_kind(table).items()}})
    return roof


# ---- sub-records: the other BASELINE configs -----------------------------------------------------
def rotating(ctx, make, n_bytes_one, want_bytes=300e6, cap=8):
    """Enough copies of an operand set that consecutive launches never find their inputs in the 126 MB L2."""
    n = max(1, min(cap, int(want_bytes // max(n_bytes_one, 1)) + 1))
    return [make() for _ in range(n)]


def bench_conv1(ctx):
    """BASELINE config 1: conv2d 3x3 s1 p1 1x56x56x64 -> 64, per-channel scales, relu, int8 out (latency), and
    its batch-256 variant (TOP/s).  Bit-exactness of this shape is a test (tests/test_gpu_parity.py)."""
    torch, pl = ctx.torch, ctx.pl
    out = {}
    for N in (1, 256):
        C = K = 64
        g = torch.Generator(device="cpu").manual_seed(1234)
        w = torch.randint(-127, 128, (K, 3, 3, C), dtype=torch.int8, generator=g).to(ctx.dev)
        wp = pl.pack_conv_weight(w)
        sf = torch.full((K,), 2e-4, device=ctx.dev)
        bf = torch.zeros(K, device=ctx.dev)
        one = N * 56 * 56 * (C + K)
        sets = rotating(ctx, lambda: (torch.randint(-127, 128, (N, 56, 56, C), dtype=torch.int8, device=ctx.dev),
                                      torch.empty((N, 56, 56, K), dtype=torch.int8, device=ctx.dev)), one) \
            if N > 1 else [(torch.randint(-127, 128, (N, 56, 56, C), dtype=torch.int8, device=ctx.dev),
                            torch.empty((N, 56, 56, K), dtype=torch.int8, device=ctx.dev))]
        fns = [lambda x=x, y=y: pl.conv2d(x, wp, K, 3, 3, sf, bf, (1, 1), (1, 1, 1, 1), act=pl.ACT_RELU, y=y)
               for x, y in sets]
        per_graph = 100 if N == 1 else 24
        for f in fns:
            f()
        torch.cuda.synchronize()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr):
            for i in range(per_graph):
                fns[i % len(fns)]()
        for _ in range(10):
            gr.replay()
        ms = ctx.timed(gr.replay, 5)
        us = ms * 1e3 / (5 * per_graph)
        ops = 2.0 * N * 56 * 56 * K * 9 * C
        byts = one + K * 9 * C + 8 * K
        out[f"n{N}"] = {"us_per_launch": us, "tops": ops / us / 1e6, "gbs": byts / us / 1e3,
                        "l2": "warm (latency case: one operand set)" if N == 1 else f"{len(sets)} rotating operand sets",
                        "timing": f"{per_graph} launches per CUDA graph, CUDA events, reference convention warmup 10 / repeats 100"}
    return out


def bench_gemm(ctx, probe):
    """BASELINE config 5, single GPU: C[M,N] = q(s[n] * (A[M,K] . B[K,N]) + b[n]), int8 out."""
    torch, pl = ctx.torch, ctx.pl
    out = []
    for (M, N, K) in ((4096, 4096, 4096), (8192, 8192, 8192), (16384, 16384, 16384), (16, 16384, 16384),
                      (1, 4096, 4096)):
        one = M * K + N * K + M * N
        scale = torch.full((N,), 1e-4, device=ctx.dev)
        bias = torch.zeros(N, device=ctx.dev)

        def make():
            return (torch.randint(-127, 128, (M, K), dtype=torch.int8, device=ctx.dev),
                    torch.randint(-127, 128, (N, K), dtype=torch.int8, device=ctx.dev),
                    torch.empty((M, N), dtype=torch.int8, device=ctx.dev))
        sets = rotating(ctx, make, one, cap=6)
        fns = [lambda a=a, b=b, c=c: pl.gemm(a, b, scale, bias, out=pl.OUT_INT8, y=c) for a, b, c in sets]
        for f in fns:
            f()
        torch.cuda.synchronize()
        # launches are captured in a CUDA graph (the host call through ctypes costs more than a skinny GEMM)
        per_graph = max(len(fns), min(24, int(2e11 / (2.0 * M * N * K + 1)) + 1))
        per_graph = (per_graph + len(fns) - 1) // len(fns) * len(fns)
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr):
            for i in range(per_graph):
                fns[i % len(fns)]()
        for _ in range(3):
            gr.replay()
        reps = max(3, min(50, int(1.5e12 / (2.0 * M * N * K * per_graph + 1))))
        ms = ctx.timed(gr.replay, reps) / (reps * per_graph)
        ops = 2.0 * M * N * K
        byts = one + 8 * N
        out.append({"shape": [M, N, K], "out": "int8", "ms": ms, "tops": ops / ms / 1e9, "gbs": byts / ms / 1e6,
                    "frac_of_probe": (ops / ms / 1e9 / probe) if probe else None,
                    "operand_sets": len(sets), "launches_per_graph": per_graph})
        del sets, fns, gr
        torch.cuda.empty_cache()
    return out


def bench_nsplit(ctx):
    """BASELINE config 5, N-split: every rank owns a column block of B; local GEMM / + NCCL all-gather of the
    output columns / fused peer store (the GEMM's TMA store warp writes the tiles into every peer's result over
    NVLink).  The gathered and the fused result are compared bit for bit with the single-GPU GEMM in-run."""
    torch, pl, ns = ctx.torch, ctx.pl, ctx.pkg.nsplit
    out = []
    shapes = [(8192, 8192, 8192)] + ([(16384, 16384, 16384)] if ctx.world >= 4 else [])
    for (M, N, K) in shapes:
        g = torch.Generator().manual_seed(1234)
        B = torch.randint(-127, 128, (K, N), dtype=torch.int8, generator=g)
        A = torch.randint(-127, 128, (M, K), dtype=torch.int8, generator=g).to(ctx.dev)
        scale = torch.full((N,), 1e-4)
        bias = torch.zeros(N)
        op = ns.NSplitGemm(pl, B.numpy(), scale.numpy(), bias.numpy(), pl.OUT_INT8)
        single = pl.gemm(A, B.t().contiguous().to(ctx.dev), scale.to(ctx.dev), bias.to(ctx.dev), out=pl.OUT_INT8)
        iters = max(5, min(100, int(1e12 / (2.0 * M * N * K / ctx.world + 1))))
        for _ in range(3):
            op.local(A)
        ms_local = ctx.timed(lambda: op.local(A), iters) / iters
        for _ in range(3):
            gathered = op(A)
        ms_nccl = ctx.timed(lambda: op(A), iters) / iters
        ok_nccl = bool(torch.equal(op(A), single))
        op.fused_setup(M)
        for _ in range(3):
            fused = op.fused(A)
        ms_fused = ctx.timed(lambda: op.fused(A), iters) / iters
        fused = op.fused(A)
        ctx.barrier()
        ok_fused = bool(torch.equal(fused, single))
        # the same through the NVSwitch multicast address: one multimem.st stream instead of G - 1 peer stores
        ms_mcast, ok_mcast = None, True
        try:
            if op.mcast_supported():
                for _ in range(3):
                    fused = op.fused_mcast(A)
                ms_mcast = ctx.timed(lambda: op.fused_mcast(A), iters) / iters
                fused = op.fused_mcast(A)
                ctx.barrier()
                ok_mcast = bool(torch.equal(fused, single))
        except Exception as exc:  # pragma: no cover - depends on the box
            sys.stderr.write(f"[bench] multicast N-split unavailable: {exc!r}\n")
            ms_mcast = None
        # fp32 results: NCCL gather against multicast stores from the epilogue registers
        f32 = {}
        try:
            opf = ns.NSplitGemm(pl, B.numpy(), scale.numpy(), bias.numpy(), pl.OUT_FP32)
            single_f = pl.gemm(A, B.t().contiguous().to(ctx.dev), scale.to(ctx.dev), bias.to(ctx.dev), out=pl.OUT_FP32)
            for _ in range(2):
                opf.local(A)
            f32["ms_local"] = ctx.timed(lambda: opf.local(A), iters) / iters
            for _ in range(2):
                opf(A)
            f32["ms_nccl_gather"] = ctx.timed(lambda: opf(A), iters) / iters
            opf.fused_setup(M)
            if opf.mcast_supported():
                for _ in range(2):
                    opf.fused_mcast(A)
                f32["ms_fused_multicast"] = ctx.timed(lambda: opf.fused_mcast(A), iters) / iters
                r_ = opf.fused_mcast(A)
                ctx.barrier()
                f32["bit_exact_vs_single_gpu"] = ctx.min_over_ranks(float(torch.equal(r_, single_f)))[0] == 1.0
            del opf, single_f
        except Exception as exc:  # pragma: no cover
            f32["error"] = repr(exc)
        ok = ctx.min_over_ranks(float(ok_nccl and ok_fused and ok_mcast))[0] == 1.0
        ops = 2.0 * M * N * K
        best = min(v for v in (ms_fused, ms_mcast) if v is not None)
        out.append({"M": M, "N": N, "K": K, "n_gpus": ctx.world, "ms_local": ms_local, "ms_nccl_gather": ms_nccl,
                    "ms_fused_peer": ms_fused, "ms_fused_multicast": ms_mcast, "tops_fused": ops / best / 1e9,
                    "bit_exact_vs_single_gpu": ok, "fp32_out": f32})
        del op, single, A, B, fused, gathered
        torch.cuda.empty_cache()
    return out


def bench_sustained(ctx, net, batch, seconds, ms_per_step):
    steps = max(50, int(seconds * 1e3 / ms_per_step))
    sampler = ClockSampler(ctx.local).start() if ctx.rank == 0 else None
    ms = ctx.timed(net.replay, steps)
    ck = sampler.stop() if sampler else None
    return {"steps": steps, "seconds": ms / 1e3, "ms_per_step": ms / steps,
            "img_s": ctx.world * batch * steps / (ms * 1e-3), "clocks": ck}


def bench_plugin(ctx, prog, net, batch, x_host, steps, warmup):
    """The same program, the same end-to-end loop, but through the reference-facing plugin surface: every op is a
    KernelLite instance from KernelRegistry run by Launch() (libpl_lite_kernels.so; lite_net.PluginNet), with the
    reference's op chain calib -> layout -> conv in front instead of the fused entry kernel."""
    import importlib
    torch = ctx.torch
    lite_net = importlib.import_module("paddle_lite_b200.lite_net")
    pnet = lite_net.PluginNet(prog, batch)
    x_dev = x_host.cuda()
    pnet.inputs[0].copy_(x_dev)
    net.input.copy_(x_dev)
    same = bool(torch.equal(pnet.run(0), net.replay()))  # logits of both paths, bit for bit
    pnet.capture()
    for _ in range(3):
        pnet.replay(0)
    ctx.barrier()
    ms = ctx.timed(lambda: pnet.replay(0), steps)
    rec = e2e_loop(ctx, pnet, x_host, steps, warmup)
    rec["input"] = rec["input"].replace("the network's", "the plugin program's")
    rec.update({"path": "KernelRegistry -> KernelLite::Launch (libpl_lite_kernels.so), one CUDA graph around the "
                        "instruction loop", "instructions_per_step": pnet.n_instructions,
                "ms_per_step_device": ms / steps, "img_s_device": ctx.world * batch * steps / (ms * 1e-3),
                "logits_identical_to_c_abi_path": same})
    del pnet
    torch.cuda.empty_cache()
    return rec


def run_b200(args):
    ctx = Ctx()
    torch = ctx.torch
    world, rank = ctx.world, ctx.rank
    batch = args.batch or default_batch(args.net)
    n0 = ctx.pl.launch_count()
    rec, prog, net, x_host, table = measure_net(ctx, args.net, batch, args.steps, args.warmup, clocks=True)
    launches_per_step = len(launch_order(net))
    ms_per_step = rec["ms_per_step"]
    roof = headline_roofline(ctx, table, prog, batch, ms_per_step)
    if args.breakdown and rank == 0:
        with open(args.breakdown, "w") as f:
            json.dump(table, f, indent=1)
    macs = ctx.nets.program_macs(prog)
    H, W, _ = prog.shapes["input"]
    line = {"metric": METRIC, "value": rec["img_s"], "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "int8", "data": "synthetic",
            "config": {"workload": workload_name(args.net, batch, H), "global_batch": batch * world,
                       "parallelism": f"batch-sharded x{world}, no collective", "cuda_graph": True,
                       "numa_node_of_rank0": ctx.numa,
                       "l2": f"per-step working set ({(ctx.nets.program_bytes(prog)[0] * batch + ctx.nets.program_bytes(prog)[1]) / 1e6:.0f} MB of activations + weights) "
                             "exceeds the 126 MB L2; consecutive steps replay the whole network, no flush needed",
                       "gmac_per_image": sum(macs.values()) / 1e9},
            "tops": rec["tops"], "clocks": rec.get("clocks"), "e2e": rec["e2e"],
            "gpu_launches": launches_per_step * args.steps, "roofline": roof, "cpu_baseline": None}

    extras = not args.no_extras
    if extras:
        line["sustained"] = bench_sustained(ctx, net, batch, 2.0, ms_per_step)
        if args.net in ("resnet50", "mobilenet_v1"):
            try:
                line["e2e_plugin"] = bench_plugin(ctx, prog, net, batch, x_host, args.steps, args.warmup)
            except Exception as exc:
                line["e2e_plugin"] = {"error": repr(exc)}
        # the same end-to-end loop with narrower host feeds (the entry kernel reads fp16 / uint8 directly)
        for feed in ("f16", "u8"):
            try:
                net = None
                torch.cuda.empty_cache()
                _, net, xh = build_net(ctx, args.net, batch, feed)
                line[f"e2e_{feed}"] = e2e_loop(ctx, net, xh, args.steps, args.warmup)
            except Exception as exc:
                line[f"e2e_{feed}"] = {"error": repr(exc)}
        net = None
        torch.cuda.empty_cache()
        nets_rec = {}
        if args.net == "resnet50":
            for name, nb, st in (("mobilenet_v1", 64, 50), ("mobilenet_v1", 1, 100), ("humanseg", 32, 50)):
                try:
                    r, _, n2, _, _ = measure_net(ctx, name, nb, st, 5)
                    nets_rec[f"{name}_b{nb}"] = r
                    del n2
                except Exception as exc:
                    nets_rec[f"{name}_b{nb}"] = {"error": repr(exc)}
                torch.cuda.empty_cache()
        line["nets"] = nets_rec
        try:
            line["conv1"] = bench_conv1(ctx)
        except Exception as exc:
            line["conv1"] = {"error": repr(exc)}
        try:
            line["gemm"] = bench_gemm(ctx, roof.get("tensor_peak_tops_probe"))
        except Exception as exc:
            line["gemm"] = {"error": repr(exc)}
        if world > 1:
            try:
                line["nsplit"] = bench_nsplit(ctx)
            except Exception as exc:
                line["nsplit"] = {"error": repr(exc)}
    line["gpu_launches_total"] = ctx.pl.launch_count() - n0
    if rank != 0:
        ctx.close()
        return 0
    if world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"], _ = cpu_reference_rate(args.net, args.cpu_seconds, steps=0)
    print(json.dumps(line))
    ctx.close()
    return 0


def main():
    args = parse()
    if args.impl == "reference":
        return run_reference(args)
    return run_b200(args)


if __name__ == "__main__":
    sys.exit(main())
