"""profiles/<tag>_results.md from gpurun_out/<tag>_results (tools/gpu_round_results.sh).  usage: make_results_md.py r01"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1] if len(sys.argv) > 1 else "r01"
O = os.path.join(ROOT, "gpurun_out", f"{tag}_results")
HBM = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6532.9


def last_json(path):
    return json.loads(open(path).read().strip().splitlines()[-1])


out = [f"# {tag}: measured results on one B200 (`tools/gpu_round_results.sh {tag}`)\n"]
out.append(open(os.path.join(O, "gpu.txt")).read().strip().replace("\n", " | ") + "\n")
probe = open(os.path.join(O, "mma_probe.txt")).read().strip().splitlines()
peak = max(float(l.split()[-2]) for l in probe if "TOPS" in l)
out.append(f"Roofs used below: HBM {HBM:.0f} GB/s (MEASURED_PEAKS.json, copy), int8 tensor {peak:.0f} TOP/s "
           "(`tools/mma_probe.py`: back-to-back `tcgen05.mma.kind::i8` on resident smem operands, all 148 SMs).\n")
out.append("## int8 tensor-pipe probe (cycles per MMA, M=128, K=32)\n\n```\n" + "\n".join(probe) + "\n```\n")

out.append("## Networks (bench.py, CUDA graph, device-resident `value`; `e2e` = pinned host fp32 in, logits out)\n")
out.append("| run | img/s (device) | ms/step | img/s (e2e) | TOP/s | igemm roofline | CPU reference img/s (cores) |")
out.append("|---|---:|---:|---:|---:|---|---:|")
for name, f in (("ResNet-50 int8 b256", "bench_resnet50_b256.json"), ("MobileNetV1 int8 b64", "bench_mobilenet_b64.json"),
                ("MobileNetV1 int8 b1", "bench_mobilenet_b1.json"),
                ("ShuffleNet-v2-like int8 b64 (SURVEY 8f rank 1 ops)", "bench_shufflenet_b64.json")):
    if not os.path.exists(os.path.join(O, f)):
        continue
    l = last_json(os.path.join(O, f))
    r = l["roofline"]
    cb = l.get("cpu_baseline")
    out.append(f"| {name} | {l['value']:.0f} | {l['ms_per_step']:.4f} | {l['e2e']['value']:.0f} | {l['tops']:.0f} | "
               f"{r['bound']}: {r['achieved']:.0f} / {r['peak']:.0f} {r['unit']} = {r['frac']:.2f} | "
               f"{'%.2f (%d)' % (cb['value'], cb['cores']) if cb else '-'} |")
out.append("")

out.append("## Config 1: single conv2d 3x3 s1 p1, 64 -> 64 channels at 56x56 (`tools/conv1_latency.py`)\n")
out.append("| shape | us / launch | TOP/s | GB/s | bit-exact vs oracle |\n|---|---:|---:|---:|---|")
for ln in open(os.path.join(O, "config1_conv.jsonl")):
    j = json.loads(ln)
    out.append(f"| {j['config']} | {j['us_per_launch']} | {j['tops']} | {j['gbs']} | {j['bit_exact_vs_oracle']} |")
out.append("")

out.append("## Layers at BASELINE shapes, batch 256, L2-cold (`tools/layer_bench.py`)\n")
out.append("| layer | ms | TOP/s | % of tensor peak | GB/s (algorithmic) | % of HBM |\n|---|---:|---:|---:|---:|---:|")
for ln in open(os.path.join(O, "layers_b256.txt")):
    f = ln.split()
    if len(f) < 5 or not f[2].startswith("ms="):
        continue
    ms = float(f[2][3:])
    tops = float(ln.split("TOPS=")[1].split()[0])
    gbs = float(ln.split("GB/s=")[1].split()[0])
    out.append(f"| {f[0]} | {ms:.4f} | {tops:.0f} | {100 * tops / peak:.0f} | {gbs:.0f} | {100 * gbs / HBM:.0f} |")
out.append("")

if os.path.exists(os.path.join(O, "channel_ops.jsonl")):
    out.append("## int8 NHWC channel movement kernels (`tools/channel_ops_bench.py`; 2 bytes per element)\n")
    out.append("| op | tensor | MB moved | us | GB/s | fraction of HBM peak |\n|---|---|---:|---:|---:|---:|")
    for ln in open(os.path.join(O, "channel_ops.jsonl")):
        if not ln.startswith("{"):
            continue
        j = json.loads(ln)
        if j["MB_moved"] < 100:
            continue  # the small humanseg shapes time the Python call, not the kernel
        out.append(f"| {j['op']} | {j['shape']} x {j['channels']} | {j['MB_moved']} | {j['us']} | {j['GB/s']} | {j['frac_hbm_peak']} |")
    out.append("")

out.append("## Config 5: int8 GEMM sweep, one GPU (`tools/gemm_sweep.py`)\n")
out.append("| M x N x K | out | ms | TOP/s | % of tensor peak | GB/s |\n|---|---|---:|---:|---:|---:|")
for ln in open(os.path.join(O, "gemm_sweep_1gpu.jsonl")):
    ln = ln.strip()
    if not ln.startswith("{"):
        continue
    j = json.loads(ln)
    M, N, K = j["shape"]
    out.append(f"| {M} x {N} x {K} | {j['out']} | {j['ms_gemm']} | {j['tops_gemm']} | {100 * j['tops_gemm'] / peak:.0f} | {j['gbs_gemm']} |")
out.append("")
open(os.path.join(ROOT, "profiles", f"{tag}_results.md"), "w").write("\n".join(out) + "\n")
print("\n".join(out)[-3000:])
