"""the ResNet-50 / MobileNet fc at batch sizes 64 / 256 (fp32 out), graph-timed like tools/gemv_time.py"""
import sys, torch
sys.path.insert(0, '.')
import __graft_entry__ as ge
pl = ge.load_package().capi
def t(M, N, K, out):
    sets = []
    for _ in range(6):
        sets.append((torch.randint(-127, 128, (M, K), dtype=torch.int8, device='cuda'),
                     torch.randint(-127, 128, (N, K), dtype=torch.int8, device='cuda'),
                     torch.empty((M, N), dtype=torch.int8 if out == pl.OUT_INT8 else torch.float32, device='cuda')))
    sc = torch.full((N,), 1e-4, device='cuda'); bs = torch.zeros(N, device='cuda')
    fns = [lambda a=a, b=b, c=c: pl.gemm(a, b, sc, bs, out=out, y=c) for a, b, c in sets]
    for f in fns: f()
    torch.cuda.synchronize()
    q = pl.last_igemm_plan()
    g = torch.cuda.CUDAGraph()
    per = 4 * len(fns)
    with torch.cuda.graph(g):
        for i in range(per): fns[i % len(fns)]()
    for _ in range(3): g.replay()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(10): g.replay()
    b.record(); torch.cuda.synchronize()
    us = a.elapsed_time(b) * 1e3 / (10 * per)
    print(f"{M}x{N}x{K} out={'i8' if out == pl.OUT_INT8 else 'f32'} {us:8.2f} us  plan BN {q['BN']} BK {q['BK']} stages {q['stages']} grid {q['grid']}", flush=True)
for shp in [(256, 1000, 2048), (256, 1000, 1024), (64, 1000, 1024), (128, 1000, 2048)]:
    t(*shp, pl.OUT_FP32)
    t(*shp, pl.OUT_INT8)
