import sys, torch
sys.path.insert(0, '.')
import __graft_entry__ as ge
pl = ge.load_package().capi
def t(M, N, K, out):
    sets = []
    n = max(1, min(6, int(3e8 // (N * K)) + 1))
    for _ in range(n):
        sets.append((torch.randint(-127, 128, (M, K), dtype=torch.int8, device='cuda'),
                     torch.randint(-127, 128, (N, K), dtype=torch.int8, device='cuda'),
                     torch.empty((M, N), dtype=torch.int8 if out == pl.OUT_INT8 else torch.float32, device='cuda')))
    sc = torch.full((N,), 1e-4, device='cuda'); bs = torch.zeros(N, device='cuda')
    fns = [lambda a=a, b=b, c=c: pl.gemm(a, b, sc, bs, out=out, y=c) for a, b, c in sets]
    for f in fns: f()
    torch.cuda.synchronize()
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
    print(f"{M}x{N}x{K} out={'i8' if out == pl.OUT_INT8 else 'f32'} {us:8.2f} us  {(N*K + M*K)/us/1e3:8.1f} GB/s", flush=True)
for shp in [(1,4096,4096),(4,4096,4096),(8,4096,4096),(1,1000,2048),(1,16384,16384),(8,16384,16384),(16,16384,16384)]:
    t(*shp, pl.OUT_INT8)
