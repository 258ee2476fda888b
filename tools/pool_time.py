"""int8 pooling at the ResNet-50 / MobileNet shapes (batch 256), L2-cold buffer rotation, CUDA events"""
import sys, torch
sys.path.insert(0, '.')
import __graft_entry__ as ge
pl = ge.load_package().capi
def t(N, H, C, k, s, pad, typ, name):
    xs = [torch.randint(-127, 128, (N, H, H, C), dtype=torch.int8, device='cuda') for _ in range(3)]
    P = (H + 2 * pad - k) // s + 1
    ys = [torch.empty((N, P, P, C), dtype=torch.int8, device='cuda') for _ in range(3)]
    for i in range(3): pl.pool2d(xs[i], (k, k), (s, s), (pad,) * 4, typ, y=ys[i])
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(30): pl.pool2d(xs[i % 3], (k, k), (s, s), (pad,) * 4, typ, y=ys[i % 3])
    b.record(); torch.cuda.synchronize()
    us = a.elapsed_time(b) * 1e3 / 30
    byts = N * H * H * C + N * P * P * C
    print(f"{name:28s} {us:8.2f} us  {byts / us / 1e3:8.1f} GB/s", flush=True)
t(256, 112, 64, 3, 2, 1, pl.POOL_MAX, "max 3x3 s2 112x112x64")
t(256, 7, 2048, 7, 1, 0, pl.POOL_AVG, "avg 7x7 global 7x7x2048")
t(256, 7, 1024, 7, 1, 0, pl.POOL_AVG, "avg 7x7 global 7x7x1024")
t(256, 56, 64, 2, 2, 0, pl.POOL_MAX, "max 2x2 s2 56x56x64")
