"""cycles per tcgen05.mma kind::i8 (M=128, K=32) as a function of N and of the operand swizzle mode"""
import ctypes as C
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge
pl = ge.load_package().capi
lib = pl.lib()
out = torch.zeros(4, dtype=torch.int32, device="cuda")
sms = torch.cuda.get_device_properties(0).multi_processor_count
clk = 1.965e9
for rb in (128, 64, 32):
    for bn in (256, 128, 64, 32, 16):
        iters = 4000
        lib.pl_i8_mma_probe(sms, 200, bn, rb, C.c_void_p(out.data_ptr()), None)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        lib.pl_i8_mma_probe(sms, iters, bn, rb, C.c_void_p(out.data_ptr()), None)
        b.record()
        torch.cuda.synchronize()
        ms = a.elapsed_time(b)
        mmas = iters * 8
        cyc = ms * 1e-3 * clk / mmas
        tops = sms * mmas * 2.0 * 128 * bn * 32 / (ms * 1e-3) / 1e12
        print(f"row_bytes={rb:3d} N={bn:3d}: {cyc:7.1f} cycles/MMA  {tops:7.1f} TOPS", flush=True)

print("CTA pairs (cta_group::2, M = 256; B half per CTA):")
for bn in (256, 128, 64):
    iters = 4000
    lib.pl_i8_mma_probe_pairs(sms, 200, bn, 128, C.c_void_p(out.data_ptr()), None)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    rc = lib.pl_i8_mma_probe_pairs(sms, iters, bn, 128, C.c_void_p(out.data_ptr()), None)
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b)
    mmas = iters * 8
    cyc = ms * 1e-3 * clk / mmas
    tops = (sms // 2) * mmas * 2.0 * 256 * bn * 32 / (ms * 1e-3) / 1e12
    print(f"pairs rc={rc} N={bn:3d}: {cyc:7.1f} cycles/MMA  {tops:7.1f} TOPS", flush=True)

for rbase in (128, 64):
    print(f"A descriptor start shifted by whole rows (halo-mode taps), row_bytes={rbase}:")
    for shift in (0, 1, 2, 3, 4, 8, 9):
        for bn in (256, 128, 64):
            iters = 4000
            rb = rbase | (shift << 8)
            lib.pl_i8_mma_probe(sms, 200, bn, rb, C.c_void_p(out.data_ptr()), None)
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            rc = lib.pl_i8_mma_probe(sms, iters, bn, rb, C.c_void_p(out.data_ptr()), None)
            b.record()
            torch.cuda.synchronize()
            ms = a.elapsed_time(b)
            mmas = iters * 8
            cyc = ms * 1e-3 * clk / mmas
            print(f"shift={shift} rc={rc} N={bn:3d}: {cyc:7.1f} cycles/MMA", flush=True)
