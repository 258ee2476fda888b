"""CPU-side checks of the drop-in boundary: the C-ABI library builds, loads and
exports every symbol include/pl_int8.h declares; host-only helpers agree with the
oracle; argument validation fails loudly (no compute calls without a GPU)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import orc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "pl_int8.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(pl_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_the_hot_path():
    names = declared_symbols()
    for must in ("pl_conv2d_i8_nhwc", "pl_dwconv2d_i8_nhwc", "pl_gemm_i8", "pl_quantize_f32_i8",
                 "pl_dequantize_i8_f32", "pl_nchw_to_nhwc", "pl_nhwc_to_nchw", "pl_pool2d_i8_nhwc",
                 "pl_add_i8", "pl_pack_conv_weight_i8", "pl_fold_scales"):
        assert must in names


def test_library_exports_every_declared_symbol(pl):
    L = pl.lib()
    for name in declared_symbols():
        assert hasattr(L, name), f"{name} declared in include/pl_int8.h but not exported"
    assert b"sm_100a" in L.pl_version()


def test_fold_scales_matches_oracle(pl):
    rng = np.random.default_rng(1)
    ws = rng.random(24, dtype=np.float32) + 0.1
    b = rng.standard_normal(24).astype(np.float32)
    for out in (orc.OUT_INT8, orc.OUT_FP32):
        for act, arg in ((orc.ACT_RELU6, 6.0), (orc.ACT_LEAKY, 0.1), (orc.ACT_NONE, 0.0)):
            want = orc.fold_scales(ws, 24, 0.013, 0.21, out, b, act, arg)
            got = pl.fold_scales(ws, 24, 0.013, 0.21, out, b, act, arg)
            assert np.array_equal(want[0], got[0]) and np.array_equal(want[1], got[1]) and want[2] == got[2]
    want = orc.fold_scales(ws[:1], 24, 0.5, 0.25, orc.OUT_INT8, None)
    got = pl.fold_scales(ws[:1], 24, 0.5, 0.25, orc.OUT_INT8, None)
    assert np.array_equal(want[0], got[0]) and np.array_equal(want[1], got[1])


def test_out_size_and_packing_helpers(pl):
    L = pl.lib()
    for args in ((224, 7, 1, 3, 3, 2), (56, 3, 1, 1, 1, 1), (17, 3, 2, 2, 2, 1), (15, 3, 1, 0, 1, 2)):
        assert L.pl_conv_out_size(*args) == orc.conv_out_size(*args)
    assert [L.pl_conv_block_k(c) for c in (3, 16, 32, 33, 64, 65, 128, 2048)] == [32, 32, 32, 64, 64, 128, 128, 128]
    assert L.pl_conv_packed_weight_bytes(64, 3, 3, 64) == 64 * 9 * 64
    assert L.pl_conv_packed_weight_bytes(32, 3, 3, 3) == 32 * 9 * 32
    assert L.pl_conv_packed_weight_bytes(80, 3, 3, 96) == 80 * 9 * 128


def test_null_and_bad_arguments_are_rejected_without_a_gpu(pl):
    L = pl.lib()
    assert L.pl_conv2d_i8_nhwc(None, None, None, None, None, None) == -1
    assert L.pl_gemm_i8(None, 0, None, 0, None, 0, 1, 1, 1, None, None) == -1
    assert L.pl_dwconv2d_i8_nhwc(None, None, None, None, None, None) == -1
    assert L.pl_quantize_f32_i8(None, None, C.c_int64(4), C.c_float(1.0), 0, None) == -1
    assert L.pl_add_i8(None, None, None, C.c_int64(4), C.c_float(1), C.c_float(1), 0, C.c_float(0), 0, 0, None) == -1
    assert L.pl_fold_scales(None, 1, 4, C.c_float(1), C.c_float(1), 0, None, None, None, 0, C.c_float(0), None) == -1


def test_product_never_imports_the_oracle():
    """The oracle is test infrastructure: nothing under paddle-lite_b200/ may reference it."""
    pkg = os.path.join(ROOT, "paddle-lite_b200")
    for dp, _, fns in os.walk(pkg):
        for fn in fns:
            if fn.endswith((".py", ".cu", ".cuh", ".cc", ".h", ".cpp")):
                txt = open(os.path.join(dp, fn), errors="ignore").read()
                assert "liboracle" not in txt and "import orc" not in txt and "oracle/" not in txt.replace(
                    "the CPU oracle", ""), f"{fn} references the oracle"
