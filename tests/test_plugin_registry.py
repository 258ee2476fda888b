"""CPU-side checks of the plugin boundary (no GPU): the kernels the north_star names
are registered under the reference's keys, and the miniature static_kernel_pick obeys
the reference's int8_out / fp32_out rule (lite/core/mir/static_kernel_pick_pass.cc:93-166)."""
import pytest


@pytest.fixture(scope="module")
def lh():
    import __graft_entry__ as ge
    return ge.load_package().lite_harness


def test_registered_kernels(lh):
    ks = set(lh.list_kernels())
    for op in ("conv2d", "depthwise_conv2d", "pool2d", "elementwise_add"):
        for alias in ("int8_out", "fp32_out"):
            assert f"{op}/{alias}/cuda/int8_t/NHWC" in ks
    for op in ("fc", "mul", "matmul"):
        for alias in ("int8out", "fp32out"):  # fc aliases as lite/kernels/arm/fc_compute.cc:368-380
            assert f"{op}/{alias}/cuda/int8_t/NCHW" in ks
    for op in ("calib", "calib_once"):
        assert f"{op}/fp32_to_int8/cuda/int8_t/NCHW" in ks and f"{op}/int8_to_fp32/cuda/int8_t/NCHW" in ks
    for op in ("layout", "layout_once"):
        assert f"{op}/int8_nchw2nhwc/cuda/int8_t/NCHW" in ks and f"{op}/nhwc2nchw/cuda/float/NCHW" in ks
    for op in ("concat", "split", "shuffle_channel", "relu", "relu6"):
        assert f"{op}/int8_nhwc/cuda/int8_t/NHWC" in ks
    for op in ("io_copy", "io_copy_once"):
        assert f"{op}/host_to_device/cuda/any/any" in ks and f"{op}/device_to_host/cuda/any/any" in ks


@pytest.mark.parametrize("op", ["conv2d", "depthwise_conv2d", "fc", "pool2d", "elementwise_add"])
def test_static_kernel_pick_output_precision_rule(lh, op):
    int8 = "int8out" if op == "fc" else "int8_out"
    fp32 = "fp32out" if op == "fc" else "fp32_out"
    key, scale = lh.pick(op, True, [True, True], [0.25, 0.25])
    assert key.split("/")[1] == int8 and scale == 0.25      # every consumer int8 -> int8 out + its scale
    key, _ = lh.pick(op, True, [True, False])
    assert key.split("/")[1] == fp32                          # one fp32 consumer -> fp32 out
    key, _ = lh.pick(op, True, [])
    assert key.split("/")[1] == fp32                          # graph output -> fp32 out
    key, _ = lh.pick(op, False, [True])
    assert key == ""                                          # no enable_int8 -> int8 kernels not eligible


def test_errors_surface_like_check_failures(lh):
    import numpy as np
    with pytest.raises(lh.LiteError):
        lh.Tensor(np.zeros((2, 2), np.float64))
