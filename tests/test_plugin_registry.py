"""CPU-side checks of the plugin boundary (no GPU): the kernels the north_star names
are registered under the reference's keys, and the miniature static_kernel_pick obeys
the reference's int8_out / fp32_out rule (lite/core/mir/static_kernel_pick_pass.cc:93-166)."""
import numpy as np
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


# ---- miniature type passes (lite/core/mir/type_cast_plan.h) -------------------
def _shufflenet_unit():
    """feed -> conv (int8) -> split -> [x1 | x2 -> 1x1 -> dw3x3 -> 1x1] -> concat -> shuffle_channel -> conv -> fetch
    (the stride-1 unit of humanseg_shfv75's ShuffleNet-v2 backbone)"""
    return [
        ("feed", [], ["image"], False),
        ("conv2d", ["image"], ["c0"], True),
        ("split", ["c0"], ["x1", "x2"], False),
        ("conv2d", ["x2"], ["b1"], True),
        ("depthwise_conv2d", ["b1"], ["b2"], True),
        ("conv2d", ["b2"], ["b3"], True),
        ("concat", ["x1", "b3"], ["cat"], False),
        ("shuffle_channel", ["cat"], ["sh"], False),
        ("conv2d", ["sh"], ["c1"], True),
        ("fetch", ["c1"], [], False),
    ]


def test_type_plan_keeps_shufflenet_unit_int8_resident(lh):
    ops = _shufflenet_unit()
    ext = lh.plan_types(ops, extension=True)
    # every op between feed and fetch runs an int8 NHWC kernel of this library
    assert ext["picked"][1] == "conv2d/int8_out/cuda/int8_t/NHWC"
    assert ext["picked"][2] == "split/int8_nhwc/cuda/int8_t/NHWC"
    assert ext["picked"][6] == "concat/int8_nhwc/cuda/int8_t/NHWC"
    assert ext["picked"][7] == "shuffle_channel/int8_nhwc/cuda/int8_t/NHWC"
    assert ext["picked"][8] == "conv2d/fp32_out/cuda/int8_t/NHWC"  # its consumer is fetch
    assert ext["inside"] == 0
    kinds = [(c[0], c[1]) for c in ext["casts"]]
    assert kinds == [("io_copy", "image"), ("calib", "image"), ("layout", "image"),
                     ("io_copy", "c1"), ("layout", "c1")] or \
        kinds == [("io_copy", "image"), ("calib", "image"), ("layout", "image"), ("io_copy", "c1")]


def test_type_plan_without_extension_round_trips(lh):
    """The reference's behaviour: split / concat / shuffle_channel have no enable_int8, so the convs
    around them fall to fp32_out and calib / layout / io_copy nodes appear inside the block."""
    ops = _shufflenet_unit()
    ref = lh.plan_types(ops, extension=False)
    assert ref["picked"][2].startswith("split/def/host")          # ARM-only kernel in the reference
    assert ref["picked"][7].startswith("shuffle_channel/def/host")
    assert ref["picked"][6] == "concat/def/cuda/float/NCHW"
    assert ref["picked"][1] == "conv2d/fp32_out/cuda/int8_t/NHWC"  # consumer split is not int8
    assert ref["inside"] >= 8
    ext = lh.plan_types(ops, extension=True)
    assert ref["bytes_per_elem"] - ext["bytes_per_elem"] >= 40     # per-element traffic of the round trips


def test_type_plan_resnet_block_and_fp32_tail(lh):
    ops = [
        ("feed", [], ["x"], False),
        ("conv2d", ["x"], ["a"], True),
        ("conv2d", ["a"], ["b"], True),
        ("elementwise_add", ["b", "a"], ["s"], True),
        ("relu", ["s"], ["r"], False),
        ("pool2d", ["r"], ["p"], True),
        ("softmax", ["p"], ["y"], False),   # fp32-only tail op
        ("fetch", ["y"], [], False),
    ]
    plan = lh.plan_types(ops, extension=True)
    assert plan["picked"][3] == "elementwise_add/int8_out/cuda/int8_t/NHWC"
    assert plan["picked"][4] == "relu/int8_nhwc/cuda/int8_t/NHWC"
    assert plan["picked"][5] == "pool2d/fp32_out/cuda/int8_t/NHWC"   # softmax consumes fp32
    assert plan["picked"][6] == "softmax/def/cuda/float/NCHW"
    inside = [c for c in plan["casts"] if c[1] not in ("x", "y")]
    assert [(c[0], c[1]) for c in inside] == [("layout", "p")]         # NHWC -> NCHW for softmax, no calib
    # a relu whose consumer is fp32 cannot stay int8: the extension must not claim it
    ops2 = [("feed", [], ["x"], False), ("conv2d", ["x"], ["a"], True), ("relu", ["a"], ["r"], False),
            ("softmax", ["r"], ["y"], False), ("fetch", ["y"], [], False)]
    plan2 = lh.plan_types(ops2, extension=True)
    assert plan2["picked"][2] == "relu/def/cuda/float/NCHW"
    assert plan2["picked"][1] == "conv2d/fp32_out/cuda/int8_t/NHWC"


def test_kernel_sources_compile_against_reference_headers():
    """The drop-in claim: the kernel class sources compile byte for byte against the reference's own
    lite/core / lite/operators headers (tools/check_dropin.sh; only the CMake-generated version.h and the
    documented MatMulParam / ActivationParam int8-config edit come from outside the reference tree)."""
    import os
    import subprocess
    ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    if not os.path.isdir("/root/reference/lite/core"):
        pytest.skip("reference tree not present on this machine")
    out = subprocess.run(["bash", os.path.join(ROOT, "tools", "check_dropin.sh"), "/root/reference"],
                         capture_output=True, text=True)
    assert out.returncode == 0, out.stdout + out.stderr
    assert out.stdout.count("ok  ") == 3


# ---- KernelGrade (lite/core/mir/static_kernel_pick_pass.h:53-215) ------------------------------------------
# enum values of lite/api/paddle_place.h
T_X86, T_CUDA, T_ARM = 2, 3, 4
P_FLOAT, P_INT8, P_ANY = 1, 2, 4
L_NCHW, L_ANY, L_NHWC = 1, 2, 3


def _grade(places, kplace, enable_int8, type_match=True):
    """the reference's arithmetic, by hand: kMax = 255, integer division per factor, best weighted place"""
    best = -1.0
    n = len(places)
    for i, (t, p, l) in enumerate(places):
        w = (n - i) / n
        s = 0
        if t == kplace[0]:
            s += 255 // 1
        if (p == kplace[1] or kplace[1] == P_ANY or p == P_ANY) and not (kplace[1] == P_INT8 and not enable_int8):
            s += 255 // 2
        if l == kplace[2] or kplace[2] == L_ANY or l == L_ANY:
            s += 255 // 4
        if not enable_int8 and type_match:
            s *= 2
        best = max(best, np.float32(w) * np.float32(s))
    return int(best)


def test_kernel_grade_scores(lh):
    import numpy as np  # noqa: F401
    cuda_i8 = (T_CUDA, P_INT8, L_NHWC)
    # one place: target + precision + layout = 255 + 127 + 63
    sc = lh.score_kernels("conv2d", [cuda_i8], True)
    assert sc and all(s == 445 for s, _ in sc)
    assert {k.split("/")[1] for _, k in sc} == {"int8_out", "fp32_out"}
    # place order is preference: the same kernels behind an x86 place score half
    sc2 = lh.score_kernels("conv2d", [(T_X86, P_FLOAT, L_NCHW), cuda_i8], True)
    assert all(s == _grade([(T_X86, P_FLOAT, L_NCHW), cuda_i8], cuda_i8, True) == 222 for s, _ in sc2)
    # three places, the kernel's own place last: weight 1/3 there, but target-only matches earlier score more
    pl3 = [(T_CUDA, P_FLOAT, L_NCHW), (T_ARM, P_FLOAT, L_NCHW), cuda_i8]
    want = _grade(pl3, cuda_i8, True)
    assert want == 255 and all(s == want for s, _ in lh.score_kernels("conv2d", pl3, True))
    # an op without enable_int8: int8 kernels get no precision score, and the sum doubles on a type match
    sc3 = lh.score_kernels("conv2d", [cuda_i8], False)
    assert all(s == (255 + 63) * 2 for s, _ in sc3)
    sc4 = lh.score_kernels("conv2d", [cuda_i8], False, {"Input": P_FLOAT})  # declared int8 input, fp32 variable
    assert all(s == 255 + 63 for s, _ in sc4)
    # kAny / kAny kernels (io_copy) are reached through the expanded places and match every place
    sc5 = lh.score_kernels("io_copy", [cuda_i8, (T_CUDA, P_FLOAT, L_NCHW)], False)
    assert sc5 and all(s == 445 * 2 for s, _ in sc5)


def test_kernel_grade_tie_order_and_pick(lh):
    """equal scores keep the candidate order of the expanded std::set<Place> (stable sort), and the int8 rule then
    takes the first candidate whose declared outputs have the wanted precision"""
    places = [(T_CUDA, P_INT8, L_NHWC), (T_CUDA, P_FLOAT, L_NCHW)]
    sc = lh.score_kernels("conv2d", places, True)
    assert [s for s, _ in sc] == sorted([s for s, _ in sc], reverse=True)
    key, scale = lh.pick("conv2d", True, [True], [0.5])
    assert key == "conv2d/int8_out/cuda/int8_t/NHWC" and scale == 0.5
    key, _ = lh.pick("conv2d", True, [False])
    assert key == "conv2d/fp32_out/cuda/int8_t/NHWC"
