"""Space-to-depth stems with two output rows per tile (IgemmParams::mpair in mode 3: TA + 1 input spans, two accumulators,
one hand-off and one 4-D TMA store of both rows).  Forced on every eligible shape (PL_S2D_PAIR=2), checked to have taken
the pair path, compared bit for bit with the oracle; shapes with an odd number of output rows or Q % 8 != 0 must fall
back to single rows and still match."""
import numpy as np
import pytest

import orc
from test_gpu_parity import dev, host

pytestmark = pytest.mark.gpu

CASES = [
    # N, H, W, C, K, R, pads, pairs expected
    (2, 64, 64, 3, 64, 7, (3, 3, 3, 3), True),     # ResNet stem geometry, P = Q = 32
    (3, 64, 224, 3, 32, 3, (1, 1, 1, 1), True),    # MobileNet stem, Q = 112
    (5, 20, 256, 4, 16, 3, (1, 1, 1, 1), True),    # Q = 128: every accumulator row is an output; odd image count
    (2, 224, 224, 3, 64, 7, (3, 3, 3, 3), True),   # the full ImageNet stem
    (3, 30, 34, 3, 32, 3, (1, 1, 1, 1), False),    # P = 15: single rows
    (1, 64, 44, 3, 64, 7, (3, 3, 3, 3), False),    # Q = 22: single rows
]


@pytest.mark.parametrize("case", CASES)
@pytest.mark.parametrize("act", ["relu", "none"])
def test_s2d_row_pairs_match_oracle(pl, monkeypatch, case, act):
    N, H, W, Cc, K, R, pads, want_pairs = case
    monkeypatch.setenv("PL_S2D_PAIR", "2")
    pl.refresh_knobs()
    try:
        rng = np.random.default_rng(N + H + K)
        g = pl.s2d_geometry(N, H, W, Cc, R, R, (2, 2), pads)
        assert g.supported
        xf = rng.uniform(-1, 1, (N, Cc, H, W)).astype(np.float32)
        w = rng.integers(-127, 128, (K, R, R, Cc), dtype=np.int8)
        ws = ((1 + np.arange(K) % 7) / (7 * 127)).astype(np.float32)
        bias = rng.uniform(-1, 1, K).astype(np.float32)
        sf, bf, _ = orc.fold_scales(ws, K, 1 / 127, Cc * R * R / 127 / 4, orc.OUT_INT8, bias)
        a_o, a_p = (orc.ACT_RELU, pl.ACT_RELU) if act == "relu" else (orc.ACT_NONE, pl.ACT_NONE)
        want = orc.conv2d(orc.nchw_to_nhwc(orc.quantize(xf, 1 / 127)), w, sf, bf, (2, 2), pads, act=a_o)
        xs = pl.quantize_nchw_to_s2d(dev(xf), 1 / 127, g)
        got = host(pl.s2d_conv2d(xs, pl.pack_s2d_conv_weight(dev(w), g), N, H, W, Cc, K, R, R, dev(sf), dev(bf), (2, 2),
                                 pads, act=a_p))
        assert bool(pl.last_igemm_plan()["tile_pairs"]) == want_pairs
        assert np.array_equal(got, want)
    finally:
        monkeypatch.undo()
        pl.refresh_knobs()
