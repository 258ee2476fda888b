"""GPU parity (-m gpu): int8 NHWC channel movement (concat / split / shuffle_channel / relu,
SURVEY 8f rank 1) through the C-ABI against the oracle's restatement of the reference's NCHW
loops (oracle/oracle.c: orc_concat_nchw, orc_split_nchw, orc_shuffle_channel_nchw, orc_relu_i8).
Shapes: the reference's own test (lite/tests/kernels/shuffle_channel_compute_test.cc: 10x16x4x4,
group 4 -- and its concat test's ragged channel counts), the ShuffleNet-v2 shapes of
humanseg_shfv75_0119_quant (C = 18 / 36 / 72 / 144, group 2) and wide 16-byte-aligned tensors.
Everything is byte movement: bit-exact."""
import numpy as np
import pytest
import torch

import orc

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).to(DEV)


def host(t):
    torch.cuda.synchronize()
    return t.cpu().numpy()


def nchw(x):
    return np.ascontiguousarray(np.moveaxis(x, -1, 1))


def nhwc(x):
    return np.ascontiguousarray(np.moveaxis(x, 1, -1))


def rand_i8(rng, shape):
    return rng.integers(-128, 128, shape, dtype=np.int8)


CONCAT_CASES = [
    ((2, 5, 7), (64, 64)),          # 16-byte chunks
    ((1, 24, 24), (36, 36)),        # humanseg: 4-byte chunks
    ((3, 4, 4), (9, 18, 3)),        # ragged: byte chunks
    ((2, 3, 5), (16, 32, 48, 16, 64, 16, 32, 32)),  # eight parts
    ((1, 12, 12), (72,)),           # single part = copy
    ((10, 4, 4), (16, 16)),
]


@pytest.mark.parametrize("case", CONCAT_CASES)
def test_concat_split_parity(pl, case):
    pix, cs = case
    rng = np.random.default_rng(sum(cs) + len(cs))
    xs = [rand_i8(rng, pix + (c,)) for c in cs]
    want = nhwc(orc.concat_nchw([nchw(x) for x in xs]))
    got = host(pl.concat([dev(x) for x in xs]))
    assert np.array_equal(got, want)
    # split is the inverse, checked against the oracle's split of the same wide tensor
    parts_want = [nhwc(p) for p in orc.split_nchw(nchw(want), list(cs))]
    parts_got = pl.split(dev(want), list(cs))
    for g, w, x in zip(parts_got, parts_want, xs):
        assert np.array_equal(host(g), w)
        assert np.array_equal(w, x)


def test_concat_from_channel_slices_and_into_a_slice(pl):
    """Parts and the wide tensor may be channel slices of wider buffers (pointer offset + pitch)."""
    rng = np.random.default_rng(5)
    a = rand_i8(rng, (2, 6, 6, 96))
    b = rand_i8(rng, (2, 6, 6, 40))
    ad, bd = dev(a), dev(b)
    y = torch.zeros((2, 6, 6, 128), dtype=torch.int8, device=DEV)
    pl.concat([ad[..., 32:64], bd[..., 4:24]], y=y[..., 16:68])
    want = np.zeros((2, 6, 6, 128), np.int8)
    want[..., 16:48] = a[..., 32:64]
    want[..., 48:68] = b[..., 4:24]
    assert np.array_equal(host(y), want)
    outs = pl.split(ad[..., 8:88], [16, 64])
    assert np.array_equal(host(outs[0]), a[..., 8:24])
    assert np.array_equal(host(outs[1]), a[..., 24:88])


SHUFFLE_CASES = [
    ((10, 4, 4), 16, 4),     # the reference's test shape
    ((2, 12, 12), 72, 2),    # humanseg, 4-byte path (C/group = 36)
    ((2, 24, 24), 36, 2),    # C/group = 18: 4-byte path
    ((1, 48, 48), 18, 2),    # C/group = 9 (odd): bytewise
    ((1, 12, 12), 144, 2),   # C/group = 72: 16-byte path
    ((2, 7, 9), 128, 2),
    ((2, 7, 9), 128, 8),
    ((3, 5, 5), 12, 3),
    ((1, 3, 3), 7, 7),
    ((1, 3, 3), 6, 1),
]


@pytest.mark.parametrize("case", SHUFFLE_CASES)
def test_shuffle_channel_parity(pl, case):
    pix, Cc, g = case
    rng = np.random.default_rng(Cc * 31 + g)
    x = rand_i8(rng, pix + (Cc,))
    want = nhwc(orc.shuffle_channel_nchw(nchw(x), g))
    got = host(pl.shuffle_channel(dev(x), g))
    assert np.array_equal(got, want)
    # known answer: channel j * g + i of the output is channel i * (C / g) + j of the input
    i, j = (g - 1), (Cc // g - 1) // 2
    assert np.array_equal(got[..., j * g + i], x[..., i * (Cc // g) + j])


def test_shuffle_channel_pitched(pl):
    rng = np.random.default_rng(9)
    x = rand_i8(rng, (2, 5, 5, 80))
    y = torch.zeros((2, 5, 5, 96), dtype=torch.int8, device=DEV)
    pl.shuffle_channel(dev(x)[..., 4:76], 2, y=y[..., 8:80])
    want = nhwc(orc.shuffle_channel_nchw(nchw(x[..., 4:76]), 2))
    got = host(y)
    assert np.array_equal(got[..., 8:80], want)
    assert not got[..., :8].any() and not got[..., 80:].any()


def test_shuffle_channel_rejects(pl):
    x = torch.zeros((1, 2, 2, 10), dtype=torch.int8, device=DEV)
    with pytest.raises(pl.PlError):
        pl.shuffle_channel(x, 3)          # C % group
    with pytest.raises(pl.PlError):
        pl.shuffle_channel(x, 2, y=x)     # in place


@pytest.mark.parametrize("n,hi,off", [(1 << 20, 127, 0), (1000003, 47, 0), (4099, 127, 3), (5, 6, 1)])
def test_relu_i8_parity(pl, n, hi, off):
    rng = np.random.default_rng(n + hi)
    x = rand_i8(rng, (n + off,))
    xd = dev(x)[off:]
    yd = torch.empty(n + off, dtype=torch.int8, device=DEV)[off:]
    pl.relu_i8(xd, hi, y=yd)
    assert np.array_equal(host(yd), orc.relu_i8(x[off:], hi))


def test_shufflenet_unit_round_trip_full_size(pl):
    """Size-independent properties at a BASELINE-sized activation (256 x 28 x 28 x 256):
    split(concat(a, b)) = (a, b); shuffle(group g) then shuffle(group C / g) is the identity;
    concat + shuffle(2) interleaves the two halves channel by channel."""
    g = torch.Generator(device=DEV).manual_seed(3)
    a = torch.randint(-128, 128, (256, 28, 28, 128), dtype=torch.int8, device=DEV, generator=g)
    b = torch.randint(-128, 128, (256, 28, 28, 128), dtype=torch.int8, device=DEV, generator=g)
    wide = pl.concat([a, b])
    pa, pb = pl.split(wide, [128, 128])
    assert torch.equal(pa, a) and torch.equal(pb, b)
    sh = pl.shuffle_channel(wide, 2)
    assert torch.equal(sh[..., 0::2], a) and torch.equal(sh[..., 1::2], b)
    back = pl.shuffle_channel(sh, 128)
    assert torch.equal(back, wide)
    r = pl.relu_i8(wide)
    assert torch.equal(r, torch.clamp(wide, min=0))


@pytest.mark.parametrize("pix,ch", [((2, 7, 9), 64), ((1, 24, 24), 18), ((3, 5, 5), 9), ((2, 4, 4), 8)])
def test_concat_shuffle2_equals_the_two_reference_ops(pl, pix, ch):
    """pl_concat_shuffle2_i8_nhwc = concat (axis C) then shuffle_channel(group 2), in one pass
    (16-byte, 4-byte and unaligned flavours)."""
    rng = np.random.default_rng(ch)
    a, b = rand_i8(rng, pix + (ch,)), rand_i8(rng, pix + (ch,))
    want = nhwc(orc.shuffle_channel_nchw(orc.concat_nchw([nchw(a), nchw(b)]), 2))
    got = host(pl.concat_shuffle2(dev(a), dev(b)))
    assert np.array_equal(got, want)
    assert np.array_equal(got[..., 0::2], a) and np.array_equal(got[..., 1::2], b)
    # sources as channel slices of one wider tensor (the stride-1 unit reads x1 from its own input)
    wide = rand_i8(rng, pix + (4 * ch,))
    wd = dev(wide)
    got = host(pl.concat_shuffle2(wd[..., :ch], wd[..., 2 * ch:3 * ch]))
    assert np.array_equal(got[..., 0::2], wide[..., :ch]) and np.array_equal(got[..., 1::2], wide[..., 2 * ch:3 * ch])


def test_concat_and_shuffle_beyond_2g_items(pl):
    """Maximum sizes: more than 2^31 work items (unaligned 1 + 2 channel parts -> byte chunks) go out in
    slabs of whole pixels; checked against torch on the device (size-independent property: concat = cat)."""
    pixels = (1 << 31) // 3 + 1000  # 3 byte-items per pixel -> just over 2^31 items
    g = torch.Generator(device=DEV).manual_seed(11)
    a = torch.randint(-128, 128, (pixels, 1), dtype=torch.int8, device=DEV, generator=g)
    b = torch.randint(-128, 128, (pixels, 2), dtype=torch.int8, device=DEV, generator=g)
    y = pl.concat([a, b])
    assert y.shape == (pixels, 3)
    assert torch.equal(y[:, :1], a) and torch.equal(y[:, 1:], b)
    pa, pb = pl.split(y, [1, 2])
    assert torch.equal(pa, a) and torch.equal(pb, b)
    del pa, pb, a, b
    sh = pl.shuffle_channel(y, 3)  # group == C: identity permutation through the bytewise path
    assert torch.equal(sh, y)
