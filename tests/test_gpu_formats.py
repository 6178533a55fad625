"""SURVEY 8(f) row f4, data formats: descriptor <-> saved-map text (okvis_ceres/src/Component.cpp:228-236, 434-437)
and FBrisk::meanValue (okvis_frontend/src/FBrisk.cpp:25-60), CUDA path through the C ABI against the oracle port."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx(mod):
    return mod.Context(0)


@pytest.fixture(scope="module")
def port():
    from oracle.refbind import Port
    return Port()


@pytest.mark.parametrize("n", [0, 1, 7, 700, 20000])
def test_to_hex_and_back(ctx, port, n):
    d = np.random.default_rng(n).integers(0, 256, (n, 48), dtype=np.uint8)
    got = ctx.descriptors_to_hex(d)
    assert got == port.descriptors_to_hex(d)
    assert got == [bytes(r).hex() for r in d]
    back = ctx.descriptors_from_hex(got)
    assert back.shape == (n, 48) and np.array_equal(back, d)
    assert np.array_equal(ctx.descriptors_from_hex([s.upper() for s in got]), d)


def test_from_hex_every_two_character_field(ctx, port):
    """All 65 536 fields, as the reference's `stringstream >> hex >> uint32_t` reads them (signs, "0x", one digit,
    white space, no digit)."""
    fields = np.arange(65536, dtype=np.uint32)
    chars = np.stack([fields >> 8, fields & 255], axis=1).astype(np.uint8).reshape(-1)
    pad = (-len(chars)) % 96
    chars = np.concatenate([chars, np.full(pad, ord("0"), np.uint8)])
    rows = [bytes(r) for r in chars.reshape(-1, 96)]
    want = port.descriptors_from_hex(rows)
    got = ctx.descriptors_from_hex(rows)
    # a field of two white-space characters ends the stream inside the sentry: the reference's `byte` stays
    # uninitialised (no value to compare with); the CUDA path defines it as 0
    ws = np.isin(chars.reshape(-1, 2), [9, 10, 11, 12, 13, 32]).all(axis=1).reshape(-1, 48)
    assert ws.sum() == 36 and not got[ws].any()
    bad = np.argwhere((got != want) & ~ws)
    assert len(bad) == 0, "first differing field: %r" % (bytes(chars[2 * (bad[0][0] * 48 + bad[0][1]):][:2]),)
    assert len({int(v) for v in want.reshape(-1)}) == 256


def test_from_hex_short_string_raises(ctx):
    with pytest.raises(IndexError):
        ctx.descriptors_from_hex(["00" * 47])


@pytest.mark.parametrize("seed", [0, 1, 2])
def test_descriptor_mean(ctx, port, seed):
    rng = np.random.default_rng(seed)
    sizes = [0, 1, 2, 3, 4, 5, 10, 11, 64, 333, 1000, 0, 2049][: 13 - seed]
    rng.shuffle(sizes)
    gs = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    centre = rng.integers(0, 256, (len(sizes), 48), dtype=np.uint8)
    d = np.zeros((gs[-1], 48), np.uint8)
    for g in range(len(sizes)):
        flips = np.packbits(rng.random((sizes[g], 384)) < 0.35, axis=1)   # ties and near-ties occur for small groups
        d[gs[g]:gs[g + 1]] = centre[g] ^ flips
    got = ctx.descriptor_mean(d, gs)
    want = port.descriptor_mean(d, gs)
    assert np.array_equal(got, want)
    bits = np.unpackbits(d, axis=1, bitorder="little")
    for g in range(len(sizes)):
        ref = np.packbits(bits[gs[g]:gs[g + 1]].sum(0) > sizes[g] // 2, bitorder="little")
        assert np.array_equal(got[g], ref)


def test_descriptor_mean_rejects_decreasing_groups(ctx, mod):
    with pytest.raises(mod.OkcError):
        ctx.descriptor_mean(np.zeros((4, 48), np.uint8), [0, 3, 2])
