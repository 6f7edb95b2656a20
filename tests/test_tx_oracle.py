"""TX direction without a GPU: the oracle's restatement of caribou_smi_generate_data against the
reference's committed behaviour (a constant test pattern: the one vector it pins), against the bit layout its
comment documents (caribou_smi.c:693-696), and the Stream::WriteSamples conversions against numpy."""
import os

import numpy as np
import pytest

GOLD = os.path.join(os.path.dirname(__file__), "golden", "tx_golden.npz")


def layout_words(iq):
    """The layout comment taken literally, byte by byte (independent of the oracle's shift sequence)."""
    i = iq[:, 0].astype(np.int32) & 0x1FFF
    q = iq[:, 1].astype(np.int32) & 0x1FFF
    b0 = 0xE0 | (i >> 8)                      # [SOF TXC CTX I12 I11 I10 I9 I8]
    b1 = (i >> 1) & 0x7F                      # [0 I7 .. I1]
    b2 = ((i & 1) << 6) | (q >> 7)            # [0 I0 Q12 .. Q7]
    b3 = q & 0x7F                             # [0 Q6 .. Q0]
    return (b0 | (b1 << 8) | (b2 << 16) | (b3 << 24)).astype(np.uint32)      # byte 0 at the lowest address


def test_reference_pattern_is_the_word_for_minus_one_zero(po):
    g = np.load(GOLD)
    w = po.tx_pack(np.array([[-1, 0]], dtype=np.int16))[0]
    for n in (1, 5, 8192, 8192 * 3 + 77):
        assert int(g["ret_%d" % n]) == n                      # caribou_smi_write returns the samples it wrote
        words = g["words_%d" % n]
        assert words.shape == (n,)
        assert np.all(words == w)
    assert w == 0x00407FFF


def test_live_reference_when_present(po):
    if not po.ref_available():
        pytest.skip("oracle/_ref not built")
    dev = po.RefDev(np.zeros(4, dtype=np.uint32), 0, 524288)
    n = 131072 * 2 + 1000
    ret, words = dev.smi_write(np.zeros((n, 2), dtype=np.int16))
    assert ret == n and words.size == n
    assert np.all(words == po.tx_pack(np.array([[-1, 0]], dtype=np.int16))[0])


def test_pack_matches_the_documented_layout(po):
    # every 13-bit code on each rail against every code of a coarse grid on the other, plus int16 values beyond 13 bits
    codes = np.arange(-4096, 4096, dtype=np.int16)
    grid = np.array([-4096, -4095, -1, 0, 1, 2, 127, 128, 255, 256, 4095], dtype=np.int16)
    iq = np.concatenate([np.stack(np.meshgrid(codes, grid, indexing="ij"), -1).reshape(-1, 2),
                         np.stack(np.meshgrid(grid, codes, indexing="ij"), -1).reshape(-1, 2),
                         np.array([[32767, -32768], [-32768, 32767], [8191, -8192], [4096, -4097]], dtype=np.int16)])
    w = po.tx_pack(iq)
    np.testing.assert_array_equal(w, layout_words(iq))
    assert np.all(w & 0x000000E0 == 0xE0)                     # SOF | TXC | CTX in byte 0
    assert np.all(w & 0x80808000 == 0)                        # the three always-zero bits (bytes 1, 2, 3)


def test_tx_rx_field_round_trip(po):
    """The 13-bit fields survive TX pack -> field extraction with sign extension (what the modem sees)."""
    rng = np.random.default_rng(5)
    iq = rng.integers(-4096, 4096, size=(4096, 2)).astype(np.int16)
    w = po.tx_pack(iq).astype(np.uint32)
    s = w.byteswap().astype(np.int64)                         # back to the reference's `s` before bswap
    i13 = ((s >> 24) & 0x1F) << 8 | ((s >> 16) & 0x7F) << 1 | ((s >> 14) & 1)
    q13 = ((s >> 8) & 0x3F) << 7 | (s & 0x7F)
    sx = lambda v: ((v ^ 0x1000) - 0x1000)
    np.testing.assert_array_equal(sx(i13), iq[:, 0])
    np.testing.assert_array_equal(sx(q13), iq[:, 1])


def test_write_conversions(po):
    rng = np.random.default_rng(6)
    x = (rng.random((5000, 2)) * 2 - 1).astype(np.float32)
    np.testing.assert_array_equal(po.tx_to_cs16(x, "CF32"), np.trunc(x * np.float32(4096.0)).astype(np.int16))
    xd = rng.random((5000, 2)) * 2 - 1
    np.testing.assert_array_equal(po.tx_to_cs16(xd, "CF64"), np.trunc(xd * 4096.0).astype(np.int16))
    b = rng.integers(-128, 128, size=(5000, 2)).astype(np.int8)
    np.testing.assert_array_equal(po.tx_to_cs16(b, "CS8"), b.astype(np.int16) << 5)
    # beyond int16 the cast keeps the low 16 bits of the int32; beyond int32 it gives 0 (x86-64 behaviour)
    edge = np.array([[8.0, -8.0], [7.99999, -8.0001], [1e9, -1e9], [524288.0, np.nan]], dtype=np.float32)
    got = po.tx_to_cs16(edge, "CF32")
    v = np.trunc(edge.astype(np.float64) * 4096.0)
    ok = np.isfinite(v) & (np.abs(v) < 2 ** 31)
    exp = np.where(ok, ((np.where(ok, v, 0).astype(np.int64) + 2 ** 15) % 2 ** 16) - 2 ** 15, 0)
    np.testing.assert_array_equal(got, exp.astype(np.int16))
