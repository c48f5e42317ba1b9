"""Known-answer tests of the reference's unit suite, restated against the oracle
(tests/unit/png-filter.test.ts, adam7.test.ts, pixel-ops.test.ts,
alpha-blending.test.ts), plus the integer identities the CUDA kernels rely on."""
import math

import numpy as np
import pytest

import oracle_py as orc

U8 = lambda *v: np.array(v, np.uint8)


# ---- png-filter.test.ts ------------------------------------------------------
def test_unfilter_none():
    assert orc.unfilter_scanline(0, U8(1, 2, 3, 4, 5), None, 1).tolist() == [1, 2, 3, 4, 5]


def test_unfilter_sub():  # :26-37
    assert orc.unfilter_scanline(1, U8(10, 20, 30, 40), None, 2).tolist() == [10, 20, 40, 60]


def test_unfilter_up():  # :39-48
    assert orc.unfilter_scanline(2, U8(5, 10, 15, 20), U8(100, 50, 80, 30), 1).tolist() == [105, 60, 95, 50]
    assert orc.unfilter_scanline(2, U8(5, 10, 15, 20), None, 1).tolist() == [5, 10, 15, 20]


def test_unfilter_average():  # :57-66
    assert orc.unfilter_scanline(3, U8(10, 20), U8(50, 30), 1).tolist() == [35, 52]


def test_unfilter_paeth_and_unknown():  # :68-84
    assert len(orc.unfilter_scanline(4, U8(5, 10), U8(100, 50), 1)) == 2
    with pytest.raises(orc.OracleError, match="Unknown filter type"):
        orc.unfilter_scanline(99, U8(1, 2, 3), None, 1)


@pytest.mark.parametrize("row,prev,bpp", [
    ([100, 150, 200, 250, 50, 75, 125, 175], [110, 140, 190, 240, 60, 80, 130, 170], 2),
    ([50, 100, 150, 200], None, 1),
    ([255, 128, 64, 200, 100, 50, 150, 75, 25], None, 3),
    ([255, 255, 255, 255, 128, 128, 128, 255], [200, 200, 200, 255, 100, 100, 100, 255], 4),
])
def test_filter_roundtrip(row, prev, bpp):  # :95-135
    pv = None if prev is None else U8(*prev)
    ft, f, _ = orc.filter_scanline(U8(*row), pv, bpp)
    assert 0 <= ft <= 4
    assert orc.unfilter_scanline(ft, f, pv, bpp).tolist() == row


def test_filter_tie_order_first_row():
    # row 0: Up == None and Paeth == Sub, the earlier candidate wins (png-filter.ts:174)
    ft, _, sums = orc.filter_scanline(U8(7, 7, 7, 7, 7, 7, 7, 7), None, 4)
    assert sums[0] == sums[2] and sums[1] == sums[4]
    assert ft in (0, 1)


def test_bytes_per_pixel():  # :11-18
    L = orc.lib()
    assert [L.orc_bytes_per_pixel(8, c) for c in (0, 2, 6, 4)] == [1, 3, 4, 2]
    assert L.orc_bytes_per_pixel(16, 0) == 2 and L.orc_bytes_per_pixel(16, 2) == 6
    assert L.orc_bytes_per_pixel(1, 0) == 1 and L.orc_bytes_per_pixel(8, 99) == -1


# ---- adam7.test.ts -------------------------------------------------------------
def test_adam7_2x2_rgba():  # :13-46
    data = U8(0, 10, 20, 30, 40, 0, 50, 60, 70, 80, 0, 90, 100, 110, 120, 130, 140, 150, 160)
    assert orc.deinterlace_adam7(data, 2, 2, 8, 6).tolist() == list(range(10, 170, 10))


def test_adam7_1bit_and_truncated():  # :48-79
    assert orc.deinterlace_adam7(U8(0, 0b10000000), 1, 1, 1, 0).tolist() == [0b10000000]
    with pytest.raises(orc.OracleError, match="Unexpected end of decompressed data"):
        orc.deinterlace_adam7(U8(0), 2, 2, 8, 6)


# ---- pixel-ops.test.ts -----------------------------------------------------------
def test_convert_packed_gray1():  # :192-198
    got = orc.convert_scanline(U8(0b10100000), 4, 1, 0, 8, 6)
    assert got.tolist() == [255, 255, 255, 255, 0, 0, 0, 255, 255, 255, 255, 255, 0, 0, 0, 255]


def test_convert_rgb16_to8():  # :200-204
    got = orc.convert_scanline(U8(0xFF, 0xFF, 0x80, 0x00, 0, 0), 1, 16, 2, 8, 6)
    assert got.tolist() == [255, round(0x8000 * 255 / 65535), 0, 255]


def test_convert_rgba8_to16():  # :205-218
    got = orc.convert_scanline(U8(1, 2, 3, 4), 1, 8, 6, 16, 6)
    assert got.tolist() == [1, 1, 2, 2, 3, 3, 4, 4]


def test_convert_errors():  # :219-225
    with pytest.raises(orc.OracleError, match="Only conversion to RGBA"):
        orc.convert_scanline(U8(1, 2, 3, 4), 1, 8, 6, 8, 2)
    with pytest.raises(orc.OracleError, match="Unsupported source color type"):
        orc.convert_scanline(U8(0), 1, 8, 3, 8, 6)


def test_scale_identities_used_by_kernels():
    """Math.round(v*toMax/fromMax) equals the integer forms in common.cuh for every input."""
    v = np.arange(65536)
    ref16 = np.array([orc.scale_sample(int(x), 16, 8) for x in v])
    assert np.array_equal(ref16, (v + 128) // 257)
    assert all(orc.scale_sample(x, 8, 16) == x * 257 for x in range(256))
    for bits, mul8, mul16 in ((1, 255, 65535), (2, 85, 21845), (4, 17, 4369)):
        for x in range(1 << bits):
            assert orc.scale_sample(x, bits, 8) == x * mul8
            assert orc.scale_sample(x, bits, 16) == x * mul16


# ---- alpha-blending.test.ts --------------------------------------------------------
def test_composite_copy_and_opaque():  # :5-39
    d = orc.composite_scanline(np.zeros(12, np.uint8), U8(255, 0, 0, 255), 1, 1, 4, False)
    assert d[4:8].tolist() == [255, 0, 0, 255]
    d = orc.composite_scanline(U8(0, 0, 255, 255), U8(255, 0, 0, 255), 0, 1, 4, True)
    assert d.tolist() == [255, 0, 0, 255]


def test_composite_semi_over_opaque():  # :41-57
    d = orc.composite_scanline(U8(0, 0, 255, 255), U8(255, 0, 0, 128), 0, 1, 4, True)
    assert d[0] in (127, 128) and d[1] == 0 and d[2] in (127, 128) and d[3] == 255


def test_composite_transparent_source():  # :59-74
    assert orc.composite_scanline(U8(100, 100, 100, 255), U8(255, 0, 0, 0), 0, 1, 4, True).tolist() == [100, 100, 100, 255]


def test_composite_semi_over_semi():  # :76-93
    d = orc.composite_scanline(U8(0, 0, 255, 128), U8(255, 0, 0, 128), 0, 1, 4, True)
    assert 169 <= d[0] <= 171 and d[1] == 0 and 84 <= d[2] <= 86 and 191 <= d[3] <= 192


def test_composite_16bit():  # :121-177
    d = orc.composite_scanline(U8(0, 0, 0, 0, 255, 255, 255, 255), U8(255, 255, 0, 0, 0, 0, 255, 255), 0, 1, 8, True)
    assert d.tolist() == [255, 255, 0, 0, 0, 0, 255, 255]
    d = orc.composite_scanline(U8(0, 0, 0, 0, 255, 255, 255, 255), U8(255, 255, 0, 0, 0, 0, 128, 0), 0, 1, 8, True)
    d = [int(v) for v in d]
    r, b, a = (d[0] << 8) | d[1], (d[4] << 8) | d[5], (d[6] << 8) | d[7]
    assert 32000 <= r <= 33500 and 32000 <= b <= 33500 and a == 65535


def test_blend_thresholds_match_integer_forms():
    """8-bit: only 255 is 'opaque', only 0 is 'transparent'.  16-bit: >= 65529 copies, <= 6 is ignored."""
    assert [a for a in range(256) if a / 255 >= 0.9999] == [255]
    assert [a for a in range(256) if not a / 255 > 0.0001] == [0]
    assert min(a for a in range(65536) if a / 65535 >= 0.9999) == 65529
    assert max(a for a in range(65536) if not a / 65535 > 0.0001) == 6
    # and the oracle agrees on the behaviour at the edges
    dst = U8(0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF)   # opaque white
    for a, copied in ((65529, True), (65528, False)):
        src = U8(0, 0, 0, 0, 0, 0, a >> 8, a & 255)              # almost opaque black
        out = orc.composite_scanline(dst, src, 0, 1, 8, True).tolist()
        assert (out == src.tolist()) == copied                      # a blend leaves alpha 65535 and colour 6..7
    for a, ignored in ((6, True), (7, False)):
        src = U8(0, 0, 0, 0, 0, 0, a >> 8, a & 255)
        out = orc.composite_scanline(dst, src, 0, 1, 8, True).tolist()
        assert (out == dst.tolist()) == ignored


def test_blend_against_python_floats():
    """The oracle's f64 blend equals the same expression evaluated by CPython doubles."""
    rng = np.random.default_rng(7)

    def js_round(x):
        r = math.floor(x)
        return r + 1 if x - r >= 0.5 else r

    for _ in range(20000):
        s = rng.integers(0, 256, 4)
        d = rng.integers(0, 256, 4)
        out = orc.composite_scanline(d.astype(np.uint8), s.astype(np.uint8), 0, 1, 4, True).tolist()
        sa = s[3] / 255
        if sa >= 0.9999:
            want = s.tolist()
        elif sa > 0.0001:
            da = d[3] / 255
            oa = sa + da * (1 - sa)
            want = [js_round(max(0, min(255, (int(s[c]) * sa + int(d[c]) * da * (1 - sa)) / oa))) for c in range(3)]
            want.append(js_round(oa * 255))
        else:
            want = d.tolist()
        assert out == want


def test_division_by_constant_is_exact():
    """csrc/common.cuh div_const(): x/255 and x/65535 through q0 = x*y, r = fma(-b, q0, x), q = fma(r, y, q0)
    with y = RN(1/b) must equal IEEE division for every integer x the blend can see (sA, dA of
    src/pixel-ops.ts:694-700).  fma is emulated exactly with rationals."""
    from fractions import Fraction

    def fma(a, b, c):
        return float(Fraction(a) * Fraction(b) + Fraction(c))   # Fraction -> float rounds to nearest even

    for b, y_hex in ((255.0, "0x1.0101010101010p-8"), (65535.0, "0x1.0001000100010p-16")):
        y = float.fromhex(y_hex)
        assert y == 1.0 / b
        for k in range(int(b) + 1):
            a = float(k)
            q0 = a * y
            q = fma(fma(-b, q0, a), y, q0)
            assert q == a / b, (b, k)
