"""JPEG half of the oracle (oracle/jpeg_oracle.c).

Parity against the reference is UNPINNED (its arithmetic lives in the un-vendored
jpeg-encoder-wasm package).  These tests pin the oracle's own
definition against an independent implementation instead: libjpeg-turbo through PIL.
The oracle follows libjpeg's colour transform, islow FDCT, quality scaling and standard
Huffman tables, so a file it writes must decode to exactly the pixels of PIL's own
4:4:4 baseline encoding of the same image -- i.e. identical quantised coefficients.
It also restates the reference's structural JPEG tests (tests/unit/jpeg-structure.test.ts).
"""
import io

import numpy as np
import pytest

import oracle_py as orc

PIL = pytest.importorskip("PIL.Image")


def _photo(rng, w, h):
    yy, xx = np.mgrid[0:h, 0:w]
    img = np.zeros((h, w, 4), np.float64)
    for c in range(3):
        img[..., c] = 128 + 90 * np.sin(xx / (7.0 + 3 * c) + c) * np.cos(yy / (11.0 - 2 * c)) + rng.normal(0, 12, (h, w))
    img[..., 3] = rng.integers(0, 256, (h, w))
    return np.clip(img, 0, 255).astype(np.uint8)


def _pil_encode(rgb, quality):
    buf = io.BytesIO()
    PIL.fromarray(rgb, "RGB").save(buf, "JPEG", quality=quality, subsampling=0, optimize=False)
    return buf.getvalue()


def _segments(data):
    """marker -> list of payloads, up to SOS."""
    out, i = {}, 2
    assert data[:2] == b"\xff\xd8"
    while i < len(data):
        assert data[i] == 0xFF
        m = data[i + 1]
        n = (data[i + 2] << 8) | data[i + 3]
        out.setdefault(m, []).append(data[i + 4:i + 2 + n])
        i += 2 + n
        if m == 0xDA:
            break
    return out


@pytest.mark.parametrize("quality", [1, 25, 50, 75, 85, 90, 95, 100])
def test_tables_match_libjpeg(quality):
    rng = np.random.default_rng(1)
    ref = _segments(_pil_encode(_photo(rng, 16, 16)[..., :3], quality))
    ours = _segments(orc.jpeg_header(16, 16, quality))
    ref_dqt = {p[0]: p[1:65] for seg in ref[0xDB] for p in [seg[k:k + 65] for k in range(0, len(seg), 65)]}
    our_dqt = {p[0]: p[1:65] for seg in ours[0xDB] for p in [seg[k:k + 65] for k in range(0, len(seg), 65)]}
    assert our_dqt == ref_dqt

    def dht(segs):
        tabs = {}
        for seg in segs:
            i = 0
            while i < len(seg):
                n = sum(seg[i + 1:i + 17])
                tabs[seg[i]] = seg[i + 1:i + 17 + n]
                i += 17 + n
        return tabs
    assert dht(ours[0xC4]) == dht(ref[0xC4])
    assert ours[0xC0][0][5:] == ref[0xC0][0][5:]      # 3 components, 1x1 sampling, same table ids


@pytest.mark.parametrize("w,h,quality", [(64, 48, 90), (33, 17, 75), (8, 8, 50), (1, 1, 85), (100, 9, 100), (250, 131, 30)])
def test_decodes_to_libjpeg_pixels(w, h, quality):
    rng = np.random.default_rng(w * 1000 + h)
    rgba = _photo(rng, w, h)
    ours = orc.jpeg_file(rgba.reshape(h, w * 4), w, h, quality)
    theirs = _pil_encode(np.ascontiguousarray(rgba[..., :3]), quality)
    a = np.asarray(PIL.open(io.BytesIO(ours)).convert("RGB"))
    b = np.asarray(PIL.open(io.BytesIO(theirs)).convert("RGB"))
    assert a.shape == (h, w, 3)
    assert np.array_equal(a, b), "oracle coefficients differ from libjpeg-turbo's"


def _pil_scan(rgb, quality):
    """libjpeg-turbo's own baseline encode of rgb (4:4:4, Annex K tables, a restart interval of one MCU row, which is the
    oracle's framing): the entropy-coded bytes between the SOS header and EOI."""
    buf = io.BytesIO()
    try:
        PIL.fromarray(rgb, "RGB").save(buf, "JPEG", quality=quality, subsampling=0, restart_marker_rows=1)
    except TypeError:
        pytest.skip("this Pillow has no restart_marker_rows")
    d = buf.getvalue()
    i = d.index(b"\xff\xda")
    return d[i + 2 + int.from_bytes(d[i + 2:i + 4], "big"):-2]


@pytest.mark.parametrize("seed", range(6))
def test_entropy_bytes_are_libjpeg_turbos(seed):
    """The strongest pin this box allows for kernel (d): with the oracle's framing (DRI = one MCU row) libjpeg-turbo's
    encoder writes exactly the oracle's entropy-coded bytes -- colour transform, islow FDCT, quantiser rounding, edge
    replication, Huffman coding, stuffing and RSTm numbering all included.  Random sizes, contents and qualities 1..100."""
    rng = np.random.default_rng(100 + seed)
    for _ in range(12):
        w, h = int(rng.integers(1, 220)), int(rng.integers(1, 130))
        quality = int(rng.choice([rng.integers(1, 101), 90, 100]))
        y, x = np.mgrid[0:h, 0:w].astype(np.float64)
        rgb = np.stack([np.clip(127 + 80 * np.sin(x / (5 + c)) * np.cos(y / (7 + c)) + rng.normal(0, rng.uniform(0, 25), (h, w)), 0, 255).astype(np.uint8)
                        for c in range(3)], -1)
        rgba = np.concatenate([rgb, np.full((h, w, 1), int(rng.integers(0, 256)), np.uint8)], -1).reshape(h, w * 4)
        assert orc.jpeg_encode_rows(rgba, w, h, 0, h, quality) == _pil_scan(np.ascontiguousarray(rgb), quality), (w, h, quality)


def test_structure_like_reference_tests():
    """tests/unit/jpeg-structure.test.ts:52-167: one SOI at 0, one EOI at the end, decodable, dims, red stays red."""
    w, h = 40, 23
    rgba = np.zeros((h, w, 4), np.uint8)
    rgba[..., 0] = 255
    rgba[..., 3] = 255
    data = orc.jpeg_file(rgba.reshape(h, w * 4), w, h, 85)
    assert data[:2] == b"\xff\xd8" and data[-2:] == b"\xff\xd9"
    assert data.count(b"\xff\xd8") == 1 and data.count(b"\xff\xd9") == 1
    img = np.asarray(PIL.open(io.BytesIO(data)).convert("RGB"))
    assert img.shape == (h, w, 3)
    assert (img[..., 0] > 200).all() and (img[..., 1] < 50).all() and (img[..., 2] < 50).all()


def test_bands_concatenate():
    """MCU rows are independent entropy segments: encoding in bands equals encoding at once."""
    rng = np.random.default_rng(3)
    w, h = 70, 45
    rgba = _photo(rng, w, h).reshape(h, w * 4)
    whole = orc.jpeg_encode_rows(rgba, w, h, 0, h, 90)
    parts = b"".join(orc.jpeg_encode_rows(rgba[a:b], w, h, a, b, 90) for a, b in ((0, 16), (16, 24), (24, 45)))
    assert parts == whole


def test_quantize_rounding():
    l = orc._jpeg_setup()
    for coef, q, want in ((0, 1, 0), (3, 1, 0), (4, 1, 1), (-4, 1, -1), (12, 1, 2), (-12, 1, -2), (127, 16, 1), (63, 16, 0),
                          (64, 16, 1), (-64, 16, -1), (-63, 16, 0)):
        assert l.orc_jpeg_quantize(coef, q) == want
