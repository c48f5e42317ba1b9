"""The JPEG decode oracle (oracle/jpeg_decode_oracle.c) against the pixels libjpeg-turbo decodes -- what the reference
gets from sharp on Node (src/decoders/jpeg-decoder.ts:196-215).  Fixtures: tests/golden/jpeg_decode, made by
tests/golden/make_jpeg_decode_golden.py with Pillow 12.2.0 / libjpeg-turbo 3.1.4.1.  CPU only."""
import os

import numpy as np
import pytest

import oracle_py as orc

HERE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "jpeg_decode")
NAMES = sorted(n[:-4] for n in os.listdir(HERE) if n.endswith(".jpg"))


def _file(name):
    with open(os.path.join(HERE, name + ".jpg"), "rb") as f:
        return f.read()


@pytest.fixture(scope="module")
def decoded():
    return np.load(os.path.join(HERE, "decoded.npz"))


@pytest.mark.parametrize("name", [n for n in NAMES if not n.startswith("p")])
def test_oracle_matches_libjpeg_turbo(name, decoded):
    want = decoded[name]
    got = orc.jpeg_decode(_file(name))
    if want.ndim == 2:
        assert got.shape == want.shape and np.array_equal(got, want)
    else:
        assert got.shape == want.shape[:2] + (4,)
        assert np.array_equal(got[..., :3], want), f"{np.count_nonzero(got[..., :3] != want)} samples differ"
        assert (got[..., 3] == 255).all()


def test_progressive_is_refused():
    with pytest.raises(orc.JpegDecodeError, match="only baseline"):
        orc.jpeg_decode(_file("p420_progressive"))


def test_header_errors_read_like_the_reference():
    """src/decoders/jpeg-decoder.ts:46-94"""
    with pytest.raises(orc.JpegDecodeError, match="Invalid JPEG: missing SOI marker"):
        orc.jpeg_decode(b"\x89PNG\r\n\x1a\n")
    with pytest.raises(orc.JpegDecodeError, match="Invalid JPEG: no SOF marker found"):
        orc.jpeg_decode(b"\xff\xd8\xff\xe0\x00\x04ab\xff\xd9")
    data = _file("c444_q90")
    sof = data.index(b"\xff\xc0")
    with pytest.raises(orc.JpegDecodeError, match="Invalid JPEG: truncated SOF marker"):
        orc.jpeg_decode(data[:sof + 6])


def test_truncated_and_corrupt_streams():
    data = _file("c420_q80")
    with pytest.raises(orc.JpegDecodeError, match="ends early"):
        orc.jpeg_decode(data[:len(data) - 300])
    rst = bytearray(_file("c420_rst3"))
    k = rst.index(b"\xff\xd1")
    rst[k + 1] = 0xD5
    with pytest.raises(orc.JpegDecodeError, match="restart marker"):
        orc.jpeg_decode(bytes(rst))


import jpeg_variants as JV


@pytest.mark.parametrize("name", sorted(JV.variants()))
def test_marker_segment_variants(name, decoded):
    """Files that differ from a fixture in their marker segments only: same pixels, or the error that names the reason."""
    data, same_as, err = JV.variants()[name]
    if err:
        with pytest.raises(orc.JpegDecodeError, match=err.replace("(", r"\(").replace(")", r"\)")):
            orc.jpeg_decode(data)
    else:
        assert np.array_equal(orc.jpeg_decode(data)[..., :3], decoded[same_as])
