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


@pytest.mark.parametrize("seed", range(4))
def test_oracle_matches_libjpeg_turbo_on_random_files(seed):
    """Where Pillow is at hand: random sizes, contents, qualities, subsamplings, optimised tables and restart intervals,
    encoded AND decoded by libjpeg-turbo, against the oracle's pixels (the GPU suite runs the device decoder through the
    same mill, tools/jpeg_decode_fuzz.py)."""
    import io
    Image = pytest.importorskip("PIL.Image")
    rng = np.random.default_rng(300 + seed)
    done = 0
    for _ in range(40):
        w, h = int(rng.integers(1, 260)), int(rng.integers(1, 200))
        gray = rng.integers(0, 5) == 0
        y, x = np.mgrid[0:h, 0:w].astype(np.float64)
        chans = [np.clip(127 + rng.uniform(10, 100) * np.sin(x / rng.uniform(2, 40) + c) * np.cos(y / rng.uniform(2, 40))
                         + rng.normal(0, rng.uniform(0, 30), (h, w)), 0, 255).astype(np.uint8) for c in range(1 if gray else 3)]
        pic = chans[0] if gray else np.stack(chans, -1)
        opts = dict(quality=int(rng.choice([rng.integers(1, 101), 75, 95, 100])))
        if not gray:
            opts["subsampling"] = int(rng.integers(0, 3))
        if rng.integers(0, 3) == 0:
            opts["optimize"] = True
        r = rng.integers(0, 4)
        if r == 1:
            opts["restart_marker_blocks"] = int(rng.integers(1, 40))
        elif r == 2:
            opts["restart_marker_rows"] = int(rng.integers(1, 4))
        buf = io.BytesIO()
        try:
            Image.fromarray(pic, "L" if gray else "RGB").save(buf, "JPEG", **opts)
        except (TypeError, OSError):   # an older Pillow without restart_marker_*, or its encoder giving up on the option set
            continue
        data = buf.getvalue()
        want = np.array(Image.open(io.BytesIO(data)))
        got = orc.jpeg_decode(data)
        assert np.array_equal(got if gray else got[..., :3], want), (w, h, gray, opts)
        done += 1
    assert done > 20
