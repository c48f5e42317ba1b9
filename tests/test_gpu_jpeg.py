"""Kernel (d) parity: GPU JPEG entropy bytes vs the oracle (bit-exact), and the
reference's structural JPEG tests (tests/unit/jpeg-structure.test.ts:52-237,
tests/integration/jpeg-output.test.ts:29-272) restated through the public API."""
import io

import numpy as np
import pytest

import helpers as H
import oracle_py as orc
import image_stitch_b200 as stitch
from image_stitch_b200 import _native
from image_stitch_b200.jpeg_container import JpegBandSplicer, jpeg_header

pytestmark = pytest.mark.gpu
PIL = pytest.importorskip("PIL.Image")


def _photo(rng, w, h):
    yy, xx = np.mgrid[0:h, 0:w]
    img = np.zeros((h, w, 4), np.float64)
    for c in range(3):
        img[..., c] = 128 + 90 * np.sin(xx / (7.0 + 3 * c) + c) * np.cos(yy / (11.0 - 2 * c)) + rng.normal(0, 12, (h, w))
    img[..., 3] = 255
    return np.clip(img, 0, 255).astype(np.uint8)


@pytest.fixture(params=["one_pass_first", "j1_to_j7"], autouse=True)
def jpeg_route(request, monkeypatch):
    """Every test on both routes of csrc/jpeg.cu: the one-pass MCU kernel (with the J1..J7 kernels behind it for
    content it cannot hold), and J1..J7 alone."""
    if request.param == "j1_to_j7":
        monkeypatch.setenv("STITCH_JPEG_ONE_PASS", "0")
    else:
        monkeypatch.delenv("STITCH_JPEG_ONE_PASS", raising=False)
    return request.param


def _gpu_entropy(ctx, rgba, w, h, quality, bands):
    job = _native.Job(ctx, w, h, 8, _native.OUT_JPEG_ENTROPY, False, bytes(4), jpeg_quality=quality)
    job.add_tile(rgba.tobytes(), width=w, height=h, bit_depth=8, color_type=6, interlace=0, filtered=0,
                 dst_x=0, dst_y=0, z_index=0, image_idx=0)
    sp = JpegBandSplicer()
    out = b""
    try:
        for a, b in bands:
            data, _ = job.run_band(a, b)
            out += sp.push(data)
    finally:
        job.close()
    return out


@pytest.mark.parametrize("w,h,quality", [(64, 48, 90), (33, 17, 75), (8, 8, 50), (1, 1, 85), (100, 9, 100),
                                         (250, 131, 30), (2051, 77, 90), (640, 480, 1)])
def test_entropy_bytes_match_oracle(ctx, w, h, quality):
    rng = np.random.default_rng(w * 31 + h)
    rgba = _photo(rng, w, h).reshape(h, w * 4)
    want = orc.jpeg_encode_rows(rgba, w, h, 0, h, quality)
    got = _gpu_entropy(ctx, rgba, w, h, quality, [(0, h)])
    assert got == want


@pytest.mark.parametrize("w,h,quality", [(64, 48, 90), (8, 8, 50), (1, 1, 85), (1024, 8, 90), (1025, 16, 90), (1032, 24, 50),
                                         (2051, 77, 75), (3000, 40, 60), (4099, 9, 85), (257, 300, 88)])
def test_one_pass_route_alone(ctx, monkeypatch, w, h, quality):
    """STITCH_JPEG_ONE_PASS=2: no second route behind the MCU kernel, so these bytes are its own.  Widths on both
    sides of the 1024-pixel chunks (the DC predictor crosses from one CTA's chunk into the next)."""
    monkeypatch.setenv("STITCH_JPEG_ONE_PASS", "2")
    rng = np.random.default_rng(w * 7 + h)
    rgba = _photo(rng, w, h).reshape(h, w * 4)
    want = orc.jpeg_encode_rows(rgba, w, h, 0, h, quality)
    assert _gpu_entropy(ctx, rgba, w, h, quality, [(0, h)]) == want


def test_one_pass_route_reports_what_it_cannot_hold(ctx, monkeypatch):
    monkeypatch.setenv("STITCH_JPEG_ONE_PASS", "2")
    rng = np.random.default_rng(12)
    w, h = 264, 16
    rgba = rng.integers(0, 256, size=(h, w * 4), dtype=np.uint8)
    with pytest.raises(_native.StitchError, match="bits in an MCU"):
        _gpu_entropy(ctx, rgba, w, h, 100, [(0, h)])


def test_smooth_and_noisy_bands_in_one_job(ctx):
    """A job whose first band the MCU kernel holds and whose second it does not (noise at quality 100)."""
    rng = np.random.default_rng(13)
    w, h = 1100, 32
    rgba = _photo(rng, w, h).reshape(h, w * 4).copy()
    rgba[16:] = rng.integers(0, 256, size=(16, w * 4), dtype=np.uint8)
    want = orc.jpeg_encode_rows(rgba, w, h, 0, h, 100)
    assert _gpu_entropy(ctx, rgba, w, h, 100, [(0, 16), (16, 32)]) == want
    assert _gpu_entropy(ctx, rgba, w, h, 100, [(0, 32)]) == want


def _mosaic_entropy(ctx, tiles, w, h, quality, bands):
    """tiles: (rgba array (th, tw, 4), dst_x, dst_y) pasted in order."""
    job = _native.Job(ctx, w, h, 8, _native.OUT_JPEG_ENTROPY, False, bytes(4), jpeg_quality=quality)
    sp = JpegBandSplicer()
    out = b""
    try:
        for k, (px, x, y, *clip) in enumerate(tiles):
            th, tw = px.shape[:2]
            extra = dict(zip(("src_off_x", "src_off_y", "vis_w", "vis_h"), clip[0])) if clip else {}
            job.add_tile(px.tobytes(), width=tw, height=th, bit_depth=8, color_type=6, interlace=0, filtered=0,
                         dst_x=x, dst_y=y, z_index=0, image_idx=k, **extra)
        for a, b in bands:
            data, _ = job.run_band(a, b)
            out += sp.push(data)
    finally:
        job.close()
    return out


@pytest.mark.parametrize("layout", ["grid", "tall_and_short", "gap", "overlap", "odd_x", "ragged_bottom", "clipped", "odd_y", "bricks"])
def test_mosaics_read_in_place_or_through_a_canvas(ctx, monkeypatch, layout):
    """An exact mosaic of RGBA8 tiles is coded straight from the tiles (no canvas), 16 bytes at a time where the rows
    allow and pixel by pixel at tile edges and odd sizes; a gap or an overlap goes through the compositor.  Either way: the oracle's bytes, and the same bytes
    with the in-place route switched off."""
    rng = np.random.default_rng(21)
    w, h = 2304, 96
    def t(tw, th):
        return _photo(rng, tw, th)
    if layout == "grid":
        tiles = [(t(1152, 48), x, y) for y in (0, 48) for x in (0, 1152)]
    elif layout == "tall_and_short":
        tiles = [(t(1000, 96), 0, 0), (t(1304, 40), 1000, 0), (t(1304, 56), 1000, 40)]
    elif layout == "gap":
        tiles = [(t(1152, 96), 0, 0), (t(1100, 96), 1152, 0)]
    elif layout == "overlap":
        tiles = [(t(1200, 96), 0, 0), (t(1152, 96), 1152, 0)]
    elif layout == "odd_x":
        tiles = [(t(1150, 96), 0, 0), (t(1154, 96), 1150, 0)]
    elif layout == "odd_y":     # tile edges inside MCU rows
        tiles = [(t(2304, 43), 0, 0), (t(1000, 53), 0, 43), (t(1304, 21), 1000, 43), (t(1304, 32), 1000, 64)]
    elif layout == "bricks":    # odd sizes everywhere: edges inside MCU rows and inside 4-pixel pieces, unaligned rows
        tiles = [(t(701, 37), 0, 0), (t(1603, 37), 701, 0), (t(1201, 59), 0, 37), (t(1103, 59), 1201, 37)]
    elif layout == "clipped":   # tiles larger than what shows of them (ClippedImageInfo.sourceOffsetX/Y), still an exact mosaic
        tiles = [(t(1200, 120), 0, 0, (16, 24, 1152, 96)), (t(1160, 96), 1152, 0, (8, 0, 1152, 96))]
    else:
        h = 93
        tiles = [(t(1152, 48), 0, 0), (t(1152, 48), 1152, 0), (t(2304, 45), 0, 48)]
    canvas = np.zeros((h, w, 4), np.uint8)
    for px, x, y, *clip in tiles:
        if clip:
            ox, oy, vw, vh = clip[0]
            px = px[oy:oy + vh, ox:ox + vw]
        canvas[y:y + px.shape[0], x:x + px.shape[1]] = px[:h - y]
    canvas = canvas.reshape(h, w * 4)
    want = orc.jpeg_encode_rows(canvas, w, h, 0, h, 85)
    bands = [(0, 48), (48, h)]
    assert _mosaic_entropy(ctx, tiles, w, h, 85, bands) == want
    assert _mosaic_entropy(ctx, tiles, w, h, 85, [(0, h)]) == want
    monkeypatch.setenv("STITCH_JPEG_DIRECT", "0")
    assert _mosaic_entropy(ctx, tiles, w, h, 85, bands) == want


def test_bands_after_an_overflow_skip_the_one_pass_route(ctx):
    """After a band the MCU kernel could not hold, the next bands go straight to J1..J7 (1, 2, 4 of them while it
    keeps happening), then the one-pass route is tried again: whatever the route, the oracle's bytes."""
    rng = np.random.default_rng(14)
    w, h = 520, 16 * 9
    rgba = _photo(rng, w, h).reshape(h, w * 4).copy()
    for k in (0, 1, 4):   # noisy bands
        rgba[16 * k:16 * k + 16] = rng.integers(0, 256, size=(16, w * 4), dtype=np.uint8)
    want = orc.jpeg_encode_rows(rgba, w, h, 0, h, 97)
    assert _gpu_entropy(ctx, rgba, w, h, 97, [(16 * k, 16 * k + 16) for k in range(9)]) == want


def test_random_noise_many_ff_bytes(ctx):
    """Noise at quality 100 produces long codes and many 0xFF bytes: stuffing and worst-case lengths."""
    rng = np.random.default_rng(11)
    w, h = 264, 72
    rgba = rng.integers(0, 256, size=(h, w * 4), dtype=np.uint8)
    want = orc.jpeg_encode_rows(rgba, w, h, 0, h, 100)
    assert want.count(b"\xff\x00") > 10
    assert _gpu_entropy(ctx, rgba, w, h, 100, [(0, h)]) == want


def test_bands_equal_whole(ctx):
    rng = np.random.default_rng(5)
    w, h = 300, 101
    rgba = _photo(rng, w, h).reshape(h, w * 4)
    whole = _gpu_entropy(ctx, rgba, w, h, 90, [(0, h)])
    parts = _gpu_entropy(ctx, rgba, w, h, 90, [(0, 8), (8, 64), (64, 96), (96, 101)])
    assert parts == whole == orc.jpeg_encode_rows(rgba, w, h, 0, h, 90)


@pytest.mark.parametrize("chunk_rows", [8, 24, 64])
def test_band_downloaded_in_sub_bands(ctx, monkeypatch, chunk_rows):
    """With a host destination a band is coded and downloaded in sub-bands (two device buffers taking turns); the bytes
    are those of the band coded at once."""
    rng = np.random.default_rng(6)
    w, h = 1300, 203
    rgba = _photo(rng, w, h).reshape(h, w * 4)
    want = orc.jpeg_encode_rows(rgba, w, h, 0, h, 80)
    monkeypatch.setenv("STITCH_JPEG_CHUNK_BYTES", str(w * 4 * chunk_rows))
    assert _gpu_entropy(ctx, rgba, w, h, 80, [(0, h)]) == want
    assert _gpu_entropy(ctx, rgba, w, h, 80, [(0, 104), (104, h)]) == want


def test_band_alignment_is_checked(ctx):
    job = _native.Job(ctx, 16, 32, 8, _native.OUT_JPEG_ENTROPY, False, bytes(4), jpeg_quality=85)
    job.add_tile(bytes(16 * 32 * 4), width=16, height=32, bit_depth=8, color_type=6, filtered=0)
    with pytest.raises(_native.StitchError, match="multiple of 8"):
        job.run_band(4, 16)
    job.close()


def test_header_matches_oracle():
    for w, h, q in ((64, 48, 90), (65535, 1, 1), (17, 9, 100)):
        assert jpeg_header(w, h, q) == orc.jpeg_header(w, h, q)
    with pytest.raises(ValueError, match="Invalid quality"):
        jpeg_header(8, 8, 0)
    with pytest.raises(ValueError, match="Invalid dimensions"):
        jpeg_header(0, 8, 50)


def test_api_jpeg_output_structure_and_pixels():
    """concatToBuffer({outputFormat:'jpeg'}): one SOI at 0, one EOI at the end, decodable, dims,
    and the decoded image equals libjpeg's decode of the oracle's file."""
    ins = [H.suite_bytes("basn2c08"), H.suite_bytes("basn6a08"), H.suite_bytes("basn0g08")]
    opts = {"inputs": ins, "layout": {"columns": 3}, "output_format": "jpeg", "jpeg_quality": 90}
    data = stitch.concat_to_buffer(opts)
    assert data[:2] == b"\xff\xd8" and data[-2:] == b"\xff\xd9"
    assert data.count(b"\xff\xd8") == 1 and data.count(b"\xff\xd9") == 1
    img = np.asarray(PIL.open(io.BytesIO(data)).convert("RGB"))
    assert img.shape == (32, 96, 3)
    _, raw = H.oracle_stitch_rows(dict(opts, output_format="png"), out_kind=1)
    want = orc.jpeg_file(np.frombuffer(raw, np.uint8).reshape(32, 96 * 4), 96, 32, 90)
    assert data == want


def test_api_jpeg_non_multiple_of_8_and_small_bands():
    ins = [{"x": 0, "y": 0, "source": H.suite_bytes("basn2c08")}, {"x": 20, "y": 11, "source": H.suite_bytes("basn6a08")}]
    opts = {"inputs": ins, "layout": {}, "output_format": "jpeg", "band_rows": 16}
    data = stitch.concat_to_buffer(opts)
    img = PIL.open(io.BytesIO(data))
    assert img.size == (52, 43)
    np.asarray(img.convert("RGB"))  # decodes
