"""The device JPEG decoder (csrc/jpeg_decode.cu) through the C ABI: byte for byte against the oracle
(oracle/jpeg_decode_oracle.c) and against the pixels libjpeg-turbo decodes (tests/golden/jpeg_decode), as a decoder
alone, as tiles of a positioned canvas, and through the public API."""
import io
import os

import numpy as np
import pytest

import helpers as H
import oracle_py as orc
import image_stitch_b200 as stitch
from image_stitch_b200 import _native
from image_stitch_b200.png_container import read_png

pytestmark = pytest.mark.gpu

HERE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "jpeg_decode")
NAMES = sorted(n[:-4] for n in os.listdir(HERE) if n.endswith(".jpg") and not n.startswith("p"))


def _file(name):
    with open(os.path.join(HERE, name + ".jpg"), "rb") as f:
        return f.read()


@pytest.fixture(scope="module")
def decoded():
    return np.load(os.path.join(HERE, "decoded.npz"))


@pytest.mark.parametrize("name", NAMES)
def test_decoder_matches_oracle_and_libjpeg_turbo(ctx, name, decoded):
    data = _file(name)
    got = ctx.decode_jpeg(data)
    want = orc.jpeg_decode(data)
    assert got.shape == want.shape
    assert np.array_equal(got, want), f"{np.count_nonzero(got != want)} samples differ from the oracle"
    gold = decoded[name]
    assert np.array_equal(got if gold.ndim == 2 else got[..., :3], gold)
    assert ctx.jpeg_info(data) == (gold.shape[1], gold.shape[0], 1 if gold.ndim == 2 else 4)


def test_larger_images_with_and_without_restart_intervals(ctx):
    """Made here with Pillow when it is installed (the GPU box has the same image); many restart intervals, many
    blocks, both chroma layouts -- against the oracle."""
    Image = pytest.importorskip("PIL.Image")
    rng = np.random.default_rng(77)
    y, x = np.mgrid[0:531, 0:777]
    pic = np.stack([(x * 3 + y) % 256, (x + y * 2) % 256, (x * y // 64) % 256], -1).astype(np.uint8)
    pic = np.clip(pic.astype(np.int16) + rng.integers(-20, 20, pic.shape), 0, 255).astype(np.uint8)
    for opts in (dict(quality=85, subsampling=2), dict(quality=92, subsampling=1, restart_marker_rows=1),
                 dict(quality=70, subsampling=0, restart_marker_blocks=7), dict(quality=60, subsampling=2, restart_marker_blocks=1)):
        buf = io.BytesIO()
        Image.fromarray(pic, "RGB").save(buf, "JPEG", **opts)
        got = ctx.decode_jpeg(buf.getvalue())
        assert np.array_equal(got, orc.jpeg_decode(buf.getvalue())), opts


@pytest.mark.parametrize("sub_bits", ["64", "256", "4096", "1000000000"])
def test_every_decoder_path_gives_the_same_pixels(ctx, monkeypatch, sub_bits):
    """Subsequence length of the speculative walk: 64 bits never settles (the one-thread-per-interval decoder takes
    over), 256 needs many rounds, 10^9 is one subsequence per restart interval."""
    monkeypatch.setenv("STITCH_JPEG_SUB_BITS", sub_bits)
    for name in ("c420_big_opt", "c420_big_rst", "g_big", "c422_rst1_opt", "c444_q100_opt", "c420_q1"):
        data = _file(name)
        assert np.array_equal(ctx.decode_jpeg(data), orc.jpeg_decode(data)), name


def test_errors(ctx):
    with pytest.raises(_native.StitchError, match="Invalid JPEG: missing SOI marker"):
        ctx.decode_jpeg(b"\x89PNG\r\n\x1a\n0000")
    with pytest.raises(_native.StitchError, match="Invalid JPEG: no SOF marker found"):
        ctx.decode_jpeg(b"\xff\xd8\xff\xe0\x00\x04ab\xff\xd9")
    with open(os.path.join(HERE, "p420_progressive.jpg"), "rb") as f:
        prog = f.read()
    with pytest.raises(_native.StitchError, match="only baseline"):
        ctx.decode_jpeg(prog)
    assert ctx.jpeg_info(prog) == (40, 40, 4)   # the header scan takes every frame type (jpeg-decoder.ts:26-40)
    data = _file("c420_q80")
    with pytest.raises(_native.StitchError, match="ends early"):
        ctx.decode_jpeg(data[:len(data) - 300])
    rst = bytearray(_file("c420_rst3"))
    k = rst.index(b"\xff\xd1")
    rst[k + 1] = 0xD5
    with pytest.raises(_native.StitchError, match="restart marker"):
        ctx.decode_jpeg(bytes(rst))
    # the context is still usable
    assert np.array_equal(ctx.decode_jpeg(data), orc.jpeg_decode(data))


@pytest.mark.parametrize("blend", [False, True])
def test_jpeg_tiles_on_a_positioned_canvas(ctx, blend):
    """JPEG tiles (colour and gray) and a translucent RGBA tile on one canvas, bands cut at odd rows."""
    rng = np.random.default_rng(5)
    cw, ch = 300, 240
    files = [("c420_big_opt", 0, 0, 0), ("c420_big_rst", 40, 30, 1), ("g_big", 90, 110, 2), ("c420_odd", 160, 170, 3)]
    tiles = []
    with _native.Job(ctx, cw, ch, 8, _native.OUT_PNG_FILTERED, blend, bytes([9, 8, 7, 200])) as job:
        for i, (name, x, y, z) in enumerate(files):
            data = _file(name)
            pix = orc.jpeg_decode(data)
            h, w = pix.shape[:2]
            vw, vh = min(w, cw - x), min(h, ch - y)
            job.add_tile_jpeg(data, dst_x=x, dst_y=y, vis_w=vw, vis_h=vh, z_index=z, image_idx=i)
            tiles.append(dict(width=w, height=h, bit_depth=8, color_type=6 if pix.ndim == 3 else 0, pixels=pix.reshape(-1),
                              dst_x=x, dst_y=y, src_off_x=0, src_off_y=0, vis_w=vw, vis_h=vh, z_index=z, image_idx=i))
        over = H.synth_pixels(rng, 120, 90, 8, 6, "any")
        job.add_tile(np.ascontiguousarray(over).reshape(-1), width=120, height=90, bit_depth=8, color_type=6, filtered=0,
                     dst_x=100, dst_y=60, z_index=9, image_idx=9)
        tiles.append(dict(width=120, height=90, bit_depth=8, color_type=6, pixels=over, dst_x=100, dst_y=60, src_off_x=0,
                          src_off_y=0, vis_w=120, vis_h=90, z_index=9, image_idx=9))
        for a, b in ((0, ch), (37, 151), (239, 240)):
            out, _ = job.run_band(a, b)
            want = orc.render_rows(cw, ch, 8, bytes([9, 8, 7, 200]), blend, tiles, a, b, 0)
            assert np.array_equal(np.asarray(out), want), (a, b)


def test_corrupt_entropy_data_surfaces_at_run_band(ctx):
    data = bytearray(_file("c444_q90"))
    sos = data.index(b"\xff\xda")
    for k in range(sos + 40, sos + 90):   # garbage in the middle of the scan
        data[k] = 0xFF if k % 2 else 0xC7
    with _native.Job(ctx, 64, 64, 8, _native.OUT_PNG_FILTERED, False, bytes(4)) as job:
        try:
            job.add_tile_jpeg(bytes(data), dst_x=0, dst_y=0)
        except _native.StitchError as e:   # the marker scan may already refuse it
            assert "JPEG" in str(e)
            return
        with pytest.raises(_native.StitchError, match="Invalid JPEG"):
            job.run_band(0, 64)


def test_public_api_with_jpeg_inputs():
    """concat_to_buffer with JPEG files among the inputs: grid and positioned, against the oracle."""
    ins = [_file("c420_q80"), H.suite_bytes("basn6a08"), _file("g_q90"), _file("c422_q85")]
    png = stitch.concat_to_buffer({"inputs": ins, "layout": {"columns": 2}})
    hdr, rows = read_png(png)
    plan, want = H.oracle_stitch_rows({"inputs": ins, "layout": {"columns": 2}})
    assert (hdr.width, hdr.height) == (plan.canvas_w, plan.canvas_h)
    assert rows == want.tobytes()
    pos = [{"x": 0, "y": 0, "source": ins[0]}, {"x": 30, "y": 20, "source": ins[1], "zIndex": 3}, {"x": 10, "y": 30, "source": ins[3]}]
    opts = {"inputs": pos, "layout": {"width": 80, "height": 70}, "backgroundColor": "#10203040"}
    png = stitch.concat_to_buffer(opts)
    _, want = H.oracle_stitch_rows(opts)
    assert read_png(png)[1] == want.tobytes()


def test_jpeg_files_in_jpeg_out():
    """JPEG tiles decoded on the device and coded again as one JPEG (the MCU kernel reads the decoded tiles in place:
    an exact mosaic of RGBA8 rows), against the two oracles chained."""
    ins = [_file("c420_q80"), _file("c420_q80"), _file("c420_q80"), _file("c420_q80")]
    opts = {"inputs": ins, "layout": {"columns": 2}, "output_format": "jpeg", "jpeg_quality": 88}
    jpg = stitch.concat_to_buffer(opts)
    plan, raw = H.oracle_stitch_rows(dict(opts, output_format="png"), out_kind=1)
    w, h = plan.canvas_w, plan.canvas_h
    assert jpg == orc.jpeg_file(np.frombuffer(raw, np.uint8).reshape(h, w * 4), w, h, 88)


def test_decoder_plugin_contract():
    """JpegBufferDecoder: getHeader + scanlines (jpeg-decoder.ts:284-347) over the device decoder."""
    from image_stitch_b200.decoders import JpegBufferDecoder, detect_image_format
    data = _file("c420_odd")
    assert detect_image_format(data) == "jpeg"
    dec = JpegBufferDecoder(data)
    h = dec.get_header()
    assert (h.width, h.height, h.channels, h.bit_depth, h.format) == (129, 65, 4, 8, "jpeg")
    rows = list(dec.scanlines())
    assert len(rows) == 65 and all(len(r) == 129 * 4 for r in rows)
    assert b"".join(rows) == orc.jpeg_decode(data).tobytes()
    g = JpegBufferDecoder(_file("g_q90"))
    assert g.get_header().channels == 1 and len(next(iter(g.scanlines()))) == 40


import jpeg_variants as JV


@pytest.mark.parametrize("name", sorted(JV.variants()))
def test_marker_segment_variants(ctx, name, decoded):
    data, same_as, err = JV.variants()[name]
    if err:
        with pytest.raises(_native.StitchError, match=err.replace("(", r"\(").replace(")", r"\)")):
            ctx.decode_jpeg(data)
    else:
        assert np.array_equal(ctx.decode_jpeg(data)[..., :3], decoded[same_as])
