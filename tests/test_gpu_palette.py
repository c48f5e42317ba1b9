"""SURVEY.md 8f.4, an EXTENSION beyond the reference: colour type 3 tiles expanded through PLTE/tRNS on the device.

The reference throws `Unsupported source color type: 3` (pixel-ops.ts:609-611) and so does the default path; what is
checked here is stitch_job_add_tile_palette against the third-party pngsuite/json goldens (63 palette files) and against
the oracle's restatement of the PNG specification's rule -- not a parity claim."""
import os
import zlib

import numpy as np
import pytest

import helpers as H
import oracle_py as orc
from image_stitch_b200 import _native, concat_to_buffer
from image_stitch_b200.decoders.png_decoder import PngBufferDecoder
from image_stitch_b200.png_container import extract_idat, parse_png_chunks, read_png

pytestmark = pytest.mark.gpu


def _palette_cases():
    return sorted(np.load(H.GOLDEN + "/pngsuite_palette_rgba.npz").files)


def _chunks(png):
    return PngBufferDecoder(png).palette()


def _raw_rgba(ctx, hdr, data, filtered, plte, trns, interlace=None):
    with _native.Job(ctx, hdr.width, hdr.height, 8, _native.OUT_RAW_RGBA, False, bytes(8)) as job:
        job.add_tile(data, width=hdr.width, height=hdr.height, bit_depth=hdr.bit_depth, color_type=3,
                     interlace=hdr.interlace_method if interlace is None else interlace, filtered=filtered,
                     plte=plte, trns=trns)
        out, _ = job.run_band(0, hdr.height)
        return np.array(out, copy=True).reshape(hdr.height, hdr.width, 4)


@pytest.mark.parametrize("stem", _palette_cases())
def test_pngsuite_palette_matches_json_goldens(ctx, stem):
    want = np.load(H.GOLDEN + "/pngsuite_palette_rgba.npz")[stem]
    png = H.suite_bytes(stem)
    hdr, stream = read_png(png)
    plte, trns = _chunks(png)
    assert np.array_equal(_raw_rgba(ctx, hdr, stream, 1, plte, trns), want)            # inflated stream
    z = extract_idat(parse_png_chunks(png))
    assert np.array_equal(_raw_rgba(ctx, hdr, z, 2, plte, trns), want)                 # the zlib stream itself (8f.2)
    idx = orc.decode_scanlines(stream, hdr.width, hdr.height, hdr.bit_depth, 3, hdr.interlace_method)
    assert np.array_equal(_raw_rgba(ctx, hdr, idx.tobytes(), 0, plte, trns, interlace=0), want)   # raw index rows
    assert np.array_equal(orc.expand_palette(idx, hdr.width, hdr.height, hdr.bit_depth, plte, trns).reshape(want.shape), want)


@pytest.mark.parametrize("bd", [1, 2, 4, 8])
@pytest.mark.parametrize("interlace", [0, 1])
def test_synthetic_palette_tiles(ctx, bd, interlace):
    rng = np.random.default_rng(40 + bd + interlace)
    for (w, h) in ((1, 1), (3, 2), (5, 7), (130, 67), (1027, 33), (4100, 9)):
        n = 1 << bd
        entries = int(rng.integers(1, n + 1))
        plte = rng.integers(0, 256, entries * 3, dtype=np.uint8).tobytes()
        trns = rng.integers(0, 256, int(rng.integers(0, entries + 1)), dtype=np.uint8).tobytes()
        idx = H.synth_pixels(rng, w, h, bd, 0, "any")      # any index, also past the end of PLTE (opaque black)
        stream = H.adam7_interlace(idx, w, h, bd, 0, rng, "random") if interlace else H.filter_image(idx, 1, "random", rng)
        want = orc.expand_palette(idx.reshape(-1), w, h, bd, plte, trns).reshape(h, w, 4)
        hdr = H.PngHeader(w, h, bd, 3, 0, 0, interlace)
        got = _raw_rgba(ctx, hdr, np.asarray(stream).tobytes(), 1, plte, trns)
        assert np.array_equal(got, want), (w, h, bd, interlace)


def test_palette_tiles_in_a_png_job_match_the_oracle(ctx):
    """Palette tiles with and without tRNS next to ordinary tiles, blended, filtered: the same bytes as the oracle fed
    with the expanded RGBA8 pixels."""
    names = ["basn3p08", "tbbn3p08", "basn6a08", "basi3p04", "tp1n3p08", "basn3p01", "tm3n3p02", "basn0g08"]
    inputs = [{"source": H.suite_bytes(n), "x": 10 * i, "y": 6 * i} for i, n in enumerate(names)]
    opts = {"inputs": inputs, "layout": {}, "enableAlphaBlending": True, "backgroundColor": "#10203080",
            "expandPalette": True}
    got = concat_to_buffer(opts)
    hdr, rows = read_png(got)
    _, want = H.oracle_stitch_rows(opts)
    assert rows == want.tobytes()
    for extra in ({"deviceInflate": True}, {"deviceInflate": True, "deviceDeflate": True}):
        _, rows2 = read_png(concat_to_buffer({**opts, **extra}))
        assert rows2 == rows
    grid = {"inputs": [H.suite_bytes(n) for n in names], "layout": {"columns": 3}, "expandPalette": True}
    _, rows = read_png(concat_to_buffer(grid))
    assert rows == H.oracle_stitch_rows(grid)[1].tobytes()


def test_without_the_option_the_reference_error_stays(ctx):
    with pytest.raises(Exception, match="Unsupported source color type: 3"):
        concat_to_buffer({"inputs": [H.suite_bytes("basn3p08")], "layout": {"columns": 1}})
    hdr, stream = read_png(H.suite_bytes("basn3p08"))
    with _native.Job(ctx, 32, 32, 8, _native.OUT_RAW_RGBA, False, bytes(8)) as job:
        with pytest.raises(_native.StitchError, match="Unsupported source color type: 3") as e:
            job.add_tile(stream, width=32, height=32, bit_depth=8, color_type=3)
        assert e.value.code == _native.E_COLOR


def test_palette_argument_checks(ctx):
    hdr, stream = read_png(H.suite_bytes("basn3p08"))
    plte, trns = _chunks(H.suite_bytes("basn3p08"))
    with _native.Job(ctx, 32, 32, 8, _native.OUT_RAW_RGBA, False, bytes(8)) as job:
        kw = dict(width=32, height=32, bit_depth=8)
        with pytest.raises(_native.StitchError, match="PLTE must hold 1 to 256 entries"):
            job.add_tile(stream, color_type=3, plte=plte[:-1], **kw)
        with pytest.raises(_native.StitchError, match="PLTE must hold 1 to 256 entries"):
            job.add_tile(stream, color_type=3, plte=bytes(3 * 257), **kw)
        with pytest.raises(_native.StitchError, match="tRNS holds 5 entries, PLTE 4"):
            job.add_tile(stream, color_type=3, plte=bytes(12), trns=bytes(5), **kw)
        with pytest.raises(_native.StitchError, match="a palette was given for color type 2"):
            job.add_tile(stream, color_type=2, plte=plte, **kw)
        with pytest.raises(_native.StitchError, match="Unsupported bit depth 16 for color type 3"):
            job.add_tile(stream, width=32, height=32, bit_depth=16, color_type=3, plte=plte)
        with pytest.raises(_native.StitchError, match="Expected 32 scanlines, decoded 31"):
            job.add_tile(stream[:-1], color_type=3, plte=plte, **kw)
