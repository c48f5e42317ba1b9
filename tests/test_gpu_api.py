"""End-to-end through the public API on the GPU: the five reference goldens
must come out byte-identical at the filtered-scanline level."""
import io
import zlib

import numpy as np
import pytest

import helpers as H
import image_stitch_b200 as stitch
from image_stitch_b200.png_container import parse_png_chunks, read_png
from test_oracle_golden import EXAMPLES

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", sorted(EXAMPLES))
def test_reference_goldens_through_api(name):
    inputs, layout, (w, h, bd) = EXAMPLES[name]
    png = stitch.concat_to_buffer({"inputs": [H.suite_bytes(n) for n in inputs], "layout": layout})
    hdr, rows = read_png(png)
    exp_hdr, exp_rows = read_png(H.expected_output(name))
    assert hdr == exp_hdr
    assert rows == exp_rows
    assert [c.type for c in parse_png_chunks(png)][0] == "IHDR" and parse_png_chunks(png)[-1].type == "IEND"


def test_config1_two_rgba_columns():
    """BASELINE config 1: two PngSuite RGBA8 PNGs, layout {columns: 2}."""
    ins = [H.suite_bytes("basn6a08")] * 2
    png = stitch.concat_to_buffer({"inputs": ins, "layout": {"columns": 2}})
    hdr, rows = read_png(png)
    assert (hdr.width, hdr.height, hdr.bit_depth, hdr.color_type) == (64, 32, 8, 6)
    _, want = H.oracle_stitch_rows({"inputs": ins, "layout": {"columns": 2}})
    assert rows == want.tobytes()


def test_positioned_api_matches_oracle():
    ins = [H.suite_bytes(n) for n in ("basn6a08", "basn2c08", "basi6a16", "basn4a08")]
    pos = [{"x": 0, "y": 0, "source": ins[0]}, {"x": 16, "y": 8, "source": ins[1], "zIndex": 5},
           {"x": -10, "y": 20, "source": ins[2]}, {"x": 40, "y": 40, "source": ins[3], "zIndex": 1}]
    for extra in ({}, {"enableAlphaBlending": False}, {"backgroundColor": "#FFFFFF80"}):
        opts = {"inputs": pos, "layout": {"width": 60, "height": 64}, **extra}
        with pytest.warns(UserWarning):
            png = stitch.concat_to_buffer(opts)
        hdr, rows = read_png(png)
        plan, want = H.oracle_stitch_rows(opts)
        assert (hdr.width, hdr.height, hdr.bit_depth) == (60, 64, 16)
        assert rows == want.tobytes(), extra


def test_stream_and_progress_and_bands():
    ins = [H.suite_bytes(n) for n in ("basn2c08", "basn0g08", "basn6a08", "basn4a08")]
    calls = []
    opts = {"inputs": ins, "layout": {"columns": 2}, "onProgress": lambda d, t: calls.append((d, t)), "band_rows": 8}
    whole = stitch.concat_to_buffer(opts)
    assert calls == [(1, 4), (2, 4), (3, 4), (4, 4)]
    streamed = stitch.concat_to_stream({**opts, "onProgress": None}).read()
    assert read_png(streamed)[1] == read_png(whole)[1]
    _, want = H.oracle_stitch_rows({"inputs": ins, "layout": {"columns": 2}})
    assert read_png(whole)[1] == want.tobytes()


def test_custom_decoder_plugin_rows():
    """A foreign ImageDecoder (scanlines() only) enters the path as raw rows (filtered = 0)."""
    rng = np.random.default_rng(0)
    pix = H.synth_pixels(rng, 20, 10, 8, 6, "any")

    class Dec:
        def get_header(self):
            return stitch.ImageHeader(20, 10, 4, 8, "jpeg")

        def scanlines(self):
            for y in range(10):
                yield pix[y].tobytes()

        def close(self):
            self.closed = True
    d = Dec()
    png = stitch.concat_to_buffer({"inputs": [d, H.suite_bytes("basn6a08")], "layout": {"columns": 2}})
    assert d.closed
    hdr, rows = read_png(png)
    tiles = [dict(width=20, height=10, bit_depth=8, color_type=6, pixels=pix, dst_x=0, dst_y=0)]
    import oracle_py as orc
    h2, s2 = read_png(H.suite_bytes("basn6a08"))
    tiles.append(dict(width=32, height=32, bit_depth=8, color_type=6, dst_x=20, dst_y=0, image_idx=1,
                      pixels=orc.decode_scanlines(s2, 32, 32, 8, 6, 0)))
    want = orc.render_rows(52, 32, 8, bytes(4), False, tiles)
    assert rows == want.tobytes()


def test_palette_input_raises_like_reference():
    with pytest.raises(stitch.StitchFailure, match="Unsupported source color type: 3"):
        stitch.concat_to_buffer({"inputs": [H.suite_bytes("basn3p08")] * 2, "layout": {"columns": 2}})


def test_png_decoder_scanlines_on_gpu():
    from image_stitch_b200.decoders import PngBufferDecoder
    import oracle_py as orc
    dec = PngBufferDecoder(H.suite_bytes("basi2c16"))
    rows = b"".join(dec.scanlines())
    h, s = read_png(H.suite_bytes("basi2c16"))
    assert rows == orc.decode_scanlines(s, h.width, h.height, h.bit_depth, h.color_type, 1).tobytes()
