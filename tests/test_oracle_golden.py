"""Pins the CPU oracle (and the host layout code) against the reference's own
golden vectors:

1. tests/utils/fixtures/expected-outputs/example1..5.png -- genuine reference
   outputs; the inflated IDAT (filter bytes included) must match byte for byte.
2. pngsuite/json goldens for every non-palette PngSuite image.
"""
import numpy as np
import pytest

import helpers as H
import oracle_py as orc

# recipes: tests/utils/generate-expected-outputs.js:29-106
EXAMPLES = {
    "example1.png": (["basn2c08", "basn0g08", "basn6a08"], {"columns": 3}, (96, 32, 8)),
    "example2.png": (["basn2c08", "basn0g08", "basn6a08"], {"rows": 3}, (32, 96, 8)),
    "example3.png": (["basn2c08", "basn0g08", "basn6a08", "basn4a08", "basn2c16", "basn0g16"], {"columns": 3}, (96, 64, 16)),
    "example4.png": (["basn0g01", "basn0g04", "basn2c08"], {"columns": 3}, (96, 32, 8)),
    "example5.png": (["basn2c08", "basn0g08", "basn6a08", "basn4a08"], {"width": 100}, (96, 64, 8)),
}


@pytest.mark.parametrize("name", sorted(EXAMPLES))
def test_oracle_reproduces_reference_output(name):
    inputs, layout, (w, h, bd) = EXAMPLES[name]
    exp_hdr, exp_rows = H.png_filtered_rows(H.expected_output(name))
    assert (exp_hdr.width, exp_hdr.height, exp_hdr.bit_depth, exp_hdr.color_type) == (w, h, bd, 6)
    plan, rows = H.oracle_stitch_rows({"inputs": [H.suite_bytes(n) for n in inputs], "layout": layout})
    assert (plan.canvas_w, plan.canvas_h, plan.out_bit_depth) == (w, h, bd)
    assert rows.tobytes() == exp_rows, f"{name}: filtered scanlines differ from the reference output"


def test_golden_filter_bytes_are_nontrivial():
    # observed in SURVEY.md 8c: Sub on first rows, Paeth elsewhere -> the selector is really exercised
    _, rows = H.png_filtered_rows(H.expected_output("example1.png"))
    a = np.frombuffer(rows, np.uint8).reshape(32, -1)
    assert set(a[:, 0].tolist()) - {0} != set()


# Third-party goldens that do NOT describe the reference's behaviour:
#  * tb*/tm*/tp1*: tRNS transparency, which the reference ignores
SKIP_PREFIXES = ("tb", "tm", "tp1")


def _json_cases():
    z = np.load(H.GOLDEN + "/pngsuite_rgba.npz")
    return sorted(k for k in z.files if not k.startswith(SKIP_PREFIXES))


@pytest.mark.parametrize("stem", _json_cases())
def test_oracle_matches_pngsuite_json(stem):
    z = np.load(H.GOLDEN + "/pngsuite_rgba.npz")
    want = z[stem]
    hdr, stream = H.inflated_idat(H.suite_bytes(stem))
    pix = orc.decode_scanlines(stream, hdr.width, hdr.height, hdr.bit_depth, hdr.color_type, hdr.interlace_method)
    n = orc.scanline_bytes(hdr.width, hdr.bit_depth, hdr.color_type)
    tbd = 16 if hdr.bit_depth == 16 else 8
    got = np.stack([orc.convert_scanline(pix[y * n:(y + 1) * n], hdr.width, hdr.bit_depth, hdr.color_type, tbd, 6)
                    for y in range(hdr.height)])
    if tbd == 16:
        g = got.reshape(hdr.height, hdr.width, 4, 2).astype(np.uint16)
        got_v = (g[..., 0] << 8) | g[..., 1]
    else:
        got_v = got.reshape(hdr.height, hdr.width, 4)
    assert np.array_equal(got_v, want)


def _palette_cases():
    return sorted(np.load(H.GOLDEN + "/pngsuite_palette_rgba.npz").files)


@pytest.mark.parametrize("stem", _palette_cases())
def test_oracle_palette_extension_matches_pngsuite_json(stem):
    """The palette expansion is an extension (the reference throws on colour type 3): its oracle
    restates the PNG specification and is pinned by the third-party goldens only."""
    from image_stitch_b200.decoders.png_decoder import PngBufferDecoder
    want = np.load(H.GOLDEN + "/pngsuite_palette_rgba.npz")[stem]
    png = H.suite_bytes(stem)
    hdr, stream = H.inflated_idat(png)
    assert hdr.color_type == 3
    plte, trns = PngBufferDecoder(png).palette()
    idx = orc.decode_scanlines(stream, hdr.width, hdr.height, hdr.bit_depth, 3, hdr.interlace_method)
    got = orc.expand_palette(idx, hdr.width, hdr.height, hdr.bit_depth, plte, trns)
    assert np.array_equal(got.reshape(want.shape), want)
