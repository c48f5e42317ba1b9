"""The PNG side of the oracle (unfilter, Adam7, sample conversion to RGBA8: SURVEY 8 rows a1-a3) against an independent
decoder, Pillow's, on random images: every 8-bit colour type, sub-byte grey, Adam7, random and adaptive filter types.
The reference's own goldens pin the oracle (tests/test_oracle_golden.py); this adds sizes and filter mixes they lack.  CPU only."""
import io

import numpy as np
import pytest

import helpers as H

Image = pytest.importorskip("PIL.Image")

FORMATS = [(8, 6), (8, 2), (8, 0), (8, 4), (1, 0), (2, 0), (4, 0)]


@pytest.mark.parametrize("seed", range(4))
def test_single_image_canvas_equals_pillows_rgba(seed):
    rng = np.random.default_rng(500 + seed)
    for _ in range(25):
        bd, ct = FORMATS[int(rng.integers(0, len(FORMATS)))]
        w, h = int(rng.integers(1, 200)), int(rng.integers(1, 120))
        pix = H.synth_pixels(rng, w, h, bd, ct, "any" if ct in (4, 6) else "opaque")
        png = H.make_png(pix.reshape(h, -1), w, h, bd, ct, interlace=int(rng.integers(0, 3) == 0),
                         mode=["adaptive", "random"][int(rng.integers(0, 2))], rng=rng)
        plan, raw = H.oracle_stitch_rows({"inputs": [png], "layout": {"columns": 1}}, out_kind=1)
        assert (plan.canvas_w, plan.canvas_h, plan.out_bit_depth) == (w, h, 8)
        want = np.array(Image.open(io.BytesIO(png)).convert("RGBA"))
        got = np.frombuffer(raw, np.uint8).reshape(h, w, 4)
        assert np.array_equal(got, want), (w, h, bd, ct)
