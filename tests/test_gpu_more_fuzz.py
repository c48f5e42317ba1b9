"""Short randomised parity runs of the device inflate against zlib (tools/inflate_fuzz.py) and of the public API against
the oracle (tools/api_fuzz.py: random PNG / JPEG inputs, layouts, blending, device codecs, PNG and JPEG output)."""
import os
import sys
import warnings

import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))

pytestmark = pytest.mark.gpu


def test_random_zlib_streams(ctx):
    import inflate_fuzz
    n = inflate_fuzz.run(6.0, 41, ctx)
    assert n > 20, "mismatch against zlib (see the output)" if n < 0 else f"only {n} streams in 6 s"


def test_random_jobs_through_the_public_api():
    import api_fuzz
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")   # clipped / invisible images are announced like the reference announces them
        n = api_fuzz.run(8.0, 42)
    assert n > 20, "mismatch against the oracle (see the output)" if n < 0 else f"only {n} jobs in 8 s"


def test_damaged_zlib_streams_are_refused_like_zlib_refuses_them(ctx):
    import inflate_corrupt_fuzz
    assert inflate_corrupt_fuzz.run(5.0, 43, ctx) > 100


def test_damaged_png_inputs_raise_like_the_oracle_side():
    import png_corrupt_fuzz
    assert png_corrupt_fuzz.run(5.0, 44) > 100
