"""A short randomised parity run of the JPEG encoder (tools/jpeg_fuzz.py): random sizes, qualities, contents with steps
and patches of noise, mosaics of randomly cut tiles, random band cuts -- every route of csrc/jpeg.cu, byte-exact
against the oracle.  (The run that found the one-pass route's first bug: entries of a block overwritten by its own code.)"""
import os
import sys

import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("seed,direct", [(11, True), (12, False), (13, True)])
def test_random_cases(ctx, monkeypatch, seed, direct):
    import jpeg_fuzz
    if not direct:
        monkeypatch.setenv("STITCH_JPEG_DIRECT", "0")
    n = jpeg_fuzz.run(6.0, seed, ctx)
    assert n > 50, "mismatch against the oracle (see the output)" if n < 0 else f"only {n} cases in 6 s"
