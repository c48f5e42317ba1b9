"""A short randomised parity run of the device JPEG decoder against libjpeg-turbo itself (Pillow on the box) and the
oracle (tools/jpeg_decode_fuzz.py): random sizes, contents, qualities, subsamplings, optimised tables, restart intervals."""
import os
import sys

import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))

pytestmark = pytest.mark.gpu
pytest.importorskip("PIL.Image")


@pytest.mark.parametrize("seed", [31, 32])
def test_random_files(ctx, seed):
    import jpeg_decode_fuzz
    n = jpeg_decode_fuzz.run(6.0, seed, ctx)
    assert n > 30, "mismatch against libjpeg-turbo (see the output)" if n < 0 else f"only {n} files in 6 s"
