"""Robustness of the device JPEG decoder (tools/jpeg_corrupt_fuzz.py): valid files truncated, bit-flipped and spliced at
random; every call comes back with pixels or a StitchError, and a valid file decodes to its pixels after each of them.
(The run that found a double free on the failure path of stitch_job_add_tile_jpeg: its error surfaced in the next call.)"""
import os
import sys

import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))

pytestmark = pytest.mark.gpu


def test_corrupted_files_are_refused_or_decoded_and_leave_no_trace(ctx):
    import jpeg_corrupt_fuzz
    assert jpeg_corrupt_fuzz.run(5.0, 51, ctx, check_every=1) > 200
