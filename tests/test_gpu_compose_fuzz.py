"""A short randomised parity run of composite + filter selection (tools/compose_fuzz.py): random canvases, tiles of
random sizes, formats, alpha kinds and positions (clipped, overlapping, with gaps), z orders, blending on and off, 8- and
16-bit output, random band cuts, on every route of the library -- byte-exact against the oracle."""
import os
import sys

import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("seed,env", [(21, {}), (22, {"STITCH_SINGLE_PASS": "2"}), (23, {"STITCH_SINGLE_PASS": "0"}), (24, {"STITCH_STRIP": "1"})])
def test_random_scenes(ctx, monkeypatch, seed, env):
    import compose_fuzz
    monkeypatch.delenv("STITCH_STRIP", raising=False)
    monkeypatch.delenv("STITCH_SINGLE_PASS", raising=False)
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    n = compose_fuzz.run(5.0, seed, ctx)
    assert n > 50, "mismatch against the oracle (see the output)" if n < 0 else f"only {n} scenes in 5 s"
