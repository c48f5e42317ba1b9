"""Device-side proof of the alpha-over fast path: blend8 against the literal f64
arithmetic of src/pixel-ops.ts:704-733 for ALL 2^32 (srcAlpha, dstAlpha, src, dst) tuples, blend16 on a 2^28
hashed sample that includes every alpha within 8 of the reference's thresholds, and 10^6 hashed pixels against
the oracle's compositeScanline."""
import numpy as np
import pytest

import oracle_py as orc

pytestmark = pytest.mark.gpu


def test_blend8_every_tuple_on_the_device(ctx):
    checked, bad, first = ctx.selftest_blend(0, 1)
    assert checked == 1 << 32
    assert bad == 0, f"{bad} mismatches, e.g. dst={first[0]:08x} src={first[1]:08x} fast={first[2]:08x} exact={first[3]:08x}"


def test_blend16_sample_on_the_device(ctx):
    checked, bad, first = ctx.selftest_blend(1, 1 << 28)
    assert checked == 1 << 28
    assert bad == 0, f"{bad} mismatches, e.g. dst={first[0]:08x}{first[1]:08x} src={first[2]:08x}{first[3]:08x}"


def test_blend8_samples_against_the_oracle(ctx):
    n = 1_000_000
    t = ctx.selftest_blend8_samples(n, seed=7)
    dest = t[:, 0].copy().view(np.uint8)
    src = t[:, 1].copy().view(np.uint8)
    want = orc.composite_scanline(dest, src, 0, n, 4, True)
    got = t[:, 2].copy().view(np.uint8)
    assert np.array_equal(np.asarray(want).reshape(-1), got)
    alphas = src.reshape(-1, 4)[:, 3]
    assert ((alphas > 0) & (alphas < 255)).mean() > 0.9     # the sample is mostly true blends
