"""Kernel (a) parity: unfilter + Adam7 on the GPU vs the oracle, bit-exact."""
import numpy as np
import pytest

import helpers as H
import oracle_py as orc
from image_stitch_b200 import _native

pytestmark = pytest.mark.gpu


def _suite_cases():
    import os
    out = []
    for n in sorted(os.listdir(os.path.join(H.GOLDEN, "pngsuite"))):
        with open(H.suite_path(n), "rb") as f:
            ct = f.read(26)[25]
        if ct != 3:
            out.append(n[:-4])
    return out


@pytest.mark.parametrize("stem", _suite_cases())
def test_pngsuite_decode_matches_oracle(ctx, stem):
    hdr, stream = H.inflated_idat(H.suite_bytes(stem))
    want = orc.decode_scanlines(stream, hdr.width, hdr.height, hdr.bit_depth, hdr.color_type, hdr.interlace_method)
    got = ctx.decode_image(hdr.width, hdr.height, hdr.bit_depth, hdr.color_type, hdr.interlace_method, stream)
    assert np.array_equal(got, want)


FORMATS = [(8, 6), (8, 2), (8, 4), (8, 0), (16, 6), (16, 2), (16, 4), (16, 0), (4, 0), (2, 0), (1, 0)]


@pytest.mark.parametrize("bd,ct", FORMATS)
@pytest.mark.parametrize("mode", ["random", "adaptive", 4, 3])
def test_synthetic_unfilter(ctx, bd, ct, mode):
    rng = np.random.default_rng(100 + bd * 7 + ct)
    for (w, h) in ((1, 1), (7, 3), (33, 65), (257, 130), (64, 1000)):
        pix = H.synth_pixels(rng, w, h, bd, ct, "any")
        bpp = max(1, {0: 1, 2: 3, 4: 2, 6: 4}[ct] * bd // 8)
        stream = H.filter_image(pix, bpp, mode, rng)
        got = ctx.decode_image(w, h, bd, ct, 0, stream.tobytes())
        assert np.array_equal(got, pix.reshape(-1)), (w, h, bd, ct, mode)
        want = orc.decode_scanlines(stream, w, h, bd, ct, 0)
        assert np.array_equal(got, want)


@pytest.mark.parametrize("bd,ct", FORMATS)
def test_synthetic_adam7(ctx, bd, ct):
    rng = np.random.default_rng(200 + bd * 7 + ct)
    for (w, h) in ((1, 1), (2, 2), (5, 3), (8, 8), (33, 17), (130, 67)):
        pix = H.synth_pixels(rng, w, h, bd, ct, "any")
        stream = H.adam7_interlace(pix, w, h, bd, ct, rng, "random")
        want = orc.deinterlace_adam7(stream, w, h, bd, ct)
        assert np.array_equal(want, pix.reshape(-1))
        got = ctx.decode_image(w, h, bd, ct, 1, stream.tobytes())
        assert np.array_equal(got, want), (w, h, bd, ct)


def test_random_bytes_all_filters(ctx):
    """Arbitrary filtered bytes (not produced by an encoder): every filter, every wrap-around."""
    rng = np.random.default_rng(5)
    for bd, ct, w, h in ((8, 6, 300, 200), (8, 2, 101, 77), (16, 6, 64, 90), (8, 0, 513, 40)):
        n = orc.scanline_bytes(w, bd, ct)
        stream = rng.integers(0, 256, size=(h, n + 1), dtype=np.uint8)
        stream[:, 0] = rng.integers(0, 5, size=h)
        want = orc.decode_scanlines(stream.reshape(-1), w, h, bd, ct, 0)
        got = ctx.decode_image(w, h, bd, ct, 0, stream.tobytes())
        assert np.array_equal(got, want)


def test_large_tile_1024(ctx):
    rng = np.random.default_rng(9)
    w = h = 1024
    pix = H.synth_pixels(rng, w, h, 8, 6, "opaque")
    stream = H.filter_image(pix, 4, "random", rng)
    got = ctx.decode_image(w, h, 8, 6, 0, stream.tobytes())
    assert np.array_equal(got, pix.reshape(-1))


def test_errors(ctx):
    with pytest.raises(_native.StitchError, match="Unknown filter type: 9") as e:
        ctx.decode_image(2, 2, 8, 6, 0, bytes([0] + [1] * 8 + [9] + [1] * 8))
    assert e.value.code == _native.E_FILTER
    with pytest.raises(_native.StitchError, match="Unsupported source color type: 3"):
        ctx.decode_image(2, 2, 8, 3, 0, bytes(6))
    with pytest.raises(_native.StitchError, match="Expected 2 scanlines, decoded 1"):
        ctx.decode_image(2, 2, 8, 6, 0, bytes(9 + 4))
    with pytest.raises(_native.StitchError, match="Unexpected end of decompressed data at pass 6, line 0"):
        ctx.decode_image(2, 2, 8, 6, 1, bytes([0, 1, 2, 3, 4]))


def test_adam7_partial_last_line_reads_zero(ctx):
    """adam7.ts:87-88: `slice` shortens a final partial line; missing bytes scatter as 0."""
    full = bytes([0, 10, 20, 30, 40, 0, 50, 60, 70, 80, 1, 90, 100, 110, 120, 130, 140, 150, 160])
    cut = full[:-3]
    want = orc.deinterlace_adam7(cut, 2, 2, 8, 6)
    got = ctx.decode_image(2, 2, 8, 6, 1, cut)
    assert np.array_equal(got, want)


@pytest.mark.parametrize("bd,ct", FORMATS)
def test_wide_and_tall_random_streams(ctx, bd, ct):
    """Several column blocks and several 32-row groups per chain (wavefront kernel seams),
    arbitrary filtered bytes with every filter type."""
    rng = np.random.default_rng(300 + bd * 7 + ct)
    for (w, h) in ((1500, 70), (130, 300), (4099, 33)):
        n = orc.scanline_bytes(w, bd, ct)
        stream = rng.integers(0, 256, size=(h, n + 1), dtype=np.uint8)
        stream[:, 0] = rng.integers(0, 5, size=h)
        want = orc.decode_scanlines(stream.reshape(-1), w, h, bd, ct, 0)
        got = ctx.decode_image(w, h, bd, ct, 0, stream.tobytes())
        assert np.array_equal(got, want), (w, h, bd, ct)


def test_many_tiles_one_job_decode(ctx):
    """Many chains of different heights in one launch (group-major unit order)."""
    rng = np.random.default_rng(77)
    job = _native.Job(ctx, 4096, 4096, 8, _native.OUT_RAW_RGBA, False, bytes(4))
    shapes = [(64, 200), (300, 31), (33, 33), (512, 130), (7, 65), (1024, 96)]
    wants, x = [], 0
    for k, (w, h) in enumerate(shapes):
        stream = rng.integers(0, 256, size=(h, w * 4 + 1), dtype=np.uint8)
        stream[:, 0] = rng.integers(0, 5, size=h)
        wants.append((x, w, h, orc.decode_scanlines(stream.reshape(-1), w, h, 8, 6, 0).reshape(h, w * 4)))
        job.add_tile(stream.tobytes(), width=w, height=h, bit_depth=8, color_type=6, interlace=0, filtered=1,
                     dst_x=x, dst_y=0, z_index=0, image_idx=k)
        x += w
    out, _ = job.run_band(0, 256)
    out = np.frombuffer(out, dtype=np.uint8).reshape(256, 4096 * 4)
    for (x0, w, h, want) in wants:
        hh = min(h, 256)
        assert np.array_equal(out[:hh, x0 * 4:(x0 + w) * 4], want[:hh])
    job.close()
