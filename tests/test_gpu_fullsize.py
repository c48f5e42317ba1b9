"""BASELINE.json configurations at their full sizes.

The oracle cannot render a 16k x 16k or 32k x 32k canvas in seconds, so each
test checks (1) row bands sampled across the canvas against the oracle,
bit-exact, and (2) a size-independent property over the whole output:
filter -> unfilter round trip (C2), bands == whole (C3), decoded JPEG == the
oracle's JPEG (C5 shape at the largest sizes JPEG allows).
"""
import io

import numpy as np
import pytest

import helpers as H
import oracle_py as orc
from image_stitch_b200 import _native, synth
from image_stitch_b200.jpeg_container import JpegBandSplicer

pytestmark = pytest.mark.gpu


def test_c2_grid_16k_sampled_bands_and_round_trip(ctx):
    """16x16 grid of 1024x1024 RGBA8 tiles (filtered streams) -> 16384x16384 filtered scanlines."""
    grid, tile, canvas = 16, 1024, 16384
    rng = np.random.default_rng(2)
    raws = [synth.smooth_rgba8(rng, tile, tile, "opaque") for _ in range(4)]
    streams = [synth.encode_filtered(r, 4) for r in raws]
    which = lambda r, c: (r * 5 + c * 3) % 4
    out = np.empty(canvas * (canvas * 4 + 1), np.uint8)
    with _native.Job(ctx, canvas, canvas, 8, _native.OUT_PNG_FILTERED, False, bytes(4)) as job:
        for r in range(grid):
            for c in range(grid):
                job.add_tile(streams[which(r, c)], width=tile, height=tile, bit_depth=8, color_type=6, filtered=1,
                             dst_x=c * tile, dst_y=r * tile, image_idx=r * grid + c)
        got, tm = job.run_band(0, canvas, out=out)  # host destination: chunked, three streams
    got = got.reshape(canvas, canvas * 4 + 1)
    # (1) sampled bands, including tile seams and both canvas edges, against the oracle
    otiles = [dict(width=tile, height=tile, bit_depth=8, color_type=6, pixels=raws[which(r, c)], dst_x=c * tile,
                   dst_y=r * tile, image_idx=r * grid + c) for r in range(grid) for c in range(grid)]
    for a, b in ((0, 12), (1018, 1030), (8189, 8195), (16372, 16384)):
        want = orc.render_rows(canvas, canvas, 8, bytes(4), False, otiles, a, b, 0)
        assert np.array_equal(got[a:b].reshape(-1), want), (a, b)
    assert set(np.unique(got[:, 0])) <= {0, 1, 2, 3, 4}
    # (2) the whole output is a valid PNG stream of the mosaic: unfilter it (kernel a, one 16384-row chain)
    back = ctx.decode_image(canvas, canvas, 8, 6, 0, out, filtered=1)
    back = np.frombuffer(back, np.uint8).reshape(canvas, canvas * 4)
    for r in range(grid):
        for c in range(grid):
            blk = back[r * tile:(r + 1) * tile, c * tile * 4:(c + 1) * tile * 4]
            assert np.array_equal(blk, raws[which(r, c)]), (r, c)


def _c3_tiles(rng, canvas, n_side, tile):
    """SURVEY.md 8d, C3 (layout: image_stitch_b200.synth.c3_layout, shared with bench.py), with pixels."""
    raws = [synth.smooth_rgba8(rng, tile, tile, k) for k in synth.C3_CONTENT]
    tiles = synth.c3_layout(rng, canvas, n_side, tile)
    for t in tiles:
        t["pixels"] = raws[t["content"]]
    return tiles


@pytest.mark.parametrize("blend,route", [(True, None), (False, None), (True, "1"), (False, "0")])
def test_c3_positioned_blend_32k_sampled_bands(ctx, capsys, monkeypatch, blend, route):
    """256 positioned overlapping 2048x2048 RGBA8 tiles on a 32768x32768 canvas, with and without zIndex alpha
    blending, on the route the library picks (blended: via the canvas; not blended: the unit kernel's plan mode) and on
    the other one (STITCH_SINGLE_PASS=1 / 0)."""
    if route is None:
        monkeypatch.delenv("STITCH_SINGLE_PASS", raising=False)
    else:
        monkeypatch.setenv("STITCH_SINGLE_PASS", route)
    canvas, tile = 32768, 2048
    rng = np.random.default_rng(3)
    tiles = _c3_tiles(rng, canvas, 16, tile)
    bg = bytes([255, 255, 255, 128])
    with _native.Job(ctx, canvas, canvas, 8, _native.OUT_PNG_FILTERED, blend, bg) as job:
        for t in tiles:
            job.add_tile(t["pixels"].reshape(-1), width=tile, height=tile, bit_depth=8, color_type=6, filtered=0,
                         dst_x=t["dst_x"], dst_y=t["dst_y"], src_off_x=t["src_off_x"], src_off_y=t["src_off_y"],
                         vis_w=t["vis_w"], vis_h=t["vis_h"], z_index=t["z_index"], image_idx=t["image_idx"])
        for a, b in ((0, 10), (2040, 2056), (16380, 16390), (32760, 32768)):
            got, _ = job.run_band(a, b)
            want = orc.render_rows(canvas, canvas, 8, bg, blend, tiles, a, b, 0)
            assert np.array_equal(np.asarray(got), want), (a, b)
        # a large band equals its pieces (size-independent property), and gives a timing for DESIGN.md
        big, tm = job.run_band(4096, 8192)
        big = np.array(big, copy=True).reshape(4096, canvas * 4 + 1)
        for a, b in ((4096, 4104), (6140, 6150), (8184, 8192)):
            part, _ = job.run_band(a, b)
            assert np.array_equal(np.asarray(part).reshape(b - a, -1), big[a - 4096:b - 4096]), (a, b)
        job.run_band(4096, 8192, download=False)          # scratch buffers grow to the band's size here
        _, tm = job.run_band(4096, 8192, download=False)
    with capsys.disabled():
        print(f"\n[c3] 4096 rows x 32768 px, blending {blend}, STITCH_SINGLE_PASS {route}: compose+filter {tm['compose_ms']:.3f} ms "
              f"({4096 * 32768 / tm['compose_ms'] / 1e3:.0f} MPix/s)")


def test_c4_mixed_formats_grid(ctx):
    """16-bit RGBA, RGB, grey+alpha, low-bit-depth grey and Adam7-interlaced tiles -> one 8-bit RGBA grid."""
    rng = np.random.default_rng(4)
    tile, n = 192, 6
    fmts = [(16, 6, 0), (8, 2, 0), (8, 4, 1), (4, 0, 0), (8, 6, 1), (16, 2, 1), (1, 0, 1), (8, 0, 0), (16, 4, 0), (2, 0, 0)]
    tiles, streams = [], []
    for i in range(n * n):
        bd, ct, il = fmts[i % len(fmts)]
        pix = H.synth_pixels(rng, tile, tile, bd, ct, ["opaque", "any"][i % 2])
        spp = {0: 1, 2: 3, 4: 2, 6: 4}[ct]
        bpp = max(1, spp * bd // 8)
        streams.append(H.adam7_interlace(pix, tile, tile, bd, ct, rng, "random") if il else H.filter_image(pix, bpp, "random", rng))
        tiles.append(dict(width=tile, height=tile, bit_depth=bd, color_type=ct, interlace=il, pixels=pix,
                          dst_x=(i % n) * tile, dst_y=(i // n) * tile, image_idx=i))
    cw = ch = tile * n
    want = orc.render_rows(cw, ch, 8, bytes(4), False, tiles, 0, ch, 0)
    with _native.Job(ctx, cw, ch, 8, _native.OUT_PNG_FILTERED, False, bytes(4)) as job:
        for t, st in zip(tiles, streams):
            job.add_tile(st, width=tile, height=tile, bit_depth=t["bit_depth"], color_type=t["color_type"],
                         interlace=t["interlace"], filtered=1, dst_x=t["dst_x"], dst_y=t["dst_y"], image_idx=t["image_idx"])
        got, _ = job.run_band(0, ch)
    assert np.array_equal(np.asarray(got), want)


def _jpeg_entropy(ctx, tiles_fn, w, h, quality, bands):
    with _native.Job(ctx, w, h, 8, _native.OUT_JPEG_ENTROPY, False, bytes(4), jpeg_quality=quality) as job:
        tiles_fn(job)
        sp, out = JpegBandSplicer(), []
        for a, b in bands:
            data, _ = job.run_band(a, b)
            out.append(sp.push(data))
    return b"".join(out)


def test_c5_jpeg_mosaic_4k_and_maximum_width(ctx):
    """Mosaic -> JPEG q=90 entropy bytes, byte-exact against the oracle; then a band at the widest canvas JPEG
    can describe (65528 = 8191 MCUs; 65536 does not fit the 16-bit SOF0 field)."""
    rng = np.random.default_rng(5)
    tile, n = 1024, 4
    raws = [synth.smooth_rgba8(rng, tile, tile, "opaque") for _ in range(3)]
    cw = ch = tile * n
    canvas = np.zeros((ch, cw * 4), np.uint8)
    for i in range(n * n):
        r, c = divmod(i, n)
        canvas[r * tile:(r + 1) * tile, c * tile * 4:(c + 1) * tile * 4] = raws[i % 3]

    def add(job):
        for i in range(n * n):
            r, c = divmod(i, n)
            job.add_tile(raws[i % 3].reshape(-1), width=tile, height=tile, bit_depth=8, color_type=6, filtered=0,
                         dst_x=c * tile, dst_y=r * tile, image_idx=i)
    got = _jpeg_entropy(ctx, add, cw, ch, 90, [(0, 1024), (1024, 3072), (3072, 4096)])
    want = orc.jpeg_encode_rows(canvas, cw, ch, 0, ch, 90)
    assert got == want

    w, h = 65528, 24
    wide = synth.smooth_rgba8(rng, w, h, "opaque")

    def add_wide(job):
        job.add_tile(wide.reshape(-1), width=w, height=h, bit_depth=8, color_type=6, filtered=0, image_idx=0)
    got = _jpeg_entropy(ctx, add_wide, w, h, 90, [(0, 8), (8, 24)])
    assert got == orc.jpeg_encode_rows(wide, w, h, 0, h, 90)
