"""Kernels (b)+(c) parity through the C ABI: normalise, place/composite,
filter selection -- bit-exact against the oracle."""
import numpy as np
import pytest

import helpers as H
import oracle_py as orc
from image_stitch_b200 import _native

pytestmark = pytest.mark.gpu


def run_job(ctx, canvas_w, canvas_h, out_bd, bg, blend, tiles, out_kind=0, bands=None, filtered_streams=None):
    """tiles: oracle-style dicts (dense pixels).  Uploads them unfiltered unless a stream is given."""
    kind = {0: _native.OUT_PNG_FILTERED, 1: _native.OUT_RAW_RGBA}[out_kind]
    outs = []
    with _native.Job(ctx, canvas_w, canvas_h, out_bd, kind, blend, bg) as job:
        for i, t in enumerate(tiles):
            st = None if filtered_streams is None else filtered_streams[i]
            data = np.ascontiguousarray(t["pixels"]).reshape(-1) if st is None else st
            job.add_tile(data, width=t["width"], height=t["height"], bit_depth=t["bit_depth"],
                         color_type=t["color_type"], interlace=0, filtered=0 if st is None else 1,
                         dst_x=t.get("dst_x", 0), dst_y=t.get("dst_y", 0), src_off_x=t.get("src_off_x", 0),
                         src_off_y=t.get("src_off_y", 0), vis_w=t.get("vis_w", t["width"]),
                         vis_h=t.get("vis_h", t["height"]), z_index=t.get("z_index", 0), image_idx=t.get("image_idx", i))
        for (a, b) in (bands or [(0, canvas_h)]):
            data, tm = job.run_band(a, b)
            outs.append(np.array(data, copy=True))
    return np.concatenate(outs)


@pytest.mark.parametrize("bpp", [1, 2, 3, 4, 6, 8])
def test_filter_rows_matches_oracle(ctx, bpp):
    rng = np.random.default_rng(bpp)
    for n_rows, row_bytes in ((1, bpp), (3, 5 * bpp), (17, 64 * bpp + bpp), (40, 1000 * bpp)):
        smooth = (np.cumsum(rng.integers(-3, 4, size=(n_rows + 1, row_bytes)), axis=1) % 256).astype(np.uint8)
        noisy = rng.integers(0, 256, size=(n_rows + 1, row_bytes), dtype=np.uint8)
        for img in (smooth, noisy):
            for use_prev in (False, True):
                rows, prev0 = img[1:], (img[0] if use_prev else None)
                got = ctx.filter_rows(rows, prev0, n_rows, row_bytes, bpp).reshape(n_rows, row_bytes + 1)
                prev = prev0
                for y in range(n_rows):
                    ft, f, _ = orc.filter_scanline(rows[y], prev, bpp)
                    assert got[y, 0] == ft, (bpp, n_rows, row_bytes, y)
                    assert np.array_equal(got[y, 1:], f)
                    prev = rows[y]


def test_filter_tie_breaks(ctx):
    rows = np.full((4, 64), 7, np.uint8)       # constant image: many equal costs
    got = ctx.filter_rows(rows, None, 4, 64, 4).reshape(4, 65)
    prev = None
    for y in range(4):
        ft, f, _ = orc.filter_scanline(rows[y], prev, 4)
        assert got[y, 0] == ft and np.array_equal(got[y, 1:], f)
        prev = rows[y]


FORMATS = [(8, 6), (8, 2), (8, 4), (8, 0), (16, 6), (16, 2), (16, 4), (16, 0), (4, 0), (2, 0), (1, 0)]


@pytest.mark.parametrize("out_bd", [8, 16])
def test_grid_mixed_formats(ctx, out_bd):
    """Config 4 shape: every accepted format normalised into one grid, with padding."""
    rng = np.random.default_rng(40 + out_bd)
    tiles, x, y, row_h = [], 0, 0, 0
    for i, (bd, ct) in enumerate(FORMATS):
        w, h = 19 + 3 * i, 11 + 2 * (i % 4)
        tiles.append(dict(width=w, height=h, bit_depth=bd, color_type=ct, pixels=H.synth_pixels(rng, w, h, bd, ct, "any"),
                          dst_x=x, dst_y=y, image_idx=i))
        x += w + (i % 3)          # gaps = background padding
        row_h = max(row_h, h)
        if i % 4 == 3:
            x, y, row_h = 0, y + row_h + 1, 0
    cw = max(t["dst_x"] + t["width"] for t in tiles) + 5
    ch = max(t["dst_y"] + t["height"] for t in tiles) + 2
    for bg in (bytes(8), bytes([10, 10, 20, 20, 30, 30, 128, 128]) if out_bd == 16 else bytes([10, 20, 30, 128])):
        for kind in (1, 0):
            want = orc.render_rows(cw, ch, out_bd, bg, False, tiles, 0, ch, kind)
            got = run_job(ctx, cw, ch, out_bd, bg, False, tiles, kind)
            assert np.array_equal(got, want), (out_bd, kind)


@pytest.mark.parametrize("out_bd", [8, 16])
@pytest.mark.parametrize("blend", [True, False])
def test_positioned_overlap_blend(ctx, out_bd, blend):
    """Config 3 shape at small scale: overlapping, clipped tiles, zIndex ties, all alpha kinds."""
    rng = np.random.default_rng(30 + out_bd + blend)
    cw, ch = 200, 150
    tiles = []
    fmts = [(8, 6), (16, 6), (8, 4), (8, 2), (16, 4)] if out_bd == 16 else [(8, 6), (8, 6), (8, 4), (8, 2), (4, 0)]
    for i in range(14):
        bd, ct = fmts[i % len(fmts)]
        w, h = int(rng.integers(20, 90)), int(rng.integers(20, 90))
        x, y = int(rng.integers(-30, cw - 10)), int(rng.integers(-30, ch - 10))
        left, top, right, bottom = max(0, x), max(0, y), min(cw, x + w), min(ch, y + h)
        mode = ["opaque", "uniform", "binary", "any"][i % 4]
        tiles.append(dict(width=w, height=h, bit_depth=bd, color_type=ct, pixels=H.synth_pixels(rng, w, h, bd, ct, mode),
                          dst_x=left, dst_y=top, src_off_x=max(0, -x), src_off_y=max(0, -y), vis_w=right - left,
                          vis_h=bottom - top, z_index=int(rng.integers(0, 4)), image_idx=i))
    bgs = [bytes(8), (bytes([255, 255] * 3 + [128, 128]) if out_bd == 16 else bytes([255, 255, 255, 128]))]
    for bg in bgs:
        for kind in (1, 0):
            want = orc.render_rows(cw, ch, out_bd, bg, blend, tiles, 0, ch, kind)
            got = run_job(ctx, cw, ch, out_bd, bg, blend, tiles, kind)
            assert np.array_equal(got, want), (out_bd, blend, kind)


def test_blend8_exhaustive_alpha_pairs(ctx):
    """Every (srcAlpha, dstAlpha) pair with varied colours: 65536 pixels per layer, two layers."""
    rng = np.random.default_rng(1)
    w, h = 256, 256
    a = np.zeros((h, w, 4), np.uint8)
    b = np.zeros((h, w, 4), np.uint8)
    a[..., :3] = rng.integers(0, 256, size=(h, w, 3))
    b[..., :3] = rng.integers(0, 256, size=(h, w, 3))
    a[..., 3] = np.arange(256)[:, None]     # bottom layer alpha = row
    b[..., 3] = np.arange(256)[None, :]     # top layer alpha = column
    tiles = [dict(width=w, height=h, bit_depth=8, color_type=6, pixels=a.reshape(h, -1), z_index=0, image_idx=0),
             dict(width=w, height=h, bit_depth=8, color_type=6, pixels=b.reshape(h, -1), z_index=1, image_idx=1)]
    for bg in (bytes(4), bytes([9, 99, 199, 255]), bytes([255, 0, 128, 77])):
        want = orc.render_rows(w, h, 8, bg, True, tiles, 0, h, 1)
        got = run_job(ctx, w, h, 8, bg, True, tiles, 1)
        assert np.array_equal(got, want)


def test_blend8_ties_and_extremes(ctx):
    """Colours 0/255/128 against every alpha: exact .5 ties happen here."""
    vals = np.array([0, 1, 127, 128, 254, 255], np.uint8)
    w, h = 256, 36
    a = np.zeros((h, w, 4), np.uint8)
    b = np.zeros((h, w, 4), np.uint8)
    for r in range(h):
        a[r, :, :3] = vals[r % 6]
        b[r, :, :3] = vals[r // 6]
        a[r, :, 3] = 255 if r % 2 else 128
    b[..., 3] = np.arange(256)[None, :]
    tiles = [dict(width=w, height=h, bit_depth=8, color_type=6, pixels=a.reshape(h, -1), z_index=0, image_idx=0),
             dict(width=w, height=h, bit_depth=8, color_type=6, pixels=b.reshape(h, -1), z_index=0, image_idx=1)]
    want = orc.render_rows(w, h, 8, bytes(4), True, tiles, 0, h, 1)
    got = run_job(ctx, w, h, 8, bytes(4), True, tiles, 1)
    assert np.array_equal(got, want)


def test_blend16_edges(ctx):
    """16-bit thresholds: alpha 0..8 and 65527..65535 (pixel-ops.ts:670-677) plus random."""
    rng = np.random.default_rng(2)
    alphas = np.concatenate([np.arange(0, 9), np.arange(65527, 65536), rng.integers(0, 65536, 238)]).astype(np.uint32)
    w, h = 256, 8

    def img(alpha_row):
        v = rng.integers(0, 65536, size=(h, w, 4)).astype(np.uint32)
        v[..., 3] = alpha_row
        out = np.zeros((h, w, 4, 2), np.uint8)
        out[..., 0] = v >> 8
        out[..., 1] = v & 255
        return out.reshape(h, -1)
    tiles = [dict(width=w, height=h, bit_depth=16, color_type=6, pixels=img(rng.permutation(alphas)[None, :]), z_index=0, image_idx=0),
             dict(width=w, height=h, bit_depth=16, color_type=6, pixels=img(alphas[None, :]), z_index=1, image_idx=1)]
    for bg in (bytes(8), bytes([255] * 8)):
        want = orc.render_rows(w, h, 16, bg, True, tiles, 0, h, 1)
        got = run_job(ctx, w, h, 16, bg, True, tiles, 1)
        assert np.array_equal(got, want)


def test_bands_equal_whole(ctx):
    """Row bands with the recomputed 1-row halo give the same bytes as one pass."""
    rng = np.random.default_rng(3)
    cw, ch = 130, 97
    tiles = []
    for i in range(6):
        w, h = 60, 50
        tiles.append(dict(width=w, height=h, bit_depth=8, color_type=6, pixels=H.synth_pixels(rng, w, h, 8, 6, "any"),
                          dst_x=int(rng.integers(0, cw - w)), dst_y=int(rng.integers(0, ch - h)), z_index=i, image_idx=i))
    want = orc.render_rows(cw, ch, 8, bytes(4), True, tiles, 0, ch, 0)
    whole = run_job(ctx, cw, ch, 8, bytes(4), True, tiles, 0)
    banded = run_job(ctx, cw, ch, 8, bytes(4), True, tiles, 0, bands=[(0, 1), (1, 16), (16, 17), (17, 64), (64, 97)])
    assert np.array_equal(whole, want) and np.array_equal(banded, want)


def test_filtered_upload_path(ctx):
    """Tiles given as inflated filtered streams go through kernel (a) inside the job."""
    rng = np.random.default_rng(4)
    tiles, streams = [], []
    for i, (bd, ct) in enumerate([(8, 6), (8, 2), (16, 6), (2, 0)]):
        w, h = 40 + i, 33
        pix = H.synth_pixels(rng, w, h, bd, ct, "any")
        bpp = max(1, {0: 1, 2: 3, 4: 2, 6: 4}[ct] * bd // 8)
        tiles.append(dict(width=w, height=h, bit_depth=bd, color_type=ct, pixels=pix, dst_x=i * 45, dst_y=0, image_idx=i))
        streams.append(H.filter_image(pix, bpp, "random", rng))
    cw, ch = 45 * 4, 33
    want = orc.render_rows(cw, ch, 16, bytes(8), False, tiles, 0, ch, 0)
    got = run_job(ctx, cw, ch, 16, bytes(8), False, tiles, 0, filtered_streams=streams)
    assert np.array_equal(got, want)


def test_empty_canvas_and_single_pixel(ctx):
    bg = bytes([1, 2, 3, 4])
    want = orc.render_rows(5, 3, 8, bg, True, [], 0, 3, 0)
    got = run_job(ctx, 5, 3, 8, bg, True, [], 0)
    assert np.array_equal(got, want)
    t = [dict(width=1, height=1, bit_depth=8, color_type=6, pixels=np.array([[200, 100, 50, 128]], np.uint8))]
    want = orc.render_rows(1, 1, 8, bg, True, t, 0, 1, 0)
    assert np.array_equal(run_job(ctx, 1, 1, 8, bg, True, t, 0), want)


@pytest.mark.parametrize("two_pass", [False, True])
@pytest.mark.parametrize("out_bd,cw,ch", [(8, 1301, 75), (8, 2048, 33), (16, 700, 50), (8, 513, 17), (16, 257, 16)])
def test_wide_canvas_multi_segment(ctx, monkeypatch, two_pass, out_bd, cw, ch):
    """Rows spanning several 2 KB segments (cross-CTA cost reduction), ragged right edges,
    bands that start inside an item; both the fused kernel and the two-pass fallback."""
    if two_pass:
        monkeypatch.setenv("STITCH_B200_TWO_PASS", "1")
    rng = np.random.default_rng(cw + ch)
    tiles = []
    for i in range(10):
        bd, ct = [(8, 6), (8, 2), (16, 6), (8, 4)][i % 4]
        w, h = int(rng.integers(40, 700)), int(rng.integers(8, 60))
        x, y = int(rng.integers(-50, cw - 20)), int(rng.integers(-10, ch - 4))
        left, top, right, bottom = max(0, x), max(0, y), min(cw, x + w), min(ch, y + h)
        tiles.append(dict(width=w, height=h, bit_depth=bd, color_type=ct,
                          pixels=H.synth_pixels(rng, w, h, bd, ct, ["opaque", "any"][i % 2]),
                          dst_x=left, dst_y=top, src_off_x=max(0, -x), src_off_y=max(0, -y), vis_w=right - left,
                          vis_h=bottom - top, z_index=int(rng.integers(0, 3)), image_idx=i))
    bg = bytes([7, 7, 9, 9, 11, 11, 200, 200]) if out_bd == 16 else bytes([7, 9, 11, 200])
    for blend in (True, False):
        want = orc.render_rows(cw, ch, out_bd, bg, blend, tiles, 0, ch, 0)
        got = run_job(ctx, cw, ch, out_bd, bg, blend, tiles, 0)
        assert np.array_equal(got, want), (out_bd, cw, ch, blend)
        cuts = sorted(set([0, 1, min(ch, 5), min(ch, 16), min(ch, 23), ch]))
        banded = run_job(ctx, cw, ch, out_bd, bg, blend, tiles, 0, bands=list(zip(cuts, cuts[1:])))
        assert np.array_equal(banded, want)


def test_filter_costs_extreme_bytes(ctx):
    """Bytes around 0x80 and wrap-arounds: the |signed byte| cost must match for every candidate."""
    rng = np.random.default_rng(11)
    w, h = 1024, 40
    pix = rng.choice(np.array([0, 1, 127, 128, 129, 254, 255], np.uint8), size=(h, w * 4))
    tiles = [dict(width=w, height=h, bit_depth=8, color_type=6, pixels=pix)]
    want = orc.render_rows(w, h, 8, bytes(4), False, tiles, 0, h, 0)
    got = run_job(ctx, w, h, 8, bytes(4), False, tiles, 0)
    assert np.array_equal(got, want)
    pix2 = rng.integers(0, 256, size=(h, w * 4), dtype=np.uint8)   # noise: near-ties between filters
    tiles = [dict(width=w, height=h, bit_depth=8, color_type=6, pixels=pix2)]
    want = orc.render_rows(w, h, 8, bytes(4), False, tiles, 0, h, 0)
    assert np.array_equal(run_job(ctx, w, h, 8, bytes(4), False, tiles, 0), want)


def test_chunked_download_seams_and_redecode(ctx, monkeypatch):
    """stitch_job_run_band with a host destination works in chunks (uploads, kernels and downloads on three
    streams, only the tiles a chunk touches are decoded before it): tiny chunks put the seams at arbitrary rows
    of filtered tiles; stitch_job_invalidate + run_band must reproduce the same bytes."""
    rng = np.random.default_rng(77)
    cw, ch = 300, 333
    tiles, streams = [], []
    for i in range(9):
        w, h = int(rng.integers(60, 140)), int(rng.integers(60, 140))
        pix = H.synth_pixels(rng, w, h, 8, 6, ["opaque", "any"][i % 2])
        streams.append(H.filter_image(pix, 4, "random", rng))
        x, y = (i % 3) * 100, (i // 3) * 111
        tiles.append(dict(width=w, height=h, bit_depth=8, color_type=6, pixels=pix, dst_x=x, dst_y=y,
                          vis_w=min(w, cw - x), vis_h=min(h, ch - y), z_index=i, image_idx=i))
    want = orc.render_rows(cw, ch, 8, bytes(4), True, tiles, 0, ch, 0)
    for chunk_bytes in ("1", "40000", "1000000000"):
        monkeypatch.setenv("STITCH_CHUNK_BYTES", chunk_bytes)
        with _native.Job(ctx, cw, ch, 8, _native.OUT_PNG_FILTERED, True, bytes(4)) as job:
            for t, st in zip(tiles, streams):
                job.add_tile(st, width=t["width"], height=t["height"], bit_depth=8, color_type=6, filtered=1,
                             dst_x=t["dst_x"], dst_y=t["dst_y"], vis_w=t["vis_w"], vis_h=t["vis_h"], z_index=t["z_index"],
                             image_idx=t["image_idx"])
            got, _ = job.run_band(0, ch)
            assert np.array_equal(np.asarray(got), want), chunk_bytes
            job.invalidate()
            again, tm = job.run_band(0, ch)
            assert np.array_equal(np.asarray(again), want), chunk_bytes
            assert tm["launches"] >= 2
