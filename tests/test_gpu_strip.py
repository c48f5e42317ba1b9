"""The composite+filter kernels (csrc/fused.cu and the row-pipelined csrc/strip.cu) against the oracle, bit-exact, on canvases wide enough
for several column strips and TMA boxes: direct tile boxes at arbitrary pixel offsets, background boxes,
composited boxes (tile edges, blending, foreign formats), the pixel handed across a strip seam, overlays whose
edges sit on 8-row / 128-pixel boundaries (the unit kernel took its predictor row from the wrong
tile there), 16-bit output, and bands cut at arbitrary rows."""
import numpy as np
import pytest

import helpers as H
import oracle_py as orc
from image_stitch_b200 import _native, synth

pytestmark = pytest.mark.gpu


@pytest.fixture(params=["unit", "unit_plan", "via_canvas", "strip"], autouse=True)
def kernel(request, monkeypatch):
    """Every case runs through every composite+filter route: the unit kernel (csrc/fused.cu) as the library picks
    its mode, the unit kernel in plan mode for every band (STITCH_SINGLE_PASS=2), busy bands through the canvas
    (STITCH_SINGLE_PASS=0), and the row-pipelined kernel (csrc/strip.cu, STITCH_STRIP=1).  The library reads the
    variables at every call."""
    monkeypatch.delenv("STITCH_STRIP", raising=False)
    monkeypatch.delenv("STITCH_SINGLE_PASS", raising=False)
    if request.param == "strip":
        monkeypatch.setenv("STITCH_STRIP", "1")
    elif request.param == "unit_plan":
        monkeypatch.setenv("STITCH_SINGLE_PASS", "2")
    elif request.param == "via_canvas":
        monkeypatch.setenv("STITCH_SINGLE_PASS", "0")
    return request.param


def _tile(rng, w, h, alpha, x, y, z, idx, canvas_w, canvas_h, bit_depth=8, color_type=6):
    if bit_depth == 8 and color_type == 6:
        pix = synth.smooth_rgba8(rng, w, h, alpha)
    else:
        pix = H.synth_pixels(rng, w, h, bit_depth, color_type, "opaque" if alpha == "opaque" else "any")
    left, top, right, bottom = max(0, x), max(0, y), min(canvas_w, x + w), min(canvas_h, y + h)
    return dict(width=w, height=h, bit_depth=bit_depth, color_type=color_type, pixels=pix, dst_x=left, dst_y=top,
                src_off_x=left - x, src_off_y=top - y, vis_w=max(0, right - left), vis_h=max(0, bottom - top),
                z_index=z, image_idx=idx)


def _render(ctx, cw, ch, out_depth, blend, bg, tiles, bands):
    got = []
    with _native.Job(ctx, cw, ch, out_depth, _native.OUT_PNG_FILTERED, blend, bg) as job:
        for t in tiles:
            if not t["vis_w"] or not t["vis_h"]:
                continue
            job.add_tile(np.ascontiguousarray(t["pixels"]).reshape(-1), width=t["width"], height=t["height"],
                         bit_depth=t["bit_depth"], color_type=t["color_type"], filtered=0, dst_x=t["dst_x"], dst_y=t["dst_y"],
                         src_off_x=t["src_off_x"], src_off_y=t["src_off_y"], vis_w=t["vis_w"], vis_h=t["vis_h"],
                         z_index=t["z_index"], image_idx=t["image_idx"])
        for a, b in bands:
            out, _ = job.run_band(a, b)
            got.append(np.array(out, copy=True))
    return got


def _check(ctx, cw, ch, out_depth, blend, bg, tiles, bands):
    got = _render(ctx, cw, ch, out_depth, blend, bg, tiles, bands)
    vis = [t for t in tiles if t["vis_w"] and t["vis_h"]]
    for (a, b), g in zip(bands, got):
        want = orc.render_rows(cw, ch, out_depth, bg, blend, vis, a, b, 0)
        if not np.array_equal(g, want):
            row = cw * (8 if out_depth == 16 else 4) + 1
            bad = np.nonzero(g != want)[0]
            raise AssertionError(f"band [{a},{b}): {bad.size} bytes differ, first at row {a + bad[0] // row} byte {bad[0] % row}")


def test_grid_of_wide_tiles_direct_boxes(ctx):
    """Plain mosaic, tile width a multiple of the box: every box is a TMA load; 3 strips at Q=1."""
    rng = np.random.default_rng(21)
    tw, th, nx, ny = 1024, 37, 5, 3
    cw, ch = tw * nx, th * ny
    tiles = [_tile(rng, tw, th, "opaque", c * tw, r * th, 0, r * nx + c, cw, ch) for r in range(ny) for c in range(nx)]
    _check(ctx, cw, ch, 8, False, bytes(4), tiles, [(0, ch), (5, 41), (36, 38)])


@pytest.mark.parametrize("blend", [False, True])
def test_tiles_at_odd_offsets_and_gaps(ctx, blend):
    """Tiles of odd widths at odd positions with gaps: direct boxes at unaligned source offsets, background boxes,
    composited edge boxes, canvas width not a multiple of anything."""
    rng = np.random.default_rng(22)
    cw, ch = 5003, 90
    tiles = []
    x = 3
    for i in range(6):
        w = int(rng.integers(500, 1100))
        tiles.append(_tile(rng, w, int(rng.integers(40, 90)), ["opaque", "uniform", "binary"][i % 3], x, int(rng.integers(0, 20)), i, i, cw, ch))
        x += w + int(rng.integers(-40, 60))
    _check(ctx, cw, ch, 8, blend, bytes([10, 200, 30, 128]), tiles, [(0, ch), (17, 60)])


@pytest.mark.parametrize("blend", [False, True])
def test_overlay_with_aligned_edges(ctx, blend):
    """A large base tile and opaque / translucent overlays whose bottom and right edges sit on 8-row and
    128-pixel boundaries (and on box and strip boundaries): the predictor row and the left pixel of what follows
    an overlay must come from the overlay."""
    rng = np.random.default_rng(23)
    cw, ch = 4608, 96
    tiles = [_tile(rng, cw, ch, "opaque", 0, 0, 0, 0, cw, ch),
             _tile(rng, 256, 32, "opaque", 256, 16, 1, 1, cw, ch),      # box-aligned, ends on row 48
             _tile(rng, 1024, 24, "opaque", 1024, 8, 1, 2, cw, ch),     # ends exactly at the strip seam x = 2048
             _tile(rng, 384, 40, "uniform", 2048, 24, 2, 3, cw, ch),    # starts at the seam, translucent
             _tile(rng, 128, 8, "opaque", 3968, 40, 3, 4, cw, ch)]
    _check(ctx, cw, ch, 8, blend, bytes(4), tiles, [(0, ch), (48, 49), (40, 72)])


def test_sixteen_bit_output_mixed_sources(ctx):
    rng = np.random.default_rng(24)
    cw, ch = 1500, 50
    tiles = [_tile(rng, 700, 50, "opaque", 0, 0, 0, 0, cw, ch, 16, 6),
             _tile(rng, 640, 45, "any", 690, 3, 1, 1, cw, ch, 16, 6),
             _tile(rng, 300, 30, "opaque", 1200, 10, 2, 2, cw, ch, 8, 2),
             _tile(rng, 256, 20, "any", 100, 5, 3, 3, cw, ch, 8, 4)]
    _check(ctx, cw, ch, 16, True, bytes([0, 1, 2, 3, 4, 5, 255, 255]), tiles, [(0, ch), (7, 33)])


def test_many_row_ranges_and_single_rows(ctx):
    """More rows than one range per CTA, and bands of one and two rows (ring shorter than its depth)."""
    rng = np.random.default_rng(25)
    cw, ch = 2300, 700
    tiles = [_tile(rng, 1200, 700, "opaque", 0, 0, 0, 0, cw, ch), _tile(rng, 1200, 650, "binary", 1100, 20, 1, 1, cw, ch)]
    _check(ctx, cw, ch, 8, True, bytes(4), tiles, [(0, ch), (0, 1), (699, 700), (350, 352), (1, 4)])
