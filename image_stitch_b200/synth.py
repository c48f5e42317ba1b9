"""Synthetic workloads of BASELINE.json / SURVEY.md 8(d), built with numpy only
(no oracle): tiles are stored as inflated, filtered scanline streams, encoded
with the usual min-sum-of-abs heuristic so the filter-type mix is realistic."""
from __future__ import annotations

import numpy as np


def smooth_rgba8(rng, w: int, h: int, alpha: str = "opaque") -> np.ndarray:
    """Smooth gradient + uniform noise +-8 per channel (C2/C3 pixel model). Returns [h, w*4] uint8."""
    yy, xx = np.mgrid[0:h, 0:w].astype(np.int32)
    img = np.empty((h, w, 4), np.uint8)
    fx, fy = rng.integers(1, 4, 3), rng.integers(1, 4, 3)
    for c in range(3):
        base = ((xx * int(fx[c]) + yy * int(fy[c])) >> 3) + int(rng.integers(0, 256))
        tri = np.abs((base % 510) - 255)
        img[..., c] = np.clip(tri + rng.integers(-8, 9, size=(h, w)), 0, 255).astype(np.uint8)
    if alpha == "opaque":
        img[..., 3] = 255
    elif alpha == "uniform":
        img[..., 3] = rng.integers(1, 255, size=(h, w))
    elif alpha == "binary":
        img[..., 3] = rng.integers(0, 2, size=(h, w)) * 255
    return img.reshape(h, w * 4)


def encode_filtered(pix: np.ndarray, bpp: int, force: int = -1, rng=None) -> np.ndarray:
    """[h, n] raw rows -> inflated IDAT stream [h*(n+1)] (PNG encoder side, vectorised over rows).
    force: -1 = min-sum-of-abs heuristic, 0..4 = that filter on every row, 5 = a random type per row (rng)."""
    h, n = pix.shape
    x = pix.astype(np.int16)
    a = np.zeros_like(x)
    a[:, bpp:] = x[:, :-bpp]
    b = np.zeros_like(x)
    b[1:] = x[:-1]
    c = np.zeros_like(x)
    c[1:, bpp:] = x[:-1, :-bpp]
    p = a + b - c
    pa, pb, pc = np.abs(p - a), np.abs(p - b), np.abs(p - c)
    pae = np.where((pa <= pb) & (pa <= pc), a, np.where(pb <= pc, b, c))
    cands = [x, x - a, x - b, x - ((a + b) >> 1), x - pae]
    cands = [(v & 0xFF).astype(np.uint8) for v in cands]
    if force == 5:
        best = (rng or np.random.default_rng(0)).integers(0, 5, size=h)
    elif force >= 0:
        best = np.full(h, force, np.int64)
    else:
        costs = np.stack([np.where(v > 127, 256 - v.astype(np.int32), v.astype(np.int32)).sum(axis=1) for v in cands])
        best = np.argmin(costs, axis=0)   # first minimum wins, like png-filter.ts:174
    out = np.empty((h, n + 1), np.uint8)
    out[:, 0] = best
    stack = np.stack(cands)               # [5, h, n]
    out[:, 1:] = stack[best, np.arange(h)]
    return out.reshape(-1)


def c3_layout(rng, canvas: int, n_side: int, tile: int):
    """n_side^2 tiles at uniform random positions in [-256, canvas - tile + 256] (about a tenth
    are clipped), zIndex a random permutation with 16 deliberate duplicates (imageIdx breaks the tie).  Tile i shows
    content i % 4 of (opaque, opaque, alpha uniform in 1..254, alpha a 0/255 mask): half of the tiles opaque, a
    quarter translucent, a quarter masked.  Returns placement dicts (no pixels; key "content")."""
    n = n_side * n_side
    z = rng.permutation(n)
    z[rng.choice(n, 16, replace=False)] = z[rng.choice(n, 16, replace=False)]
    tiles = []
    for i in range(n):
        x = int(rng.integers(-256, canvas - tile + 257))
        y = int(rng.integers(-256, canvas - tile + 257))
        left, top, right, bottom = max(0, x), max(0, y), min(canvas, x + tile), min(canvas, y + tile)
        tiles.append(dict(width=tile, height=tile, bit_depth=8, color_type=6, content=i % 4, dst_x=left, dst_y=top,
                          src_off_x=left - x, src_off_y=top - y, vis_w=right - left, vis_h=bottom - top,
                          z_index=int(z[i]), image_idx=i))
    return tiles


C3_CONTENT = ("opaque", "opaque", "uniform", "binary")
