"""Randomised parity run of composite + filter selection against the oracle: random canvases, tiles of random sizes,
formats, alpha kinds, positions (clipped at every edge, overlapping, leaving gaps), z orders, blending on and off, 8- and
16-bit output, random band cuts -- through the route the environment selects (STITCH_SINGLE_PASS, STITCH_STRIP).
python tools/compose_fuzz.py [seconds] [seed].  GPU box only."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import helpers as H
import oracle_py as orc
from image_stitch_b200 import _native, synth

FORMATS = [(8, 6), (8, 6), (8, 6), (8, 2), (8, 0), (8, 4), (16, 6), (16, 2), (16, 0), (16, 4)]


def make_tile(rng, w, h, alpha, x, y, z, idx, cw, ch, bit_depth, color_type):
    if bit_depth == 8 and color_type == 6:
        pix = synth.smooth_rgba8(rng, w, h, alpha)
    else:
        pix = H.synth_pixels(rng, w, h, bit_depth, color_type, "opaque" if alpha == "opaque" else "any")
    left, top, right, bottom = max(0, x), max(0, y), min(cw, x + w), min(ch, y + h)
    return dict(width=w, height=h, bit_depth=bit_depth, color_type=color_type, pixels=pix, dst_x=left, dst_y=top,
                src_off_x=left - x, src_off_y=top - y, vis_w=max(0, right - left), vis_h=max(0, bottom - top),
                z_index=z, image_idx=idx)


def run(budget, seed, context=None):
    """Random scenes for `budget` seconds; returns their number, or -1 after printing the first mismatch."""
    rng = np.random.default_rng(seed)
    ctx = context or _native.Context(0)
    n, t_end = 0, time.time() + budget
    while time.time() < t_end:
        cw = int(rng.choice([rng.integers(1, 200), rng.integers(200, 1400), rng.integers(1400, 3300)]))
        ch = int(rng.choice([rng.integers(1, 30), rng.integers(30, 200)]))
        out_depth = 16 if rng.integers(0, 5) == 0 else 8
        blend = bool(rng.integers(0, 2))
        bg = bytes(int(v) for v in rng.integers(0, 256, 8)) if rng.integers(0, 2) else bytes(8)
        tiles = []
        aligned = rng.integers(0, 3) == 0   # now and then: edges on 8-row / 128-pixel boundaries
        for i in range(int(rng.integers(1, 10))):
            w = int(rng.choice([rng.integers(1, 40), rng.integers(40, cw + 60)]))
            h = int(rng.choice([rng.integers(1, 12), rng.integers(12, ch + 40)]))
            x, y = int(rng.integers(-w + 1, cw)), int(rng.integers(-h + 1, ch))
            if aligned:
                x, y, w, h = x // 128 * 128, y // 8 * 8, max(128, w // 128 * 128), max(8, h // 8 * 8)
            bd, ct = FORMATS[int(rng.integers(0, len(FORMATS)))]
            alpha = ["opaque", "uniform", "binary"][int(rng.integers(0, 3))]
            tiles.append(make_tile(rng, w, h, alpha, x, y, int(rng.integers(0, 4)), i, cw, ch, bd, ct))
        vis = [t for t in tiles if t["vis_w"] and t["vis_h"]]
        cuts = [0]
        while cuts[-1] < ch:
            cuts.append(min(ch, cuts[-1] + int(rng.integers(1, max(2, ch)))))
        bands = list(zip(cuts[:-1], cuts[1:]))
        got = []
        with _native.Job(ctx, cw, ch, out_depth, _native.OUT_PNG_FILTERED, blend, bg) as job:
            for t in vis:
                job.add_tile(np.ascontiguousarray(t["pixels"]).reshape(-1), width=t["width"], height=t["height"],
                             bit_depth=t["bit_depth"], color_type=t["color_type"], filtered=0, dst_x=t["dst_x"], dst_y=t["dst_y"],
                             src_off_x=t["src_off_x"], src_off_y=t["src_off_y"], vis_w=t["vis_w"], vis_h=t["vis_h"],
                             z_index=t["z_index"], image_idx=t["image_idx"])
            for a, b in bands:
                out, _ = job.run_band(a, b)
                got.append(np.array(out, copy=True))
        for (a, b), g in zip(bands, got):
            want = orc.render_rows(cw, ch, out_depth, bg, blend, vis, a, b, 0)
            if not np.array_equal(g, want):
                row = cw * (8 if out_depth == 16 else 4) + 1
                bad = np.nonzero(g != want)[0]
                print(f"MISMATCH: canvas {cw}x{ch} depth {out_depth} blend {blend} bg {bg.hex()} band [{a},{b}): {bad.size} bytes differ, "
                      f"first at row {a + bad[0] // row} byte {bad[0] % row}; tiles: "
                      + "; ".join(f"{t['width']}x{t['height']} {t['bit_depth']}/{t['color_type']} at ({t['dst_x']},{t['dst_y']}) off ({t['src_off_x']},{t['src_off_y']}) vis {t['vis_w']}x{t['vis_h']} z{t['z_index']}" for t in vis))
                return -1
        n += 1
    return n


if __name__ == "__main__":
    n_ok = run(float(sys.argv[1]) if len(sys.argv) > 1 else 60.0, int(sys.argv[2]) if len(sys.argv) > 2 else 1)
    if n_ok < 0:
        sys.exit(1)
    print(f"{n_ok} random scenes byte-exact against the oracle")
