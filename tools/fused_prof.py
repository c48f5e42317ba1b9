"""Composite+filter launch alone on the bench workloads, inputs resident (for ncu and quick timings).

  python tools/fused_prof.py c2 [runs]            16x16 grid of 1024^2 tiles -> 16384^2
  python tools/fused_prof.py c3 [runs] [y0 y1]    256 positioned 2048^2 tiles on 32768^2, blending, rows [y0,y1)
  python tools/fused_prof.py c3nb ...             the same without alpha blending
GPU box only."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from image_stitch_b200 import _native, synth  # noqa: E402


def c2(runs):
    grid, tile, canvas = 16, 1024, 16384
    rng = np.random.default_rng(2)
    raws = [synth.smooth_rgba8(rng, tile, tile, "opaque") for _ in range(4)]
    ctx = _native.Context(0)
    with _native.Job(ctx, canvas, canvas, 8, _native.OUT_PNG_FILTERED, False, bytes(4)) as job:
        for r in range(grid):
            for c in range(grid):
                job.add_tile(raws[(r * 5 + c * 3) % 4].reshape(-1), width=tile, height=tile, bit_depth=8, color_type=6,
                             filtered=0, dst_x=c * tile, dst_y=r * tile, image_idx=r * grid + c)
        ms = [job.run_band(0, canvas, download=False)[1]["compose_ms"] for _ in range(runs)]
    print("c2 compose+filter ms:", [round(m, 3) for m in ms], f"best {canvas * canvas / min(ms) / 1e3:.0f} MPix/s "
          f"{8 * canvas * canvas / min(ms) / 1e6:.0f} GB/s", flush=True)
    ctx.close()


def c3(runs, blend, y0, y1):
    from test_gpu_fullsize import _c3_tiles
    canvas, tile = 32768, 2048
    tiles = _c3_tiles(np.random.default_rng(3), canvas, 16, tile)
    ctx = _native.Context(0)
    with _native.Job(ctx, canvas, canvas, 8, _native.OUT_PNG_FILTERED, blend, bytes([255, 255, 255, 128])) as job:
        for t in tiles:
            job.add_tile(t["pixels"].reshape(-1), width=tile, height=tile, bit_depth=8, color_type=6, filtered=0,
                         dst_x=t["dst_x"], dst_y=t["dst_y"], src_off_x=t["src_off_x"], src_off_y=t["src_off_y"],
                         vis_w=t["vis_w"], vis_h=t["vis_h"], z_index=t["z_index"], image_idx=t["image_idx"])
        ms = [job.run_band(y0, y1, download=False)[1]["compose_ms"] for _ in range(runs)]
    px = (y1 - y0) * canvas
    print(f"c3 blend={blend} rows [{y0},{y1}) compose+filter ms:", [round(m, 3) for m in ms],
          f"best {px / min(ms) / 1e3:.0f} MPix/s {8 * px / min(ms) / 1e6:.0f} GB/s", flush=True)
    ctx.close()


if __name__ == "__main__":
    what = sys.argv[1] if len(sys.argv) > 1 else "c2"
    runs = int(sys.argv[2]) if len(sys.argv) > 2 else 4
    if what == "c2":
        c2(runs)
    else:
        y0 = int(sys.argv[3]) if len(sys.argv) > 4 else 4096
        y1 = int(sys.argv[4]) if len(sys.argv) > 4 else 8192
        c3(runs, what == "c3", y0, y1)
