"""C3 shape (256 positioned overlapping 2048x2048 RGBA8 tiles on a 32768-wide canvas): times the
composite+filter-select launch for a 4096-row band, alpha blending on and off.  GPU box only."""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from image_stitch_b200 import _native
from test_gpu_fullsize import _c3_tiles

canvas, tile = 32768, 2048
tiles = _c3_tiles(np.random.default_rng(3), canvas, 16, tile)
ctx = _native.Context(0)
for blend in (True, False):
    with _native.Job(ctx, canvas, canvas, 8, _native.OUT_PNG_FILTERED, blend, bytes([255, 255, 255, 128])) as job:
        for t in tiles:
            job.add_tile(t["pixels"].reshape(-1), width=tile, height=tile, bit_depth=8, color_type=6, filtered=0,
                         dst_x=t["dst_x"], dst_y=t["dst_y"], src_off_x=t["src_off_x"], src_off_y=t["src_off_y"],
                         vis_w=t["vis_w"], vis_h=t["vis_h"], z_index=t["z_index"], image_idx=t["image_idx"])
        ms = []
        got, tm = job.run_band(4096, 8192)          # with download first, like the test
        print("  with download:", {k: round(v, 3) for k, v in tm.items()})
        for _ in range(4):
            _, tm = job.run_band(4096, 8192, download=False)
            ms.append(tm["compose_ms"])
        px = 4096 * canvas
        print("  all runs (ms):", [round(m, 3) for m in ms])
        print(f"blend={blend}: compose+filter {min(ms):.3f} ms for 4096 x 32768 px = {px / min(ms) / 1e3:.0f} MPix/s", flush=True)
ctx.close()
