"""Device JPEG decode of tiles: wall time from the first add_tile_jpeg to the composited band, for a
4 x 4 mosaic of 2048^2 4:2:0 q=85 tiles, with and without restart intervals.  GPU box only (needs Pillow to make the files).

  python tools/jpeg_decode_bench.py"""
import io
import os
import sys
import time

import numpy as np
from PIL import Image

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from image_stitch_b200 import _native, synth  # noqa: E402

tile, grid = 2048, 4
rng = np.random.default_rng(8)
pic = synth.smooth_rgba8(rng, tile, tile, "opaque").reshape(tile, tile, 4)[..., :3]
ctx = _native.Context(0)
for label, opts in (("no restart markers", dict()), ("restart every MCU row", dict(restart_marker_rows=1)),
                    ("restart every 8 MCUs", dict(restart_marker_blocks=8))):
    buf = io.BytesIO()
    Image.fromarray(np.ascontiguousarray(pic), "RGB").save(buf, "JPEG", quality=85, subsampling=2, **opts)
    data = buf.getvalue()
    t0 = time.perf_counter()
    Image.open(io.BytesIO(data)).load()
    cpu_ms = (time.perf_counter() - t0) * 1e3
    best = None
    for _ in range(3):
        with _native.Job(ctx, tile * grid, tile * grid, 8, _native.OUT_PNG_FILTERED, False, bytes(4)) as job:
            t0 = time.perf_counter()
            for k in range(grid * grid):
                job.add_tile_jpeg(data, dst_x=(k % grid) * tile, dst_y=(k // grid) * tile, image_idx=k)
            t1 = time.perf_counter()
            job.run_band(0, tile * grid, download=False)
            t2 = time.perf_counter()
        ms = ((t1 - t0) * 1e3, (t2 - t0) * 1e3)
        best = ms if best is None or ms[1] < best[1] else best
    px = grid * grid * tile * tile
    print(f"{label:24s} file {len(data) / 1e6:5.2f} MB: 16 tiles queued in {best[0]:7.2f} ms, decoded + composited + filtered at "
          f"{best[1]:8.2f} ms ({px / best[1] / 1e3:7.0f} MPix/s); libjpeg-turbo, one tile, one core: {cpu_ms:6.1f} ms "
          f"({tile * tile / cpu_ms / 1e3:5.0f} MPix/s)", flush=True)
ctx.close()
