"""C5 shape: grid mosaic -> JPEG q=90 entropy bytes on the GPU (one 16384 x 16384 band per run), device timings.
python tools/jpeg_bench.py [tile]   (tile = 1023: rows that are not 16-byte aligned, tile edges inside MCU rows)"""
import sys, os, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from image_stitch_b200 import _native, synth

grid, tile = 16, (int(sys.argv[1]) if len(sys.argv) > 1 else 1024)
canvas = grid * tile
rng = np.random.default_rng(5)
raws = [synth.smooth_rgba8(rng, tile, tile, "opaque") for _ in range(4)]
sigma = float(os.environ.get("JPEG_BENCH_NOISE", "0"))   # gaussian noise on top: JPEG_BENCH_NOISE=6 is a grainy photograph, 12 a very noisy one
if sigma > 0:
    for r in raws:
        rgb = r.reshape(tile, tile, 4)[..., :3]
        rgb[...] = np.clip(rgb.astype(np.float32) + rng.normal(0, sigma, rgb.shape), 0, 255).astype(np.uint8)
quality = int(os.environ.get("JPEG_BENCH_QUALITY", "90"))
ctx = _native.Context(0)
with _native.Job(ctx, canvas, canvas, 8, _native.OUT_JPEG_ENTROPY, False, bytes(4), jpeg_quality=quality) as job:
    for r in range(grid):
        for c in range(grid):
            job.add_tile(raws[(r * 5 + c * 3) % 4].reshape(-1), width=tile, height=tile, bit_depth=8, color_type=6,
                         filtered=0, dst_x=c * tile, dst_y=r * tile, image_idx=r * grid + c)
    for rows in (canvas, 2048 // 8 * 8):
        best = None
        for _ in range(4):
            t0 = time.perf_counter()
            _, tm = job.run_band(0, rows, download=False)
            dt = (time.perf_counter() - t0) * 1e3
            best = tm if best is None or tm["compose_ms"] < best["compose_ms"] else best
        bpp = 8.0 * job.band_device()[3] / (rows * canvas)
        print(f"{rows} rows ({bpp:.2f} bits/pixel): compose+jpeg {best['compose_ms']:.3f} ms on device ({rows * canvas / best['compose_ms'] / 1e3:.0f} MPix/s), "
              f"launches {best['launches']}, wall {dt:.2f} ms", flush=True)
ctx.close()
