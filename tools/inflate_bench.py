"""Times inflate on the device alone: N zlib streams of 1024x1024 RGBA8 tiles (level-6 host zlib of
the filtered rows), one warp per stream.  python tools/inflate_bench.py [tiles] [size]"""
import sys, os, time, zlib
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from image_stitch_b200 import _native, synth

n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
T = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
ctx = _native.Context(0)
rng = np.random.default_rng(2)
distinct = []
for i in range(min(n, 8)):
    st = bytes(synth.encode_filtered(synth.smooth_rgba8(rng, T, T, "opaque"), 4))
    t0 = time.perf_counter(); z = zlib.compress(st, 6); t1 = time.perf_counter()
    t2 = time.perf_counter(); zlib.decompress(z); t3 = time.perf_counter()
    distinct.append((st, z))
print(f"tile {T}x{T}: {len(st)} B -> {len(z)} B (ratio {len(st)/len(z):.2f}); host zlib deflate {len(st)/(t1-t0)/1e6:.0f} MB/s, inflate {len(st)/(t3-t2)/1e6:.0f} MB/s (1 thread)")
G = int(np.ceil(np.sqrt(n)))
job = _native.Job(ctx, T * G, T * G, 8, _native.OUT_PNG_FILTERED, False, bytes(4))
raw = 0
for i in range(n):
    st, z = distinct[i % len(distinct)]
    job.add_tile(z, width=T, height=T, bit_depth=8, color_type=6, filtered=2, dst_x=(i % G) * T, dst_y=(i // G) * T, image_idx=i)
    raw += len(st)
best = 1e9
for it in range(4):
    tm = job.inflate(force=True)
    best = min(best, tm["decode_ms"])
print(f"inflate {n} streams: {best:.2f} ms, {raw / best / 1e6:.1f} GB/s of inflated bytes, {raw / n / best / 1e3:.1f} MB/s per stream")
out, tm = job.run_band(0, min(T * G, 64))
print("band ok", tm["launches"])
