"""The bench's JPEG section (32768^2 mosaic, q = 90) with the entropy bytes downloaded, for several sub-band sizes
(STITCH_JPEG_CHUNK_BYTES): python tools/jpeg_d2h.py.  GPU box only."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench
from image_stitch_b200 import _native, synth

ctx = _native.Context(0)
dev = torch.device("cuda:0")
rng = np.random.default_rng(5)
T, S = bench.C3_TILE, bench.C3_SIDE
raws = [torch.from_numpy(synth.smooth_rgba8(rng, T, T, "opaque").reshape(-1)).to(dev) for _ in range(4)]
bufs = [raws[i % 4].clone() for i in range(S * S)]
with _native.Job(ctx, bench.C3_CANVAS, bench.C3_CANVAS, 8, _native.OUT_JPEG_ENTROPY, False, bytes(4), jpeg_quality=90) as job:
    for i, b in enumerate(bufs):
        job.add_tile(None, device_ptr=b.data_ptr(), length=b.numel(), width=T, height=T, bit_depth=8, color_type=6,
                     filtered=0, dst_x=(i % S) * T, dst_y=(i // S) * T, image_idx=i)
    job.run_band(0, bench.C3_CANVAS, download=False)
    n_bytes = job.band_device()[3]
    host = torch.empty(n_bytes + 64, dtype=torch.uint8, pin_memory=True)
    for mb in (64, 128, 256, 512, 1024, 8192):
        os.environ["STITCH_JPEG_CHUNK_BYTES"] = str(mb << 20)
        job.run_band(0, bench.C3_CANVAS, out=host)
        ts = []
        for _ in range(3):
            t0 = time.perf_counter()
            job.run_band(0, bench.C3_CANVAS, out=host)
            ts.append((time.perf_counter() - t0) * 1e3)
        print(f"sub-bands of {mb} MB of pixels: {min(ts):.2f} ms wall with the {n_bytes / 1e6:.0f} MB downloaded", flush=True)
