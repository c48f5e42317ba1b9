"""Kernel (a) by filter type ("variants forced to each single filter type"): 256 tiles of
1024 x 1024 RGBA8, streams resident in HBM, stitch_job_decode(force) timed on the device.  GPU box only."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from image_stitch_b200 import _native, synth

tile, n = 1024, 256
rng = np.random.default_rng(2)
pix = [synth.smooth_rgba8(rng, tile, tile, "opaque") for _ in range(4)]
ctx = _native.Context(0)
names = {-1: "heuristic (1 Sub + 1023 Average rows)", 0: "None", 1: "Sub", 2: "Up", 3: "Average", 4: "Paeth", 5: "random per row"}
for force in (-1, 0, 1, 2, 3, 4, 5):
    streams = [torch.from_numpy(synth.encode_filtered(p, 4, force, rng)).cuda() for p in pix]
    with _native.Job(ctx, tile * 16, tile * 16, 8, _native.OUT_PNG_FILTERED, False, bytes(4)) as job:
        for k in range(n):
            s = streams[k % 4]
            job.add_tile(None, device_ptr=s.data_ptr(), length=s.numel(), width=tile, height=tile, bit_depth=8, color_type=6,
                         filtered=1, dst_x=(k % 16) * tile, dst_y=(k // 16) * tile, image_idx=k)
        ms = [job.decode(force=True)["decode_ms"] for _ in range(6)]
    best = min(ms[1:])
    print(f"{names[force]:40s} {best:.3f} ms  {n * tile * tile * 8 / best / 1e6:7.0f} GB/s", flush=True)
ctx.close()
