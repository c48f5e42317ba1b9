"""Kernel (a) by pixel format: 64 tiles of 1024 x 1024, random filter type per row, streams resident in HBM."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from image_stitch_b200 import _native, synth

tile, n = 1024, 64
rng = np.random.default_rng(4)
ctx = _native.Context(0)
for name, bd, ct, spp in (("RGBA8", 8, 6, 4), ("RGBA16", 16, 6, 4), ("RGB8", 8, 2, 3), ("RGB16", 16, 2, 3), ("GA8", 8, 4, 2),
                          ("GA16", 16, 4, 2), ("gray8", 8, 0, 1), ("gray16", 16, 0, 1)):
    bpp = spp * bd // 8
    rb = tile * bpp
    pix = [rng.integers(0, 256, size=(tile, rb), dtype=np.uint8) for _ in range(2)]
    for force, fname in ((5, "random"), (3, "Average")):
        streams = [torch.from_numpy(synth.encode_filtered(p, bpp, force, rng)).cuda() for p in pix]
        with _native.Job(ctx, tile * 8, tile * 8, 16 if bd == 16 else 8, _native.OUT_PNG_FILTERED, False, bytes(8)) as job:
            for k in range(n):
                s = streams[k % 2]
                job.add_tile(None, device_ptr=s.data_ptr(), length=s.numel(), width=tile, height=tile, bit_depth=bd,
                             color_type=ct, filtered=1, dst_x=(k % 8) * tile, dst_y=(k // 8) * tile, image_idx=k)
            ms = min(job.decode(force=True)["decode_ms"] for _ in range(5))
        print(f"{name:7s} bpp={bpp} {fname:8s} {ms:.3f} ms  {2 * n * tile * rb / ms / 1e6:7.0f} GB/s  {n * tile * tile / ms / 1e3:7.0f} MPix/s", flush=True)
ctx.close()
