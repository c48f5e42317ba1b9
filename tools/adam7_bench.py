"""Kernel (a) on Adam7-interlaced tiles: 64 tiles of 1024 x 1024, 7 pass chains each + the gather."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
from image_stitch_b200 import _native, synth

XS, YS, DX, DY = [0, 4, 0, 2, 0, 1, 0], [0, 0, 4, 0, 2, 0, 1], [8, 8, 4, 4, 2, 2, 1], [8, 8, 8, 4, 4, 2, 2]


def interlace(pix, w, h, bpp, rng):
    src = pix.reshape(h, w, bpp)
    parts = []
    for p in range(7):
        sub = np.ascontiguousarray(src[YS[p]::DY[p], XS[p]::DX[p]]).reshape(-1, (-(-(w - XS[p]) // DX[p])) * bpp)
        parts.append(synth.encode_filtered(sub, bpp, 5, rng))
    return np.concatenate(parts)


tile, n = 1024, 64
rng = np.random.default_rng(4)
ctx = _native.Context(0)
for name, bd, ct, spp in (("RGBA8", 8, 6, 4), ("RGB8", 8, 2, 3), ("RGBA16", 16, 6, 4), ("gray8", 8, 0, 1)):
    bpp = spp * bd // 8
    pix = [rng.integers(0, 256, size=(tile, tile * bpp), dtype=np.uint8) for _ in range(2)]
    streams = [torch.from_numpy(interlace(p, tile, tile, bpp, rng)).cuda() for p in pix]
    with _native.Job(ctx, tile * 8, tile * 8, 16 if bd == 16 else 8, _native.OUT_PNG_FILTERED, False, bytes(8)) as job:
        for k in range(n):
            s = streams[k % 2]
            job.add_tile(None, device_ptr=s.data_ptr(), length=s.numel(), width=tile, height=tile, bit_depth=bd,
                         color_type=ct, interlace=1, filtered=1, dst_x=(k % 8) * tile, dst_y=(k // 8) * tile, image_idx=k)
        ms = min(job.decode(force=True)["decode_ms"] for _ in range(5))
    print(f"{name:7s} Adam7 {ms:.3f} ms  {n * tile * tile / ms / 1e3:7.0f} MPix/s", flush=True)
ctx.close()
