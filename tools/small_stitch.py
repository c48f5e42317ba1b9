import sys, os
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
import numpy as np
from image_stitch_b200 import _native, synth
ctx = _native.Context(0)
rng = np.random.default_rng(1)
G=int(sys.argv[1]) if len(sys.argv)>1 else 4; T=int(sys.argv[2]) if len(sys.argv)>2 else 256
pix = synth.smooth_rgba8(rng, T, T, "opaque"); st = synth.encode_filtered(pix, 4)
job = _native.Job(ctx, T*G, T*G, 8, _native.OUT_PNG_FILTERED, False, bytes(4))
for r in range(G):
    for c in range(G):
        job.add_tile(st, width=T, height=T, bit_depth=8, color_type=6, interlace=0, filtered=1, dst_x=c*T, dst_y=r*T, z_index=0, image_idx=r*G+c, length=len(st))
for i in range(3):
    out, tm = job.run_band(0, T*G)
print("ok", tm)
