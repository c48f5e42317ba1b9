"""Host cost of queueing the 256 tiles of the bench's end-to-end step (stitch_job_add_tile from pinned buffers): about
12 us per tile = 3.1 ms before run_band can start, of which the Python side (descriptor, address) is 0.5 ms and five CUDA
runtime calls per tile (two stream-ordered allocations, the copy, event create + record) the rest.  GPU box only."""
import sys, time
import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from image_stitch_b200 import _native, synth
T, G = 1024, 16
C = T * G
rng = np.random.default_rng(2)
st = torch.from_numpy(synth.encode_filtered(synth.smooth_rgba8(rng, T, T, "opaque"), 4))
ctx = _native.Context(0)
host_in = torch.empty((G * G, st.numel()), dtype=torch.uint8, pin_memory=True)
for k in range(G * G): host_in[k].copy_(st)
for rep in range(3):
    torch.cuda.synchronize()
    j = _native.Job(ctx, C, C, 8, _native.OUT_PNG_FILTERED, False, bytes(4))
    t0 = time.perf_counter()
    for k in range(G * G):
        j.add_tile(host_in[k], length=st.numel(), width=T, height=T, bit_depth=8, color_type=6, filtered=1,
                   dst_x=(k % G) * T, dst_y=(k // G) * T, image_idx=k)
    t1 = time.perf_counter()
    torch.cuda.synchronize(); t2 = time.perf_counter()
    j.close()
    print(f"256 add_tile: {1e3*(t1-t0):.2f} ms issue, uploads done at {1e3*(t2-t0):.2f} ms")
t0 = time.perf_counter()
for k in range(256):
    td = _native.TileDesc(width=T, height=T, bit_depth=8, color_type=6, interlace=0, filtered=1, dst_x=0, dst_y=0, src_off_x=0, src_off_y=0, vis_w=T, vis_h=T, z_index=0, image_idx=k)
print(f"256 TileDesc: {1e3*(time.perf_counter()-t0):.2f} ms")
t0 = time.perf_counter()
for k in range(256): a = _native._addr(host_in[k])
print(f"256 _addr(tensor row): {1e3*(time.perf_counter()-t0):.2f} ms")
