"""Timeline of one end-to-end step (C2, pinned host buffers both ways): STITCH_TRACE=1 makes run_band print when
each chunk was rendered and when it had left the device."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from image_stitch_b200 import _native, synth
T, G = 1024, 16
C = T * G
rng = np.random.default_rng(2)
st = torch.from_numpy(synth.encode_filtered(synth.smooth_rgba8(rng, T, T, "opaque"), 4))
ctx = _native.Context(0)
host_out = torch.empty(C * (C * 4 + 1), dtype=torch.uint8, pin_memory=True)
host_in = torch.empty((G * G, st.numel()), dtype=torch.uint8, pin_memory=True)
for k in range(G * G):
    host_in[k].copy_(st)
for i in range(3):
    if i == 2: os.environ["STITCH_TRACE"] = "1"
    torch.cuda.synchronize(); t0 = time.perf_counter()
    j = _native.Job(ctx, C, C, 8, _native.OUT_PNG_FILTERED, False, bytes(4))
    for r in range(G):
        for c in range(G):
            k = r * G + c
            j.add_tile(host_in[k], length=st.numel(), width=T, height=T, bit_depth=8, color_type=6, filtered=1,
                       dst_x=c * T, dst_y=r * T, image_idx=k)
    out, tm = j.run_band(0, C, out=host_out)
    j.close(); torch.cuda.synchronize()
    print("e2e", round((time.perf_counter() - t0) * 1e3, 2), "ms")
