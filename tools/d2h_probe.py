"""How fast does a finished band leave the device?  C2 canvas with the inputs resident: run_band with a pinned host
destination (chunked 2-D copies) against one flat copy of the same bytes."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from image_stitch_b200 import _native, synth
T, G = 1024, 16
C = T * G
rng = np.random.default_rng(2)
st = torch.from_numpy(synth.encode_filtered(synth.smooth_rgba8(rng, T, T, "opaque"), 4))
dev_in = st.cuda()
ctx = _native.Context(0)
host_out = torch.empty(C * (C * 4 + 1), dtype=torch.uint8, pin_memory=True)
job = _native.Job(ctx, C, C, 8, _native.OUT_PNG_FILTERED, False, bytes(4))
for r in range(G):
    for c in range(G):
        job.add_tile(None, device_ptr=dev_in.data_ptr(), length=st.numel(), width=T, height=T, bit_depth=8, color_type=6,
                     filtered=1, dst_x=c * T, dst_y=r * T, image_idx=r * G + c)
for i in range(3):
    t0 = time.perf_counter()
    out, tm = job.run_band(0, C, out=host_out)
    dt = time.perf_counter() - t0
    print("run_band to host:", round(dt * 1e3, 2), "ms wall", {k: round(v, 2) for k, v in tm.items()})
flat = torch.empty(C * (C * 4 + 1), dtype=torch.uint8, device="cuda")
for i in range(3):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    host_out.copy_(flat, non_blocking=True); torch.cuda.synchronize()
    print("flat D2H:", round((time.perf_counter() - t0) * 1e3, 2), "ms")

# ---- the e2e step of bench.py: uploads + kernels + downloads ----
host_in = torch.empty((G * G, st.numel()), dtype=torch.uint8, pin_memory=True)
for k in range(G * G):
    host_in[k].copy_(st)
def e2e(download, chunk=None):
    if chunk: os.environ["STITCH_CHUNK_BYTES"] = str(chunk)
    else: os.environ.pop("STITCH_CHUNK_BYTES", None)
    best = None
    for i in range(3):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        j = _native.Job(ctx, C, C, 8, _native.OUT_PNG_FILTERED, False, bytes(4))
        for r in range(G):
            for c in range(G):
                k = r * G + c
                j.add_tile(host_in[k], length=st.numel(), width=T, height=T, bit_depth=8, color_type=6, filtered=1,
                           dst_x=c * T, dst_y=r * T, image_idx=k)
        t1 = time.perf_counter()
        if download: out, tm = j.run_band(0, C, out=host_out)
        else: out, tm = j.run_band(0, C, download=False)
        j.close(); torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        if best is None or dt < best[0]: best = (dt, t1 - t0, tm)
    print(f"e2e download={download} chunk={chunk}: {best[0]*1e3:.2f} ms wall (add_tile calls {best[1]*1e3:.2f} ms)", {k: round(v, 2) for k, v in best[2].items()})
e2e(False)
e2e(True)
e2e(True, 16 << 20)
e2e(True, 256 << 20)
e2e(True, 2 << 30)

# ---- does the band download overlap with an unrelated big upload? ----
big_h = torch.empty(1 << 30, dtype=torch.uint8, pin_memory=True)
big_d = torch.empty(1 << 30, dtype=torch.uint8, device="cuda")
s_up = torch.cuda.Stream()
os.environ.pop("STITCH_CHUNK_BYTES", None)
for i in range(3):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    with torch.cuda.stream(s_up):
        big_d.copy_(big_h, non_blocking=True)
    out, tm = job.run_band(0, C, out=host_out)
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    print(f"resident band download + unrelated 1 GiB upload: band done {1e3*(t1-t0):.2f} ms, all done {1e3*(time.perf_counter()-t0):.2f} ms")
