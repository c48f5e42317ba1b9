"""The bench's end-to-end step (256 tile uploads from pinned host memory + run_band into a pinned host buffer) with the
upload buffers allocated write-combined (cudaHostAllocWriteCombined) against plain page-locked memory.  GPU box only."""
import ctypes, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from image_stitch_b200 import _native, synth

T, G = 1024, 16
C = T * G
rng = np.random.default_rng(2)
st = np.asarray(synth.encode_filtered(synth.smooth_rgba8(rng, T, T, "opaque"), 4))
n = st.size
ctx = _native.Context(0)
import glob
_cands = glob.glob(os.path.join(os.path.dirname(torch.__file__), "..", "nvidia", "cuda_runtime", "lib", "libcudart.so*")) + ["libcudart.so.12", "libcudart.so"]
for _c in _cands:
    try:
        rt = ctypes.CDLL(_c); break
    except OSError:
        continue
rt.cudaHostAlloc.argtypes = [ctypes.POINTER(ctypes.c_void_p), ctypes.c_size_t, ctypes.c_uint]
rt.cudaFreeHost.argtypes = [ctypes.c_void_p]
out_row = C * 4 + 1
host_out = torch.empty(out_row * C, dtype=torch.uint8, pin_memory=True)


def alloc(flags):
    if flags is None:
        t = torch.empty((G * G, n), dtype=torch.uint8, pin_memory=True)
        return t.numpy(), t
    ptr = ctypes.c_void_p()
    err = rt.cudaHostAlloc(ctypes.byref(ptr), G * G * n, flags)
    assert int(err) == 0, err
    buf = (ctypes.c_uint8 * (G * G * n)).from_address(ptr.value)
    return np.frombuffer(buf, np.uint8).reshape(G * G, n), ptr


for name, flags in (("page-locked", None), ("write-combined", 4), ("page-locked", None), ("write-combined", 4)):
    arr, keep = alloc(flags)
    for k in range(G * G):
        arr[k] = st
    ts = []
    for rep in range(6):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        with _native.Job(ctx, C, C, 8, _native.OUT_PNG_FILTERED, False, bytes(4)) as j:
            for k in range(G * G):
                j.add_tile(arr[k], length=n, width=T, height=T, bit_depth=8, color_type=6, filtered=1,
                           dst_x=(k % G) * T, dst_y=(k // G) * T, image_idx=k)
            j.run_band(0, C, out=host_out)
        ts.append((time.perf_counter() - t0) * 1e3)
    print(f"{name}: {min(ts[1:]):.2f} ms best, {np.median(ts[1:]):.2f} ms median per step", flush=True)
    if flags is not None:
        rt.cudaFreeHost(keep)
