import sys, os, time, zlib, cProfile, pstats
sys.path.insert(0, "/root/repo")
import numpy as np
import image_stitch_b200 as stitch
from image_stitch_b200 import synth
from image_stitch_b200.png_container import PNG_SIGNATURE, PngHeader, create_chunk, create_ihdr, create_iend, serialize_chunk
G, T = 16, 1024
rng = np.random.default_rng(2)
files = []
for i in range(4):
    st = bytes(synth.encode_filtered(synth.smooth_rgba8(rng, T, T, "opaque"), 4))
    files.append(PNG_SIGNATURE + serialize_chunk(create_ihdr(PngHeader(T, T, 8, 6, 0, 0, 0))) + serialize_chunk(create_chunk("IDAT", zlib.compress(st, 6))) + serialize_chunk(create_iend()))
inputs = [files[i % 4] for i in range(G * G)]
for fmt in ("png", "jpeg"):
    opts = {"inputs": inputs, "layout": {"columns": G}, "device_inflate": True, "device_deflate": True}
    if fmt == "jpeg": opts.update(output_format="jpeg", jpeg_quality=90); opts.pop("device_deflate")
    stitch.concat_to_buffer(opts)
    pr = cProfile.Profile(); pr.enable()
    t0 = time.perf_counter(); out = stitch.concat_to_buffer(opts); dt = time.perf_counter() - t0
    pr.disable()
    print(fmt, f"{dt*1e3:.0f} ms")
    pstats.Stats(pr).sort_stats("cumulative").print_stats(22)
