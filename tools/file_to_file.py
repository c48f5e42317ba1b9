"""PNG files in, PNG file out through the public API (concat_to_buffer): host zlib both ways (the reference's
arrangement) against inflate and deflate on the device.  python tools/file_to_file.py [grid] [tile]"""
import sys, os, time, zlib
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import image_stitch_b200 as stitch
from image_stitch_b200 import synth
from image_stitch_b200.png_container import PNG_SIGNATURE, PngHeader, create_chunk, create_ihdr, create_iend, read_png, serialize_chunk

_pos = [a for a in sys.argv[1:] if not a.startswith("--")]
G = int(_pos[0]) if len(_pos) > 0 else 8
T = int(_pos[1]) if len(_pos) > 1 else 1024
rng = np.random.default_rng(2)
files = []
for i in range(4):
    st = bytes(synth.encode_filtered(synth.smooth_rgba8(rng, T, T, "opaque"), 4))
    png = PNG_SIGNATURE + serialize_chunk(create_ihdr(PngHeader(T, T, 8, 6, 0, 0, 0))) + \
        serialize_chunk(create_chunk("IDAT", zlib.compress(st, 6))) + serialize_chunk(create_iend())
    files.append(png)
inputs = [files[i % 4] for i in range(G * G)]
print(f"{G*G} PNG tiles of {T}x{T} ({sum(len(f) for f in inputs)/1e6:.0f} MB of files) -> {G*T}x{G*T} PNG")
results = {}
for name, extra in (("device inflate + deflate", {"device_inflate": True, "device_deflate": True}),
                    ("device deflate only", {"device_deflate": True}),
                    ("host zlib both ways", {})):
    if name == "host zlib both ways" and G * T > 4096 and "--all" not in sys.argv:
        print(f"{name}: skipped above 4096 px (pass --all)"); continue
    best = 1e9
    for rep in range(2):
        t0 = time.perf_counter()
        out = stitch.concat_to_buffer(dict({"inputs": inputs, "layout": {"columns": G}}, **extra))
        best = min(best, time.perf_counter() - t0)
    results[name] = out
    print(f"{name}: {best*1e3:.0f} ms, {len(out)/1e6:.1f} MB out, {G*T*G*T/best/1e6:.0f} MPix/s")
if "--jpeg" in sys.argv:   # the same files to one JPEG (q = 90): the device inflates, unfilters and codes; the host splices the bands
    for name, extra in (("device inflate", {"device_inflate": True}),):
        best = 1e9
        for rep in range(3):
            t0 = time.perf_counter()
            jpg = stitch.concat_to_buffer(dict({"inputs": inputs, "layout": {"columns": G}, "output_format": "jpeg", "jpeg_quality": 90}, **extra))
            best = min(best, time.perf_counter() - t0)
        print(f"JPEG, {name}: {best*1e3:.0f} ms, {len(jpg)/1e6:.1f} MB out, {G*T*G*T/best/1e6:.0f} MPix/s")
rows = [read_png(o)[1] for o in results.values()]
assert all(r == rows[0] for r in rows), "outputs differ"
print("all outputs decode to the same filtered rows")
