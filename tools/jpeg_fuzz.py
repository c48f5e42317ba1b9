"""Randomised parity run of the JPEG encoder against the oracle: random canvas sizes, qualities, contents (smooth,
steps, patches of noise: MCUs that do not fit their columns), mosaics of randomly sized tiles (read in place or through
the canvas), random band cuts.  python tools/jpeg_fuzz.py [seconds] [seed].  GPU box only."""
import io, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import oracle_py as orc
from image_stitch_b200 import _native
from image_stitch_b200.jpeg_container import JpegBandSplicer
try:
    from PIL import Image as _pil
except ImportError:
    _pil = None

rng = None
ctx = None


def content(w, h):
    yy, xx = np.mgrid[0:h, 0:w]
    img = np.zeros((h, w, 4), np.float64)
    for c in range(3):
        img[..., c] = 128 + rng.uniform(20, 100) * np.sin(xx / rng.uniform(5, 60) + c) * np.cos(yy / rng.uniform(5, 60))
    kind = rng.integers(0, 4)
    if kind >= 1:   # steps: rectangles of other levels
        for _ in range(int(rng.integers(1, 6))):
            x0, y0 = int(rng.integers(0, w)), int(rng.integers(0, h))
            img[y0:y0 + int(rng.integers(1, 64)), x0:x0 + int(rng.integers(1, 256)), :3] += rng.uniform(-120, 120, 3)
    if kind >= 2:   # patches of noise
        for _ in range(int(rng.integers(1, 4))):
            x0, y0 = int(rng.integers(0, w)), int(rng.integers(0, h))
            ph, pw = int(rng.integers(1, 40)), int(rng.integers(1, 300))
            sl = img[y0:y0 + ph, x0:x0 + pw, :3]
            sl[...] = rng.integers(0, 256, sl.shape)
    if kind == 3:
        img[..., :3] += rng.normal(0, rng.uniform(1, 14), (h, w, 3))
    img[..., 3] = rng.integers(0, 256)
    return np.clip(img, 0, 255).astype(np.uint8)


def cuts(total, lo, step):
    out, a = [0], 0
    while a < total:
        a = min(total, a + int(rng.integers(lo, lo + step)))
        out.append(a)
    return out


def run(budget, seed, context=None):
    """Random cases for `budget` seconds; returns their number, or -1 after printing the first mismatch."""
    global rng, ctx
    rng = np.random.default_rng(seed)
    ctx = context or _native.Context(0)
    n = 0
    t_end = time.time() + budget
    while time.time() < t_end:
        w = int(rng.choice([rng.integers(1, 64), rng.integers(64, 1500), rng.integers(1500, 4200)]))
        h = int(rng.choice([rng.integers(1, 40), rng.integers(40, 260)]))
        q = int(rng.choice([rng.integers(1, 101), 90, 95, 100]))
        canvas = content(w, h)
        # a mosaic of tiles cut from the canvas at random edges (multiples of 4 or not), so the expected canvas is known
        xs = cuts(w, 1, 1400) if rng.integers(0, 2) else [0, w]
        if rng.integers(0, 2):
            xs = sorted(set([0, w] + [min(w, (x + 3) // 4 * 4) for x in xs]))
        ys = cuts(h, 1, 120) if rng.integers(0, 2) else [0, h]
        if rng.integers(0, 2):
            ys = sorted(set([0, h] + [min(h, (y + 7) // 8 * 8) for y in ys]))
        job = _native.Job(ctx, w, h, 8, _native.OUT_JPEG_ENTROPY, False, bytes(4), jpeg_quality=q)
        try:
            k = 0
            for ya, yb in zip(ys[:-1], ys[1:]):
                for xa, xb in zip(xs[:-1], xs[1:]):
                    tile = np.ascontiguousarray(canvas[ya:yb, xa:xb])
                    job.add_tile(tile.tobytes(), width=xb - xa, height=yb - ya, bit_depth=8, color_type=6, interlace=0, filtered=0,
                                 dst_x=xa, dst_y=ya, z_index=0, image_idx=k)
                    k += 1
            bands = [b for b in cuts((h + 7) // 8, 1, 12)]
            bands = [min(h, 8 * b) for b in bands]
            sp, got = JpegBandSplicer(), b""
            for a, b in zip(bands[:-1], bands[1:]):
                if a == b:
                    continue
                data, _ = job.run_band(a, b)
                got += sp.push(data)
        finally:
            job.close()
        want = orc.jpeg_encode_rows(canvas.reshape(h, w * 4), w, h, 0, h, q)
        if n % 4 == 0 and _pil is not None:   # and libjpeg-turbo itself: its encoder, with the same framing, writes these very bytes
            buf = io.BytesIO()
            try:
                _pil.fromarray(np.ascontiguousarray(canvas[..., :3]), "RGB").save(buf, "JPEG", quality=q, subsampling=0, restart_marker_rows=1)
                d = buf.getvalue()
                i = d.index(b"\xff\xda")
                if d[i + 2 + int.from_bytes(d[i + 2:i + 4], "big"):-2] != want:
                    print(f"ORACLE differs from libjpeg-turbo: w={w} h={h} q={q}")
                    return -1
            except (TypeError, OSError):
                pass
        if got != want:
            print(f"MISMATCH: w={w} h={h} q={q} xs={xs} ys={ys} bands={bands} ({len(got)} vs {len(want)} bytes)")
            first = next((i for i in range(min(len(got), len(want))) if got[i] != want[i]), min(len(got), len(want)))
            print(f"first difference at byte {first}")
            os.makedirs("gpurun_out", exist_ok=True)
            np.save("gpurun_out/fuzz_fail.npy", canvas)
            with open("gpurun_out/fuzz_fail.txt", "w") as f:
                f.write(repr(dict(w=w, h=h, q=q, xs=xs, ys=ys, bands=bands)))
            return -1
        n += 1
    return n


if __name__ == "__main__":
    n_ok = run(float(sys.argv[1]) if len(sys.argv) > 1 else 60.0, int(sys.argv[2]) if len(sys.argv) > 2 else 1)
    if n_ok < 0:
        sys.exit(1)
    print(f"{n_ok} random cases byte-exact against the oracle")
