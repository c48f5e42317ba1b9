"""Randomised parity run of the device JPEG decoder against libjpeg-turbo ITSELF (through Pillow, when the box has it)
and against the oracle: random sizes, contents, qualities, subsamplings (4:4:4 / 4:2:2 / 4:2:0 / gray), optimised
Huffman tables, restart intervals.  python tools/jpeg_decode_fuzz.py [seconds] [seed].  GPU box only."""
import io, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import oracle_py as orc
from image_stitch_b200 import _native


def run(budget, seed, context=None):
    """Random files for `budget` seconds; returns their number, or -1 after printing the first mismatch."""
    from PIL import Image
    rng = np.random.default_rng(seed)
    ctx = context or _native.Context(0)
    n, t_end = 0, time.time() + budget
    while time.time() < t_end:
        w = int(rng.choice([rng.integers(1, 20), rng.integers(20, 300), rng.integers(300, 1800)]))
        h = int(rng.choice([rng.integers(1, 20), rng.integers(20, 300), rng.integers(300, 900)]))
        gray = rng.integers(0, 5) == 0
        y, x = np.mgrid[0:h, 0:w].astype(np.float64)
        chans = []
        for c in range(1 if gray else 3):
            a = 127 + rng.uniform(10, 100) * np.sin(x / rng.uniform(2, 40) + c) * np.cos(y / rng.uniform(2, 40))
            kind = rng.integers(0, 3)
            if kind >= 1:
                a += rng.normal(0, rng.uniform(0.5, 30), size=(h, w))
            if kind == 2:
                a[(x > w * rng.uniform(0, 1)) & (y < h * rng.uniform(0, 1))] = rng.uniform(0, 255)
            chans.append(np.clip(a, 0, 255).astype(np.uint8))
        pic = chans[0] if gray else np.stack(chans, axis=-1)
        opts = dict(quality=int(rng.choice([rng.integers(1, 101), 75, 85, 95, 100])))
        if not gray:
            opts["subsampling"] = int(rng.integers(0, 3))
        if rng.integers(0, 3) == 0:
            opts["optimize"] = True
        r = rng.integers(0, 4)
        if r == 1:
            opts["restart_marker_blocks"] = int(rng.integers(1, 40))
        elif r == 2:
            opts["restart_marker_rows"] = int(rng.integers(1, 4))
        buf = io.BytesIO()
        try:
            Image.fromarray(pic, "L" if gray else "RGB").save(buf, "JPEG", **opts)
        except TypeError:   # an older Pillow without restart_marker_*
            opts.pop("restart_marker_blocks", None); opts.pop("restart_marker_rows", None)
            buf = io.BytesIO()
            Image.fromarray(pic, "L" if gray else "RGB").save(buf, "JPEG", **opts)
        except OSError:     # Pillow's encoder gives up on some option sets ("broken data stream"): not our case
            continue
        data = buf.getvalue()
        want = np.array(Image.open(io.BytesIO(data)))
        got = ctx.decode_jpeg(data)
        got = got if gray else got[..., :3]
        if got.shape != want.shape or not np.array_equal(got, want):
            bad = np.argwhere(got != want) if got.shape == want.shape else []
            print(f"MISMATCH against libjpeg-turbo: {w}x{h} gray={gray} {opts}: {len(bad)} samples differ, first at {bad[0] if len(bad) else None}")
            os.makedirs("gpurun_out", exist_ok=True)
            open("gpurun_out/jpeg_decode_fuzz_fail.jpg", "wb").write(data)
            return -1
        if n % 8 == 0:   # the oracle too, now and then (it is the slower one)
            o = orc.jpeg_decode(data)
            o = o if gray else o[..., :3]
            if not np.array_equal(o, want):
                print(f"ORACLE differs from libjpeg-turbo: {w}x{h} gray={gray} {opts}")
                return -1
        n += 1
    return n


if __name__ == "__main__":
    n_ok = run(float(sys.argv[1]) if len(sys.argv) > 1 else 60.0, int(sys.argv[2]) if len(sys.argv) > 2 else 1)
    if n_ok < 0:
        sys.exit(1)
    print(f"{n_ok} random JPEG files decoded to libjpeg-turbo's pixels")
