"""Robustness run of the device JPEG decoder: valid files from tests/golden/jpeg_decode truncated, bit-flipped and spliced
at random.  Every call must come back -- pixels or a StitchError -- and the decoder must still decode a valid file after
it.  python tools/jpeg_corrupt_fuzz.py [seconds] [seed]  (run it under `timeout`).  GPU box only."""
import glob, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from image_stitch_b200 import _native


def run(budget, seed, context=None, check_every=50):
    rng = np.random.default_rng(seed)
    ctx = context or _native.Context(0)
    gold = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "jpeg_decode")
    files = [open(f, "rb").read() for f in sorted(glob.glob(os.path.join(gold, "*.jpg"))) if "progressive" not in f]
    good = ctx.decode_jpeg(files[0]).copy()
    n = errors = 0
    t_end = time.time() + budget
    while time.time() < t_end:
        d = bytearray(files[int(rng.integers(0, len(files)))])
        kind = int(rng.integers(0, 4))
        if kind == 0:
            d = d[:int(rng.integers(2, len(d)))]
        elif kind == 1:
            for _ in range(int(rng.integers(1, 6))):
                i = int(rng.integers(2, len(d)))
                d[i] ^= 1 << int(rng.integers(0, 8))
        elif kind == 2:
            i = int(rng.integers(2, len(d)))
            d[i:i + int(rng.integers(1, 40))] = bytes(rng.integers(0, 256, int(rng.integers(1, 40)), dtype=np.uint8))
        else:
            o = files[int(rng.integers(0, len(files)))]
            i, j = int(rng.integers(2, len(d))), int(rng.integers(2, len(o)))
            d = d[:i] + o[j:]
        try:
            ctx.decode_jpeg(bytes(d))
        except _native.StitchError:
            errors += 1
        n += 1
        if check_every and n % check_every == 0:
            try:
                same = np.array_equal(ctx.decode_jpeg(files[0]), good)
                why = "its pixels differ"
            except _native.StitchError as e:
                same, why = False, str(e)
            if not same:
                print(f"after corrupted file {n} (kind {kind}, {len(d)} bytes) a valid file no longer decodes: {why}")
                os.makedirs("gpurun_out", exist_ok=True)
                open("gpurun_out/jpeg_corrupt_fail.jpg", "wb").write(bytes(d))
                return -1
    print(f"{n} corrupted files: {errors} refused, {n - errors} decoded to something; the decoder is still sound")
    return n


if __name__ == "__main__":
    sys.exit(0 if run(float(sys.argv[1]) if len(sys.argv) > 1 else 20.0, int(sys.argv[2]) if len(sys.argv) > 2 else 1, None, int(sys.argv[3]) if len(sys.argv) > 3 else 50) >= 0 else 1)
