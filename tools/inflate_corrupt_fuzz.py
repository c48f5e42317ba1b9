"""Robustness run of the device inflate: valid zlib streams truncated, bit-flipped and overwritten at random.  Every call
must come back -- bytes or a StitchError with zlib's kind of message -- and a valid stream must inflate correctly after
each.  Where zlib itself accepts the damaged stream, the bytes must be zlib's.  python tools/inflate_corrupt_fuzz.py
[seconds] [seed]  (run it under `timeout`).  GPU box only."""
import os, sys, time, zlib
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from image_stitch_b200 import _native
from inflate_fuzz import content


def run(budget, seed, context=None):
    rng = np.random.default_rng(seed)
    ctx = context or _native.Context(0)
    ref_raw = content(rng, 50000).tobytes()
    ref_z = zlib.compress(ref_raw, 6)
    n = refused = 0
    t_end = time.time() + budget
    while time.time() < t_end:
        raw = content(rng, int(rng.integers(1, 60000))).tobytes()
        z = bytearray(zlib.compress(raw, int(rng.integers(0, 10))))
        kind = int(rng.integers(0, 3))
        if kind == 0:
            z = z[:int(rng.integers(0, len(z)))]
        elif kind == 1:
            for _ in range(int(rng.integers(1, 4))):
                z[int(rng.integers(0, len(z)))] ^= 1 << int(rng.integers(0, 8))
        else:
            i = int(rng.integers(0, len(z)))
            z[i:i + int(rng.integers(1, 30))] = bytes(rng.integers(0, 256, int(rng.integers(1, 30)), dtype=np.uint8))
        try:
            want = zlib.decompress(bytes(z))
        except zlib.error:
            want = None
        try:
            got = bytes(ctx.inflate(bytes(z), len(raw) + 64))
        except _native.StitchError:
            got = None
            refused += 1
        if want is not None and got is not None and got != want and len(want) <= len(raw) + 64:
            print(f"zlib accepts the damaged stream ({len(want)} bytes) and the device returns other bytes ({len(got)})")
            return -1
        if want is not None and got is None and len(want) <= len(raw) + 64:
            print("zlib accepts the damaged stream, the device refuses it")
            return -1
        n += 1
        if bytes(ctx.inflate(ref_z, len(ref_raw) + 16)) != ref_raw:
            print(f"after damaged stream {n} a valid stream no longer inflates correctly")
            return -1
    print(f"{n} damaged streams: {refused} refused; the device inflate is still sound")
    return n


if __name__ == "__main__":
    sys.exit(0 if run(float(sys.argv[1]) if len(sys.argv) > 1 else 20.0, int(sys.argv[2]) if len(sys.argv) > 2 else 1) >= 0 else 1)
