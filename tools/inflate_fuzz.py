"""Randomised parity run of the device inflate against zlib: random contents (text-like, filtered-row-like residuals,
runs, noise, mixtures), sizes, levels, strategies, window sizes and memory levels.  python tools/inflate_fuzz.py [seconds] [seed]"""
import os, sys, time, zlib
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from image_stitch_b200 import _native


def content(rng, n):
    kind = int(rng.integers(0, 6))
    if kind == 0:     # residual-like: small signed values around 0
        return (rng.normal(0, rng.uniform(0.5, 20), n).astype(np.int64) & 0xFF).astype(np.uint8)
    if kind == 1:     # noise
        return rng.integers(0, 256, n, dtype=np.uint8)
    if kind == 2:     # long runs
        v = rng.integers(0, 256, max(1, n // int(rng.integers(1, 400)) + 1), dtype=np.uint8)
        return np.repeat(v, int(rng.integers(1, 400)))[:n]
    if kind == 3:     # text-like: few symbols, repeated phrases
        words = [bytes(rng.integers(97, 123, int(rng.integers(1, 12)), dtype=np.uint8)) for _ in range(int(rng.integers(2, 200)))]
        out = bytearray()
        while len(out) < n:
            out += words[int(rng.integers(0, len(words)))] + b" "
        return np.frombuffer(bytes(out[:n]), np.uint8)
    if kind == 4:     # periodic with drift (far matches)
        p = int(rng.integers(1, 40000))
        base = rng.integers(0, 256, p, dtype=np.uint8)
        reps = np.tile(base, n // p + 1)[:n].copy()
        idx = rng.integers(0, n, max(1, n // 500))
        reps[idx] = rng.integers(0, 256, idx.size, dtype=np.uint8)
        return reps
    parts = [content(rng, max(1, n // 3)) for _ in range(3)]
    return np.concatenate(parts)[:n]


def run(budget, seed, context=None):
    rng = np.random.default_rng(seed)
    ctx = context or _native.Context(0)
    n, t_end = 0, time.time() + budget
    while time.time() < t_end:
        size = int(rng.choice([rng.integers(0, 300), rng.integers(300, 70000), rng.integers(70000, 1500000)]))
        raw = content(rng, size).tobytes() if size else b""
        level = int(rng.integers(0, 10))
        strategy = int(rng.choice([zlib.Z_DEFAULT_STRATEGY, zlib.Z_FILTERED, zlib.Z_HUFFMAN_ONLY, zlib.Z_RLE, zlib.Z_FIXED]))
        wbits = int(rng.integers(9, 16))
        mem = int(rng.integers(1, 10))
        co = zlib.compressobj(level, zlib.DEFLATED, wbits, mem, strategy)
        z = co.compress(raw)
        if rng.integers(0, 4) == 0 and len(raw) > 10:   # a sync flush in the middle
            co2 = zlib.compressobj(level, zlib.DEFLATED, wbits, mem, strategy)
            cut = int(rng.integers(1, len(raw)))
            z = co2.compress(raw[:cut]) + co2.flush(zlib.Z_SYNC_FLUSH) + co2.compress(raw[cut:]) + co2.flush()
        else:
            z += co.flush()
        got = ctx.inflate(z, len(raw) + 16)
        if bytes(got) != raw:
            print(f"MISMATCH: {len(raw)} bytes, level {level} strategy {strategy} wbits {wbits} memLevel {mem}: got {len(got)} bytes")
            os.makedirs("gpurun_out", exist_ok=True)
            open("gpurun_out/inflate_fuzz_fail.z", "wb").write(z)
            return -1
        n += 1
    return n


if __name__ == "__main__":
    n_ok = run(float(sys.argv[1]) if len(sys.argv) > 1 else 60.0, int(sys.argv[2]) if len(sys.argv) > 2 else 1)
    if n_ok < 0:
        sys.exit(1)
    print(f"{n_ok} random zlib streams inflated to the original bytes")
