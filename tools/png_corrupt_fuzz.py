"""Robustness run of the PNG input path through the public API: inputs whose inflated streams carry unknown filter types or
end early, whose zlib streams are damaged (with the inflate on the host or on the device).  A damaged input must raise --
the reference's errors -- or give exactly what the oracle gives for it, and a sound job must come out right after each.
python tools/png_corrupt_fuzz.py [seconds] [seed]  (run it under `timeout`).  GPU box only."""
import os, sys, time, warnings, zlib
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import helpers as H
import image_stitch_b200 as stitch
from image_stitch_b200.png_container import PngHeader, build_png, create_chunk, create_ihdr, create_iend, read_png

FORMATS = [(8, 6), (8, 2), (8, 0), (8, 4), (16, 6), (16, 0)]


def run(budget, seed):
    rng = np.random.default_rng(seed)
    good_in = [H.suite_bytes("basn6a08"), H.suite_bytes("basn2c08")]
    good_opts = {"inputs": good_in, "layout": {"columns": 2}}
    _, good_want = H.oracle_stitch_rows(good_opts)
    n = raised = 0
    t_end = time.time() + budget
    while time.time() < t_end:
        bd, ct = FORMATS[int(rng.integers(0, len(FORMATS)))]
        w, h = int(rng.integers(1, 300)), int(rng.integers(1, 120))
        pix = H.synth_pixels(rng, w, h, bd, ct, "any" if ct in (4, 6) else "opaque")
        spp = {0: 1, 2: 3, 4: 2, 6: 4}[ct]
        stream = bytearray(H.filter_image(pix.reshape(h, -1), max(1, spp * bd // 8), "random", rng).tobytes())
        row = len(stream) // h
        kind = int(rng.integers(0, 4))
        if kind == 0:
            stream[row * int(rng.integers(0, h))] = int(rng.integers(5, 256))          # unknown filter type
        elif kind == 1:
            stream = stream[:int(rng.integers(0, len(stream)))]                         # fewer scanlines than promised
        z = bytearray(zlib.compress(bytes(stream), 6))
        if kind == 2:
            z[int(rng.integers(2, len(z)))] ^= 1 << int(rng.integers(0, 8))              # a damaged zlib stream
        elif kind == 3:
            z = z[:int(rng.integers(2, len(z)))]
        png = build_png([create_ihdr(PngHeader(w, h, bd, ct, 0, 0, 0)), create_chunk("IDAT", bytes(z)), create_iend()])
        opts = {"inputs": [png, png], "layout": {"columns": 2}}
        if rng.integers(0, 2):
            opts["device_inflate"] = True
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            lenient = False
            try:
                _, want = H.oracle_stitch_rows(dict(opts, device_inflate=False))
            except Exception as e:
                want = None
                # The one known difference: a stream that is sound up to its last row and fails only the Adler-32 check.
                # Host zlib reads to the end and refuses it; the device stops, like the reference's early break
                # (src/decoders/png-decoder.ts:164-167), once the image's rows are there, and never sees the trailer.
                lenient = opts.get("device_inflate", False) and "incorrect data check" in str(e)
            why = ""
            try:
                got = read_png(stitch.concat_to_buffer(opts))[1]
            except Exception as e:
                got, why = None, f" ({e})"
                raised += 1
            if lenient and got is not None:
                n += 1
                continue
            # The other known difference, the other way round: a stream cut or damaged BEHIND the image's last byte.  The host
            # inflates what there is and has every row; the device meets the damage in the window it decodes ahead and
            # refuses the stream ("Inflate error: ...") rather than trust bytes made in that window.
            if want is not None and got is None and opts.get("device_inflate", False) and "Inflate error" in why:
                n += 1
                continue
            if (want is None) != (got is None) or (want is not None and got != want.tobytes()):
                print(f"damaged input kind {kind} ({w}x{h} {bd}/{ct}, device_inflate={opts.get('device_inflate', False)}): "
                      f"the oracle side {'raised' if want is None else 'gave rows'}, the API {'raised' if got is None else 'gave rows'}{why}")
                os.makedirs("gpurun_out", exist_ok=True)
                open("gpurun_out/png_corrupt_fail.png", "wb").write(png)
                open("gpurun_out/png_corrupt_fail_stream.bin", "wb").write(bytes(stream))
                return -1
            if read_png(stitch.concat_to_buffer(good_opts))[1] != good_want.tobytes():
                print(f"after damaged input {n} a sound job comes out wrong")
                return -1
        n += 1
    print(f"{n} damaged PNG inputs: {raised} raised, like the oracle side; sound jobs stay right")
    return n


if __name__ == "__main__":
    sys.exit(0 if run(float(sys.argv[1]) if len(sys.argv) > 1 else 20.0, int(sys.argv[2]) if len(sys.argv) > 2 else 1) >= 0 else 1)
