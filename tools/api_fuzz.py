"""Randomised parity run of the PUBLIC API (concat_to_buffer) against the oracle: random PNG inputs (formats, bit depths,
Adam7, filter types), now and then a JPEG file, grid and positioned layouts, blending, backgrounds, device inflate /
deflate, PNG and JPEG output.  python tools/api_fuzz.py [seconds] [seed].  GPU box only."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import helpers as H
import oracle_py as orc
import image_stitch_b200 as stitch
from image_stitch_b200.png_container import read_png

FORMATS = [(8, 6), (8, 6), (8, 2), (8, 0), (8, 4), (16, 6), (16, 2), (16, 0), (16, 4), (1, 0), (2, 0), (4, 0)]
JPEGS = ["c420_q80", "c444_q90", "c422_q85", "g_q90", "c420_odd"]


def run(budget, seed):
    rng = np.random.default_rng(seed)
    n, t_end = 0, time.time() + budget
    gold = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "jpeg_decode")
    while time.time() < t_end:
        inputs = []
        for _ in range(int(rng.integers(1, 7))):
            if rng.integers(0, 8) == 0:
                with open(os.path.join(gold, JPEGS[int(rng.integers(0, len(JPEGS)))] + ".jpg"), "rb") as f:
                    inputs.append(f.read())
                continue
            bd, ct = FORMATS[int(rng.integers(0, len(FORMATS)))]
            w = int(rng.choice([rng.integers(1, 30), rng.integers(30, 700)]))
            h = int(rng.choice([rng.integers(1, 30), rng.integers(30, 200)]))
            pix = H.synth_pixels(rng, w, h, bd, ct, "any" if ct in (4, 6) else "opaque")
            inputs.append(H.make_png(pix.reshape(h, -1), w, h, bd, ct, interlace=int(rng.integers(0, 4) == 0),
                                     mode=["adaptive", "random"][int(rng.integers(0, 2))], rng=rng))
        opts = {}
        if rng.integers(0, 2):
            opts["inputs"] = inputs
            opts["layout"] = {"columns": int(rng.integers(1, len(inputs) + 1))} if rng.integers(0, 2) else {"rows": int(rng.integers(1, len(inputs) + 1))}
        else:
            opts["inputs"] = [{"x": int(rng.integers(0, 300)), "y": int(rng.integers(0, 120)), "source": s,
                               **({"zIndex": int(rng.integers(0, 4))} if rng.integers(0, 2) else {})} for s in inputs]
            opts["layout"] = {"width": int(rng.integers(1, 900)), "height": int(rng.integers(1, 260))} if rng.integers(0, 2) else {}
            opts["enableAlphaBlending"] = bool(rng.integers(0, 2))
        if rng.integers(0, 2):
            opts["backgroundColor"] = "#" + "".join(f"{int(v):02x}" for v in rng.integers(0, 256, 4))
        jpeg = rng.integers(0, 4) == 0
        if jpeg:
            opts.update(output_format="jpeg", jpeg_quality=int(rng.integers(1, 101)))
        else:
            if rng.integers(0, 2):
                opts["device_deflate"] = True
        if rng.integers(0, 2):
            opts["device_inflate"] = True
        try:
            plan, want = H.oracle_stitch_rows(dict(opts, output_format="png"), out_kind=1 if jpeg else 0)
        except Exception as e:   # what the host rejects (mixed bit depths and the like), it must reject here too
            try:
                stitch.concat_to_buffer(opts)
            except Exception:
                n += 1
                continue
            print(f"the oracle side raised {e!r}, the API did not: {opts.get('layout')}")
            return -1
        got = stitch.concat_to_buffer(opts)
        if jpeg:
            if plan.out_bit_depth != 8:
                n += 1
                continue
            w, h = plan.canvas_w, plan.canvas_h
            ok = got == orc.jpeg_file(np.frombuffer(want, np.uint8).reshape(h, w * 4), w, h, opts["jpeg_quality"])
        else:
            ok = read_png(got)[1] == want.tobytes()
        if not ok:
            print(f"MISMATCH: {len(inputs)} inputs, options { {k: v for k, v in opts.items() if k != 'inputs'} }")
            return -1
        n += 1
    return n


if __name__ == "__main__":
    n_ok = run(float(sys.argv[1]) if len(sys.argv) > 1 else 60.0, int(sys.argv[2]) if len(sys.argv) > 2 else 1)
    if n_ok < 0:
        sys.exit(1)
    print(f"{n_ok} random jobs through the public API byte-exact against the oracle")
