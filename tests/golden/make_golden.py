#!/usr/bin/env python3
"""Builds tests/golden/ from the read-only reference checkout (/root/reference).

Run here (authoring container) only; the GPU box never sees /root/reference.
What is copied are FIXTURES, not reference source code:
  * pngsuite/png/*.png          PngSuite corpus (third-party, freely redistributable test images)
  * expected-outputs/example1..5.png   genuine reference outputs
        (tests/utils/fixtures/expected-outputs, recipes in tests/utils/generate-expected-outputs.js:29-106)
  * pngsuite_rgba.npz           the lunapaint/pngsuite JSON goldens (pngsuite/json/*.json) as arrays
  * pngsuite_palette_rgba.npz   the same for the colour type 3 files: the reference throws on those, the goldens pin
                                the palette EXTENSION and nothing else
"""
import json
import os
import shutil
import struct
import sys

import numpy as np

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))


def png_ihdr(path):
    with open(path, "rb") as f:
        d = f.read(33)
    return struct.unpack(">IIBBBBB", d[16:29])


def main():
    png_src = os.path.join(REF, "pngsuite", "png")
    png_dst = os.path.join(HERE, "pngsuite")
    os.makedirs(png_dst, exist_ok=True)
    names = sorted(n for n in os.listdir(png_src) if n.endswith(".png") and not n.startswith("x"))
    for n in names:
        shutil.copyfile(os.path.join(png_src, n), os.path.join(png_dst, n))
    exp_src = os.path.join(REF, "tests", "utils", "fixtures", "expected-outputs")
    exp_dst = os.path.join(HERE, "expected-outputs")
    os.makedirs(exp_dst, exist_ok=True)
    for n in sorted(os.listdir(exp_src)):
        shutil.copyfile(os.path.join(exp_src, n), os.path.join(exp_dst, n))
    # JSON goldens: flat [r,g,b,a,...]; 8-bit values for <=8-bit sources, 16-bit values for 16-bit sources.
    arrays, palette = {}, {}
    jdir = os.path.join(REF, "pngsuite", "json")
    for n in names:
        stem = n[:-4]
        jp = os.path.join(jdir, stem + ".json")
        if not os.path.exists(jp):
            continue
        w, h, bd, ct, _, _, _ = png_ihdr(os.path.join(png_src, n))
        with open(jp) as f:
            vals = json.load(f)
        a = np.asarray(vals, dtype=np.uint16 if bd == 16 else np.uint8)
        if a.size != w * h * 4:
            print("skip (size)", stem, a.size, w, h)
            continue
        # the reference throws on palette images: kept apart from the parity goldens
        (palette if ct == 3 else arrays)[stem] = a.reshape(h, w, 4)
    np.savez_compressed(os.path.join(HERE, "pngsuite_rgba.npz"), **arrays)
    np.savez_compressed(os.path.join(HERE, "pngsuite_palette_rgba.npz"), **palette)
    print(f"{len(names)} PNGs, {len(arrays)} JSON goldens, {len(palette)} palette goldens (extension)")


if __name__ == "__main__":
    sys.exit(main())
