"""Makes tests/golden/jpeg_decode: small baseline JPEG files and the pixels libjpeg-turbo decodes from them with
its default settings (JDCT_ISLOW, fancy upsampling) -- what the reference gets from sharp on Node
(src/decoders/jpeg-decoder.ts:196-215).  Run in the build container (Pillow 12.2.0, libjpeg-turbo 3.1.4.1):

    python tests/golden/make_jpeg_decode_golden.py

Writes <name>.jpg per case and decoded.npz (name -> HxWx3 RGB or HxW gray array as Pillow returns it)."""
import io
import os

import numpy as np
from PIL import Image, features

HERE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "jpeg_decode")


def picture(rng, w, h, gray=False):
    """Smooth gradients, a few hard edges and noise: all coefficient magnitudes and run lengths get exercised."""
    y, x = np.mgrid[0:h, 0:w].astype(np.float64)
    chans = []
    for c in range(1 if gray else 3):
        a = 127 + 90 * np.sin(x / (3.0 + c) + c) * np.cos(y / (4.0 + 2 * c)) + 0.4 * (x + y) * (1 if c != 1 else -1)
        a += rng.normal(0, 12, size=(h, w))
        a[(x > w * 0.6) & (y < h * 0.4)] = 250 - 60 * c
        a[(x < w * 0.2) & (y > h * 0.7)] = 5 + 40 * c
        chans.append(np.clip(a, 0, 255).astype(np.uint8))
    return chans[0] if gray else np.stack(chans, axis=-1)


CASES = [
    # name, width, height, gray, save options
    ("c444_q90", 37, 29, False, dict(quality=90, subsampling=0)),
    ("c422_q85", 50, 33, False, dict(quality=85, subsampling=1)),
    ("c420_q80", 61, 45, False, dict(quality=80, subsampling=2)),
    ("c420_rst3", 70, 52, False, dict(quality=75, subsampling=2, restart_marker_blocks=3)),
    ("c444_rstrow", 48, 40, False, dict(quality=92, subsampling=0, restart_marker_rows=1)),
    ("c422_rst1_opt", 45, 23, False, dict(quality=60, subsampling=1, restart_marker_blocks=1, optimize=True)),
    ("g_q90", 40, 31, True, dict(quality=90)),
    ("g_rst2", 33, 17, True, dict(quality=70, restart_marker_blocks=2)),
    ("c420_w3", 3, 3, False, dict(quality=90, subsampling=2)),      # downsampled width 2: box replication
    ("c420_w4", 4, 5, False, dict(quality=90, subsampling=2)),
    ("c420_w5", 5, 5, False, dict(quality=90, subsampling=2)),      # downsampled width 3: fancy
    ("c422_w4", 4, 9, False, dict(quality=90, subsampling=1)),
    ("c422_w6", 6, 2, False, dict(quality=90, subsampling=1)),
    ("c444_1x1", 1, 1, False, dict(quality=95, subsampling=0)),
    ("c420_q1", 64, 48, False, dict(quality=1, subsampling=2)),     # coarsest tables
    ("c444_q100_opt", 32, 32, False, dict(quality=100, subsampling=0, optimize=True)),
    ("c420_big_opt", 256, 200, False, dict(quality=75, subsampling=2, optimize=True)),
    ("c420_big_rst", 256, 200, False, dict(quality=88, subsampling=2, restart_marker_rows=2)),
    ("g_big", 200, 120, True, dict(quality=50, optimize=True)),
    ("c420_odd", 129, 65, False, dict(quality=85, subsampling=2)),
    ("p420_progressive", 40, 40, False, dict(quality=85, subsampling=2, progressive=True)),   # must be refused
]


def main():
    os.makedirs(HERE, exist_ok=True)
    rng = np.random.default_rng(20260)
    decoded = {}
    for name, w, h, gray, opts in CASES:
        img = Image.fromarray(picture(rng, w, h, gray), "L" if gray else "RGB")
        buf = io.BytesIO()
        img.save(buf, "JPEG", **opts)
        data = buf.getvalue()
        with open(os.path.join(HERE, name + ".jpg"), "wb") as f:
            f.write(data)
        back = Image.open(io.BytesIO(data))
        back.load()
        assert back.mode == ("L" if gray else "RGB"), back.mode
        decoded[name] = np.asarray(back).copy()
    np.savez_compressed(os.path.join(HERE, "decoded.npz"), **decoded)
    with open(os.path.join(HERE, "VERSIONS.txt"), "w") as f:
        import PIL
        f.write(f"Pillow {PIL.__version__}, libjpeg-turbo {features.version('libjpeg_turbo')}\n")
    print(len(CASES), "cases,", sum(os.path.getsize(os.path.join(HERE, n)) for n in os.listdir(HERE)), "bytes")


if __name__ == "__main__":
    main()
