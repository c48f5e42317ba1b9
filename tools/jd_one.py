"""One 2048^2 4:2:0 JPEG tile through the device decoder twice (for ncu launch lists).  Needs Pillow; GPU box only."""
import io
import os
import sys

import numpy as np
from PIL import Image

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from image_stitch_b200 import _native, synth  # noqa: E402

tile = 2048
pic = synth.smooth_rgba8(np.random.default_rng(8), tile, tile, "opaque").reshape(tile, tile, 4)[..., :3]
buf = io.BytesIO()
Image.fromarray(np.ascontiguousarray(pic), "RGB").save(buf, "JPEG", quality=85, subsampling=2)
ctx = _native.Context(0)
for _ in range(2):
    ctx.decode_jpeg(buf.getvalue())
