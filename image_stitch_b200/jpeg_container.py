"""JPEG container pieces on the host (the role of src/jpeg-encoder.ts:96-191).

The reference's JpegEncoder.header() yields nothing, the first encode_strip()
output carries SOI + tables and finish() carries EOI
(src/jpeg-encoder.ts:123-143,155-191).  Here the markers come from
stitch_jpeg_header() in libstitch_b200 and the entropy-coded MCU rows from the
GPU (csrc/jpeg.cu).  The restart interval is one MCU row, so every band the
library returns is a byte-aligned run of complete segments ending in RSTm:
bands are spliced by plain concatenation, in ascending row order.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _native

EOI = b"\xff\xd9"
_INFO = C.sizeof(_native.JpegBandInfo)


def jpeg_header(width: int, height: int, quality: int) -> bytes:
    """SOI, APP0 (JFIF), DQT, SOF0, DHT, DRI, SOS."""
    if width <= 0 or height <= 0:
        raise ValueError(f"Invalid dimensions: {width}x{height}")                    # src/jpeg-encoder.ts:109-111
    if quality < 1 or quality > 100:
        raise ValueError(f"Invalid quality: {quality} (must be 1-100)")               # :112-114
    if width > 65535 or height > 65535:
        raise ValueError(f"Invalid dimensions: {width}x{height} (JPEG stores sizes in 16 bits)")
    lib = _native.load()
    buf = np.zeros(1024, np.uint8)
    n = C.c_size_t()
    rc = lib.stitch_jpeg_header(width, height, quality, buf.ctypes.data, buf.size, C.byref(n))
    if rc != _native.STITCH_OK:
        raise _native.StitchError(rc, f"stitch_jpeg_header failed ({rc})")
    return buf[: n.value].tobytes()


class JpegBandSplicer:
    """Joins the per-band results of stitch_job_run_band (STITCH_OUT_JPEG_ENTROPY)."""

    def __init__(self) -> None:
        self.bytes_out = 0

    def push(self, band) -> bytes:
        data = memoryview(np.ascontiguousarray(band)).cast("B")
        info = _native.JpegBandInfo.from_buffer_copy(bytes(data[:_INFO]))
        body = bytes(data[_INFO:])
        if info.n_bits != len(body) * 8:
            raise _native.StitchError(_native.E_STATE, "JPEG band header does not match its payload")
        self.bytes_out += len(body)
        return body

    def finish(self) -> bytes:
        return EOI
