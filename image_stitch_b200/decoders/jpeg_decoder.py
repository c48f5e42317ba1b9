"""JPEG decoder plugin (src/decoders/jpeg-decoder.ts).

The reference parses the frame header itself (parseJpegHeader, :46-94) and hands the file to a third-party decoder
-- sharp (libjpeg-turbo) on Node, jpeg-js or the browser elsewhere (:196-281) -- then yields RGBA8 or gray8 rows
(:317-347).  Here the header scan is the same host code, and the pixels are made by libstitch_b200's device decoder
(csrc/jpeg_decode.cu), which redoes libjpeg-turbo's default decode bit for bit: ``jpeg_file()``
hands the file to the engine (it crosses the bus compressed), ``scanlines()`` runs the same kernels through
stitch_decode_jpeg.  No sample is decoded on the CPU.
"""
from __future__ import annotations

from typing import Iterator, Optional, Tuple

from ..types import ImageHeader

# SOF markers (jpeg-decoder.ts:26-40): every frame type, so that the size of a file the device decoder will refuse
# is still reported like the reference reports it
JPEG_SOF_MARKERS = (0xC0, 0xC1, 0xC2, 0xC3, 0xC5, 0xC6, 0xC7, 0xC9, 0xCA, 0xCB, 0xCD, 0xCE, 0xCF)


def parse_jpeg_header(data) -> Tuple[int, int, int]:
    """(width, height, components) from the first SOF marker (jpeg-decoder.ts:46-94, same messages)."""
    data = bytes(data)
    if len(data) < 2 or data[0] != 0xFF or data[1] != 0xD8:
        raise ValueError("Invalid JPEG: missing SOI marker")
    offset = 2
    while offset < len(data) - 1:
        if data[offset] != 0xFF:
            offset += 1
            continue
        marker = data[offset + 1]
        offset += 2
        if marker in (0xFF, 0x00):
            continue
        if marker in JPEG_SOF_MARKERS:
            if offset + 7 > len(data):
                raise ValueError("Invalid JPEG: truncated SOF marker")
            height = (data[offset + 3] << 8) | data[offset + 4]
            width = (data[offset + 5] << 8) | data[offset + 6]
            return width, height, data[offset + 7]
        if offset + 2 > len(data):
            break
        offset += (data[offset] << 8) | data[offset + 1]
    raise ValueError("Invalid JPEG: no SOF marker found")


class JpegBufferDecoder:
    """JpegBufferDecoder / JpegFileDecoder (jpeg-decoder.ts:284-380)."""

    def __init__(self, data=None, path: Optional[str] = None, options=None):
        self._data = None if data is None else bytes(data)
        self._path = path
        self._header: Optional[ImageHeader] = None

    def _bytes(self) -> bytes:
        if self._data is None:
            with open(self._path, "rb") as f:
                self._data = f.read()
        return self._data

    def get_header(self) -> ImageHeader:
        if self._header is None:
            w, h, comps = parse_jpeg_header(self._bytes())
            # :304-312: grayscale stays one channel, everything else is delivered as RGBA
            self._header = ImageHeader(w, h, 1 if comps == 1 else 4, 8, "jpeg", {"components": comps})
        return self._header

    def jpeg_file(self) -> bytes:
        """The file itself: the engine adds it as a tile that the device decodes (stitch_job_add_tile_jpeg)."""
        return self._bytes()

    def scanlines(self) -> Iterator[bytes]:
        from .. import engine
        pixels = engine.decode_jpeg(self._bytes())
        for y in range(pixels.shape[0]):
            yield pixels[y].tobytes()

    def close(self) -> None:
        if self._path is not None:
            self._data = None


class JpegFileDecoder(JpegBufferDecoder):
    def __init__(self, path: str, options=None):
        super().__init__(None, path, options)


class _JpegPlugin:
    format = "jpeg"

    def create(self, input, options=None):
        if isinstance(input, str):
            return JpegFileDecoder(input, options)
        if isinstance(input, (bytes, bytearray, memoryview)):
            return JpegBufferDecoder(input, None, options)
        raise TypeError("Unsupported JPEG input type for decoder plugin")   # jpeg-decoder.ts:411


jpeg_decoder = _JpegPlugin()
