"""PNG decoder plugin (src/decoders/png-decoder.ts).

Container parsing and zlib inflate run on the host, exactly as in the
reference.  What the reference does next on the CPU -- unfilterScanline per
row, deinterlaceAdam7 for interlaced files (png-decoder.ts:63-229) -- is
kernel (a) of libstitch_b200: ``inflated()`` hands the still-filtered stream
to the engine, and ``scanlines()`` (the plugin contract for foreign callers)
runs the same kernel through stitch_decode_image.  No pixel byte is unfiltered
on the CPU.
"""
from __future__ import annotations

from typing import Iterator, Optional, Tuple

from ..png_container import extract_idat, parse_png_chunks, read_png, parse_png_header, scanline_bytes
from ..types import ImageHeader, PngHeader

_CHANNELS = {0: 1, 2: 3, 3: 3, 4: 2, 6: 4}  # palette reported as 3 channels: png-decoder.ts:32-34


def png_header_to_image_header(h: PngHeader) -> ImageHeader:
    if h.color_type not in _CHANNELS:
        raise ValueError(f"Unsupported PNG color type: {h.color_type}")
    return ImageHeader(h.width, h.height, _CHANNELS[h.color_type], h.bit_depth, "png", {
        "colorType": h.color_type, "compressionMethod": h.compression_method,
        "filterMethod": h.filter_method, "interlaceMethod": h.interlace_method})


class PngBufferDecoder:
    """PngBufferDecoder / PngFileDecoder (png-decoder.ts:236-420): headers are
    read from the first chunk only; pixel data is touched lazily."""

    def __init__(self, data=None, path: Optional[str] = None):
        self._data = None if data is None else bytes(data)
        self._path = path
        self._header: Optional[PngHeader] = None
        self._inflated: Optional[bytes] = None

    def _bytes(self) -> bytes:
        if self._data is None:
            with open(self._path, "rb") as f:
                self._data = f.read()
        return self._data

    def png_header(self) -> PngHeader:
        if self._header is None:
            if self._data is None and self._path is not None:
                with open(self._path, "rb") as f:
                    self._header = parse_png_header(f.read(64))
            else:
                self._header = parse_png_header(self._bytes())
        return self._header

    def get_header(self) -> ImageHeader:
        return png_header_to_image_header(self.png_header())

    def inflated(self) -> Tuple[PngHeader, bytes]:
        """Inflated, still filtered IDAT stream: where north_star starts the GPU path."""
        if self._inflated is None:
            header, stream = read_png(self._bytes())
            self._header = header
            self._inflated = stream
        return self.png_header(), self._inflated

    def deflated(self) -> Tuple[PngHeader, bytes]:
        """The zlib stream itself (IDAT payloads concatenated, png-decoder.ts:100-117 before the
        decompression stream): with the `device_inflate` option the GPU path starts here."""
        chunks = parse_png_chunks(self._bytes())
        self._header = parse_png_header(self._bytes())
        return self._header, extract_idat(chunks)

    def palette(self) -> Tuple[Optional[bytes], bytes]:
        """(PLTE body, tRNS body) of a colour type 3 file, for the `expand_palette` extension;
        the reference never reads these chunks."""
        plte, trns = None, b""
        for c in parse_png_chunks(self._bytes()):
            if c.type == "PLTE" and plte is None:
                plte = bytes(c.data)
            elif c.type == "tRNS":
                trns = bytes(c.data)
            elif c.type == "IDAT":
                break
        return plte, trns

    def scanlines(self) -> Iterator[bytes]:
        from .. import engine
        header, stream = self.inflated()
        rows = engine.decode_image(header, stream)
        n = scanline_bytes(header.width, header.bit_depth, header.color_type)
        for y in range(header.height):
            yield rows[y * n:(y + 1) * n].tobytes()

    def close(self) -> None:
        self._inflated = None
        if self._path is not None:
            self._data = None


class PngFileDecoder(PngBufferDecoder):
    def __init__(self, path: str):
        super().__init__(None, path)


class _PngPlugin:
    format = "png"

    def create(self, input, options=None):
        if isinstance(input, str):
            return PngFileDecoder(input)
        if isinstance(input, (bytes, bytearray, memoryview)):
            return PngBufferDecoder(input)
        raise TypeError("Unsupported input type for PNG decoder")


png_decoder = _PngPlugin()
