"""PNG container I/O on the host (out of the GPU hot path).

Chunk reader with CRC check and IHDR parse (src/png-parser.ts), chunk
builder / IHDR / IEND / buildPng (src/png-writer.ts), CRC32 and big-endian
helpers (src/utils.ts), IDAT concatenation + inflate
(src/png-decompress.ts, src/streaming-inflate.ts).  zlib itself is the
runtime's, as in the reference.
"""
from __future__ import annotations

import struct
import zlib
from typing import List, NamedTuple, Tuple

from .types import PngHeader

PNG_SIGNATURE = bytes([137, 80, 78, 71, 13, 10, 26, 10])


class PngChunk(NamedTuple):
    length: int
    type: str
    data: bytes
    crc: int


def png_crc32(data: bytes) -> int:
    return zlib.crc32(data) & 0xFFFFFFFF


def is_png_signature(data) -> bool:
    return len(data) >= 8 and bytes(data[:8]) == PNG_SIGNATURE


def get_samples_per_pixel(color_type: int) -> int:
    """src/utils.ts:92-101"""
    try:
        return {0: 1, 2: 3, 3: 1, 4: 2, 6: 4}[color_type]
    except KeyError:
        raise ValueError(f"Unknown color type: {color_type}") from None


def get_bytes_per_pixel(bit_depth: int, color_type: int) -> int:
    """src/png-filter.ts:186-211"""
    return (get_samples_per_pixel(color_type) * bit_depth + 7) // 8


def scanline_bytes(width: int, bit_depth: int, color_type: int) -> int:
    return (width * bit_depth * get_samples_per_pixel(color_type) + 7) // 8


def parse_png_chunks(data) -> List[PngChunk]:
    """src/png-parser.ts:17-86 -- same error strings.  Bodies of 64 KiB and more (IDAT of a large tile) are views
    into `data`, not copies."""
    if not isinstance(data, (bytes, memoryview)):
        data = bytes(data)
    view = memoryview(data)
    if not is_png_signature(data):
        raise ValueError("Invalid PNG signature")
    off = 8
    chunks: List[PngChunk] = []
    n = len(data)
    while off < n:
        if off + 12 > n:
            raise ValueError("Incomplete PNG chunk")
        (length,) = struct.unpack_from(">I", data, off)
        ctype = bytes(view[off + 4: off + 8])
        off += 8
        if off + length + 4 > n:
            raise ValueError("Incomplete PNG chunk data")
        body = view[off: off + length] if length >= 65536 else bytes(view[off: off + length])
        off += length
        (crc,) = struct.unpack_from(">I", data, off)
        off += 4
        name = ctype.decode("latin-1")
        if zlib.crc32(body, zlib.crc32(ctype)) & 0xFFFFFFFF != crc:   # CRC over type + data, without joining them
            raise ValueError(f"CRC mismatch for chunk {name}")
        chunks.append(PngChunk(length, name, body, crc))
    return chunks


def parse_ihdr(chunk: PngChunk) -> PngHeader:
    """src/png-parser.ts:91-109"""
    if chunk.type != "IHDR":
        raise ValueError("Not an IHDR chunk")
    if len(chunk.data) != 13:
        raise ValueError("Invalid IHDR chunk length")
    w, h, bd, ct, cm, fm, im = struct.unpack(">IIBBBBB", chunk.data)
    return PngHeader(w, h, bd, ct, cm, fm, im)


def parse_png_header(data) -> PngHeader:
    """src/png-parser.ts:114-142: reads only the first chunk."""
    data = bytes(data[:64]) if len(data) > 64 else bytes(data)
    if not is_png_signature(data):
        raise ValueError("Invalid PNG signature")
    if len(data) < 8 + 12:
        raise ValueError("Incomplete PNG chunk")
    (length,) = struct.unpack_from(">I", data, 8)
    ctype = data[12:16]
    if 16 + length + 4 > len(data):
        raise ValueError("Incomplete PNG chunk data")
    body = data[16:16 + length]
    (crc,) = struct.unpack_from(">I", data, 16 + length)
    if png_crc32(ctype + body) != crc:
        raise ValueError(f"CRC mismatch for chunk {ctype.decode('latin-1')}")
    if ctype != b"IHDR":
        raise ValueError("First chunk must be IHDR")
    return parse_ihdr(PngChunk(length, "IHDR", body, crc))


def extract_idat(chunks: List[PngChunk]) -> bytes:
    parts = [c.data for c in chunks if c.type == "IDAT"]
    return parts[0] if len(parts) == 1 else b"".join(parts)   # one IDAT: the chunk body itself, no copy


def inflate_idat(compressed: bytes) -> bytes:
    """Host inflate of the concatenated IDAT payload (src/png-decompress.ts)."""
    d = zlib.decompressobj()
    out = d.decompress(compressed)
    out += d.flush()
    return out


def read_png(data) -> Tuple[PngHeader, bytes]:
    """Header + inflated (still filtered) scanline stream of a PNG file."""
    chunks = parse_png_chunks(data)
    if not chunks or chunks[0].type != "IHDR":
        raise ValueError("First chunk must be IHDR")
    header = parse_ihdr(chunks[0])
    return header, inflate_idat(extract_idat(chunks))


def create_chunk(ctype: str, data: bytes) -> PngChunk:
    """src/png-writer.ts:12-30"""
    tb = ctype.encode("latin-1")
    if len(tb) != 4:
        raise ValueError("Chunk type must be exactly 4 characters")
    data = bytes(data)
    return PngChunk(len(data), ctype, data, png_crc32(tb + data))


def serialize_chunk(chunk: PngChunk) -> bytes:
    """src/png-writer.ts:35-56"""
    return struct.pack(">I", chunk.length) + chunk.type.encode("latin-1") + bytes(chunk.data) + struct.pack(">I", chunk.crc)


def create_ihdr(header: PngHeader) -> PngChunk:
    """src/png-writer.ts:61-73"""
    return create_chunk("IHDR", struct.pack(
        ">IIBBBBB", header.width, header.height, header.bit_depth, header.color_type,
        header.compression_method, header.filter_method, header.interlace_method))


def create_iend() -> PngChunk:
    return create_chunk("IEND", b"")


def build_png(chunks: List[PngChunk]) -> bytes:
    """src/png-writer.ts:85-107"""
    return PNG_SIGNATURE + b"".join(serialize_chunk(c) for c in chunks)
