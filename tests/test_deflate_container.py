"""Host side of SURVEY.md 8(f).1: zlib framing, Adler-32 / CRC-32 combination (no GPU needed)."""
import os
import zlib

import numpy as np

from image_stitch_b200.deflate_container import adler32_combine, crc32_combine, idat_chunk
from image_stitch_b200.png_container import parse_png_chunks, PNG_SIGNATURE


def test_crc32_combine_matches_zlib_on_concatenations():
    rng = np.random.default_rng(0)
    for la, lb in [(0, 0), (0, 5), (5, 0), (1, 1), (4, 1000), (1000, 4), (65536, 65537), (3, 1 << 20)]:
        a, b = rng.bytes(la), rng.bytes(lb)
        assert crc32_combine(zlib.crc32(a), zlib.crc32(b), lb) == zlib.crc32(a + b), (la, lb)


def test_adler32_combine_matches_zlib_on_concatenations():
    rng = np.random.default_rng(1)
    for la, lb in [(0, 0), (0, 5), (5, 0), (70000, 3), (3, 70000), (1 << 20, 1 << 19)]:
        a, b = rng.bytes(la), rng.bytes(lb)
        assert adler32_combine(zlib.adler32(a), zlib.adler32(b), lb) == zlib.adler32(a + b), (la, lb)


def test_idat_chunk_is_a_valid_png_chunk():
    body = os.urandom(5000)
    raw = idat_chunk(b"\x78\x01", body, zlib.crc32(body), b"\x01\x00\x00\xff\xff")
    (chunk,) = parse_png_chunks(PNG_SIGNATURE + raw)   # parse_png_chunks verifies the CRC
    assert chunk.type == "IDAT" and chunk.data == b"\x78\x01" + body + b"\x01\x00\x00\xff\xff"
