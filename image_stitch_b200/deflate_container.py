"""zlib framing around the deflate blocks that libstitch_b200 produces on the device
(STITCH_OUT_PNG_DEFLATE): the role of StreamingDeflator's output side (src/streaming-deflate.ts:130-185).

Every band is a run of complete, byte-aligned, non-final deflate blocks, so the stream is
  78 01 | band | band | ... | 01 00 00 FF FF | Adler-32 (big endian)
with the Adler-32 of all raw rows combined from the per-band values (RFC 1950; zlib's adler32_combine).
The CRC-32 of each band's compressed bytes comes from the device as well, so an IDAT chunk's check value
(src/png-writer.ts:12-30: CRC over type and data) is crc32_combine(crc32(b"IDAT" + prefix), band crc, length):
no host pass over the data.
"""
from __future__ import annotations

import ctypes as C
import zlib

import numpy as np

from . import _native

ZLIB_HEADER = b"\x78\x01"            # deflate, 32 KiB window, no preset dictionary, check bits
FINAL_BLOCK = b"\x01\x00\x00\xff\xff"  # BFINAL = 1, stored, zero bytes
_M = 65521
_INFO = C.sizeof(_native.DeflateBandInfo)


def adler32_combine(adler1: int, adler2: int, len2: int) -> int:
    """Adler-32 of A+B from adler32(A), adler32(B) and len(B)."""
    a1, b1 = adler1 & 0xFFFF, adler1 >> 16
    a2, b2 = adler2 & 0xFFFF, adler2 >> 16
    a = (a1 + a2 - 1) % _M
    b = (b1 + b2 + (len2 % _M) * (a1 - 1)) % _M
    return (b << 16) | a


_CRC_POLY = 0xEDB88320


def _crc_mulmod(a: int, b: int) -> int:
    """a(x) * b(x) mod P over GF(2), bits reflected as in the CRC register (bit 31 is x^0)."""
    p = 0
    for j in range(32):
        if (a >> (31 - j)) & 1:
            p ^= b
        b = (b >> 1) ^ (_CRC_POLY if b & 1 else 0)
    return p


def crc32_combine(crc1: int, crc2: int, len2: int) -> int:
    """zlib.crc32(A + B) from zlib.crc32(A), zlib.crc32(B) and len(B): crc1 * x^(8 len2) + crc2."""
    p, sq, n = 0x80000000, 0x80000000 >> 8, len2   # 1, x^8
    while n:
        if n & 1:
            p = _crc_mulmod(p, sq)
        sq = _crc_mulmod(sq, sq)
        n >>= 1
    return _crc_mulmod(crc1, p) ^ crc2


def idat_chunk(prefix: bytes, body: bytes, body_crc: int, suffix: bytes = b"") -> bytes:
    """A serialised IDAT chunk holding prefix + body + suffix, its CRC derived from the device's CRC of body."""
    crc = crc32_combine(zlib.crc32(b"IDAT" + prefix), body_crc, len(body))
    if suffix:
        crc = crc32_combine(crc, zlib.crc32(suffix), len(suffix))
    n = len(prefix) + len(body) + len(suffix)
    return b"".join((n.to_bytes(4, "big"), b"IDAT", prefix, body, suffix, crc.to_bytes(4, "big")))


def idat_parts(prefix: bytes, body, body_crc: int, suffix: bytes = b""):
    """The same chunk as (head, body, tail) without copying the body (hundreds of MB per band): the stream yields the
    three parts one after the other."""
    crc = crc32_combine(zlib.crc32(b"IDAT" + prefix), body_crc, len(body))
    if suffix:
        crc = crc32_combine(crc, zlib.crc32(suffix), len(suffix))
    n = len(prefix) + len(body) + len(suffix)
    return n.to_bytes(4, "big") + b"IDAT" + prefix, body, suffix + crc.to_bytes(4, "big")


class DeflateBandSplicer:
    """Joins the per-band results of stitch_job_run_band (STITCH_OUT_PNG_DEFLATE), in ascending row order."""

    def __init__(self) -> None:
        self.adler = 1
        self.raw_bytes = 0
        self.bytes_out = 0
        self.last_crc32 = 0   # CRC-32 of the body push() returned last (computed on the device)

    def push(self, band, copy: bool = True):
        """The band's deflate blocks; with copy=False a view into `band` (which must then stay alive and unchanged)."""
        data = memoryview(np.ascontiguousarray(band)).cast("B")
        info = _native.DeflateBandInfo.from_buffer_copy(bytes(data[:_INFO]))
        body = data[_INFO:_INFO + info.n_bytes]
        if copy:
            body = bytes(body)
        if len(body) != info.n_bytes:
            raise _native.StitchError(_native.E_STATE, "deflate band header does not match its payload")
        self.adler = adler32_combine(self.adler, info.adler, info.n_raw)
        self.last_crc32 = info.crc32
        self.raw_bytes += info.n_raw
        self.bytes_out += len(body)
        return body

    def finish(self) -> bytes:
        return FINAL_BLOCK + self.adler.to_bytes(4, "big")
