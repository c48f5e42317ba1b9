"""Magic-byte format detection (src/decoders/format-detection.ts:10-137)."""
from __future__ import annotations

_HEIC_BRANDS = {"heic", "heix", "hevc", "hevx", "heim", "heis", "hevm", "hevs", "mif1", "msf1"}


def detect_image_format(b) -> str:
    b = bytes(b[:32])
    if len(b) < 4:
        return "unknown"
    if len(b) >= 8 and b[:8] == bytes([0x89, 0x50, 0x4E, 0x47, 0x0D, 0x0A, 0x1A, 0x0A]):
        return "png"
    if b[0] == 0xFF and b[1] == 0xD8 and b[2] == 0xFF:
        return "jpeg"
    if len(b) >= 12 and b[4:8] == b"ftyp":
        if b[8:12].decode("latin-1") in _HEIC_BRANDS:
            return "heic"
        if len(b) >= 20:
            for i in range(16, min(len(b) - 3, 32), 4):
                if b[i:i + 4].decode("latin-1") in _HEIC_BRANDS:
                    return "heic"
    return "unknown"


def read_magic_bytes(inp) -> bytes:
    if isinstance(inp, (bytes, bytearray, memoryview)):
        return bytes(inp[:32])
    if isinstance(inp, str):
        try:
            with open(inp, "rb") as f:
                return f.read(32)
        except OSError as e:
            raise OSError(f"Failed to read file for format detection: {e}") from e
    raise TypeError("Unsupported input type for format detection")


def detect_format(inp) -> str:
    return detect_image_format(read_magic_bytes(inp))


def validate_format(fmt: str) -> None:
    if fmt == "unknown":
        raise ValueError("Unknown or unsupported image format. Supported formats: PNG, JPEG, HEIC")
