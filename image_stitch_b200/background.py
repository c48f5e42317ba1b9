"""Background colour handling (host side, once per job).

parse_background_color -- src/pixel-ops.ts:8-89
rgba_to_color_type     -- src/pixel-ops.ts:94-167
get_transparent_color  -- src/pixel-ops.ts:255-287
"""
from __future__ import annotations

import math
from typing import Sequence, Tuple, Union

_NAMED = {
    "black": (0, 0, 0, 255), "white": (255, 255, 255, 255), "red": (255, 0, 0, 255),
    "green": (0, 255, 0, 255), "blue": (0, 0, 255, 255), "yellow": (255, 255, 0, 255),
    "cyan": (0, 255, 255, 255), "magenta": (255, 0, 255, 255), "gray": (128, 128, 128, 255),
    "grey": (128, 128, 128, 255),
}


def _js_round(x: float) -> int:
    r = math.floor(x)
    return int(r + 1) if x - r >= 0.5 else int(r)


def _is_int(v) -> bool:
    return isinstance(v, int) and not isinstance(v, bool) or (isinstance(v, float) and v.is_integer())


def _parse_hex_pair(s: str) -> float:
    # JS parseInt(s, 16): longest valid hex prefix, NaN when there is none
    digits = ""
    for ch in s:
        if ch in "0123456789abcdefABCDEF":
            digits += ch
        else:
            break
    return float(int(digits, 16)) if digits else float("nan")


def parse_background_color(color: Union[None, str, Sequence[int]]) -> Tuple[int, int, int, int]:
    if color is None or color == "transparent":
        return (0, 0, 0, 0)
    if isinstance(color, (list, tuple)):
        if len(color) == 3:
            if not all(_is_int(v) and 0 <= v <= 255 for v in color):
                raise ValueError("RGB color values must be integers between 0 and 255")
            r, g, b = (int(v) for v in color)
            return (r, g, b, 255)
        if len(color) == 4:
            if not all(_is_int(v) and 0 <= v <= 255 for v in color):
                raise ValueError("RGBA color values must be integers between 0 and 255")
            r, g, b, a = (int(v) for v in color)
            return (r, g, b, a)
        raise ValueError("Color array must have 3 (RGB) or 4 (RGBA) values")
    if not isinstance(color, str):
        raise ValueError(f"Unsupported color format: {color}. Use hex (#RRGGBB), RGB array [r,g,b], or named color")
    lower = color.lower()
    if lower in _NAMED:
        return _NAMED[lower]
    if color.startswith("#"):
        hx = color[1:]
        a = 255.0
        if len(hx) in (3, 4):
            r = _parse_hex_pair(hx[0] * 2)
            g = _parse_hex_pair(hx[1] * 2)
            b = _parse_hex_pair(hx[2] * 2)
            if len(hx) == 4:
                a = _parse_hex_pair(hx[3] * 2)
        elif len(hx) in (6, 8):
            r = _parse_hex_pair(hx[0:2])
            g = _parse_hex_pair(hx[2:4])
            b = _parse_hex_pair(hx[4:6])
            if len(hx) == 8:
                a = _parse_hex_pair(hx[6:8])
        else:
            raise ValueError(f"Invalid hex color format: {color}. Expected #RGB, #RGBA, #RRGGBB, or #RRGGBBAA")
        if any(math.isnan(v) for v in (r, g, b, a)):
            raise ValueError(f"Invalid hex color: {color}")
        return (int(r), int(g), int(b), int(a))
    raise ValueError(f"Unsupported color format: {color}. Use hex (#RRGGBB), RGB array [r,g,b], or named color")


def rgba_to_color_type(rgba: Sequence[int], color_type: int, bit_depth: int) -> bytes:
    r, g, b, a = rgba

    def scale(v: int) -> int:
        if bit_depth == 16:
            return _js_round((v * 0xFFFF) / 0xFF)
        if bit_depth == 8:
            return v
        max_val = (1 << bit_depth) - 1
        return _js_round((v * max_val) / 0xFF)

    def w16(v: int) -> bytes:
        return bytes([(v >> 8) & 0xFF, v & 0xFF])

    if color_type == 0:
        gray = scale(_js_round(0.299 * r + 0.587 * g + 0.114 * b))
        return w16(gray) if bit_depth == 16 else bytes([gray & 0xFF])
    if color_type == 2:
        if bit_depth == 16:
            return w16(scale(r)) + w16(scale(g)) + w16(scale(b))
        return bytes([scale(r) & 0xFF, scale(g) & 0xFF, scale(b) & 0xFF])
    if color_type == 4:
        gray = scale(_js_round(0.299 * r + 0.587 * g + 0.114 * b))
        alpha = scale(a)
        if bit_depth == 16:
            return w16(gray) + w16(alpha)
        return bytes([gray & 0xFF, alpha & 0xFF])
    if color_type == 6:
        if bit_depth == 16:
            return w16(scale(r)) + w16(scale(g)) + w16(scale(b)) + w16(scale(a))
        return bytes([scale(r) & 0xFF, scale(g) & 0xFF, scale(b) & 0xFF, scale(a) & 0xFF])
    raise ValueError(f"Unsupported color type: {color_type}")


def get_transparent_color(color_type: int, bit_depth: int, background_color=None) -> bytes:
    if background_color is not None:
        return rgba_to_color_type(parse_background_color(background_color), color_type, bit_depth)
    bps = 2 if bit_depth == 16 else 1
    if color_type == 0:
        return bytes(bps)
    if color_type == 2:
        return bytes(3 * bps)
    if color_type == 4:
        return bytes(4 if bit_depth == 16 else 2)
    if color_type == 6:
        return bytes(8 if bit_depth == 16 else 4)
    raise ValueError(f"Unsupported color type: {color_type}")
