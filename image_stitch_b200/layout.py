"""Layout resolution on the host: O(#tiles) integer work, produces the placed
tile rectangles the kernels consume.

Grid mode:        calculate_layout / calculate_pixel_based_layout
                  -- src/image-concat-core.ts:132-204, :209-261
Positioned mode:  calculate_canvas_size, clip_images_to_canvas,
                  build_scanline_index -- src/positioned-layout.ts:80-242
Common format:    determine_common_format -- src/pixel-ops.ts:293-307
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Callable, Dict, List, Optional, Sequence, Tuple

from .types import ImageHeader, Layout, PngHeader


def image_header_to_png_header(header: ImageHeader) -> PngHeader:
    """src/image-concat-core.ts:47-71: colour type from the channel count."""
    try:
        color_type = {1: 0, 2: 4, 3: 2, 4: 6}[header.channels]
    except KeyError:
        raise ValueError(f"Unsupported channel count: {header.channels}") from None
    return PngHeader(header.width, header.height, header.bit_depth, color_type, 0, 0, 0)


def determine_common_format(headers: Sequence[PngHeader]) -> Tuple[int, int]:
    bit_depth = 8
    for h in headers:
        if h.bit_depth == 16:
            bit_depth = 16
    return bit_depth, 6


@dataclass
class GridLayout:
    grid: List[List[int]]
    row_heights: List[int]
    col_widths: List[List[int]]
    total_width: int
    total_height: int


def calculate_pixel_based_layout(headers: Sequence[PngHeader], max_width=None, max_height=None,
                                 fixed_columns=None, fixed_rows=None) -> List[List[int]]:
    grid: List[List[int]] = []
    row: List[int] = []
    row_w = 0
    row_h = 0
    total_h = 0
    for i, h in enumerate(headers):
        exceed_w = bool(max_width) and (row_w + h.width > max_width)
        exceed_c = bool(fixed_columns) and (len(row) >= fixed_columns)
        if (exceed_w or exceed_c) and row:
            if bool(max_height) and (total_h + row_h + h.height > max_height):
                break
            grid.append(row)
            total_h += row_h
            row = [i]
            row_w = h.width
            row_h = h.height
        else:
            row.append(i)
            row_w += h.width
            row_h = max(row_h, h.height)
        if fixed_rows and len(grid) >= fixed_rows and not row:
            break
    if row:
        grid.append(row)
    return grid


def calculate_layout(headers: Sequence[PngHeader], layout: Layout) -> GridLayout:
    n = len(headers)
    if layout.columns and not layout.height:
        cols = layout.columns
        rows = math.ceil(n / cols)
        grid = [[(r * cols + c) if (r * cols + c) < n else -1 for c in range(cols)] for r in range(rows)]
    elif layout.rows and not layout.width:
        rows = layout.rows
        cols = math.ceil(n / rows)
        grid = [[(c * rows + r) if (c * rows + r) < n else -1 for c in range(cols)] for r in range(rows)]
    elif layout.width or layout.height:
        grid = calculate_pixel_based_layout(headers, layout.width, layout.height, layout.columns, layout.rows)
    else:
        grid = [list(range(n))]
    row_heights: List[int] = []
    col_widths: List[List[int]] = []
    for row in grid:
        max_h = 0
        widths: List[int] = []
        for idx in row:
            if idx >= 0:
                max_h = max(max_h, headers[idx].height)
                widths.append(headers[idx].width)
            else:
                widths.append(0)
        row_heights.append(max_h)
        col_widths.append(widths)
    total_h = sum(row_heights)
    total_w = max((sum(w) for w in col_widths), default=0)
    return GridLayout(grid, row_heights, col_widths, total_w, total_h)


# ---------------------------------------------------------------- positioned


@dataclass
class PositionedImageInfo:
    image_idx: int
    x: int
    y: int
    z_index: int
    width: int
    height: int
    current_scanline: int = 0


@dataclass
class ClippedImageInfo:
    image_idx: int
    original_x: int
    original_y: int
    original_width: int
    original_height: int
    clipped_x: int
    clipped_y: int
    clipped_width: int
    clipped_height: int
    source_offset_x: int
    source_offset_y: int
    fully_clipped: bool


@dataclass
class ScanlineIntersection:
    image_idx: int
    local_y: int
    start_x: int
    end_x: int
    z_index: int


def calculate_canvas_size(images: Sequence[dict], explicit_width: Optional[int] = None,
                          explicit_height: Optional[int] = None) -> Tuple[int, int]:
    if explicit_width is not None and explicit_height is not None:
        return explicit_width, explicit_height
    max_right = 0
    max_bottom = 0
    for img in images:
        max_right = max(max_right, img["x"] + img["width"])
        max_bottom = max(max_bottom, img["y"] + img["height"])
    w = explicit_width if explicit_width is not None else max(1, max_right)
    h = explicit_height if explicit_height is not None else max(1, max_bottom)
    return w, h


def _default_logger(message: str) -> None:
    import warnings
    warnings.warn(message, stacklevel=3)


def clip_images_to_canvas(positions: Sequence[dict], headers: Sequence[PngHeader], canvas_width: int,
                          canvas_height: int, logger: Callable[[str], None] = _default_logger):
    clipped: List[ClippedImageInfo] = []
    positioned: List[PositionedImageInfo] = []
    for i, pos in enumerate(positions):
        x, y = pos["x"], pos["y"]
        w, h = headers[i].width, headers[i].height
        left, top = max(0, x), max(0, y)
        right, bottom = min(canvas_width, x + w), min(canvas_height, y + h)
        is_clipped = x < 0 or y < 0 or x + w > canvas_width or y + h > canvas_height
        fully = right <= left or bottom <= top
        if is_clipped:
            clipped.append(ClippedImageInfo(
                i, x, y, w, h, left, top, 0 if fully else right - left, 0 if fully else bottom - top,
                max(0, -x), max(0, -y), fully))
            if fully:
                logger(f"Image #{i + 1} is completely outside canvas bounds: position=({x}, {y}), "
                       f"size=({w}×{h}), canvas=({canvas_width}×{canvas_height}). "
                       f"Image will not be rendered.")
            else:
                parts = []
                if x < 0:
                    parts.append(f"left by {-x}px")
                if y < 0:
                    parts.append(f"top by {-y}px")
                if x + w > canvas_width:
                    parts.append(f"right by {x + w - canvas_width}px")
                if y + h > canvas_height:
                    parts.append(f"bottom by {y + h - canvas_height}px")
                logger(f"Image #{i + 1} clipped ({', '.join(parts)}): original=({x}, {y}, {w}×{h}), "
                       f"visible=({left}, {top}, {right - left}×{bottom - top}), "
                       f"canvas=({canvas_width}×{canvas_height})")
        if not fully:
            z = pos.get("z_index")
            positioned.append(PositionedImageInfo(i, left, top, i if z is None else z, right - left, bottom - top, 0))
    return clipped, positioned


def build_scanline_index(positioned: Sequence[PositionedImageInfo], canvas_height: int) -> Dict[int, List[ScanlineIntersection]]:
    """src/positioned-layout.ts:201-242.  The GPU engine does not need the
    per-row map (it sorts tiles once by (z_index, image_idx)); it is kept for
    API parity and as the statement of the ordering contract."""
    index: Dict[int, List[ScanlineIntersection]] = {}
    for y in range(canvas_height):
        hits = [ScanlineIntersection(p.image_idx, y - p.y, p.x, p.x + p.width, p.z_index)
                for p in positioned if p.y <= y < p.y + p.height]
        hits.sort(key=lambda s: (s.z_index, s.image_idx))
        if hits:
            index[y] = hits
    return index
