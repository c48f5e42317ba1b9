"""Host-side engine: turns ConcatOptions into a placed-tile plan and drives
libstitch_b200 band by band.

This is the Python restatement of the *orchestration* half of
src/image-concat-core.ts (CoreStreamingConcatenator.stream :927-1003,
streamGridMode :1008-1094, streamPositionedMode :1099-1217): input collection,
decoder creation, header reading, common-format choice, layout resolution,
background colour, progress callbacks and error wrapping.  The per-row loops
(:389-685) are NOT here: they are the CUDA kernels behind the C ABI.
"""
from __future__ import annotations

import os
import threading
from dataclasses import dataclass, field
from typing import Callable, Dict, Iterator, List, Optional, Sequence, Tuple

from . import _native
from .background import get_transparent_color
from .decoders.decoder_factory import (create_decoders, extract_positions, has_positioned_images,
                                       validate_positioned_inputs)
from .decoders.plugin_registry import get_default_decoder_plugins
from .layout import (calculate_canvas_size, calculate_layout, clip_images_to_canvas, determine_common_format,
                     image_header_to_png_header)
from .png_container import get_bytes_per_pixel, scanline_bytes
from .types import ConcatOptions, PngHeader


class StitchFailure(Exception):
    """createStitchError (src/image-concat-core.ts:21-28)."""

    def __init__(self, message: str, cause: Optional[BaseException] = None):
        super().__init__(f"Failed to stitch images: {message}")
        self.__cause__ = cause


@dataclass
class TilePlacement:
    image_idx: int
    dst_x: int
    dst_y: int
    src_off_x: int
    src_off_y: int
    vis_w: int
    vis_h: int
    z_index: int


@dataclass
class JobPlan:
    mode: str                     # "grid" | "positioned"
    canvas_w: int
    canvas_h: int
    out_bit_depth: int
    out_color_type: int
    output_format: str            # "png" | "jpeg"
    jpeg_quality: int
    alpha_blend: bool
    background: bytes
    headers: List[PngHeader]
    placements: List[TilePlacement]
    warnings: List[str] = field(default_factory=list)
    device_deflate: bool = False  # PNG only: bands leave the device as deflate blocks


_ctx_lock = threading.Lock()
_contexts: Dict[int, _native.Context] = {}


def default_device() -> int:
    if "STITCH_B200_DEVICE" in os.environ:
        return int(os.environ["STITCH_B200_DEVICE"])
    return int(os.environ.get("LOCAL_RANK", "0"))


def get_context(device: Optional[int] = None) -> _native.Context:
    dev = default_device() if device is None else device
    with _ctx_lock:
        ctx = _contexts.get(dev)
        if ctx is None or not ctx.handle:
            ctx = _native.Context(dev)
            _contexts[dev] = ctx
        return ctx


def decode_image(header: PngHeader, stream, device: Optional[int] = None):
    """Kernel (a) alone: what ImageDecoder.scanlines() yields, as one array."""
    return get_context(device).decode_image(header.width, header.height, header.bit_depth, header.color_type,
                                            header.interlace_method, stream, filtered=1)


def decode_jpeg(data, device: Optional[int] = None):
    """The device JPEG decoder alone: (H, W, 4) RGBA8 or (H, W) gray8."""
    return get_context(device).decode_jpeg(data)


def collect_inputs(options: ConcatOptions) -> list:
    """stream() :929-949 -- arrays, iterables and async iterables are all accepted."""
    inputs = options.inputs
    if inputs is None:
        raise ValueError("At least one input image is required")
    if hasattr(inputs, "__aiter__"):
        import asyncio

        async def _drain():
            return [i async for i in inputs]
        inputs = asyncio.run(_drain())
    else:
        inputs = list(inputs)
    if len(inputs) == 0:
        raise ValueError("At least one input image is required")
    return inputs


def plan_job(options: ConcatOptions, inputs: list, headers: List[PngHeader],
             logger: Optional[Callable[[str], None]] = None) -> JobPlan:
    """Layout + format decisions of streamGridMode / streamPositionedMode."""
    target_bd, target_ct = determine_common_format(headers)
    output_format = options.output_format or "png"
    final_bd = 8 if output_format == "jpeg" else target_bd      # :1024-1027, :1145-1148
    final_ct = 6 if output_format == "jpeg" else target_ct
    background = get_transparent_color(final_ct, final_bd, options.background_color)
    quality = 85 if options.jpeg_quality is None else options.jpeg_quality  # :1054
    warnings: List[str] = []
    log = logger or warnings.append
    placements: List[TilePlacement] = []
    if has_positioned_images(inputs):
        validate_positioned_inputs(inputs)
        positions = extract_positions(inputs)
        layout = options.layout
        cw, ch = calculate_canvas_size(
            [{"x": p["x"], "y": p["y"], "width": headers[i].width, "height": headers[i].height}
             for i, p in enumerate(positions)], layout.width, layout.height)
        clipped, positioned = clip_images_to_canvas(positions, headers, cw, ch, log)
        clip_by_idx = {c.image_idx: c for c in clipped}
        for p in positioned:
            c = clip_by_idx.get(p.image_idx)
            placements.append(TilePlacement(p.image_idx, p.x, p.y, c.source_offset_x if c else 0,
                                            c.source_offset_y if c else 0, p.width, p.height, p.z_index))
        alpha = options.enable_alpha_blending is not False      # :1170
        return JobPlan("positioned", cw, ch, final_bd, final_ct, output_format, quality, alpha, background,
                       headers, placements, warnings)
    layout = options.layout
    if not layout.columns and not layout.rows and not layout.width and not layout.height:
        raise ValueError("Grid mode requires layout: columns, rows, width, or height")   # :1017
    g = calculate_layout(headers, layout)
    y = 0
    for r, row in enumerate(g.grid):
        x = 0
        for c, idx in enumerate(row):
            if idx >= 0:
                h = headers[idx]
                placements.append(TilePlacement(idx, x, y, 0, 0, h.width, h.height, 0))
            x += g.col_widths[r][c]
        y += g.row_heights[r]
    return JobPlan("grid", g.total_width, g.total_height, final_bd, final_ct, output_format, quality, False,
                   background, headers, placements, warnings)


@dataclass
class TileSource:
    header: PngHeader     # the header the pixel bytes follow (true PNG header for inflated streams)
    data: object          # bytes-like
    filtered: int         # 0 dense rows, 1 inflated stream, 2 zlib stream, 3 JPEG file
    interlace: int
    plte: Optional[bytes] = None   # expand_palette extension: chunk bodies of a colour type 3 input
    trns: bytes = b""


def read_tile_source(decoder, declared: PngHeader, image_idx: int, device_inflate: bool = False,
                     expand_palette: bool = False) -> TileSource:
    """Fetches one input's pixel bytes.  Decoders offering ``inflated()`` hand over the
    still-filtered stream (GPU unfilter); any other ImageDecoder is drained
    through scanlines() with the reference's row checks (image-concat-core.ts:427-448)."""
    def with_palette(src: TileSource) -> TileSource:
        if expand_palette and src.header.color_type == 3 and hasattr(decoder, "palette"):
            src.plte, src.trns = decoder.palette()
            if src.plte is None:
                raise StitchFailure(f"input #{image_idx + 1} has colour type 3 but no PLTE chunk")
        return src

    if hasattr(decoder, "jpeg_file"):          # filtered == 3: the file itself, decoded on the device
        return TileSource(declared, decoder.jpeg_file(), 3, 0)
    if device_inflate and hasattr(decoder, "deflated"):
        hdr, zstream = decoder.deflated()      # filtered == 2: the device inflates
        return with_palette(TileSource(hdr, zstream, 2, 1 if hdr.interlace_method == 1 else 0))
    if hasattr(decoder, "inflated"):
        hdr, stream = decoder.inflated()
        return with_palette(TileSource(hdr, stream, 1, 1 if hdr.interlace_method == 1 else 0))
    expected = scanline_bytes(declared.width, declared.bit_depth, declared.color_type)
    rows = []
    for n, row in enumerate(decoder.scanlines()):
        if n >= declared.height:
            break
        if len(row) != expected:
            bits = declared.bit_depth * {0: 1, 2: 3, 4: 2, 6: 4}[declared.color_type]
            actual = 0 if bits == 0 else (len(row) * 8) / bits
            actual_s = f"{int(actual)}px" if float(actual).is_integer() else f"{actual:.2f}px"
            raise StitchFailure(
                f"dimension mismatch for input #{image_idx + 1}. Expected {declared.width}px wide scanline "
                f"({expected} raw bytes) but decoder produced {actual_s} ({len(row)} raw bytes).")
        rows.append(bytes(row))
    if len(rows) < declared.height:
        raise StitchFailure(
            f"dimension mismatch for input #{image_idx + 1}. Expected {declared.height}px tall image but "
            f"decoder ended after {len(rows)}px.")
    return TileSource(declared, b"".join(rows), 0, 0)


def split_rows(y0: int, y1: int, parts: int, align: int = 8) -> List[Tuple[int, int]]:
    """Contiguous row ranges, edges aligned (JPEG MCU rows)."""
    total = y1 - y0
    out = []
    prev = y0
    for k in range(1, parts + 1):
        edge = y1 if k == parts else y0 + (total * k // parts) // align * align
        edge = max(edge, prev)
        if edge > prev:
            out.append((prev, edge))
        prev = edge
    return out


class BandRunner:
    """One device's share of a job: a stitch_job restricted to rows [y0,y1)
    holding only the tiles that intersect it (plus the 1-row PNG halo)."""

    def __init__(self, plan: JobPlan, sources: Dict[int, TileSource], device: int, y0: int, y1: int):
        self.plan, self.y0, self.y1 = plan, y0, y1
        self.ctx = get_context(device)
        kind = _native.OUT_JPEG_ENTROPY if plan.output_format == "jpeg" else (
            _native.OUT_PNG_DEFLATE if plan.device_deflate else _native.OUT_PNG_FILTERED)
        self.job = _native.Job(self.ctx, plan.canvas_w, plan.canvas_h, plan.out_bit_depth, kind, plan.alpha_blend,
                               plan.background, plan.jpeg_quality, band=(y0, y1))
        halo0 = max(0, y0 - 1)
        self.tiles: List[TilePlacement] = []
        for p in plan.placements:
            if p.dst_y < y1 and p.dst_y + p.vis_h > halo0:
                s = sources[p.image_idx]
                if s.filtered == 3:
                    self.job.add_tile_jpeg(s.data, dst_x=p.dst_x, dst_y=p.dst_y, src_off_x=p.src_off_x, src_off_y=p.src_off_y,
                                           vis_w=p.vis_w, vis_h=p.vis_h, z_index=p.z_index, image_idx=p.image_idx)
                    self.tiles.append(p)
                    continue
                self.job.add_tile(s.data, width=s.header.width, height=s.header.height, bit_depth=s.header.bit_depth,
                                  color_type=s.header.color_type, interlace=s.interlace, filtered=s.filtered,
                                  dst_x=p.dst_x, dst_y=p.dst_y, src_off_x=p.src_off_x, src_off_y=p.src_off_y,
                                  vis_w=p.vis_w, vis_h=p.vis_h, z_index=p.z_index, image_idx=p.image_idx,
                                  plte=s.plte, trns=s.trns)
                self.tiles.append(p)

    def run(self, a: int, b: int):
        return self.job.run_band(a, b)

    def close(self):
        self.job.close()


def default_band_rows(plan: JobPlan) -> int:
    row = plan.canvas_w * (8 if plan.out_bit_depth == 16 else 4) + 1
    rows = max(8, (256 << 20) // row // 8 * 8)
    return rows


def run_plan(plan: JobPlan, sources: Dict[int, TileSource], devices: Optional[Sequence[int]] = None,
             band_rows: Optional[int] = None,
             on_band: Optional[Callable[[int, int], None]] = None) -> Iterator[Tuple[int, int, object, dict]]:
    """Yields (y0, y1, band bytes, timings) in ascending row order.  With several
    devices each one owns a contiguous share of the canvas rows and works in
    its own thread; no data-path collective is involved."""
    devs = list(devices) if devices else [default_device()]
    shares = split_rows(0, plan.canvas_h, len(devs))
    step = band_rows or default_band_rows(plan)
    if plan.output_format == "jpeg":
        step = max(8, step // 8 * 8)
    if len(shares) == 1:
        runner = BandRunner(plan, sources, devs[0], 0, plan.canvas_h)
        try:
            for a in range(0, plan.canvas_h, step):
                b = min(plan.canvas_h, a + step)
                data, tm = runner.run(a, b)
                yield a, b, data, tm
                if on_band:
                    on_band(a, b)
        finally:
            runner.close()
        return
    from concurrent.futures import ThreadPoolExecutor
    runners = [BandRunner(plan, sources, devs[i], s[0], s[1]) for i, s in enumerate(shares)]

    def work(r: BandRunner):
        return [(a, min(r.y1, a + step)) + r.run(a, min(r.y1, a + step)) for a in range(r.y0, r.y1, step)]

    try:
        with ThreadPoolExecutor(max_workers=len(runners)) as pool:
            futures = [pool.submit(work, r) for r in runners]
            for f in futures:  # ascending band order
                for a, b, data, tm in f.result():
                    yield a, b, data, tm
                    if on_band:
                        on_band(a, b)
    finally:
        for r in runners:
            r.close()
