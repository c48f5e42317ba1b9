"""One process per GPU: the canvas is sharded into contiguous row bands
(SURVEY.md section 8e).  There is NO data-path collective: rank k composites and
encodes rows [y0_k, y1_k) from the tiles that intersect them (plus the one
predictor row above, recomputed rather than exchanged) and the bands are
handed, in ascending order, to the container writer on rank 0.  The only
communication is that final hand-off (a gather of byte strings), which stands
in for "D2H into pinned buffers, consumed in order" when ranks are separate
processes.

Launch with ``torchrun --nproc-per-node N`` (backend nccl on GPUs; the host
logic is exercised on CPU with gloo in tests/test_distributed_gloo.py).
"""
from __future__ import annotations

import zlib
from typing import Callable, Dict, List, Optional, Tuple

import numpy as np

from . import engine
from .engine import JobPlan, TilePlacement, TileSource
from .png_container import PNG_SIGNATURE, create_chunk, create_iend, create_ihdr, serialize_chunk
from .types import PngHeader

BandRenderer = Callable[[JobPlan, Dict[int, TileSource], int, int], bytes]


def rank_rows(canvas_h: int, world: int, align: int = 8) -> List[Tuple[int, int]]:
    """Row band of every rank (edges multiples of `align`, so JPEG MCU rows never straddle ranks).
    Ranks beyond the number of aligned bands get an empty range."""
    shares = engine.split_rows(0, canvas_h, world, align)
    return shares + [(canvas_h, canvas_h)] * (world - len(shares))


def placements_for_band(plan: JobPlan, y0: int, y1: int) -> List[TilePlacement]:
    """Tiles a rank has to hold: everything that touches rows [y0-1, y1) (1-row PNG predictor halo)."""
    halo0 = max(0, y0 - 1)
    return [p for p in plan.placements if p.dst_y < y1 and p.dst_y + p.vis_h > halo0]


def gpu_band(plan: JobPlan, sources: Dict[int, TileSource], y0: int, y1: int, device: Optional[int] = None) -> bytes:
    """Default renderer: libstitch_b200 on this rank's device."""
    runner = engine.BandRunner(plan, sources, engine.default_device() if device is None else device, y0, y1)
    try:
        step = engine.default_band_rows(plan)
        parts = []
        for a in range(y0, y1, step):
            data, _ = runner.run(a, min(y1, a + step))
            if plan.output_format == "jpeg":
                from .jpeg_container import JpegBandSplicer
                parts.append(JpegBandSplicer().push(data))
            else:
                parts.append(np.ascontiguousarray(data).tobytes())
        return b"".join(parts)
    finally:
        runner.close()


def gather_bytes(local: bytes, dst: int = 0) -> Optional[List[bytes]]:
    """Variable-length gather to `dst` over the default process group (gloo or nccl)."""
    import torch
    import torch.distributed as dist
    world, rank = dist.get_world_size(), dist.get_rank()
    dev = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend() == "nccl" else torch.device("cpu")
    sizes = [torch.zeros(1, dtype=torch.int64, device=dev) for _ in range(world)]
    dist.all_gather(sizes, torch.tensor([len(local)], dtype=torch.int64, device=dev))
    cap = max(1, int(max(int(s.item()) for s in sizes)))
    buf = torch.zeros(cap, dtype=torch.uint8, device=dev)
    if local:
        buf[: len(local)] = torch.frombuffer(bytearray(local), dtype=torch.uint8).to(dev)
    out = [torch.zeros(cap, dtype=torch.uint8, device=dev) for _ in range(world)]
    dist.all_gather(out, buf)   # all_gather works on both backends; only dst keeps the result
    if rank != dst:
        return None
    return [out[r][: int(sizes[r].item())].cpu().numpy().tobytes() for r in range(world)]


def write_container(plan: JobPlan, bands: List[bytes]) -> bytes:
    """Rank 0: bands (ascending rows) -> PNG or JPEG file, like api.StreamingConcatenator."""
    if plan.output_format == "jpeg":
        from .jpeg_container import EOI, jpeg_header
        return jpeg_header(plan.canvas_w, plan.canvas_h, plan.jpeg_quality) + b"".join(bands) + EOI
    ihdr = serialize_chunk(create_ihdr(PngHeader(plan.canvas_w, plan.canvas_h, plan.out_bit_depth, plan.out_color_type, 0, 0, 0)))
    idat = serialize_chunk(create_chunk("IDAT", zlib.compress(b"".join(bands), 6)))
    return PNG_SIGNATURE + ihdr + idat + serialize_chunk(create_iend())


def stitch_rank_band(options, render: Optional[BandRenderer] = None):
    """Plans the job on every rank (cheap, deterministic), renders this rank's band.
    Returns (plan, (y0, y1), band bytes)."""
    import torch.distributed as dist
    from .api import StreamingConcatenator
    from .decoders.decoder_factory import create_decoders
    from .decoders.plugin_registry import get_default_decoder_plugins
    from .layout import image_header_to_png_header
    rank, world = dist.get_rank(), dist.get_world_size()
    sc = StreamingConcatenator(options)
    opts = sc.options
    inputs = engine.collect_inputs(opts)
    plugins = list(opts.decoders) if opts.decoders is not None else get_default_decoder_plugins()
    decoders = create_decoders(inputs, opts.decoder_options or {}, plugins)
    try:
        headers = [image_header_to_png_header(d.get_header()) for d in decoders]
        plan = engine.plan_job(opts, inputs, headers, logger=lambda m: None)
        y0, y1 = rank_rows(plan.canvas_h, world)[rank]
        if y1 <= y0:
            return plan, (y0, y1), b""
        needed = sorted({p.image_idx for p in placements_for_band(plan, y0, y1)})
        sources = {i: engine.read_tile_source(decoders[i], headers[i], i, bool(opts.device_inflate), bool(opts.expand_palette)) for i in needed}
        band = (render or gpu_band)(plan, sources, y0, y1)
        return plan, (y0, y1), band
    finally:
        for d in decoders:
            try:
                d.close()
            except Exception:
                pass


def concat_to_buffer_distributed(options, render: Optional[BandRenderer] = None) -> Optional[bytes]:
    """concatToBuffer over all ranks of the default process group: the file on rank 0, None elsewhere."""
    plan, _, band = stitch_rank_band(options, render)
    bands = gather_bytes(band, 0)
    if bands is None:
        return None
    return write_container(plan, bands)
