"""One process per GPU: the canvas is sharded into contiguous row bands
(SURVEY.md section 8e).  There is NO data-path collective: rank k composites and
encodes rows [y0_k, y1_k) from the tiles that intersect them (plus the one
predictor row above, recomputed rather than exchanged) and the bands are
handed, in ascending order, to the container writer on rank 0.  The only
communication is that final hand-off: a gather TO RANK 0 ONLY of container-ready
pieces (every rank deflates its own band, or passes on the device's deflate
blocks) over a host-memory (gloo) group -- no NCCL collective, no padding --,
which stands in for "D2H into pinned buffers, consumed in order" when ranks are
separate processes.

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


IDAT_PIECE = 8 << 20   # bytes of compressed data per IDAT chunk written by rank 0


def gpu_band(plan: JobPlan, sources: Dict[int, TileSource], y0: int, y1: int, device: Optional[int] = None):
    """Default renderer: libstitch_b200 on this rank's device.  Returns the band's rows as the device left them:
    filtered scanlines (bytes), JPEG entropy segments (bytes), or -- with device_deflate -- a list of
    (deflate blocks, crc32 of them, adler32, raw length) per chunk."""
    runner = engine.BandRunner(plan, sources, engine.default_device() if device is None else device, y0, y1)
    try:
        step = engine.default_band_rows(plan)
        parts = []
        for a in range(y0, y1, step):
            data, _ = runner.run(a, min(y1, a + step))
            if plan.output_format == "jpeg":
                from .jpeg_container import JpegBandSplicer
                parts.append(JpegBandSplicer().push(data))
            elif plan.device_deflate:
                from .deflate_container import DeflateBandSplicer
                sp = DeflateBandSplicer()
                body = sp.push(data)
                parts.append((body, sp.last_crc32, sp.adler, sp.raw_bytes))
            else:
                parts.append(np.ascontiguousarray(data).tobytes())
        return parts if plan.device_deflate and plan.output_format != "jpeg" else b"".join(parts)
    finally:
        runner.close()


def band_to_pieces(plan: JobPlan, band) -> dict:
    """What a rank hands to rank 0: container-ready pieces of ITS rows, so that rank 0 only frames and writes.
    PNG: raw deflate data ending on a byte boundary (sync flush), so that the ranks' pieces concatenate into one
    zlib stream, with the Adler-32 and length of the rows they encode -- every rank deflates its own band (level 6,
    like streaming-deflate.ts:57-66) in parallel, or passes on the device's deflate blocks with the device's CRC-32.
    JPEG: the band's entropy-coded segments (restart intervals make them concatenate)."""
    if plan.output_format == "jpeg":
        return {"kind": "jpeg", "pieces": [bytes(band)] if band else []}
    if isinstance(band, list):   # device deflate: (blocks, crc32, adler32, raw length) per chunk
        from .deflate_container import adler32_combine
        adler, raw = 1, 0
        for _, _, a, n in band:
            adler = adler32_combine(adler, a, n)
            raw += n
        return {"kind": "deflate", "pieces": [bytes(b) for b, _, _, _ in band], "crcs": [c for _, c, _, _ in band], "adler": adler, "raw": raw}
    raw = bytes(band)
    if not raw:
        return {"kind": "deflate", "pieces": [], "crcs": None, "adler": 1, "raw": 0}
    co = zlib.compressobj(6, zlib.DEFLATED, -15)
    z = co.compress(raw) + co.flush(zlib.Z_SYNC_FLUSH)
    return {"kind": "deflate", "pieces": [z[o:o + IDAT_PIECE] for o in range(0, len(z), IDAT_PIECE)], "crcs": None,
            "adler": zlib.adler32(raw), "raw": len(raw)}


_host_group = None


def host_group():
    """The process group the hand-off to rank 0 uses: host memory to host memory (gloo), whatever the default
    backend is -- the bands are already in host memory and no GPU should see them again."""
    global _host_group
    import torch.distributed as dist
    if dist.get_backend() == "gloo":
        return None
    if _host_group is None:
        _host_group = dist.new_group(backend="gloo")
    return _host_group


def gather_pieces(local, dst: int = 0) -> Optional[list]:
    """Gather of one object per rank to `dst` ONLY, in rank order (= ascending rows): no padding, no copy to the other
    ranks, no NCCL."""
    import torch.distributed as dist
    world, rank = dist.get_world_size(), dist.get_rank()
    out = [None] * world if rank == dst else None
    dist.gather_object(local, out, dst=dst, group=host_group())
    return out


def gather_bytes(local: bytes, dst: int = 0) -> Optional[List[bytes]]:
    """Variable-length gather of byte strings to `dst` (see gather_pieces)."""
    return gather_pieces(bytes(local), dst)


def write_container(plan: JobPlan, parts: List[dict]) -> bytes:
    """Rank 0: the ranks' pieces (ascending rows) -> PNG or JPEG file, like api.StreamingConcatenator: several IDAT
    chunks (a chunk holds at most 2^31 - 1 bytes), one zlib stream whose Adler-32 is combined from the ranks'."""
    if plan.output_format == "jpeg":
        from .jpeg_container import EOI, jpeg_header
        return jpeg_header(plan.canvas_w, plan.canvas_h, plan.jpeg_quality) + b"".join(b"".join(p["pieces"]) for p in parts) + EOI
    from .deflate_container import FINAL_BLOCK, ZLIB_HEADER, adler32_combine, idat_chunk
    out = [PNG_SIGNATURE, serialize_chunk(create_ihdr(PngHeader(plan.canvas_w, plan.canvas_h, plan.out_bit_depth, plan.out_color_type, 0, 0, 0)))]
    head, adler = ZLIB_HEADER, 1
    for p in parts:
        adler = adler32_combine(adler, p["adler"], p["raw"])
        for i, piece in enumerate(p["pieces"]):
            if p["crcs"] is not None:   # the device's CRC-32 of the piece: the chunk CRC needs no pass over the bytes
                out.append(idat_chunk(head, piece, p["crcs"][i]))
            else:
                out.append(serialize_chunk(create_chunk("IDAT", head + piece)))
            head = b""
    out.append(serialize_chunk(create_chunk("IDAT", head + FINAL_BLOCK + adler.to_bytes(4, "big"))))
    out.append(serialize_chunk(create_iend()))
    return b"".join(out)


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
        plan.device_deflate = bool(opts.device_deflate) and plan.output_format != "jpeg"
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
    parts = gather_pieces(band_to_pieces(plan, band), 0)
    if parts is None:
        return None
    return write_container(plan, parts)
