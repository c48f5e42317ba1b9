"""Public API: same entry points and option meanings as image-stitch.

concat_to_buffer / concat_to_stream / concat_streaming / concat_to_file /
StreamingConcatenator  <->  src/image-concat.ts:32-52,
src/image-concat-core.ts:279-285,927,1475-1511.

Container writing stays on the host like the reference's L1 (SURVEY.md
section 1): zlib deflate with the reference's batching
(src/image-concat-core.ts:336-383), IDAT/IHDR/IEND chunks (src/png-writer.ts),
JPEG markers (src/jpeg-encoder.ts).  Every pixel operation runs on the GPU
through libstitch_b200.
"""
from __future__ import annotations

import io
import os
import zlib
from typing import Dict, Iterator, List, Optional, Union

import numpy as np

from . import engine
from ._native import StitchError
from .decoders.decoder_factory import create_decoders
from .decoders.plugin_registry import get_default_decoder_plugins
from .decoders.png_decoder import PngBufferDecoder
from .engine import JobPlan, StitchFailure, TileSource
from .layout import image_header_to_png_header
from .png_container import PNG_SIGNATURE, create_chunk, create_iend, create_ihdr, serialize_chunk
from .types import ConcatOptions, PngHeader


class StreamingConcatenator:
    """CoreStreamingConcatenator (src/image-concat-core.ts:279-1470)."""

    def __init__(self, options: Union[ConcatOptions, dict]):
        opts = ConcatOptions.coerce(options)
        if opts.inputs is None or (isinstance(opts.inputs, (list, tuple)) and len(opts.inputs) == 0):
            raise ValueError("At least one input image is required")   # :287-293
        self.options = opts
        self.last_timings: List[dict] = []
        self.plan: Optional[JobPlan] = None

    # -- progress (createProgressTracker :1401-1428) -------------------------
    def _progress(self, plan: JobPlan, total_inputs: int):
        cb = self.options.on_progress
        if cb is None:
            return None
        last_row: Dict[int, int] = {}
        for p in plan.placements:
            last_row[p.image_idx] = p.dst_y + p.vis_h
        pending = sorted(last_row.items(), key=lambda kv: kv[1])
        state = {"done": 0, "pos": 0}

        def on_band(y0: int, y1: int) -> None:
            while state["pos"] < len(pending) and pending[state["pos"]][1] <= y1:
                state["pos"] += 1
                state["done"] += 1
                cb(state["done"], total_inputs)
        return on_band

    def stream(self) -> Iterator[bytes]:
        opts = self.options
        inputs = engine.collect_inputs(opts)
        plugins = list(opts.decoders) if opts.decoders is not None else get_default_decoder_plugins()
        decoders = create_decoders(inputs, opts.decoder_options or {}, plugins)
        try:
            headers: List[PngHeader] = [image_header_to_png_header(d.get_header()) for d in decoders]
            plan = engine.plan_job(opts, inputs, headers, logger=_warn)
            plan.device_deflate = bool(opts.device_deflate) and plan.output_format != "jpeg"
            self.plan = plan
            needed = sorted({p.image_idx for p in plan.placements})
            sources: Dict[int, TileSource] = {}

            def read(idx: int) -> TileSource:
                try:
                    return engine.read_tile_source(decoders[idx], headers[idx], idx, bool(opts.device_inflate),
                                                   bool(opts.expand_palette))
                except StitchFailure:
                    raise
                except Exception as e:  # decoder failures surface as stitch errors (:21-28)
                    raise StitchFailure(f"unable to read input #{idx + 1}: {e}", e) from e

            # The built-in PNG decoders spend their time in zlib (chunk CRC, host inflate), which releases the GIL:
            # they are read on a few threads.  Foreign decoders are drained one after the other, as the reference
            # does.  Errors surface in input order either way.
            threaded = [i for i in needed if isinstance(decoders[i], PngBufferDecoder)]
            if len(threaded) > 1:
                from concurrent.futures import ThreadPoolExecutor
                with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1, len(threaded))) as pool:
                    futures = {i: pool.submit(read, i) for i in threaded}
                    for idx in needed:
                        sources[idx] = futures[idx].result() if idx in futures else read(idx)
            else:
                for idx in needed:
                    sources[idx] = read(idx)
            on_band = self._progress(plan, len(inputs))
            try:
                if plan.output_format == "jpeg":
                    yield from self._stream_jpeg(plan, sources, on_band)
                else:
                    yield from self._stream_png(plan, sources, on_band)
            except StitchError as e:
                raise StitchFailure(str(e), e) from e
        finally:
            for d in decoders:   # :997-1002
                try:
                    d.close()
                except Exception:
                    pass

    # -- PNG (streamCompressedData :309-384, streamPositionedPngData :1222-1314) --
    def _stream_png(self, plan: JobPlan, sources, on_band) -> Iterator[bytes]:
        bpp = 8 if plan.out_bit_depth == 16 else 4
        yield PNG_SIGNATURE
        yield serialize_chunk(create_ihdr(PngHeader(plan.canvas_w, plan.canvas_h, plan.out_bit_depth,
                                                    plan.out_color_type, 0, 0, 0)))
        scanline_size = plan.canvas_w * bpp + 1
        if plan.device_deflate:
            # the device already produced deflate blocks: one IDAT per band, zlib framing around them
            # (the chunk CRC comes from the device's CRC-32 of the band: no host pass over the bytes)
            from .deflate_container import ZLIB_HEADER, DeflateBandSplicer, idat_parts
            splicer = DeflateBandSplicer()
            head = ZLIB_HEADER
            for y0, y1, data, tm in engine.run_plan(plan, sources, self.options.devices, self.options.band_rows, on_band):
                self.last_timings.append(tm)
                body = splicer.push(data, copy=False)      # a view: every band has its own buffer
                yield from idat_parts(head, body, splicer.last_crc32)
                head = b""
            yield serialize_chunk(create_chunk("IDAT", head + splicer.finish()))
            yield serialize_chunk(create_iend())
            return
        batch_rows = max(50, (1 << 20) // scanline_size)          # :336-344
        deflater = zlib.compressobj(6)                              # level 6 (:341)
        rows_in_batch = 0
        for y0, y1, data, tm in engine.run_plan(plan, sources, self.options.devices, self.options.band_rows, on_band):
            self.last_timings.append(tm)
            buf = memoryview(np.ascontiguousarray(data)).cast("B")
            n_rows = y1 - y0
            r = 0
            while r < n_rows:
                take = min(n_rows - r, batch_rows - rows_in_batch)
                out = deflater.compress(buf[r * scanline_size:(r + take) * scanline_size])
                r += take
                rows_in_batch += take
                if rows_in_batch == batch_rows:                    # periodic Z_SYNC_FLUSH (:361-364)
                    out += deflater.flush(zlib.Z_SYNC_FLUSH)
                    rows_in_batch = 0
                if out:
                    yield serialize_chunk(create_chunk("IDAT", out))
        tail = deflater.flush()
        if tail:
            yield serialize_chunk(create_chunk("IDAT", tail))
        yield serialize_chunk(create_iend())

    # -- JPEG (streamJpegData :837-920, streamPositionedJpegData :1319-1399) ----
    def _stream_jpeg(self, plan: JobPlan, sources, on_band) -> Iterator[bytes]:
        from .jpeg_container import JpegBandSplicer, jpeg_header
        yield jpeg_header(plan.canvas_w, plan.canvas_h, plan.jpeg_quality)
        splicer = JpegBandSplicer()
        for y0, y1, data, tm in engine.run_plan(plan, sources, self.options.devices, self.options.band_rows, on_band):
            self.last_timings.append(tm)
            chunk = splicer.push(data)
            if chunk:
                yield chunk
        yield splicer.finish()


def _warn(message: str) -> None:
    import warnings
    warnings.warn(message, stacklevel=2)   # console.warn in the reference (:1127-1132)


def concat_streaming(options) -> Iterator[bytes]:
    """coreConcatStreaming (:1475-1480)"""
    return StreamingConcatenator(options).stream()


def concat_to_buffer(options) -> bytes:
    """concatToBuffer (src/image-concat.ts:34-36) / concatCore (:1484-1503)"""
    return b"".join(concat_streaming(options))


class _GeneratorReader(io.RawIOBase):
    def __init__(self, gen: Iterator[bytes]):
        self._gen, self._buf = gen, b""

    def readable(self) -> bool:
        return True

    def readinto(self, b) -> int:
        while not self._buf:
            try:
                self._buf = next(self._gen)
            except StopIteration:
                return 0
        n = min(len(b), len(self._buf))
        b[:n] = self._buf[:n]
        self._buf = self._buf[n:]
        return n


def concat_to_stream(options) -> io.BufferedReader:
    """concatToStream (src/image-concat.ts:38-41): a readable stream, chunks in file order."""
    return io.BufferedReader(_GeneratorReader(concat_streaming(options)))


def concat_to_file(options, path: str) -> int:
    n = 0
    with open(path, "wb") as f:
        for chunk in concat_streaming(options):
            f.write(chunk)
            n += len(chunk)
    return n
