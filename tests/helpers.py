"""Shared test helpers: golden fixtures, oracle-side stitching, synthetic tiles."""
from __future__ import annotations

import os
import zlib

import numpy as np

import oracle_py as orc
from image_stitch_b200 import engine
from image_stitch_b200.decoders.decoder_factory import create_decoders
from image_stitch_b200.layout import image_header_to_png_header
from image_stitch_b200.png_container import parse_png_chunks, read_png, scanline_bytes
from image_stitch_b200.types import ConcatOptions, PngHeader

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def suite_path(name: str) -> str:
    return os.path.join(GOLDEN, "pngsuite", name if name.endswith(".png") else name + ".png")


def suite_bytes(name: str) -> bytes:
    with open(suite_path(name), "rb") as f:
        return f.read()


def expected_output(name: str) -> bytes:
    with open(os.path.join(GOLDEN, "expected-outputs", name), "rb") as f:
        return f.read()


def inflated_idat(png_bytes: bytes):
    return read_png(png_bytes)


def oracle_plan(options):
    """Host planning (product code) + pixel sources decoded by the ORACLE."""
    opts = ConcatOptions.coerce(options)
    inputs = engine.collect_inputs(opts)
    decoders = create_decoders(inputs, opts.decoder_options or {}, opts.decoders)
    headers = [image_header_to_png_header(d.get_header()) for d in decoders]
    plan = engine.plan_job(opts, inputs, headers, logger=lambda m: None)
    tiles = []
    for p in plan.placements:
        dec = decoders[p.image_idx]
        if hasattr(dec, "inflated"):
            h, stream = dec.inflated()
            pix = orc.decode_scanlines(stream, h.width, h.height, h.bit_depth, h.color_type, h.interlace_method)
            if h.color_type == 3 and opts.expand_palette:   # extension (oracle.h): the tile enters as RGBA8
                plte, trns = dec.palette()
                pix = orc.expand_palette(pix, h.width, h.height, h.bit_depth, plte, trns).reshape(-1)
                h = PngHeader(h.width, h.height, 8, 6, h.compression_method, h.filter_method, h.interlace_method)
        elif hasattr(dec, "jpeg_file"):   # JPEG input: the oracle's own decoder (oracle/jpeg_decode_oracle.c)
            h = headers[p.image_idx]
            pix = orc.jpeg_decode(dec.jpeg_file()).reshape(-1)
        else:
            h = headers[p.image_idx]
            pix = np.frombuffer(b"".join(dec.scanlines()), np.uint8)
        tiles.append(dict(width=h.width, height=h.height, bit_depth=h.bit_depth, color_type=h.color_type, pixels=pix,
                          dst_x=p.dst_x, dst_y=p.dst_y, src_off_x=p.src_off_x, src_off_y=p.src_off_y,
                          vis_w=p.vis_w, vis_h=p.vis_h, z_index=p.z_index, image_idx=p.image_idx))
    return plan, tiles


def oracle_stitch_rows(options, out_kind=0, y0=0, y1=None):
    plan, tiles = oracle_plan(options)
    rows = orc.render_rows(plan.canvas_w, plan.canvas_h, plan.out_bit_depth, plan.background, plan.alpha_blend,
                           tiles, y0, y1, out_kind)
    return plan, rows


def png_filtered_rows(png_bytes: bytes):
    """(header, inflated IDAT) of an output PNG: the filtered scanlines incl. filter bytes."""
    return read_png(png_bytes)


# ---- synthetic inputs -----------------------------------------

def synth_pixels(rng, w, h, bit_depth, color_type, alpha_mode="opaque"):
    """Dense unfiltered rows of a smooth-gradient + noise image in the given PNG format."""
    spp = {0: 1, 2: 3, 4: 2, 6: 4}[color_type]
    maxv = (1 << bit_depth) - 1
    yy, xx = np.mgrid[0:h, 0:w]
    chans = []
    for c in range(spp):
        base = ((xx * (3 + c) + yy * (5 - c)) % 256) / 255.0
        noise = rng.integers(-8, 9, size=(h, w)) / 255.0
        chans.append(np.clip(base + noise, 0, 1))
    img = np.stack(chans, axis=-1)
    has_alpha = color_type in (4, 6)
    if has_alpha:
        if alpha_mode == "opaque":
            img[..., -1] = 1.0
        elif alpha_mode == "uniform":
            img[..., -1] = rng.integers(1, 255, size=(h, w)) / 255.0
        elif alpha_mode == "binary":
            img[..., -1] = rng.integers(0, 2, size=(h, w)).astype(float)
        elif alpha_mode == "any":
            img[..., -1] = rng.integers(0, 256, size=(h, w)) / 255.0
    vals = np.rint(img * maxv).astype(np.uint32)
    if bit_depth == 16:
        out = np.zeros((h, w, spp, 2), np.uint8)
        out[..., 0] = vals >> 8
        out[..., 1] = vals & 0xFF
        return out.reshape(h, -1)
    if bit_depth == 8:
        return vals.astype(np.uint8).reshape(h, -1)
    ppb = 8 // bit_depth
    rb = (w * bit_depth + 7) // 8
    packed = np.zeros((h, rb), np.uint8)
    v = vals[..., 0]
    for x in range(w):
        packed[:, x // ppb] |= (v[:, x] << ((ppb - 1 - x % ppb) * bit_depth)).astype(np.uint8)
    return packed


def filter_image(pixels2d, bpp, mode="adaptive", rng=None):
    """Dense rows -> inflated IDAT stream.  mode: 'adaptive' uses the oracle's
    own filterScanline (a9), an int 0..4 forces one type, 'random' mixes."""
    h, n = pixels2d.shape
    out = np.zeros((h, n + 1), np.uint8)
    prev = None
    for y in range(h):
        row = np.ascontiguousarray(pixels2d[y])
        if mode == "adaptive":
            ft, f, _ = orc.filter_scanline(row, prev, bpp)
        else:
            ft = int(rng.integers(0, 5)) if mode == "random" else int(mode)
            f = forced_filter(row, prev, bpp, ft)
        out[y, 0] = ft
        out[y, 1:] = f
        prev = row
    return out.reshape(-1)


def forced_filter(row, prev, bpp, ft):
    n = len(row)
    x = row.astype(np.int32)
    a = np.zeros(n, np.int32)
    a[bpp:] = x[:-bpp] if bpp < n else 0
    b = np.zeros(n, np.int32) if prev is None else prev.astype(np.int32)
    c = np.zeros(n, np.int32)
    if prev is not None and bpp < n:
        c[bpp:] = b[:-bpp]
    if ft == 0:
        p = np.zeros(n, np.int32)
    elif ft == 1:
        p = a
    elif ft == 2:
        p = b
    elif ft == 3:
        p = (a + b) >> 1
    else:
        pp = a + b - c
        pa, pb, pc = np.abs(pp - a), np.abs(pp - b), np.abs(pp - c)
        p = np.where((pa <= pb) & (pa <= pc), a, np.where(pb <= pc, b, c))
    return ((x - p) & 0xFF).astype(np.uint8)


def adam7_interlace(pixels2d, w, h, bit_depth, color_type, rng=None, mode="random"):
    """Dense rows -> inflated Adam7 stream (7 passes, each filtered on its own chain)."""
    spp = {0: 1, 2: 3, 4: 2, 6: 4}[color_type]
    bpp = max(1, spp * bit_depth // 8)
    xs, ys = [0, 4, 0, 2, 0, 1, 0], [0, 0, 4, 0, 2, 0, 1]
    dx, dy = [8, 8, 4, 4, 2, 2, 1], [8, 8, 8, 4, 4, 2, 2]
    chunks = []
    for p in range(7):
        pw = max(0, -(-(w - xs[p]) // dx[p]))
        ph = max(0, -(-(h - ys[p]) // dy[p]))
        if pw == 0 or ph == 0:
            continue
        rb = (pw * bit_depth * spp + 7) // 8
        sub = np.zeros((ph, rb), np.uint8)
        if bit_depth >= 8:
            src = pixels2d.reshape(h, w, bpp)
            sub = src[ys[p]::dy[p], xs[p]::dx[p]].reshape(ph, -1)[:, :rb].copy()
        else:
            ppb = 8 // bit_depth
            mask = (1 << bit_depth) - 1
            for px in range(pw):
                fx = xs[p] + px * dx[p]
                v = (pixels2d[ys[p]::dy[p], fx // ppb] >> ((ppb - 1 - fx % ppb) * bit_depth)) & mask
                sub[:, px // ppb] |= (v << ((ppb - 1 - px % ppb) * bit_depth)).astype(np.uint8)
        chunks.append(filter_image(np.ascontiguousarray(sub), bpp, mode, rng))
    return np.concatenate(chunks) if chunks else np.zeros(0, np.uint8)


def make_png(pixels2d, w, h, bit_depth, color_type, interlace=0, mode="adaptive", rng=None) -> bytes:
    """A complete PNG file around synthetic pixels (for API-level tests)."""
    from image_stitch_b200.png_container import build_png, create_chunk, create_ihdr, create_iend
    spp = {0: 1, 2: 3, 4: 2, 6: 4}[color_type]
    bpp = max(1, spp * bit_depth // 8)
    if interlace:
        stream = adam7_interlace(pixels2d, w, h, bit_depth, color_type, rng, mode if mode != "adaptive" else "adaptive")
    else:
        stream = filter_image(pixels2d, bpp, mode, rng)
    hdr = PngHeader(w, h, bit_depth, color_type, 0, 0, interlace)
    return build_png([create_ihdr(hdr), create_chunk("IDAT", zlib.compress(stream.tobytes(), 6)), create_iend()])
