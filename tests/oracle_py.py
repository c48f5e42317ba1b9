"""ctypes wrapper over oracle/liboracle.so -- the CPU restatement of the
reference (TEST INFRASTRUCTURE; see oracle/oracle.h).  Only tests/, smoke()
and bench.py's cpu_baseline / --impl reference legs import this module."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_ORACLE_DIR = os.path.join(_ROOT, "oracle")
_LIB = os.path.join(_ORACLE_DIR, "liboracle.so")

ORC_OK, ORC_E_FILTER, ORC_E_COLOR, ORC_E_TRUNCATED, ORC_E_TARGET = 0, -4, -5, -6, -9


class OrcTile(C.Structure):
    _fields_ = [
        ("width", C.c_uint32), ("height", C.c_uint32), ("bit_depth", C.c_uint32), ("color_type", C.c_uint32),
        ("pixels", C.c_void_p), ("dst_x", C.c_int32), ("dst_y", C.c_int32), ("src_off_x", C.c_uint32),
        ("src_off_y", C.c_uint32), ("vis_w", C.c_uint32), ("vis_h", C.c_uint32), ("z_index", C.c_int32),
        ("image_idx", C.c_uint32),
    ]


def build(force: bool = False) -> str:
    srcs = [os.path.join(_ORACLE_DIR, f) for f in ("oracle.c", "jpeg_oracle.c", "jpeg_decode_oracle.c", "oracle.h", "Makefile")]
    stale = not os.path.exists(_LIB) or any(os.path.getmtime(s) > os.path.getmtime(_LIB) for s in srcs if os.path.exists(s))
    if force or stale:
        subprocess.run(["make", "-C", _ORACLE_DIR, "-B" if force else "-s"], check=True, capture_output=True)
    return _LIB


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        l = C.CDLL(build())
        vp, sz, i32p = C.c_void_p, C.c_size_t, C.POINTER(C.c_int)
        l.orc_paeth_predictor.argtypes = [C.c_int] * 3
        l.orc_bytes_per_pixel.argtypes = [C.c_int, C.c_int]
        l.orc_scanline_bytes.argtypes = [C.c_uint32, C.c_int, C.c_int]
        l.orc_scanline_bytes.restype = sz
        l.orc_unfilter_scanline.argtypes = [C.c_int, vp, vp, C.c_int, sz, vp]
        l.orc_filter_scanline.argtypes = [vp, vp, C.c_int, sz, vp, vp, vp]
        l.orc_deinterlace_adam7.argtypes = [vp, sz, C.c_uint32, C.c_uint32, C.c_int, C.c_int, vp, i32p, i32p, i32p]
        l.orc_decode_scanlines.argtypes = [vp, sz, C.c_uint32, C.c_uint32, C.c_int, C.c_int, C.c_int, vp,
                                           C.POINTER(C.c_uint32), i32p]
        l.orc_scale_sample.argtypes = [C.c_uint32, C.c_int, C.c_int]
        l.orc_scale_sample.restype = C.c_uint32
        l.orc_convert_scanline.argtypes = [vp, C.c_uint32, C.c_int, C.c_int, C.c_int, C.c_int, vp]
        l.orc_expand_palette_scanline.argtypes = [vp, C.c_uint32, C.c_int, vp, sz, vp, sz, vp]
        l.orc_expand_palette_scanline.restype = None
        l.orc_composite_scanline.argtypes = [vp, vp, C.c_uint32, C.c_uint32, C.c_int, C.c_int]
        l.orc_composite_scanline.restype = None
        l.orc_render_rows.argtypes = [C.c_uint32, C.c_uint32, C.c_int, vp, C.c_int, C.POINTER(OrcTile), sz,
                                      C.c_uint32, C.c_uint32, C.c_int, vp, sz]
        l.orc_render_rows.restype = C.c_int64
        _lib = l
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _u8(a):
    return np.ascontiguousarray(np.frombuffer(a, dtype=np.uint8) if isinstance(a, (bytes, bytearray, memoryview)) else a, dtype=np.uint8)


class OracleError(Exception):
    def __init__(self, code, message):
        super().__init__(message)
        self.code = code


def scanline_bytes(w, bd, ct):
    return lib().orc_scanline_bytes(w, bd, ct)


def unfilter_scanline(ft, scanline, prev, bpp):
    s = _u8(scanline)
    pv = None if prev is None else _u8(prev)
    out = np.zeros(len(s), np.uint8)
    rc = lib().orc_unfilter_scanline(int(ft), _p(s), _p(pv), bpp, len(s), _p(out))
    if rc == ORC_E_FILTER:
        raise OracleError(rc, f"Unknown filter type: {ft}")
    return out


def filter_scanline(row, prev, bpp):
    r = _u8(row)
    pv = None if prev is None else _u8(prev)
    out = np.zeros(len(r), np.uint8)
    scratch = np.zeros(4 * len(r) + 8, np.uint8)
    sums = np.zeros(5, np.uint64)
    ft = lib().orc_filter_scanline(_p(r), _p(pv), bpp, len(r), _p(out), _p(scratch), _p(sums))
    return ft, out, sums


def deinterlace_adam7(data, w, h, bd, ct):
    d = _u8(data)
    out = np.zeros(max(1, h * scanline_bytes(w, bd, ct)), np.uint8)
    ep, el, ef = C.c_int(0), C.c_int(0), C.c_int(0)
    rc = lib().orc_deinterlace_adam7(_p(d), len(d), w, h, bd, ct, _p(out), C.byref(ep), C.byref(el), C.byref(ef))
    if rc == ORC_E_TRUNCATED:
        raise OracleError(rc, f"Unexpected end of decompressed data at pass {ep.value}, line {el.value}")
    if rc == ORC_E_FILTER:
        raise OracleError(rc, f"Unknown filter type: {ef.value}")
    if rc:
        raise OracleError(rc, f"oracle error {rc}")
    return out[: h * scanline_bytes(w, bd, ct)]


def decode_scanlines(data, w, h, bd, ct, interlace):
    d = _u8(data)
    n = scanline_bytes(w, bd, ct)
    out = np.zeros(max(1, h * n), np.uint8)
    rows, ef = C.c_uint32(0), C.c_int(0)
    if interlace:
        return deinterlace_adam7(d, w, h, bd, ct)
    rc = lib().orc_decode_scanlines(_p(d), len(d), w, h, bd, ct, interlace, _p(out), C.byref(rows), C.byref(ef))
    if rc == ORC_E_TRUNCATED:
        raise OracleError(rc, f"Expected {h} scanlines, decoded {rows.value}")
    if rc == ORC_E_FILTER:
        raise OracleError(rc, f"Unknown filter type: {ef.value}")
    if rc:
        raise OracleError(rc, f"oracle error {rc}")
    return out[: h * n]


def scale_sample(v, fb, tb):
    return lib().orc_scale_sample(v, fb, tb)


def convert_scanline(src, width, sbd, sct, tbd, tct):
    s = _u8(src)
    out = np.zeros(max(len(s), width * (8 if tbd == 16 else 4)), np.uint8)
    rc = lib().orc_convert_scanline(_p(s), width, sbd, sct, tbd, tct, _p(out))
    if rc == ORC_E_TARGET:
        raise OracleError(rc, "Only conversion to RGBA (color type 6) is supported")
    if rc == ORC_E_COLOR:
        raise OracleError(rc, f"Unsupported source color type: {sct}")
    if sbd == tbd and sct == tct:
        return out[: len(s)]
    return out[: width * (8 if tbd == 16 else 4)]


def expand_palette(pix, w, h, bd, plte, trns=b""):
    """Extension (oracle.h): h rows of palette indices -> (h, w*4) RGBA8."""
    n = scanline_bytes(w, bd, 3)
    src = _u8(pix)
    p, t = _u8(plte), _u8(trns if trns else b"\0")
    out = np.zeros((h, w * 4), np.uint8)
    for y in range(h):
        row = np.ascontiguousarray(src[y * n:(y + 1) * n])
        lib().orc_expand_palette_scanline(_p(row), w, bd, _p(p), len(plte), _p(t), len(trns) if trns else 0, _p(out[y]))
    return out


def composite_scanline(dest, source, start_x, source_width, bpp, blend):
    d = _u8(dest).copy()
    s = _u8(source)
    lib().orc_composite_scanline(_p(d), _p(s), start_x, source_width, bpp, 1 if blend else 0)
    return d


def render_rows(canvas_w, canvas_h, out_bit_depth, background, alpha_blend, tiles, y0=0, y1=None, out_kind=0):
    """tiles: list of dicts with width,height,bit_depth,color_type,pixels(np.uint8 dense rows),
    dst_x,dst_y,src_off_x,src_off_y,vis_w,vis_h,z_index,image_idx."""
    y1 = canvas_h if y1 is None else y1
    arr = (OrcTile * max(1, len(tiles)))()
    keep = []
    for i, t in enumerate(tiles):
        px = _u8(t["pixels"])
        keep.append(px)
        arr[i] = OrcTile(t["width"], t["height"], t["bit_depth"], t["color_type"], px.ctypes.data,
                         t.get("dst_x", 0), t.get("dst_y", 0), t.get("src_off_x", 0), t.get("src_off_y", 0),
                         t.get("vis_w", t["width"]), t.get("vis_h", t["height"]), t.get("z_index", 0),
                         t.get("image_idx", i))
    bpp = 8 if out_bit_depth == 16 else 4
    row = canvas_w * bpp + (1 if out_kind == 0 else 0)
    out = np.zeros(max(1, row * (y1 - y0)), np.uint8)
    bg = np.frombuffer(bytes(background)[:8].ljust(8, b"\0"), np.uint8).copy()
    n = lib().orc_render_rows(canvas_w, canvas_h, out_bit_depth, _p(bg), 1 if alpha_blend else 0, arr, len(tiles),
                              y0, y1, out_kind, _p(out), out.size)
    if n == ORC_E_COLOR:
        raise OracleError(int(n), "Unsupported source color type")
    if n < 0:
        raise OracleError(int(n), f"oracle error {n}")
    return out[:n]


# ---- JPEG half (oracle/jpeg_oracle.c) ---------------------------------------------------
def _jpeg_setup():
    l = lib()
    if getattr(l, "_jpeg_ready", False):
        return l
    vp, sz = C.c_void_p, C.c_size_t
    l.orc_jpeg_quant_tables.argtypes = [C.c_int, vp, vp]
    l.orc_jpeg_quant_tables.restype = None
    l.orc_jpeg_coefficients.argtypes = [vp, sz, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int, vp]
    l.orc_jpeg_coefficients.restype = C.c_int64
    l.orc_jpeg_encode_rows.argtypes = [vp, sz, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int, vp, sz]
    l.orc_jpeg_encode_rows.restype = C.c_int64
    l.orc_jpeg_header.argtypes = [C.c_uint32, C.c_uint32, C.c_int, vp, sz]
    l.orc_jpeg_header.restype = C.c_int64
    l.orc_jpeg_quantize.argtypes = [C.c_int, C.c_int]
    l._jpeg_ready = True
    return l


def jpeg_quant_tables(quality):
    l = _jpeg_setup()
    a, b = np.zeros(64, np.uint16), np.zeros(64, np.uint16)
    l.orc_jpeg_quant_tables(int(quality), _p(a), _p(b))
    return a, b


def jpeg_header(width, height, quality):
    l = _jpeg_setup()
    buf = np.zeros(1024, np.uint8)
    n = l.orc_jpeg_header(width, height, int(quality), _p(buf), buf.size)
    if n < 0:
        raise OracleError(int(n), "invalid JPEG dimensions")
    return buf[:n].tobytes()


def jpeg_encode_rows(rgba_rows, width, height, y0, y1, quality):
    """rgba_rows: uint8 array [(y1-y0), width*4] = canvas rows y0..y1-1. Returns the entropy bytes (with RSTm)."""
    l = _jpeg_setup()
    r = np.ascontiguousarray(rgba_rows, dtype=np.uint8).reshape(y1 - y0, width * 4)
    cap = ((width + 7) // 8) * ((y1 - y0 + 7) // 8) * 3 * 260 + 64
    out = np.zeros(cap, np.uint8)
    n = l.orc_jpeg_encode_rows(_p(r), width * 4, width, height, y0, y1, int(quality), _p(out), cap)
    if n < 0:
        raise OracleError(int(n), f"oracle error {n}")
    return out[:n].tobytes()


def jpeg_coefficients(rgba_rows, width, height, y0, y1, quality):
    l = _jpeg_setup()
    r = np.ascontiguousarray(rgba_rows, dtype=np.uint8).reshape(y1 - y0, width * 4)
    n = ((width + 7) // 8) * ((y1 - y0 + 7) // 8) * 3 * 64
    out = np.zeros(n, np.int16)
    l.orc_jpeg_coefficients(_p(r), width * 4, width, height, y0, y1, int(quality), _p(out))
    return out.reshape(-1, 3, 64)


def jpeg_file(rgba, width, height, quality):
    """Whole JPEG file by the oracle: header + entropy segments + EOI."""
    r = np.ascontiguousarray(rgba, dtype=np.uint8).reshape(height, width * 4)
    return jpeg_header(width, height, quality) + jpeg_encode_rows(r, width, height, 0, height, quality) + b"\xff\xd9"


# ---- JPEG decode (oracle/jpeg_decode_oracle.c) ----------------------------------------------
def _jdec_setup():
    l = lib()
    if getattr(l, "_jdec_ready", False):
        return l
    vp, sz = C.c_void_p, C.c_size_t
    l.jdec_info.argtypes = [vp, sz, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
    l.jdec_info.restype = C.c_int
    l.jdec_decode.argtypes = [vp, sz, vp, sz]
    l.jdec_decode.restype = C.c_int
    l.jdec_message.argtypes = [C.c_int]
    l.jdec_message.restype = C.c_char_p
    l._jdec_ready = True
    return l


class JpegDecodeError(Exception):
    pass


def jpeg_decode(data: bytes):
    """Baseline JPEG -> (H, W, 4) RGBA8 or (H, W) gray8, as libjpeg-turbo's default path gives it; raises
    JpegDecodeError with the decoder's message."""
    l = _jdec_setup()
    buf = np.frombuffer(bytes(data), dtype=np.uint8)
    w, h, ch = C.c_uint32(), C.c_uint32(), C.c_uint32()
    rc = l.jdec_info(_p(buf), buf.size, C.byref(w), C.byref(h), C.byref(ch))
    if rc:
        raise JpegDecodeError(l.jdec_message(rc).decode())
    out = np.zeros(h.value * w.value * ch.value, dtype=np.uint8)
    rc = l.jdec_decode(_p(buf), buf.size, _p(out), out.size)
    if rc:
        raise JpegDecodeError(l.jdec_message(rc).decode())
    return out.reshape(h.value, w.value, 4) if ch.value == 4 else out.reshape(h.value, w.value)
