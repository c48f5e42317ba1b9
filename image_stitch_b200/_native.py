"""ctypes binding over include/stitch_b200.h (libstitch_b200.so).

This is the Python twin of the N-API addon described in INTEGRATION.md.  There
is no CPU fallback: if the shared library is missing, or no CUDA device is
visible, every entry point raises.
"""
from __future__ import annotations

import ctypes as C
import weakref
import os
from typing import Optional

_LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "libstitch_b200.so")

STITCH_OK = 0
E_JPEG = -11
E_INVALID, E_CUDA, E_NOMEM, E_FILTER, E_COLOR, E_TRUNCATED, E_STATE, E_CAPACITY, E_INFLATE, E_NODEVICE = (
    -1, -2, -3, -4, -5, -6, -7, -8, -9, -10)
OUT_PNG_FILTERED, OUT_RAW_RGBA, OUT_JPEG_ENTROPY, OUT_PNG_DEFLATE = 0, 1, 2, 3

EXPORTS = [
    "stitch_abi_version", "stitch_ctx_create", "stitch_ctx_destroy", "stitch_ctx_set_stream",
    "stitch_last_error", "stitch_job_create", "stitch_job_destroy", "stitch_job_add_tile",
    "stitch_job_add_tile_device", "stitch_job_add_tile_palette", "stitch_job_decode", "stitch_job_invalidate", "stitch_job_run_band", "stitch_job_band_device",
    "stitch_job_status", "stitch_decode_image", "stitch_filter_rows", "stitch_jpeg_header",
    "stitch_job_inflate", "stitch_inflate", "stitch_filter_rows_noalloc",
    "stitch_selftest_blend", "stitch_selftest_blend8_samples",
    "stitch_jpeg_info", "stitch_job_add_tile_jpeg", "stitch_decode_jpeg",
]


class StitchError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(message)
        self.code = code


class NativeLibraryMissing(ImportError):
    pass


class JobDesc(C.Structure):
    _fields_ = [
        ("canvas_w", C.c_uint32), ("canvas_h", C.c_uint32), ("out_bit_depth", C.c_uint32),
        ("out_kind", C.c_uint32), ("alpha_blend", C.c_uint32), ("jpeg_quality", C.c_uint32),
        ("background", C.c_uint8 * 8), ("band_y0", C.c_uint32), ("band_y1", C.c_uint32),
        ("reserved", C.c_uint32 * 4),
    ]


class TileDesc(C.Structure):
    _fields_ = [
        ("width", C.c_uint32), ("height", C.c_uint32), ("bit_depth", C.c_uint32), ("color_type", C.c_uint32),
        ("interlace", C.c_uint32), ("filtered", C.c_uint32), ("dst_x", C.c_int32), ("dst_y", C.c_int32),
        ("src_off_x", C.c_uint32), ("src_off_y", C.c_uint32), ("vis_w", C.c_uint32), ("vis_h", C.c_uint32),
        ("z_index", C.c_int32), ("image_idx", C.c_uint32),
    ]


class Timings(C.Structure):
    _fields_ = [
        ("h2d_ms", C.c_float), ("decode_ms", C.c_float), ("compose_ms", C.c_float), ("d2h_ms", C.c_float),
        ("total_ms", C.c_float), ("launches", C.c_uint32), ("reserved", C.c_uint32),
    ]

    def as_dict(self) -> dict:
        return {k: getattr(self, k) for k, _ in self._fields_ if k != "reserved"}


class DeflateBandInfo(C.Structure):
    _fields_ = [("n_bytes", C.c_uint64), ("n_raw", C.c_uint64), ("adler", C.c_uint32), ("crc32", C.c_uint32)]


class JpegBandInfo(C.Structure):
    _fields_ = [("n_bits", C.c_uint64), ("last_dc", C.c_int32 * 3), ("reserved", C.c_uint32)]


_lib: Optional[C.CDLL] = None


def load() -> C.CDLL:
    """Loads libstitch_b200.so; raises NativeLibraryMissing when it was not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(_LIB_PATH):
        raise NativeLibraryMissing(
            f"{_LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(there is no CPU fallback for the stitch kernels)")
    lib = C.CDLL(_LIB_PATH)
    vp, u8p, sz = C.c_void_p, C.POINTER(C.c_uint8), C.c_size_t
    lib.stitch_abi_version.restype = C.c_int
    lib.stitch_ctx_create.argtypes = [C.c_int, C.POINTER(vp)]
    lib.stitch_ctx_destroy.argtypes = [vp]
    lib.stitch_ctx_destroy.restype = None
    lib.stitch_ctx_set_stream.argtypes = [vp, vp]
    lib.stitch_last_error.argtypes = [vp]
    lib.stitch_last_error.restype = C.c_char_p
    lib.stitch_job_create.argtypes = [vp, C.POINTER(JobDesc), C.POINTER(vp)]
    lib.stitch_job_destroy.argtypes = [vp]
    lib.stitch_job_destroy.restype = None
    lib.stitch_job_add_tile.argtypes = [vp, C.POINTER(TileDesc), vp, sz]
    lib.stitch_job_add_tile_device.argtypes = [vp, C.POINTER(TileDesc), vp, sz]
    lib.stitch_job_add_tile_palette.argtypes = [vp, C.POINTER(TileDesc), vp, sz, vp, sz, vp, sz]
    lib.stitch_job_decode.argtypes = [vp, C.c_int, C.POINTER(Timings)]
    lib.stitch_job_invalidate.argtypes = [vp]
    lib.stitch_job_run_band.argtypes = [vp, C.c_uint32, C.c_uint32, vp, sz, C.POINTER(sz), C.POINTER(Timings)]
    lib.stitch_job_band_device.argtypes = [vp, C.POINTER(vp), C.POINTER(sz), C.POINTER(sz), C.POINTER(sz)]
    lib.stitch_job_status.argtypes = [vp]
    lib.stitch_decode_image.argtypes = [vp, C.POINTER(TileDesc), vp, sz, vp, sz, C.POINTER(sz)]
    lib.stitch_filter_rows.argtypes = [vp, vp, vp, C.c_uint32, sz, C.c_uint32, vp, sz]
    lib.stitch_jpeg_info.argtypes = [vp, vp, sz, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
    lib.stitch_job_add_tile_jpeg.argtypes = [vp, C.POINTER(TileDesc), vp, sz]
    lib.stitch_decode_jpeg.argtypes = [vp, vp, sz, vp, sz, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
    lib.stitch_jpeg_header.argtypes = [C.c_uint32, C.c_uint32, C.c_uint32, vp, sz, C.POINTER(sz)]
    lib.stitch_job_inflate.argtypes = [vp, C.c_int, C.POINTER(Timings)]
    lib.stitch_inflate.argtypes = [vp, vp, sz, vp, sz, C.POINTER(sz)]
    lib.stitch_selftest_blend.argtypes = [vp, C.c_uint32, C.c_uint64, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), C.POINTER(C.c_uint32)]
    lib.stitch_selftest_blend8_samples.argtypes = [vp, C.c_uint32, C.c_uint32, vp]
    lib.stitch_filter_rows_noalloc.argtypes = [vp, vp, sz, vp, C.c_uint32, sz, C.c_uint32, vp, sz]
    for name in EXPORTS:  # every symbol include/stitch_b200.h declares must be exported
        getattr(lib, name)
    _lib = lib
    return lib


def _addr(buf) -> int:
    """Address of a bytes-like / numpy / ctypes buffer without copying."""
    if buf is None:
        return 0
    if isinstance(buf, int):
        return buf
    if hasattr(buf, "__array_interface__"):
        return buf.__array_interface__["data"][0]
    if hasattr(buf, "data_ptr"):
        return buf.data_ptr()
    if isinstance(buf, bytes):
        return C.cast(C.c_char_p(buf), C.c_void_p).value or 0
    if isinstance(buf, (bytearray, memoryview)):
        mv = memoryview(buf)
        if not len(mv):
            return 0
        if mv.readonly:   # e.g. a chunk body inside a PNG file's bytes: numpy gives the address without a copy
            import numpy as np
            return np.frombuffer(mv, np.uint8).__array_interface__["data"][0]
        return C.addressof(C.c_uint8.from_buffer(mv))
    return C.addressof(buf)


class Context:
    """One stitch_ctx: one CUDA device, one stream."""

    def __init__(self, device: int = 0):
        self.lib = load()
        self.handle = C.c_void_p()
        rc = self.lib.stitch_ctx_create(int(device), C.byref(self.handle))
        if rc == E_NODEVICE:
            raise StitchError(rc, "No CUDA device available: libstitch_b200 has no CPU fallback")
        if rc != STITCH_OK:
            raise StitchError(rc, f"stitch_ctx_create failed ({rc})")
        self.device = device
        self._jobs = weakref.WeakSet()   # jobs still open: they hold device memory of this context

    def close(self) -> None:
        if self.handle:
            for job in list(self._jobs):   # a job left open (an exception mid-band) must not outlive its context
                job.close()
            self.lib.stitch_ctx_destroy(self.handle)
            self.handle = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def last_error(self) -> str:
        msg = self.lib.stitch_last_error(self.handle)
        return msg.decode("utf-8", "replace") if msg else ""

    def check(self, rc: int) -> None:
        if rc != STITCH_OK:
            raise StitchError(rc, self.last_error() or f"libstitch_b200 error {rc}")

    def set_stream(self, cuda_stream: int) -> None:
        self.check(self.lib.stitch_ctx_set_stream(self.handle, C.c_void_p(cuda_stream)))

    # ---- stand-alone pieces -------------------------------------------------
    def decode_image(self, width, height, bit_depth, color_type, interlace, data, filtered=1):
        """Kernel (a) alone: inflated stream -> dense unfiltered rows (numpy uint8)."""
        import numpy as np
        td = TileDesc(width=width, height=height, bit_depth=bit_depth, color_type=color_type,
                      interlace=interlace, filtered=filtered)
        spp = {0: 1, 2: 3, 3: 1, 4: 2, 6: 4}.get(color_type, 1)
        rb = (width * bit_depth * spp + 7) // 8
        out = np.empty(max(1, rb * height), dtype=np.uint8)
        n = C.c_size_t()
        data = bytes(data) if isinstance(data, memoryview) else data
        rc = self.lib.stitch_decode_image(self.handle, C.byref(td), _addr(data), len(data), _addr(out), out.size, C.byref(n))
        self.check(rc)
        return out[: n.value]

    def jpeg_info(self, data):
        """(width, height, channels) of a JPEG file without decoding (parseJpegHeader, jpeg-decoder.ts:46-94)."""
        data = bytes(data) if isinstance(data, memoryview) else data
        w, h, ch = C.c_uint32(), C.c_uint32(), C.c_uint32()
        self.check(self.lib.stitch_jpeg_info(self.handle, _addr(data), len(data), C.byref(w), C.byref(h), C.byref(ch)))
        return w.value, h.value, ch.value

    def decode_jpeg(self, data):
        """The device JPEG decoder alone: (H, W, 4) RGBA8 or (H, W) gray8 (numpy uint8)."""
        import numpy as np
        data = bytes(data) if isinstance(data, memoryview) else data
        w, h, ch = self.jpeg_info(data)
        out = np.empty(max(1, w * h * ch), dtype=np.uint8)
        rc = self.lib.stitch_decode_jpeg(self.handle, _addr(data), len(data), _addr(out), out.size, C.byref(C.c_uint32()),
                                         C.byref(C.c_uint32()), C.byref(C.c_uint32()))
        self.check(rc)
        return out[: w * h * ch].reshape((h, w, 4) if ch == 4 else (h, w))

    def inflate(self, zdata, capacity: int):
        """Inflate alone: one zlib stream -> at most `capacity` bytes (numpy uint8)."""
        import numpy as np
        zdata = bytes(zdata) if isinstance(zdata, memoryview) else zdata
        out = np.empty(max(1, capacity), dtype=np.uint8)
        n = C.c_size_t()
        rc = self.lib.stitch_inflate(self.handle, _addr(zdata), len(zdata), _addr(out), capacity, C.byref(n))
        self.check(rc)
        return out[: n.value]

    def filter_rows(self, rows, prev_row, n_rows, row_bytes, bpp):
        """Kernel (c) alone.  Returns n_rows*(1+row_bytes) bytes."""
        import numpy as np
        rows = np.ascontiguousarray(rows, dtype=np.uint8)
        prev = None if prev_row is None else np.ascontiguousarray(prev_row, dtype=np.uint8)
        out = np.empty(n_rows * (row_bytes + 1), dtype=np.uint8)
        rc = self.lib.stitch_filter_rows(self.handle, _addr(rows), _addr(prev), n_rows, row_bytes, bpp, _addr(out), out.size)
        self.check(rc)
        return out


    def selftest_blend(self, mode: int, amount: int = 0):
        """(checked, mismatches, first offending tuple) of the device-side blend proof (stitch_selftest_blend)."""
        chk, bad, first = C.c_uint64(), C.c_uint64(), (C.c_uint32 * 4)()
        self.check(self.lib.stitch_selftest_blend(self.handle, mode, amount, C.byref(chk), C.byref(bad), first))
        return chk.value, bad.value, tuple(first)

    def selftest_blend8_samples(self, n: int, seed: int = 1):
        import numpy as np
        out = np.empty((n, 3), dtype=np.uint32)
        self.check(self.lib.stitch_selftest_blend8_samples(self.handle, n, seed, _addr(out)))
        return out

    def filter_rows_noalloc(self, d_rows: int, in_pitch: int, d_prev_row: int, n_rows: int, row_bytes: int, bpp: int,
                            d_out: int, out_pitch: int) -> None:
        """Kernel (c) with DEVICE pointers and the library's static scratch: allocates nothing (operators' addition)."""
        self.check(self.lib.stitch_filter_rows_noalloc(self.handle, C.c_void_p(d_rows), in_pitch, C.c_void_p(d_prev_row or None),
                                                       n_rows, row_bytes, bpp, C.c_void_p(d_out), out_pitch))


class Job:
    """One stitch_job (one output canvas, or one row band of it)."""

    def __init__(self, ctx: Context, canvas_w: int, canvas_h: int, out_bit_depth: int = 8,
                 out_kind: int = OUT_PNG_FILTERED, alpha_blend: bool = False, background: bytes = b"",
                 jpeg_quality: int = 85, band=None):
        self.ctx = ctx
        self.lib = ctx.lib
        d = JobDesc(canvas_w=canvas_w, canvas_h=canvas_h, out_bit_depth=out_bit_depth, out_kind=out_kind,
                    alpha_blend=1 if alpha_blend else 0, jpeg_quality=jpeg_quality)
        bg = bytes(background)[:8].ljust(8, b"\0")
        for i in range(8):
            d.background[i] = bg[i]
        if band is not None:
            d.band_y0, d.band_y1 = band
        self.desc = d
        self.handle = C.c_void_p()
        ctx.check(self.lib.stitch_job_create(ctx.handle, C.byref(d), C.byref(self.handle)))
        self.bpp = 8 if out_bit_depth == 16 else 4
        self.out_kind = out_kind
        self.canvas_w, self.canvas_h = canvas_w, canvas_h
        self._keep = []
        self.last_rows = 0
        ctx._jobs.add(self)

    def close(self) -> None:
        if self.handle:
            if self.ctx.handle:
                self.lib.stitch_job_destroy(self.handle)
            self.handle = C.c_void_p()
            self._keep.clear()
            self.ctx._jobs.discard(self)

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def add_tile(self, data, *, width, height, bit_depth, color_type, interlace=0, filtered=1, dst_x=0, dst_y=0,
                 src_off_x=0, src_off_y=0, vis_w=None, vis_h=None, z_index=0, image_idx=0, device_ptr=None,
                 length=None, plte=None, trns=b"") -> None:
        td = TileDesc(width=width, height=height, bit_depth=bit_depth, color_type=color_type, interlace=interlace,
                      filtered=filtered, dst_x=dst_x, dst_y=dst_y, src_off_x=src_off_x, src_off_y=src_off_y,
                      vis_w=width if vis_w is None else vis_w, vis_h=height if vis_h is None else vis_h,
                      z_index=z_index, image_idx=image_idx)
        if device_ptr is not None:
            rc = self.lib.stitch_job_add_tile_device(self.handle, C.byref(td), C.c_void_p(device_ptr), length)
        else:
            if isinstance(data, memoryview) and not data.c_contiguous:
                data = bytes(data)
            n = (data.nbytes if isinstance(data, memoryview) else len(data)) if length is None else length
            if plte is not None:   # 8f.4 extension: colour type 3 with its PLTE / tRNS chunk bodies
                plte, trns = bytes(plte), bytes(trns or b"")
                rc = self.lib.stitch_job_add_tile_palette(self.handle, C.byref(td), _addr(data), n, _addr(plte), len(plte),
                                                          _addr(trns) if trns else None, len(trns))
            else:
                rc = self.lib.stitch_job_add_tile(self.handle, C.byref(td), _addr(data), n)
            self._keep.append(data)   # a pinned source is copied asynchronously: it must outlive the upload (until close)
        self.ctx.check(rc)

    def add_tile_jpeg(self, data, *, dst_x=0, dst_y=0, src_off_x=0, src_off_y=0, vis_w=None, vis_h=None, z_index=0,
                      image_idx=0) -> None:
        """8f.4 extension: a baseline JPEG file as a tile, decoded on the device (csrc/jpeg_decode.cu)."""
        data = bytes(data) if isinstance(data, (memoryview, bytearray)) else data
        w, h, _ = self.ctx.jpeg_info(data)
        td = TileDesc(width=w, height=h, bit_depth=8, color_type=6, interlace=0, filtered=0, dst_x=dst_x, dst_y=dst_y,
                      src_off_x=src_off_x, src_off_y=src_off_y, vis_w=w if vis_w is None else vis_w,
                      vis_h=h if vis_h is None else vis_h, z_index=z_index, image_idx=image_idx)
        self.ctx.check(self.lib.stitch_job_add_tile_jpeg(self.handle, C.byref(td), _addr(data), len(data)))
        self._keep.append(data)

    def decode(self, force: bool = False) -> dict:
        t = Timings()
        self.ctx.check(self.lib.stitch_job_decode(self.handle, 1 if force else 0, C.byref(t)))
        return t.as_dict()

    def inflate(self, force: bool = False) -> dict:
        """Inflates the tiles that were added as zlib streams (filtered=2); timings of that step alone."""
        t = Timings()
        self.ctx.check(self.lib.stitch_job_inflate(self.handle, 1 if force else 0, C.byref(t)))
        return t.as_dict()

    def invalidate(self) -> None:
        """The next run_band unfilters its tiles again (inputs stay resident)."""
        self.ctx.check(self.lib.stitch_job_invalidate(self.handle))

    def band_bytes(self, y0: int, y1: int) -> int:
        row = self.canvas_w * self.bpp + (1 if self.out_kind in (OUT_PNG_FILTERED, OUT_PNG_DEFLATE) else 0)
        if self.out_kind == OUT_PNG_DEFLATE:   # worst case: every 32 KiB block stored, 10 bytes of framing each
            raw = row * (y1 - y0)
            return C.sizeof(DeflateBandInfo) + raw + (raw // 32768 + 1) * 10 + (raw // (1 << 20) + 2) * 16 + 64
        if self.out_kind == OUT_JPEG_ENTROPY:
            # worst case of baseline Huffman coding plus the band header
            mcus = ((self.canvas_w + 7) // 8) * ((y1 - y0 + 7) // 8)
            return C.sizeof(JpegBandInfo) + mcus * 3 * 256 + 64
        return row * (y1 - y0)

    def run_band(self, y0: int, y1: int, out=None, download: bool = True):
        """Returns (numpy uint8 array or None, timings dict)."""
        import numpy as np
        t = Timings()
        n = C.c_size_t()
        if download and out is None:
            out = np.empty(self.band_bytes(y0, y1), dtype=np.uint8)
        if download:
            size = out.numel() * out.element_size() if hasattr(out, "numel") else (out.nbytes if hasattr(out, "nbytes") else len(out))
            rc = self.lib.stitch_job_run_band(self.handle, y0, y1, _addr(out), size, C.byref(n), C.byref(t))
        else:
            rc = self.lib.stitch_job_run_band(self.handle, y0, y1, None, 0, C.byref(n), C.byref(t))
        self.ctx.check(rc)
        self.last_rows = y1 - y0
        res = out[: n.value] if download else None
        return res, t.as_dict()

    def band_device(self):
        ptr, pitch, lead, rb = C.c_void_p(), C.c_size_t(), C.c_size_t(), C.c_size_t()
        self.ctx.check(self.lib.stitch_job_band_device(self.handle, C.byref(ptr), C.byref(pitch), C.byref(lead), C.byref(rb)))
        return ptr.value, pitch.value, lead.value, rb.value
