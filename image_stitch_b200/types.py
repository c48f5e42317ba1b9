"""Public types of the host side.

Mirrors the reference's option and plugin surface so that a user of
image-stitch finds the same names and meanings (snake_case for Python):

* ``ConcatOptions``      -- src/types.ts:43-144
* ``PngHeader``          -- src/types.ts:16-24
* ``ImageHeader``        -- src/decoders/types.ts:9-22
* ``ImageDecoder``       -- src/decoders/types.ts:35-71
* ``DecoderPlugin``      -- src/decoders/types.ts:165-173
* ``PositionedImage``    -- src/decoders/types.ts:120-129
* ``ImageSource``        -- src/decoders/types.ts:139-146
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Any, Callable, Iterable, Iterator, Optional, Protocol, Sequence, Union, runtime_checkable


@dataclass
class PngHeader:
    width: int
    height: int
    bit_depth: int
    color_type: int
    compression_method: int = 0
    filter_method: int = 0
    interlace_method: int = 0


@dataclass
class ImageHeader:
    width: int
    height: int
    channels: int
    bit_depth: int
    format: str = "unknown"
    metadata: Optional[dict] = None


@runtime_checkable
class ImageDecoder(Protocol):
    """Decoder contract (src/decoders/types.ts:35-71).

    ``scanlines()`` yields one *unfiltered* row per step, exactly
    ``ceil(width*bit_depth*samples/8)`` bytes long.  A decoder may additionally
    offer ``inflated()`` returning ``(PngHeader, bytes)`` -- the inflated,
    still filtered IDAT stream -- which lets the engine start the GPU path
    where BASELINE.json's north_star starts it.
    """

    def get_header(self) -> ImageHeader: ...
    def scanlines(self) -> Iterator[bytes]: ...
    def close(self) -> None: ...


@runtime_checkable
class DecoderPlugin(Protocol):
    format: str

    def create(self, input: Any, options: Optional[dict] = None) -> ImageDecoder: ...


@dataclass
class PositionedImage:
    x: int
    y: int
    source: Any
    z_index: Optional[int] = None


@dataclass
class ImageSource:
    width: int
    height: int
    factory: Callable[[], Union[bytes, bytearray, memoryview]]


ImageInput = Union[str, bytes, bytearray, memoryview, ImageDecoder, PositionedImage, ImageSource]


@dataclass
class Layout:
    columns: Optional[int] = None
    rows: Optional[int] = None
    width: Optional[int] = None
    height: Optional[int] = None


@dataclass
class ConcatOptions:
    inputs: Iterable[ImageInput]
    layout: Union[Layout, dict] = field(default_factory=Layout)
    decoder_options: Optional[dict] = None
    decoders: Optional[Sequence[DecoderPlugin]] = None
    output_format: str = "png"
    jpeg_quality: Optional[int] = None
    background_color: Any = None
    enable_alpha_blending: Optional[bool] = None
    on_progress: Optional[Callable[[int, int], None]] = None
    # additive (not in the reference): which CUDA devices run the row bands
    devices: Optional[Sequence[int]] = None
    band_rows: Optional[int] = None
    # additive: DEFLATE the filtered rows on the device; the IDAT bytes then differ from the
    # host's zlib output but inflate to the same rows
    device_deflate: bool = False
    # additive: PNG inputs cross the bus as their zlib streams and are inflated on the device;
    # same pixels, and zlib's error messages when a stream is corrupt
    device_inflate: bool = False
    # additive, BEYOND the reference: colour type 3 inputs are expanded through PLTE/tRNS on the
    # device.  Off by default: the reference fails with `Unsupported source color type: 3`, and so does this build
    expand_palette: bool = False

    @staticmethod
    def coerce(options: Union["ConcatOptions", dict]) -> "ConcatOptions":
        if isinstance(options, ConcatOptions):
            opts = options
        else:
            known = {
                "inputs", "layout", "decoder_options", "decoders", "output_format", "jpeg_quality",
                "background_color", "enable_alpha_blending", "on_progress", "devices", "band_rows",
                "device_deflate", "device_inflate", "expand_palette",
            }
            camel = {
                "decoderOptions": "decoder_options", "outputFormat": "output_format",
                "jpegQuality": "jpeg_quality", "backgroundColor": "background_color",
                "enableAlphaBlending": "enable_alpha_blending", "onProgress": "on_progress",
                "deviceDeflate": "device_deflate", "deviceInflate": "device_inflate",
                "expandPalette": "expand_palette",
            }
            kw = {}
            for k, v in dict(options).items():
                k = camel.get(k, k)
                if k not in known:
                    raise TypeError(f"Unknown option: {k}")
                kw[k] = v
            if "inputs" not in kw:
                kw["inputs"] = None
            opts = ConcatOptions(**kw)
        if isinstance(opts.layout, dict):
            opts.layout = Layout(**opts.layout)
        elif opts.layout is None:
            opts.layout = Layout()
        return opts
