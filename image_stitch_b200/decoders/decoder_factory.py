"""Decoder selection (src/decoders/decoder-factory.ts:116-321)."""
from __future__ import annotations

from typing import Iterator, List, Optional, Sequence

from ..types import ImageHeader, ImageSource, PositionedImage
from .format_detection import detect_format, validate_format
from .plugin_registry import get_default_decoder_plugins


def is_positioned_image(inp) -> bool:
    if isinstance(inp, PositionedImage):
        return True
    return isinstance(inp, dict) and "x" in inp and "y" in inp and "source" in inp and \
        isinstance(inp["x"], (int, float)) and isinstance(inp["y"], (int, float))


def _as_positioned(inp) -> PositionedImage:
    if isinstance(inp, PositionedImage):
        return inp
    return PositionedImage(inp["x"], inp["y"], inp["source"], inp.get("z_index", inp.get("zIndex")))


def is_image_source(inp) -> bool:
    return isinstance(inp, ImageSource) or (isinstance(inp, dict) and all(k in inp for k in ("factory", "width", "height")))


def _is_decoder(obj) -> bool:
    return all(callable(getattr(obj, n, None)) for n in ("get_header", "scanlines", "close"))


class LazyImageDecoder:
    """decoder-factory.ts:36-82: header from the declared size, data on demand."""

    def __init__(self, source, options, plugins):
        self.source = source if isinstance(source, ImageSource) else ImageSource(source["width"], source["height"], source["factory"])
        self.options, self.plugins = options, plugins
        self._inner = None

    def get_header(self) -> ImageHeader:
        return ImageHeader(self.source.width, self.source.height, 4, 8, "unknown")

    def _decoder(self):
        if self._inner is None:
            self._inner = create_decoder(self.source.factory(), self.options, self.plugins)
        return self._inner

    def scanlines(self) -> Iterator[bytes]:
        dec = self._decoder()
        try:
            yield from dec.scanlines()
        finally:
            dec.close()
            self._inner = None

    def close(self) -> None:
        pass


def extract_source(inp):
    return _as_positioned(inp).source if is_positioned_image(inp) else inp


def create_decoder(inp, options: Optional[dict] = None, plugins: Optional[Sequence] = None):
    source = extract_source(inp)
    if _is_decoder(source):
        return source
    available = list(plugins) if plugins else get_default_decoder_plugins()
    if is_image_source(source):
        return LazyImageDecoder(source, options or {}, available)
    if isinstance(source, (str, bytes, bytearray, memoryview)):
        fmt = detect_format(source)
        validate_format(fmt)
        plugin = next((p for p in available if p.format == fmt), None)
        if plugin is None:
            raise ValueError(f'No decoder registered for format "{fmt}". Provide a matching plugin via options.decoders.')
        return plugin.create(source, options or {})
    raise TypeError("Unsupported input type. Expected string (file path), Uint8Array, ArrayBuffer, Blob, "
                    "ImageDecoder instance, or PositionedImage")


def create_decoders(inputs, options=None, plugins=None) -> List:
    return [create_decoder(i, options, plugins) for i in inputs]


def has_positioned_images(inputs) -> bool:
    return any(is_positioned_image(i) for i in inputs)


def extract_positions(inputs):
    out = []
    for i in inputs:
        if is_positioned_image(i):
            p = _as_positioned(i)
            out.append({"x": p.x, "y": p.y, "z_index": p.z_index})
        else:
            out.append(None)
    return out


def validate_positioned_inputs(inputs) -> None:
    n_pos = sum(1 for i in inputs if is_positioned_image(i))
    if 0 < n_pos < len(inputs):
        raise ValueError("Cannot mix positioned and non-positioned images. All inputs must be PositionedImage "
                         f"objects or none can be. Found {n_pos} positioned and {len(inputs) - n_pos} "
                         "non-positioned images.")
