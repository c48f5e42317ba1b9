"""Default decoder plugin list (src/decoders/plugin-registry.ts)."""
from __future__ import annotations

from typing import List, Optional

from .jpeg_decoder import jpeg_decoder
from .png_decoder import png_decoder

_default: Optional[list] = None


def get_default_decoder_plugins() -> List:
    global _default
    if not _default:
        _default = [png_decoder, jpeg_decoder]
    return _default


def set_default_decoder_plugins(plugins) -> None:
    global _default
    _default = list(plugins)


def clear_default_decoder_plugins() -> None:
    global _default
    _default = None
