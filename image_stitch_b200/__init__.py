"""image_stitch_b200 -- B200-native pixel pipeline behind image-stitch's API.

(The task names the package ``image-stitch_b200``; Python cannot import a
hyphenated name, so the directory is ``image_stitch_b200``.)
"""
from .api import (StreamingConcatenator, concat_streaming, concat_to_buffer, concat_to_file, concat_to_stream)
from .engine import StitchFailure
from .types import ConcatOptions, DecoderPlugin, ImageDecoder, ImageHeader, ImageSource, Layout, PngHeader, PositionedImage

__all__ = [
    "StreamingConcatenator", "concat_streaming", "concat_to_buffer", "concat_to_file", "concat_to_stream",
    "StitchFailure", "ConcatOptions", "DecoderPlugin", "ImageDecoder", "ImageHeader", "ImageSource", "Layout",
    "PngHeader", "PositionedImage",
]
