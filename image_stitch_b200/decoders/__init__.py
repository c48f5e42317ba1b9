from .decoder_factory import (create_decoder, create_decoders, extract_positions, has_positioned_images,
                              is_positioned_image, validate_positioned_inputs)
from .format_detection import detect_format, detect_image_format, validate_format
from .plugin_registry import clear_default_decoder_plugins, get_default_decoder_plugins, set_default_decoder_plugins
from .png_decoder import PngBufferDecoder, PngFileDecoder, png_decoder
from .jpeg_decoder import JpegBufferDecoder, JpegFileDecoder, jpeg_decoder, parse_jpeg_header
