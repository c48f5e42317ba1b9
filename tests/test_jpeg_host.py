"""Host side of the JPEG input path (no GPU): the header scan mirrors parseJpegHeader (src/decoders/jpeg-decoder.ts:46-94),
the plugin is selected by magic bytes, the C ABI exports the three entry points."""
import os

import pytest

from image_stitch_b200.decoders import (JpegBufferDecoder, create_decoder, detect_image_format, get_default_decoder_plugins,
                                        parse_jpeg_header)

HERE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "jpeg_decode")


def _file(name):
    with open(os.path.join(HERE, name + ".jpg"), "rb") as f:
        return f.read()


def test_header_scan():
    assert parse_jpeg_header(_file("c420_q80")) == (61, 45, 3)
    assert parse_jpeg_header(_file("g_q90")) == (40, 31, 1)
    assert parse_jpeg_header(_file("p420_progressive")) == (40, 40, 3)   # SOF2 is a frame header too (:26-40)
    with pytest.raises(ValueError, match="Invalid JPEG: missing SOI marker"):
        parse_jpeg_header(b"\x00\x01\x02")
    with pytest.raises(ValueError, match="Invalid JPEG: no SOF marker found"):
        parse_jpeg_header(b"\xff\xd8\xff\xe0\x00\x04ab\xff\xd9")
    data = _file("c444_q90")
    with pytest.raises(ValueError, match="Invalid JPEG: truncated SOF marker"):
        parse_jpeg_header(data[:data.index(b"\xff\xc0") + 6])


def test_plugin_selection_and_header():
    data = _file("c422_q85")
    assert detect_image_format(data) == "jpeg"
    dec = create_decoder(data, {}, get_default_decoder_plugins())
    assert isinstance(dec, JpegBufferDecoder)
    h = dec.get_header()
    assert (h.width, h.height, h.channels, h.bit_depth) == (50, 33, 4, 8)
    assert dec.jpeg_file() == data
    assert create_decoder(_file("g_q90"), {}, get_default_decoder_plugins()).get_header().channels == 1
