"""The reference's own JPEG decoder tests (tests/integration/jpeg-decoder.test.ts), case for case, over the device
decoder.  createTestJpeg (a solid-colour JPEG) is made here by this library's own JPEG output path (kernel d), so the
files also exercise encoder -> decoder on the GPU; every decode is compared with the oracle as well."""
import os

import numpy as np
import pytest

import helpers as H
import oracle_py as orc
import image_stitch_b200 as stitch
from image_stitch_b200.decoders import JpegBufferDecoder, JpegFileDecoder

pytestmark = pytest.mark.gpu


def create_test_jpeg(width, height, color):
    """tests/utils/image-fixtures.ts createTestJpeg: a width x height image of one RGBA colour, as JPEG."""
    pix = np.tile(np.asarray(color, np.uint8), (height, width, 1))
    png = H.make_png(pix.reshape(height, width * 4), width, height, 8, 6)
    return stitch.concat_to_buffer({"inputs": [png], "layout": {"columns": 1}, "outputFormat": "jpeg", "jpegQuality": 90})


def _rows(dec):
    return [np.frombuffer(r, np.uint8) for r in dec.scanlines()]


# ---- Header Parsing (:15-54) ----
def test_parses_jpeg_header_correctly():
    dec = JpegBufferDecoder(create_test_jpeg(64, 48, (255, 128, 64, 255)))
    h = dec.get_header()
    assert (h.format, h.width, h.height, h.bit_depth, h.channels) == ("jpeg", 64, 48, 8, 4)
    dec.close()


@pytest.mark.parametrize("w,h,color", [(32, 32, (100, 150, 200, 255)), (100, 50, (200, 100, 50, 255))])
def test_handles_square_images_and_aspect_ratios(w, h, color):
    hd = JpegBufferDecoder(create_test_jpeg(w, h, color)).get_header()
    assert (hd.width, hd.height) == (w, h)


# ---- Scanline Iteration (:56-118) ----
def test_yields_correct_number_of_scanlines():
    data = create_test_jpeg(16, 12, (255, 0, 0, 255))
    rows = _rows(JpegBufferDecoder(data))
    assert len(rows) == 12 and all(r.size == 16 * 4 for r in rows)
    assert np.array_equal(np.concatenate(rows), orc.jpeg_decode(data).reshape(-1))


def test_scanlines_contain_valid_pixel_data():
    rows = _rows(JpegBufferDecoder(create_test_jpeg(8, 8, (100, 150, 200, 255))))
    assert rows and all(r.any() for r in rows)
    px = np.stack(rows).reshape(8, 8, 4)
    assert np.abs(px[..., :3].astype(int) - np.array([100, 150, 200])).max() <= 4 and (px[..., 3] == 255).all()


def test_can_call_scanlines_twice():
    dec = JpegBufferDecoder(create_test_jpeg(8, 8, (255, 255, 0, 255)))
    assert len(_rows(dec)) == 8 and len(_rows(dec)) == 8
    dec.close()


# ---- Memory Management (:120-155) ----
def test_close_releases_and_new_decoder_after_closing():
    data = create_test_jpeg(16, 16, (128, 128, 128, 255))
    d1 = JpegBufferDecoder(data)
    h1 = d1.get_header()
    _rows(d1)
    d1.close()
    d1.close()   # multiple close calls do not error
    d2 = JpegBufferDecoder(data)
    h2 = d2.get_header()
    assert (h1.width, h1.height) == (h2.width, h2.height)
    d2.close()


# ---- File-based Input (:157-186) ----
def test_decodes_jpeg_from_file_path(tmp_path):
    path = os.path.join(tmp_path, "test.jpg")
    with open(path, "wb") as f:
        f.write(create_test_jpeg(24, 16, (50, 100, 150, 255)))
    dec = JpegFileDecoder(path)
    h = dec.get_header()
    assert (h.width, h.height, h.format) == (24, 16, "jpeg")
    assert len(_rows(dec)) == 16
    dec.close()


# ---- Decoder Options (:188-224) ----
@pytest.mark.parametrize("options", [{"useImageDecoderAPI": False, "preferWasm": True}, {}, None])
def test_decoder_options(options):
    data = create_test_jpeg(8, 8, (1, 2, 3, 255))
    dec = JpegBufferDecoder(data, None, options) if options is not None else JpegBufferDecoder(data)
    assert dec.get_header().format == "jpeg"
    dec.close()


# ---- Color Variations (:226-279) ----
@pytest.mark.parametrize("color,channel", [((255, 0, 0, 255), 0), ((0, 255, 0, 255), 1), ((0, 0, 255, 255), 2)])
def test_decodes_primary_colours(color, channel):
    rows = _rows(JpegBufferDecoder(create_test_jpeg(8, 8, color)))
    assert any(r[channel::4].mean() > 100 for r in rows)
