"""Host logic: layout, clipping, z-order, background colours, container I/O.
Restates tests/unit/positioned-layout.test.ts, pixel-ops.test.ts:230-390,
png-parser/png-writer tests of the reference."""
import zlib

import pytest

import helpers as H
from image_stitch_b200 import engine
from image_stitch_b200.background import get_transparent_color, parse_background_color
from image_stitch_b200.layout import (build_scanline_index, calculate_canvas_size, calculate_layout,
                                      clip_images_to_canvas, determine_common_format)
from image_stitch_b200.png_container import (build_png, create_chunk, create_ihdr, create_iend, parse_png_chunks,
                                             parse_png_header, read_png)
from image_stitch_b200.types import ConcatOptions, Layout, PngHeader

HDR = lambda w, h, bd=8, ct=6: PngHeader(w, h, bd, ct)


def test_grid_columns_and_rows():
    hs = [HDR(32, 32)] * 5
    g = calculate_layout(hs, Layout(columns=2))
    assert g.grid == [[0, 1], [2, 3], [4, -1]] and (g.total_width, g.total_height) == (64, 96)
    g = calculate_layout(hs, Layout(rows=2))
    assert g.grid == [[0, 2, 4], [1, 3, -1]] and (g.total_width, g.total_height) == (96, 64)


def test_grid_variable_sizes_and_width_wrap():
    hs = [HDR(10, 5), HDR(20, 8), HDR(7, 3)]
    g = calculate_layout(hs, Layout(columns=3))
    assert g.col_widths == [[10, 20, 7]] and g.row_heights == [8] and g.total_width == 37
    g = calculate_layout([HDR(32, 32)] * 4, Layout(width=100))
    assert g.grid == [[0, 1, 2], [3]] and (g.total_width, g.total_height) == (96, 64)
    g = calculate_layout([HDR(32, 32)] * 6, Layout(width=64, height=64))
    assert g.grid == [[0, 1], [2, 3]]


def test_common_format():
    assert determine_common_format([HDR(1, 1, 8, 0), HDR(1, 1, 8, 2)]) == (8, 6)
    assert determine_common_format([HDR(1, 1, 8, 6), HDR(1, 1, 16, 0)]) == (16, 6)


def test_canvas_size():
    imgs = [dict(x=0, y=0, width=100, height=100), dict(x=150, y=50, width=100, height=100)]
    assert calculate_canvas_size(imgs) == (250, 150)
    assert calculate_canvas_size(imgs, 500, 400) == (500, 400)
    assert calculate_canvas_size(imgs, 300, None) == (300, 150)
    assert calculate_canvas_size([dict(x=-50, y=-50, width=10, height=10)]) == (1, 1)


def test_clipping_offsets_and_messages():
    msgs = []
    clipped, pos = clip_images_to_canvas([dict(x=-10, y=-20), dict(x=90, y=90), dict(x=200, y=0), dict(x=5, y=5)],
                                         [HDR(50, 50)] * 4, 100, 100, msgs.append)
    assert [c.image_idx for c in clipped] == [0, 1, 2]
    c0 = clipped[0]
    assert (c0.clipped_x, c0.clipped_y, c0.clipped_width, c0.clipped_height) == (0, 0, 40, 30)
    assert (c0.source_offset_x, c0.source_offset_y) == (10, 20)
    assert (clipped[1].clipped_width, clipped[1].clipped_height, clipped[1].source_offset_x) == (10, 10, 0)
    assert clipped[2].fully_clipped and [p.image_idx for p in pos] == [0, 1, 3]
    assert "left by 10px, top by 20px" in msgs[0] and "completely outside canvas bounds" in msgs[2]
    assert [p.z_index for p in pos] == [0, 1, 3]  # zIndex defaults to the input index


def test_scanline_index_z_order():
    _, pos = clip_images_to_canvas([dict(x=0, y=0, z_index=5), dict(x=0, y=0, z_index=1), dict(x=0, y=0, z_index=5)],
                                   [HDR(4, 2)] * 3, 4, 2, lambda m: None)
    idx = build_scanline_index(pos, 2)
    assert [s.image_idx for s in idx[0]] == [1, 0, 2]


def test_background_parsing():
    assert parse_background_color(None) == (0, 0, 0, 0) and parse_background_color("transparent") == (0, 0, 0, 0)
    assert parse_background_color("#F00") == (255, 0, 0, 255) and parse_background_color("#FF000080") == (255, 0, 0, 128)
    assert parse_background_color("White") == (255, 255, 255, 255) and parse_background_color([1, 2, 3]) == (1, 2, 3, 255)
    for bad in ("#12345", "#GG0000", "rgb(1,2,3)", [1, 2], [1, 2, 300]):
        with pytest.raises(ValueError):
            parse_background_color(bad)
    assert get_transparent_color(6, 16, "white") == b"\xff" * 8
    assert get_transparent_color(6, 8) == b"\0" * 4 and get_transparent_color(6, 16) == b"\0" * 8
    assert get_transparent_color(6, 16, [1, 2, 3, 4]) == bytes([1, 1, 2, 2, 3, 3, 4, 4])


def test_container_roundtrip_and_errors():
    hdr = PngHeader(2, 1, 8, 6, 0, 0, 0)
    png = build_png([create_ihdr(hdr), create_chunk("IDAT", zlib.compress(b"\0" + bytes(8))), create_iend()])
    assert parse_png_header(png) == hdr
    h2, stream = read_png(png)
    assert h2 == hdr and stream == b"\0" + bytes(8)
    assert [c.type for c in parse_png_chunks(png)] == ["IHDR", "IDAT", "IEND"]
    bad = bytearray(png)
    bad[30] ^= 1
    with pytest.raises(ValueError, match="CRC mismatch"):
        parse_png_chunks(bytes(bad))
    with pytest.raises(ValueError, match="Invalid PNG signature"):
        parse_png_chunks(b"notapng!" + png[8:])


def test_plan_grid_and_positioned():
    ins = [H.suite_bytes("basn2c08"), H.suite_bytes("basn0g08"), H.suite_bytes("basn6a08")]
    opts = ConcatOptions.coerce({"inputs": ins, "layout": {"columns": 3}})
    plan, tiles = H.oracle_plan(opts)
    assert (plan.mode, plan.canvas_w, plan.canvas_h, plan.alpha_blend) == ("grid", 96, 32, False)
    assert [(t["dst_x"], t["dst_y"]) for t in tiles] == [(0, 0), (32, 0), (64, 0)]
    pos = [{"x": -8, "y": 4, "source": ins[0]}, {"x": 20, "y": 0, "source": ins[2], "zIndex": -1}]
    plan, tiles = H.oracle_plan({"inputs": pos, "layout": {"width": 40, "height": 30}})
    assert (plan.mode, plan.alpha_blend, plan.canvas_w, plan.canvas_h) == ("positioned", True, 40, 30)
    assert (tiles[0]["src_off_x"], tiles[0]["vis_w"], tiles[0]["vis_h"]) == (8, 24, 26)
    assert (tiles[1]["vis_w"], tiles[1]["vis_h"], tiles[1]["z_index"]) == (20, 30, -1)
    with pytest.raises(ValueError, match="Cannot mix positioned"):
        H.oracle_plan({"inputs": [pos[0], ins[1]], "layout": {}})
    with pytest.raises(ValueError, match="Grid mode requires layout"):
        H.oracle_plan({"inputs": ins, "layout": {}})
    with pytest.raises(ValueError, match="At least one input"):
        H.oracle_plan({"inputs": [], "layout": {"columns": 1}})


def test_split_rows_alignment():
    parts = engine.split_rows(0, 1000, 8)
    assert parts[0][0] == 0 and parts[-1][1] == 1000
    assert all(a % 8 == 0 for a, _ in parts) and all(p[1] == q[0] for p, q in zip(parts, parts[1:]))
    assert engine.split_rows(0, 5, 8) == [(0, 5)]


def test_device_codec_options_and_deflated_source():
    """Additive options and the decoder method that hands over the zlib stream itself."""
    import zlib
    from image_stitch_b200 import engine
    from image_stitch_b200.decoders.png_decoder import PngBufferDecoder
    from image_stitch_b200.types import ConcatOptions
    o = ConcatOptions.coerce({"inputs": [b""], "layout": {"columns": 1}, "deviceInflate": True, "deviceDeflate": True})
    assert o.device_inflate is True and o.device_deflate is True
    assert ConcatOptions.coerce({"inputs": [b""], "layout": {"columns": 1}}).device_inflate is False
    with pytest.raises(TypeError):
        ConcatOptions.coerce({"inputs": [], "device_unflate": True})
    data = H.suite_bytes("basi6a08")
    dec = PngBufferDecoder(data)
    hdr, z = dec.deflated()
    hdr2, inflated = dec.inflated()
    assert (hdr.width, hdr.height, hdr.interlace_method) == (32, 32, 1) and hdr == hdr2
    assert zlib.decompress(z) == inflated                      # the stream the device inflates is the IDAT payload
    src = engine.read_tile_source(dec, hdr, 0, device_inflate=True)
    assert (src.filtered, src.interlace, bytes(src.data)) == (2, 1, z)
    src = engine.read_tile_source(dec, hdr, 0)
    assert (src.filtered, bytes(src.data)) == (1, inflated)

    class Foreign:  # a decoder without deflated(): the option falls back to scanlines()
        def scanlines(self):
            return iter([bytes(hdr.width * 4)] * hdr.height)
    src = engine.read_tile_source(Foreign(), hdr, 0, device_inflate=True)
    assert src.filtered == 0 and len(src.data) == hdr.width * 4 * hdr.height


def test_expand_palette_option_and_source():
    from image_stitch_b200 import engine
    from image_stitch_b200.decoders.png_decoder import PngBufferDecoder
    from image_stitch_b200.layout import image_header_to_png_header
    o = ConcatOptions.coerce({"inputs": [b""], "layout": {"columns": 1}, "expandPalette": True})
    assert o.expand_palette is True
    assert ConcatOptions.coerce({"inputs": [b""], "layout": {"columns": 1}}).expand_palette is False
    png = H.suite_bytes("tbbn3p08")
    dec = PngBufferDecoder(png)
    hdr = image_header_to_png_header(dec.get_header())
    plte, trns = dec.palette()
    assert len(plte) % 3 == 0 and 0 < len(trns) <= len(plte) // 3
    src = engine.read_tile_source(dec, hdr, 0, expand_palette=True)
    assert src.header.color_type == 3 and src.plte == plte and src.trns == trns and src.filtered == 1
    src = engine.read_tile_source(dec, hdr, 0, device_inflate=True, expand_palette=True)
    assert src.filtered == 2 and src.plte == plte
    assert engine.read_tile_source(dec, hdr, 0).plte is None            # default: the reference's error follows
    dec2 = PngBufferDecoder(H.suite_bytes("basn2c08"))
    assert engine.read_tile_source(dec2, image_header_to_png_header(dec2.get_header()), 0, expand_palette=True).plte is None
