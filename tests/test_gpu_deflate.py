"""DEFLATE on the device.  The bytes need not equal any host zlib's; the contract is that
the stream inflates (with the host's zlib, Adler-32 checked by it) to exactly the filtered rows that
STITCH_OUT_PNG_FILTERED produces, for every block shape: Huffman-coded, stored, single-symbol, short tail."""
import io
import zlib

import numpy as np
import pytest

import helpers as H
import image_stitch_b200 as stitch
from image_stitch_b200 import _native, synth
from image_stitch_b200.deflate_container import FINAL_BLOCK, ZLIB_HEADER, DeflateBandSplicer

pytestmark = pytest.mark.gpu


def _both(ctx, pix, w, h, bands, out_bd=8):
    """(filtered rows, zlib stream assembled from deflate bands) for one raw RGBA tile covering the canvas."""
    rows = []
    with _native.Job(ctx, w, h, out_bd, _native.OUT_PNG_FILTERED, False, bytes(8)) as job:
        job.add_tile(pix.reshape(-1), width=w, height=h, bit_depth=out_bd, color_type=6, filtered=0)
        got, _ = job.run_band(0, h)
        rows = bytes(np.asarray(got))
    sp = DeflateBandSplicer()
    z = ZLIB_HEADER
    with _native.Job(ctx, w, h, out_bd, _native.OUT_PNG_DEFLATE, False, bytes(8)) as job:
        job.add_tile(pix.reshape(-1), width=w, height=h, bit_depth=out_bd, color_type=6, filtered=0)
        for a, b in bands:
            data, _ = job.run_band(a, b)
            body = sp.push(data)
            # the device's CRC-32 of the band's compressed bytes is what zlib computes over them on the host
            assert sp.last_crc32 == zlib.crc32(body), (a, b, len(body))
            z += body
    z += sp.finish()
    assert sp.raw_bytes == len(rows)
    return rows, z


@pytest.mark.parametrize("w,h", [(1, 1), (7, 3), (64, 48), (333, 257), (2048, 129), (8191, 5)])
def test_smooth_images_round_trip(ctx, w, h):
    rng = np.random.default_rng(w * 7 + h)
    pix = synth.smooth_rgba8(rng, w, h, "uniform")
    rows, z = _both(ctx, pix, w, h, [(0, h)])
    assert zlib.decompress(z) == rows          # zlib verifies the Adler-32 trailer as well
    if w * h >= 64 * 48:
        assert len(z) < len(rows)              # residuals compress


def test_noise_falls_back_to_stored_blocks(ctx):
    rng = np.random.default_rng(1)
    w, h = 512, 200
    pix = rng.integers(0, 256, size=(h, w * 4), dtype=np.uint8)
    rows, z = _both(ctx, pix, w, h, [(0, h)])
    assert zlib.decompress(z) == rows
    assert len(z) <= len(rows) + (len(rows) // 32768 + 1) * 10 + 16


def test_constant_image_single_symbol_blocks(ctx):
    w, h = 1024, 64
    pix = np.zeros((h, w * 4), np.uint8)
    rows, z = _both(ctx, pix, w, h, [(0, h)])
    assert zlib.decompress(z) == rows
    assert len(z) < len(rows) // 6             # one bit per byte plus headers


def test_bands_concatenate(ctx):
    rng = np.random.default_rng(9)
    w, h = 700, 301
    pix = synth.smooth_rgba8(rng, w, h, "opaque")
    rows, z = _both(ctx, pix, w, h, [(0, 10), (10, 11), (11, 200), (200, 301)])
    assert zlib.decompress(z) == rows


def test_sixteen_bit_output(ctx):
    rng = np.random.default_rng(16)
    w, h = 300, 90
    pix = rng.integers(0, 256, size=(h, w * 8), dtype=np.uint8)
    pix[:, 1::2] = 0                            # smooth-ish high bytes only
    rows, z = _both(ctx, pix, w, h, [(0, h)], out_bd=16)
    assert zlib.decompress(z) == rows


def test_api_device_deflate_png_decodes_to_same_pixels():
    PIL = pytest.importorskip("PIL.Image")
    ins = [H.suite_bytes("basn2c08"), H.suite_bytes("basn6a08"), H.suite_bytes("basn0g08"), H.suite_bytes("basn6a08")]
    opts = {"inputs": ins, "layout": {"columns": 2}}
    ref = stitch.concat_to_buffer(opts)
    dev = stitch.concat_to_buffer(dict(opts, device_deflate=True))
    assert dev[:8] == ref[:8]
    a = np.asarray(PIL.open(io.BytesIO(ref)).convert("RGBA"))
    b = np.asarray(PIL.open(io.BytesIO(dev)).convert("RGBA"))
    assert a.shape == b.shape and np.array_equal(a, b)
    # and the filtered rows inside the IDAT stream are the same bytes
    from image_stitch_b200.png_container import read_png
    assert read_png(ref)[1] == read_png(dev)[1]


def test_c2_band_ratio_and_rate(ctx, capsys):
    """One tile row of C2 (1024 rows x 16384 px): ratio against host zlib level 6 and device time."""
    rng = np.random.default_rng(2)
    tile = 1024
    raws = [synth.smooth_rgba8(rng, tile, tile, "opaque") for _ in range(4)]
    with _native.Job(ctx, tile * 16, tile, 8, _native.OUT_PNG_DEFLATE, False, bytes(4)) as job:
        for c in range(16):
            job.add_tile(raws[c % 4].reshape(-1), width=tile, height=tile, bit_depth=8, color_type=6, filtered=0,
                         dst_x=c * tile, image_idx=c)
        job.run_band(0, tile)
        data, tm = job.run_band(0, tile)
    sp = DeflateBandSplicer()
    body = sp.push(data)
    assert sp.last_crc32 == zlib.crc32(body)   # ~35 MB: several hundred CTAs of the CRC kernel and its tail
    z = ZLIB_HEADER + body + sp.finish()
    rows = zlib.decompress(z)
    assert len(rows) == tile * (tile * 16 * 4 + 1)
    with capsys.disabled():
        print(f"\n[deflate] {len(rows) / 1e6:.1f} MB -> {len(z) / 1e6:.1f} MB (ratio {len(rows) / len(z):.2f}); "
              f"composite+filter+deflate {tm['compose_ms']:.3f} ms on the device")


def test_chunked_download_of_deflate_bands(ctx, monkeypatch):
    """Host destination: the band is rendered, compressed and downloaded chunk by chunk (two device buffers
    taking turns); the chunks' blocks and Adler-32 values must join into one valid stream."""
    rng = np.random.default_rng(31)
    w, h = 600, 555
    pix = synth.smooth_rgba8(rng, w, h, "uniform")
    monkeypatch.setenv("STITCH_CHUNK_BYTES", "100000")   # ~41 rows per chunk -> 14 chunks
    rows, z = _both(ctx, pix, w, h, [(0, 300), (300, h)])
    assert zlib.decompress(z) == rows
