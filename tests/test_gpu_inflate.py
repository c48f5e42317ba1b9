"""inflate on the device.  The checker is the host's zlib (the algorithm behind both of the
reference's back ends, DecompressionStream and pako: src/streaming-inflate.ts:23-76): every stream must inflate to
exactly zlib.decompress's bytes, for every block type and table shape, and every corruption zlib rejects must
be rejected with zlib's message."""
import zlib

import numpy as np
import pytest

import helpers as H
import image_stitch_b200 as stitch
from image_stitch_b200 import _native, synth
from image_stitch_b200.png_container import extract_idat, parse_png_chunks, parse_png_header, read_png

pytestmark = pytest.mark.gpu


def _z(data, level=6, strategy=zlib.Z_DEFAULT_STRATEGY, wbits=15, memlevel=8):
    c = zlib.compressobj(level, zlib.DEFLATED, wbits, memlevel, strategy)
    return c.compress(bytes(data)) + c.flush()


def _check(ctx, raw, z):
    got = ctx.inflate(z, len(raw))
    assert bytes(got) == bytes(raw)


def _corpus(rng):
    smooth = synth.encode_filtered(synth.smooth_rgba8(rng, 256, 200, "uniform"), 4)
    text = (b"the quick brown fox jumps over the lazy dog. " * 3000)[: 100_000]
    return {
        "empty": b"",
        "one": b"\x07",
        "short": b"abc",
        "zeros": bytes(200_000),                                   # distance 1, length 258 runs
        "period3": b"abc" * 50_000,                                # overlap with distance < length
        "period40k": bytes(rng.integers(0, 256, 30_000, dtype=np.uint8)) * 4,   # distances near 32 KiB
        "noise": bytes(rng.integers(0, 256, 150_000, dtype=np.uint8)),
        "low_entropy": bytes(rng.integers(0, 4, 120_000, dtype=np.uint8)),     # 2-3 bit codes: three literals per step
        "skewed": bytes(np.minimum(rng.geometric(0.02, 200_000), 255).astype(np.uint8)),  # long codes for rare bytes
        "text": text,
        "filtered_rows": bytes(smooth),
    }


@pytest.mark.parametrize("level", [0, 1, 6, 9])
def test_levels_match_zlib(ctx, level):
    rng = np.random.default_rng(level)
    for name, raw in _corpus(rng).items():
        _check(ctx, raw, _z(raw, level))


@pytest.mark.parametrize("strategy", [zlib.Z_FIXED, zlib.Z_HUFFMAN_ONLY, zlib.Z_RLE, zlib.Z_FILTERED])
def test_strategies_match_zlib(ctx, strategy):
    rng = np.random.default_rng(11)
    for name, raw in _corpus(rng).items():
        _check(ctx, raw, _z(raw, 6, strategy))


def _lz_mix(rng, n, alphabet, p_copy, max_dist, max_len):
    """Bytes made like an LZ77 stream: literals from a small alphabet, copies at random distances and lengths (short
    distances give overlapping copies and chains of copies of copies inside one window of the device's walk)."""
    out = bytearray(rng.integers(0, alphabet, 64, dtype=np.uint8).tobytes())
    while len(out) < n:
        if rng.random() < p_copy:
            d = int(rng.integers(1, min(max_dist, len(out)) + 1))
            ln = int(rng.integers(3, max_len + 1))
            for _ in range(ln):
                out.append(out[-d])
        else:
            out += rng.integers(0, alphabet, int(rng.integers(1, 12)), dtype=np.uint8).tobytes()
    return bytes(out[:n])


@pytest.mark.parametrize("seed", range(12))
def test_random_lz_mixes_match_zlib(ctx, seed):
    """Copies of copies at every distance scale: inside a lane's own bytes, across lanes of one window, across
    windows, near the 32 KiB limit; matches from 3 to 258 bytes; alphabets from 2 symbols (three literals per table
    entry) to 256 (long codes)."""
    rng = np.random.default_rng(1000 + seed)
    alphabet = [2, 4, 16, 64, 256, 256][seed % 6]
    p_copy = [0.05, 0.3, 0.6, 0.9][seed % 4]
    max_dist = [4, 40, 300, 3000, 32768, 32768][(seed // 2) % 6]
    max_len = [4, 12, 64, 258][(seed // 3) % 4]
    raw = _lz_mix(rng, 60_000 + 7919 * seed, alphabet, p_copy, max_dist, max_len)
    for level in (1, 6, 9):
        _check(ctx, raw, _z(raw, level))
    _check(ctx, raw, _z(raw, 6, zlib.Z_FIXED))
    _check(ctx, raw, _z(raw, 9, zlib.Z_DEFAULT_STRATEGY, 15, 1))   # memLevel 1: a new block every ~128 symbols


def test_small_window_and_memlevel(ctx):
    rng = np.random.default_rng(5)
    raw = _corpus(rng)["filtered_rows"]
    for wbits, mem in [(9, 1), (12, 4), (15, 9)]:
        _check(ctx, raw, _z(raw, 6, wbits=wbits, memlevel=mem))    # memlevel 1: a block every ~128 symbols


def test_sync_flushed_stream_with_many_blocks(ctx):
    """What the reference's own writer produces: Z_SYNC_FLUSH every batch (empty stored blocks inside the stream)."""
    rng = np.random.default_rng(6)
    c = zlib.compressobj(6)
    raw = b""
    z = b""
    for i in range(40):
        part = bytes(rng.integers(0, 32, int(rng.integers(1, 9000)), dtype=np.uint8))
        raw += part
        z += c.compress(part) + c.flush(zlib.Z_SYNC_FLUSH if i % 3 else zlib.Z_FULL_FLUSH)
    z += c.flush()
    _check(ctx, raw, z)


def test_large_stream(ctx):
    rng = np.random.default_rng(7)
    raw = bytes(synth.encode_filtered(synth.smooth_rgba8(rng, 1024, 1024, "opaque"), 4))
    z = _z(raw)
    assert 1.2 < len(raw) / len(z)
    _check(ctx, raw, z)


def test_capacity_smaller_than_stream(ctx):
    raw = bytes(range(256)) * 100
    with pytest.raises(_native.StitchError) as e:
        ctx.inflate(_z(raw), 1000)
    assert e.value.code == _native.E_CAPACITY


def _bits(*fields):
    """LSB-first bit packer: (value, nbits) pairs -> bytes."""
    acc = n = 0
    for v, k in fields:
        acc |= v << n
        n += k
    return acc.to_bytes((n + 7) // 8, "little")


_CASES = [
    ("incorrect header check", b"\x78\x9d" + _z(b"abc")[2:]),
    ("unknown compression method", b"\x77\x09" + _z(b"abc")[2:]),      # 0x7709 % 31 == 0, CM = 7
    ("invalid window size", b"\x88\x1c" + _z(b"abc")[2:]),             # CINFO = 8
    ("invalid block type", b"\x78\x01" + _bits((1, 1), (3, 2)) + bytes(8)),
    ("invalid stored block lengths", b"\x78\x01" + _bits((1, 1), (0, 2)) + b"\x05\x00\x00\x00" + bytes(9)),
    ("too many length or distance symbols", b"\x78\x01" + _bits((1, 1), (2, 2), (30, 5), (0, 5), (0, 4)) + bytes(16)),
    ("invalid code lengths set", b"\x78\x01" + _bits((1, 1), (2, 2), (0, 5), (0, 5), (15, 4), *[(1, 3)] * 19) + bytes(16)),
    ("invalid distance too far back", b"\x78\x01" + _bits((1, 1), (1, 2), (0b1000000, 7), (0, 5)) + bytes(8)),   # length 3, distance 1 at output position 0
    ("invalid literal/length code", b"\x78\x01" + _bits((1, 1), (1, 2), (0b01100011, 8)) + bytes(8)),   # fixed code 11000110 = symbol 286
    ("invalid distance code", b"\x78\x01" + _bits((1, 1), (1, 2), (0x0c, 8), (0b1000000, 7), (0b01111, 5)) + bytes(8)),  # literal 0, length 3, distance code 30
]


@pytest.mark.parametrize("message,stream", _CASES, ids=[c[0] for c in _CASES])
def test_corrupt_streams_fail_like_zlib(ctx, message, stream):
    with pytest.raises(zlib.error) as ze:
        zlib.decompress(stream)
    assert message in str(ze.value)
    with pytest.raises(_native.StitchError) as e:
        ctx.inflate(stream, 1000)
    assert e.value.code == _native.E_INFLATE
    assert str(e.value) == f"Inflate error: {message}"


def test_wrong_adler_and_truncation(ctx):
    raw = bytes(range(256)) * 40
    z = bytearray(_z(raw))
    z[-1] ^= 1
    with pytest.raises(_native.StitchError) as e:
        ctx.inflate(bytes(z), len(raw))
    assert str(e.value) == "Inflate error: incorrect data check"
    for cut in (1, 4, 5, len(z) // 2, len(z) - 3):
        with pytest.raises(_native.StitchError) as e:
            ctx.inflate(bytes(_z(raw)[:-cut]), len(raw))
        assert e.value.code == _native.E_INFLATE, cut
    # a flipped bit in the middle either breaks the code stream or the check value: never silently accepted
    rng = np.random.default_rng(3)
    base = _z(bytes(rng.integers(0, 64, 50_000, dtype=np.uint8)))
    for _ in range(40):
        z = bytearray(base)
        z[int(rng.integers(2, len(z) - 4))] ^= 1 << int(rng.integers(0, 8))
        with pytest.raises(_native.StitchError):
            ctx.inflate(bytes(z), 50_000)


# ---- through the job: tiles added as zlib streams -------------------------------------------------------------

def _suite_names():
    return ["basn0g01", "basn0g02", "basn0g04", "basn0g08", "basn0g16", "basn2c08", "basn2c16", "basn4a08", "basn4a16",
            "basn6a08", "basn6a16", "basi0g01", "basi0g08", "basi2c08", "basi2c16", "basi4a16", "basi6a08", "basi6a16"]


@pytest.mark.parametrize("name", _suite_names())
def test_pngsuite_tiles_from_idat(ctx, name):
    """decode from the zlib stream == decode from the host-inflated stream (which the oracle tests pin)."""
    data = H.suite_bytes(name)
    hdr, inflated = read_png(data)
    z = extract_idat(parse_png_chunks(data))
    want = ctx.decode_image(hdr.width, hdr.height, hdr.bit_depth, hdr.color_type, hdr.interlace_method, inflated, filtered=1)
    got = ctx.decode_image(hdr.width, hdr.height, hdr.bit_depth, hdr.color_type, hdr.interlace_method, z, filtered=2)
    assert bytes(got) == bytes(want)


def test_job_with_many_zlib_tiles_equals_inflated_tiles(ctx):
    rng = np.random.default_rng(9)
    T, G = 96, 5
    tiles = [synth.encode_filtered(synth.smooth_rgba8(rng, T, T, "uniform"), 4) for _ in range(G * G)]
    outs = []
    for mode in (1, 2):
        with _native.Job(ctx, T * G, T * G, 8, _native.OUT_PNG_FILTERED, False, bytes(8)) as job:
            for i, st in enumerate(tiles):
                data = bytes(st) if mode == 1 else _z(st, [1, 6, 9][i % 3])
                job.add_tile(data, width=T, height=T, bit_depth=8, color_type=6, filtered=mode,
                             dst_x=(i % G) * T, dst_y=(i // G) * T, image_idx=i)
            rows = []
            for a in range(0, T * G, 64):    # bands: tiles are inflated when the first band that needs them runs
                got, _ = job.run_band(a, min(T * G, a + 64))
                rows.append(bytes(np.asarray(got)))
            outs.append(b"".join(rows))
    assert outs[0] == outs[1]


def test_c2_size_job_from_zlib_streams(ctx):
    """BASELINE config 2 at full size with every tile entering as its zlib stream: the 16384 x 16384 band is
    byte-identical to the one made from host-inflated streams, and inflating again gives the same bytes."""
    import hashlib
    grid, tile, canvas = 16, 1024, 16384
    rng = np.random.default_rng(2)
    streams = [bytes(synth.encode_filtered(synth.smooth_rgba8(rng, tile, tile, "opaque"), 4)) for _ in range(4)]
    zs = [_z(st, lvl) for st, lvl in zip(streams, (6, 1, 9, 6))]
    which = lambda r, c: (r * 5 + c * 3) % 4
    out = np.empty(canvas * (canvas * 4 + 1), np.uint8)
    digests = []
    for mode in (1, 2):
        with _native.Job(ctx, canvas, canvas, 8, _native.OUT_PNG_FILTERED, False, bytes(4)) as job:
            for r in range(grid):
                for c in range(grid):
                    k = which(r, c)
                    job.add_tile(streams[k] if mode == 1 else zs[k], width=tile, height=tile, bit_depth=8, color_type=6,
                                 filtered=mode, dst_x=c * tile, dst_y=r * tile, image_idx=r * grid + c)
            got, _ = job.run_band(0, canvas, out=out)
            digests.append(hashlib.sha256(memoryview(got)).hexdigest())
            if mode == 2:
                job.inflate(force=True)          # every stream again, in place
                job.invalidate()
                got, _ = job.run_band(0, canvas, out=out)
                digests.append(hashlib.sha256(memoryview(got)).hexdigest())
    assert digests[0] == digests[1] == digests[2]


def test_job_reports_inflate_and_row_count_errors(ctx):
    rng = np.random.default_rng(10)
    T = 64
    st = bytes(synth.encode_filtered(synth.smooth_rgba8(rng, T, T, "uniform"), 4))
    bad = bytearray(_z(st))
    bad[len(bad) // 2] ^= 0x55
    with _native.Job(ctx, T, T, 8, _native.OUT_PNG_FILTERED, False, bytes(8)) as job:
        job.add_tile(bytes(bad), width=T, height=T, bit_depth=8, color_type=6, filtered=2)
        with pytest.raises(_native.StitchError) as e:
            job.run_band(0, T)
        assert e.value.code == _native.E_INFLATE and str(e.value).startswith("Inflate error: ")
    short = _z(st[: (T * 4 + 1) * 10 + 7])     # ten complete rows and a bit
    with _native.Job(ctx, T, T, 8, _native.OUT_PNG_FILTERED, False, bytes(8)) as job:
        job.add_tile(short, width=T, height=T, bit_depth=8, color_type=6, filtered=2)
        with pytest.raises(_native.StitchError) as e:
            job.run_band(0, T)
        assert e.value.code == _native.E_TRUNCATED
        assert str(e.value) == f"Expected {T} scanlines, decoded 10"     # png-decoder.ts:209-211


# ---- through the public API ---------------------------------------------------------------------------------------

@pytest.mark.parametrize("opts", [
    {"layout": {"columns": 3}},
    {"layout": {"width": 100}},
    {"layout": {"columns": 2}, "output_format": "jpeg", "jpeg_quality": 90},
    {"layout": {"columns": 3}, "device_deflate": True},
])
def test_concat_to_buffer_with_device_inflate(opts):
    names = ["basn2c08", "basn0g08", "basn6a08", "basi4a08", "basn2c16", "basi0g16", "basn0g01", "basi6a16"]
    ins = [H.suite_bytes(n) for n in names]
    want = stitch.concat_to_buffer(dict(opts, inputs=ins))
    got = stitch.concat_to_buffer(dict(opts, inputs=ins, device_inflate=True))
    if opts.get("output_format") == "jpeg":
        assert got == want
    else:
        assert read_png(got) == read_png(want)      # same header, same filtered rows (deflate bytes may differ)


def test_positioned_with_device_inflate_and_corrupt_input():
    pos = [{"x": 0, "y": 0, "source": H.suite_bytes("basn2c08")}, {"x": 10, "y": 6, "source": H.suite_bytes("basi6a08")}]
    want = stitch.concat_to_buffer({"inputs": pos, "layout": {}})
    got = stitch.concat_to_buffer({"inputs": pos, "layout": {}, "deviceInflate": True})
    assert read_png(got) == read_png(want)
    # corrupt the zlib stream inside the container (chunk CRC fixed up): the reference reports pako's message
    from image_stitch_b200.png_container import build_png, create_chunk
    chunks = parse_png_chunks(H.suite_bytes("basn6a08"))
    out = []
    for c in chunks:
        if c.type == "IDAT":
            d = bytearray(c.data)
            d[len(d) // 2] ^= 0xff
            c = create_chunk("IDAT", bytes(d))
        out.append(c)
    with pytest.raises(Exception) as e:
        stitch.concat_to_buffer({"inputs": [build_png(out)], "layout": {"columns": 1}, "device_inflate": True})
    assert "Failed to stitch images" in str(e.value) and ("Inflate error" in str(e.value) or "scanlines" in str(e.value))
