"""Byte surgery on the JPEG fixtures: files that differ in their marker segments only (used by the oracle test and the
GPU test of the JPEG input path)."""
import os
import struct

HERE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "jpeg_decode")


def fixture(name):
    with open(os.path.join(HERE, name + ".jpg"), "rb") as f:
        return f.read()


def segments(data):
    """[(marker, payload bytes)] up to SOS, and everything from the SOS marker on."""
    out, o = [], 2
    while True:
        assert data[o] == 0xFF
        m = data[o + 1]
        if m == 0xDA:
            return out, data[o:]
        n = struct.unpack(">H", data[o + 2:o + 4])[0]
        out.append((m, data[o + 4:o + 2 + n]))
        o += 2 + n


def build(segs, tail, fill=0):
    out = bytearray(b"\xff\xd8")
    for m, payload in segs:
        out += b"\xff" * (1 + fill) + bytes([m]) + struct.pack(">H", len(payload) + 2) + payload
    return bytes(out + tail)


ADOBE = lambda transform: (0xEE, b"Adobe" + struct.pack(">HHH", 100, 0, 0) + bytes([transform]))


def variants():
    """name -> (file, same pixels as fixture name or None, expected error substring or None)"""
    base = fixture("c420_q80")
    segs, tail = segments(base)
    no_jfif = [s for s in segs if s[0] != 0xE0]
    exif_with_thumbnail = (0xE1, b"Exif\0\0" + fixture("c444_1x1"))          # a whole JPEG inside APP1
    comment = (0xFE, b"\xff\xd8\xff\xc0 not a marker \xff\xda")
    split_tables = []                                                          # every DHT / DQT table in a segment of its own
    for m, p in segs:
        if m == 0xC4:
            q = 0
            while q < len(p):
                n = 17 + sum(p[q + 1:q + 17])
                split_tables.append((m, p[q:q + n]))
                q += n
        elif m == 0xDB:
            for q in range(0, len(p), 65):
                split_tables.append((m, p[q:q + 65]))
        else:
            split_tables.append((m, p))
    tables_after_sof = [s for s in segs if s[0] == 0xC0] and ([s for s in segs if s[0] not in (0xC4, 0xDB)] + [s for s in segs if s[0] in (0xC4, 0xDB)])
    rgb_ids = []
    for m, p in no_jfif:
        if m == 0xC0:
            p = bytearray(p)
            p[6], p[9], p[12] = ord("R"), ord("G"), ord("B")
            p = bytes(p)
        rgb_ids.append((m, p))
    rgb_tail = bytearray(tail)
    rgb_tail[5], rgb_tail[7], rgb_tail[9] = ord("R"), ord("G"), ord("B")      # component selectors of the scan header
    return {
        "fill_bytes_before_markers": (build(segs, tail, fill=3), "c420_q80", None),
        "exif_with_embedded_jpeg": (build([exif_with_thumbnail] + segs, tail), "c420_q80", None),
        "comment_with_marker_bytes": (build(segs + [comment], tail), "c420_q80", None),
        "one_table_per_segment": (build(split_tables, tail), "c420_q80", None),
        "tables_after_the_frame_header": (build(tables_after_sof, tail), "c420_q80", None),
        "adobe_ycck_flag_1_without_jfif": (build(no_jfif + [ADOBE(1)], tail), "c420_q80", None),
        "no_jfif_no_adobe": (build(no_jfif, tail), "c420_q80", None),
        "trailing_garbage_after_eoi": (base + b"\x00\xff\xd0garbage", "c420_q80", None),
        "adobe_transform_0_is_rgb": (build(no_jfif + [ADOBE(0)], tail), None, "must be YCbCr"),
        "component_ids_rgb_without_jfif": (build(rgb_ids, bytes(rgb_tail)), None, "must be YCbCr"),
        "restart_markers_without_dri": (build([s for s in segments(fixture("c420_rst3"))[0] if s[0] != 0xDD], segments(fixture("c420_rst3"))[1]),
                                        None, "restart marker|ends early"),
        "twelve_bit_precision": (build([(m, (b"\x0c" + p[1:]) if m == 0xC0 else p) for m, p in segs], tail), None, "precision must be 8"),
        "four_components": (build([(m, (p[:5] + b"\x04" + p[6:] + b"\x04\x11\x00") if m == 0xC0 else p) for m, p in segs], tail), None, "1 (gray) or 3"),
    }
