"""Row-band sharding across ranks on CPU: world_size 2 and 3, gloo backend.
Each rank renders only its band (with its 1-row halo and only the tiles that touch it); the
concatenation gathered on rank 0 must equal the whole canvas rendered at once."""
import json
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

import helpers as H
import oracle_py as orc
from image_stitch_b200 import distributed as D

HERE = os.path.dirname(os.path.abspath(__file__))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _run(case, world, tmp_path, extra_env=None, gpu=False):
    out = str(tmp_path / f"{case}_{world}.bin")
    port = _free_port()
    procs = []
    for r in range(world):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE=str(world), LOCAL_RANK=str(r), MASTER_ADDR="127.0.0.1",
                   MASTER_PORT=str(port), **(extra_env or {}))
        if not gpu:
            env["CUDA_VISIBLE_DEVICES"] = ""
        procs.append(subprocess.Popen([sys.executable, os.path.join(HERE, "dist_worker.py"), case, out], env=env,
                                      stdout=subprocess.PIPE, stderr=subprocess.STDOUT))
    for p in procs:
        o, _ = p.communicate(timeout=240)
        assert p.returncode == 0, o.decode()[-2000:]
    with open(out, "rb") as f:
        data = f.read()
    with open(out + ".json") as f:
        meta = json.load(f)
    return data, meta


def _inputs():
    return [H.suite_bytes(n) for n in ("basn2c08", "basn6a08", "basn0g08", "basn4a08", "basi6a08", "basn2c08")]


@pytest.mark.parametrize("world", [2, 3])
def test_grid_png_bands(world, tmp_path):
    data, meta = _run("grid_png", world, tmp_path)
    _, want = H.oracle_stitch_rows({"inputs": _inputs(), "layout": {"columns": 2}})
    assert data == want.tobytes()
    assert meta["rows"][0][0] == 0 and meta["rows"][-1][1] == meta["h"]


def test_positioned_png_bands(tmp_path):
    data, _ = _run("positioned_png", 2, tmp_path)
    opts = {"inputs": [{"x": 7 * i, "y": 13 * i, "source": s} for i, s in enumerate(_inputs())], "layout": {}}
    _, want = H.oracle_stitch_rows(opts)
    assert data == want.tobytes()


def test_jpeg_bands(tmp_path):
    data, meta = _run("jpeg", 2, tmp_path)
    opts = {"inputs": _inputs(), "layout": {"columns": 2}}
    _, raw = H.oracle_stitch_rows(opts, out_kind=1)
    w, h = meta["w"], meta["h"]
    assert data == orc.jpeg_encode_rows(np.frombuffer(raw, np.uint8).reshape(h, w * 4), w, h, 0, h, 90)


def test_rank_rows_properties():
    for h in (1, 7, 8, 9, 64, 100, 4096, 32768):
        for world in (1, 2, 3, 4, 8):
            rows = D.rank_rows(h, world)
            assert len(rows) == world
            assert rows[0][0] == 0 and max(b for _, b in rows) == h
            prev = 0
            for a, b in rows:
                assert a <= b and (a == prev or a == b == h)
                if b > a:
                    assert a % 8 == 0 and (b % 8 == 0 or b == h)
                    prev = b


@pytest.mark.parametrize("case", ["grid_png", "positioned_png"])
def test_whole_file_through_rank0(case, tmp_path):
    """concat_to_buffer_distributed: every rank deflates its own band (sync-flushed raw deflate), rank 0 frames the
    pieces into IDAT chunks with the combined Adler-32: the file must be a valid PNG of the whole canvas."""
    from image_stitch_b200.png_container import read_png
    data, _ = _run(case, 3, tmp_path, {"STITCH_DIST_FILE": "1"})
    opts = ({"inputs": _inputs(), "layout": {"columns": 2}} if case == "grid_png" else
            {"inputs": [{"x": 7 * i, "y": 13 * i, "source": s} for i, s in enumerate(_inputs())], "layout": {}})
    _, want = H.oracle_stitch_rows(opts)
    header, rows = read_png(data)   # checks every chunk CRC and the zlib trailer
    assert rows == want.tobytes()
    assert data.count(b"IDAT") >= 3


def test_whole_jpeg_through_rank0(tmp_path):
    data, _ = _run("jpeg", 2, tmp_path, {"STITCH_DIST_FILE": "1"})
    opts = {"inputs": _inputs(), "layout": {"columns": 2}}
    plan, raw = H.oracle_stitch_rows(opts, out_kind=1)
    w, h = plan.canvas_w, plan.canvas_h
    assert data == orc.jpeg_file(np.frombuffer(raw, np.uint8).reshape(h, w * 4), w, h, 90)
