"""The distributed product path on GPUs (it had only ever run with the oracle as renderer):
two ranks, each rendering its row band with libstitch_b200 (two GPUs when the box has them, else both ranks share the
one GPU), pieces gathered to rank 0 over the host group, container written there; and one process driving two
device contexts from threads (ConcatOptions.devices), which is what needs the per-device kernel set-up."""
import os

import numpy as np
import pytest

import helpers as H
import image_stitch_b200 as stitch
from image_stitch_b200.png_container import read_png
from test_distributed_gloo import _inputs, _run

pytestmark = pytest.mark.gpu


def _opts(case):
    if case == "positioned_png":
        return {"inputs": [{"x": 7 * i, "y": 13 * i, "source": s} for i, s in enumerate(_inputs())], "layout": {}}
    return {"inputs": _inputs(), "layout": {"columns": 2}}


@pytest.mark.parametrize("case,deflate", [("grid_png", False), ("positioned_png", False), ("grid_png", True), ("positioned_png", True)])
def test_two_ranks_png_file(case, deflate, tmp_path):
    env = {"STITCH_DIST_FILE": "1", "STITCH_DIST_GPU": "1"}
    if deflate:
        env["STITCH_DIST_DEVICE_DEFLATE"] = "1"
    data, _ = _run(case, 2, tmp_path, env, gpu=True)
    _, want = H.oracle_stitch_rows(_opts(case))
    _, rows = read_png(data)
    assert rows == want.tobytes()


def test_two_ranks_jpeg_equals_one_device(tmp_path):
    data, _ = _run("jpeg", 2, tmp_path, {"STITCH_DIST_FILE": "1", "STITCH_DIST_GPU": "1"}, gpu=True)
    one = stitch.concat_to_buffer({"inputs": _inputs(), "layout": {"columns": 2}, "output_format": "jpeg", "jpeg_quality": 90})
    assert data == one


@pytest.mark.parametrize("deflate", [False, True])
def test_one_process_two_device_contexts(deflate):
    """ConcatOptions.devices: one thread per device context.  With two GPUs the second context runs on cuda:1
    (per-device shared-memory opt-in, occupancy, CRC tables); with one GPU both shares run on cuda:0 from two threads."""
    import torch
    devs = [0, 1] if torch.cuda.device_count() >= 2 else [0, 0]
    opts = {"inputs": [H.suite_bytes(n) for n in ("basn2c08", "basn6a08", "basn0g08", "basn4a08")] * 4, "layout": {"columns": 4},
            "devices": devs, "device_deflate": deflate, "band_rows": 16}
    png = stitch.concat_to_buffer(opts)
    _, rows = read_png(png)
    _, want = H.oracle_stitch_rows({"inputs": opts["inputs"], "layout": {"columns": 4}})
    assert rows == want.tobytes()
