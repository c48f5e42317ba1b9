"""The C-ABI library loads on a CPU-only box, exports every symbol
include/stitch_b200.h declares, and refuses to run without a GPU."""
import ctypes
import os
import re

import pytest

from image_stitch_b200 import _native

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "stitch_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(stitch_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_exported():
    lib = _native.load()
    names = _declared()
    assert len(names) >= 16
    for n in names:
        assert hasattr(lib, n), f"{n} declared in stitch_b200.h but not exported"
    assert sorted(_native.EXPORTS) == names
    assert lib.stitch_abi_version() == 1


def test_struct_layouts_match_header():
    assert ctypes.sizeof(_native.JobDesc) == 56
    assert ctypes.sizeof(_native.TileDesc) == 56
    assert ctypes.sizeof(_native.Timings) == 28
    assert ctypes.sizeof(_native.JpegBandInfo) == 24


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(_native.StitchError) as e:
        _native.Context(0)
    assert e.value.code == _native.E_NODEVICE
    import image_stitch_b200 as s
    import helpers as H
    with pytest.raises(Exception):
        s.concat_to_buffer({"inputs": [H.suite_bytes("basn6a08")] * 2, "layout": {"columns": 2}})
