"""stitch_filter_rows_noalloc (operators' addition, no counterpart in the reference): kernel (c) with the caller's
device buffers and the library's static __device__ scratch.  Checked against the existing entry point
stitch_filter_rows (bit-exact: integer work) on two calls in a row -- the second call is smaller and has other
data, so sums left over from the first call would show -- and against its documented limits.  Collected last
(file name) so that a failure here cannot hide another test under `pytest -x`."""
import numpy as np
import pytest

from image_stitch_b200 import _native

pytestmark = pytest.mark.gpu


def _rows(rng, n_rows, row_bytes):
    base = rng.integers(0, 256, size=(1, row_bytes), dtype=np.int64)
    drift = np.cumsum(rng.integers(-3, 4, size=(n_rows, row_bytes)), axis=0)
    return ((base + drift) & 0xFF).astype(np.uint8)


def _run(ctx, rows, prev, bpp, check_free):
    import torch
    dev = torch.device("cuda", ctx.device)
    n_rows, row_bytes = rows.shape
    in_pitch, out_pitch = row_bytes + 5, row_bytes + 1 + 7      # pitches need no alignment
    d_in = torch.zeros((n_rows, in_pitch), dtype=torch.uint8, device=dev)
    d_in[:, :row_bytes] = torch.from_numpy(rows).to(dev)
    d_prev = torch.from_numpy(prev).to(dev) if prev is not None else None
    d_out = torch.full((n_rows, out_pitch), 0xEE, dtype=torch.uint8, device=dev)
    torch.cuda.synchronize(dev)
    free0 = torch.cuda.mem_get_info(dev)[0]
    ctx.filter_rows_noalloc(d_in.data_ptr(), in_pitch, d_prev.data_ptr() if d_prev is not None else 0, n_rows, row_bytes, bpp,
                            d_out.data_ptr(), out_pitch)
    if check_free:   # not on the first call of a process: loading the kernels' module takes device memory once
        assert torch.cuda.mem_get_info(dev)[0] == free0, "the call changed the amount of free device memory"
    got = d_out.cpu().numpy()
    assert (got[:, row_bytes + 1:] == 0xEE).all(), "bytes past a row were written"
    return got[:, :row_bytes + 1].reshape(-1)


def test_noalloc_matches_filter_rows_twice(ctx):
    rng = np.random.default_rng(77)
    for k, (n_rows, row_bytes, bpp, with_prev) in enumerate(((61, 20011, 4, True), (23, 9000, 8, False), (40, 33000, 4, True))):
        rows = _rows(rng, n_rows, row_bytes)
        prev = _rows(rng, 1, row_bytes)[0] if with_prev else None
        want = ctx.filter_rows(rows, prev, n_rows, row_bytes, bpp)
        got = _run(ctx, rows, prev, bpp, check_free=k > 0)
        assert np.array_equal(got, np.asarray(want)), (n_rows, row_bytes, bpp)
        assert set(np.unique(got.reshape(n_rows, -1)[:, 0])) <= {0, 1, 2, 3, 4}


def test_noalloc_rejects_sizes_beyond_the_scratch(ctx):
    import torch
    dev = torch.device("cuda", ctx.device)
    buf = torch.zeros(64, dtype=torch.uint8, device=dev)
    with pytest.raises(_native.StitchError) as e:
        ctx.filter_rows_noalloc(buf.data_ptr(), 4, 0, 16385, 4, 4, buf.data_ptr(), 5)
    assert e.value.code == _native.E_CAPACITY and "16384" in str(e.value)
    with pytest.raises(_native.StitchError) as e:
        ctx.filter_rows_noalloc(buf.data_ptr(), (1 << 25) + 1, 0, 1, (1 << 25) + 1, 4, buf.data_ptr(), (1 << 25) + 2)
    assert e.value.code == _native.E_CAPACITY
    with pytest.raises(_native.StitchError) as e:
        ctx.filter_rows_noalloc(0, 4, 0, 1, 4, 4, buf.data_ptr(), 5)
    assert e.value.code == _native.E_INVALID
