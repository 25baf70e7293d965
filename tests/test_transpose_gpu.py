"""GPU parity: tiled transpose through the C ABI vs the oracle (bit-exact)."""
import numpy as np
import pytest

import oracle
from conftest import dev, dev_empty, host

pytestmark = pytest.mark.gpu


def run_transpose(ctx, a, ld_in=None, ld_out=None):
    rows, cols = a.shape
    ld_in = cols if ld_in is None else ld_in
    ld_out = rows if ld_out is None else ld_out
    buf = np.zeros((rows, ld_in), dtype=a.dtype)
    buf[:, :cols] = a
    d_in = dev(buf)
    d_out = dev(np.full((cols, ld_out), 123, dtype=a.dtype))
    fn = "cme213_transpose_b32" if a.dtype.itemsize == 4 else "cme213_transpose_b64"
    ctx.call(fn, d_in, d_out, rows, cols, ld_in, ld_out)
    return host(d_out).reshape(cols, ld_out)


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("shape", [(64, 64), (128, 256), (4, 4), (100, 36), (1, 1), (3, 5), (257, 129),
                                   (1031, 2053), (2048, 1024), (8, 4096)])
def test_transpose_matches_oracle(ctx, shape, dtype):
    rng = np.random.default_rng(shape[0] * 7919 + shape[1])
    a = rng.standard_normal(shape).astype(dtype)
    got = run_transpose(ctx, a)
    assert np.array_equal(got.view(np.uint8), oracle.transpose(a).view(np.uint8))


@pytest.mark.parametrize("dtype", [np.int32, np.float64])
def test_transpose_with_pitches_and_ragged_columns(ctx, dtype):
    # the ELL conversion shape: [rows x K] staged with pitch Kp -> [K x rows] inside a wider array
    rng = np.random.default_rng(1)
    for rows, cols, ld_in, ld_out in [(256, 7, 8, 320), (64, 3, 4, 64), (512, 16, 16, 1024), (60, 5, 9, 61)]:
        a = rng.integers(-1000, 1000, (rows, cols)).astype(dtype)
        got = run_transpose(ctx, a, ld_in, ld_out)
        assert np.array_equal(got[:, :rows], a.T)
        assert np.all(got[:, rows:] == 123)  # padding of the output pitch is never written


def test_transpose_rejects_bad_arguments(ctx, cme):
    d = dev(np.zeros(16, np.float32))
    with pytest.raises(cme.Cme213Error) as e:
        ctx.call("cme213_transpose_b32", d, d, 4, 4, 4, 4)  # in place
    assert e.value.code == cme.capi.ERR_INVALID
    with pytest.raises(cme.Cme213Error):
        ctx.call("cme213_transpose_b32", d, dev(np.zeros(16, np.float32)), 4, 4, 3, 4)  # ld_in < cols
    ctx.call("cme213_transpose_b32", None, None, 0, 5, 5, 0)  # empty is fine


@pytest.mark.parametrize("dtype,n", [(np.float32, 16384), (np.float64, 16384)])
def test_transpose_full_size_properties(ctx, dtype, n):
    # BASELINE config C3 at full size: involution + sampled direct check (the oracle would take minutes)
    import torch
    tdt = torch.float32 if dtype == np.float32 else torch.float64
    a = torch.empty((n, n), dtype=tdt, device="cuda")
    name = "cme213_fill_uniform_f32" if dtype == np.float32 else "cme213_fill_uniform_f64"
    ctx.call(name, a, n * n, 77)
    b = torch.empty_like(a)
    fn = "cme213_transpose_b32" if dtype == np.float32 else "cme213_transpose_b64"
    ctx.call(fn, a, b, n, n, n, n)
    rng = np.random.default_rng(2)
    r = torch.from_numpy(rng.integers(0, n, 100000)).cuda()
    c = torch.from_numpy(rng.integers(0, n, 100000)).cuda()
    assert torch.equal(b[c, r], a[r, c])
    assert torch.equal(b[:, 5], a[5, :]) and torch.equal(b[n - 1, :], a[:, n - 1])
    c2 = torch.empty_like(a)
    ctx.call(fn, b, c2, n, n, n, n)
    assert torch.equal(c2, a)
