"""GPU parity: DIA SpMV and the device-side CSR -> DIA conversion (diagonal discovery by mark + scan + compaction)
through the C ABI vs the oracle and vs the CSR kernel."""
import ctypes as C

import numpy as np
import pytest

import oracle
from oracle import generators as g
from conftest import dev, dev_empty, host, golden

pytestmark = pytest.mark.gpu


def sfx(dt):
    return "f64" if np.dtype(dt) == np.float64 else "f32"


def gpu_diagonals(ctx, ptr, col, cols, max_diags=64):
    d_off = dev_empty(max(max_diags, 1), np.int32)
    n = C.c_int()
    ctx.call("cme213_csr_diagonals", len(ptr) - 1, cols, dev(ptr), dev(col), max_diags, d_off, C.byref(n))
    return n.value, host(d_off)[:min(n.value, max_diags)]


def gpu_to_dia(ctx, ptr, col, val, offs, pitch):
    d_dat = dev_empty(len(offs) * pitch, val.dtype)
    miss = C.c_int64(-1)
    ctx.call("cme213_csr_to_dia_" + sfx(val.dtype), len(ptr) - 1, dev(ptr), dev(col), dev(val), len(offs), dev(offs), pitch,
             d_dat, C.byref(miss))
    return d_dat, miss.value


def gpu_dia(ctx, rows, cols, pitch, offs, d_dat, x, y_off=0):
    d_y = dev_empty(rows + y_off, x.dtype)
    ctx.call("cme213_spmv_dia_" + sfx(x.dtype), rows, cols, pitch, len(offs), dev(offs), d_dat, dev(x),
             d_y.data_ptr() + y_off * x.dtype.itemsize)
    return host(d_y)[y_off:]


def test_bell_garland_fig1(ctx):
    f = golden("bell_fig2.npz")
    ptr, idx, val = f["csr_ptr"], f["csr_idx"], f["csr_val"]
    n, offs = gpu_diagonals(ctx, ptr, idx, 4)
    assert n == 3 and list(offs) == [-2, 0, 1]
    d_dat, miss = gpu_to_dia(ctx, ptr, idx, val, offs, 4)
    assert miss == 0 and list(host(d_dat)) == [0, 0, 5, 6, 1, 2, 3, 4, 7, 8, 9, 0]   # SC09 Fig. 3 "data"
    assert list(gpu_dia(ctx, 4, 4, 4, offs, d_dat, np.array([1, 2, 3, 4], np.float32))) == [15, 28, 50, 28]
    assert gpu_to_dia(ctx, ptr, idx, val, np.array([0, 1], np.int32), 4)[1] == 2
    assert gpu_diagonals(ctx, ptr, idx, 4, max_diags=2)[0] == 3                       # count is exact even when truncated


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("shape", [(1, 1, 1), (5, 1, 1), (8, 8, 1), (9, 7, 5), (32, 16, 8), (33, 31, 6), (64, 64, 17)])
def test_laplacian_dia_equals_csr_bit_exact(ctx, cme, dtype, shape):
    nx, ny, nz = shape
    rows = nx * ny * nz
    ptr, col, val = g.laplace3d(nx, ny, nz, dtype=dtype)
    x = np.random.default_rng(rows).uniform(-1, 1, rows).astype(dtype)
    n, offs = gpu_diagonals(ctx, ptr, col, rows)
    want_n, want_offs = oracle.csr_diagonals(ptr, col, rows)
    assert n == want_n and np.array_equal(offs, want_offs)
    want = oracle.csr_spmv(ptr, col, val, x)
    for pitch in (int(cme.load().cme213_ell_pitch(rows)), rows, rows + 1):   # vector path, exact, odd (scalar path)
        d_dat, miss = gpu_to_dia(ctx, ptr, col, val, offs, pitch)
        assert miss == 0
        assert np.array_equal(host(d_dat), oracle.csr_to_dia(ptr, col, val, offs, pitch)[0])
        got = gpu_dia(ctx, rows, rows, pitch, offs, d_dat, x)
        assert np.array_equal(got, want)
        assert np.array_equal(got, oracle.dia_spmv(rows, rows, pitch, offs, host(d_dat), x))
    # the CSR kernel gives the same bits
    d_y = dev_empty(rows, dtype)
    ctx.call("cme213_spmv_csr_" + sfx(dtype), 0, rows, dev(ptr), dev(col), dev(val), dev(x), d_y, 0)
    assert np.array_equal(host(d_y), want)


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_rectangular_banded_and_unaligned(ctx, dtype):
    # random band matrix, more columns than rows, diagonals at odd offsets, unaligned y and x
    rows, cols = 3001, 3500
    rng = np.random.default_rng(8)
    offs = np.array([-1700, -33, -2, -1, 0, 1, 3, 64, 499, 2100], np.int32)
    ptr, col, val = [0], [], []
    for r in range(rows):
        for o in offs:
            c = r + int(o)
            if 0 <= c < cols and rng.random() < 0.8:
                col.append(c); val.append(rng.uniform(-1, 1))
        ptr.append(len(col))
    ptr, col, val = np.array(ptr, np.int32), np.array(col, np.int32), np.array(val, dtype)
    x = rng.uniform(-1, 1, cols + 1).astype(dtype)
    n, got_offs = gpu_diagonals(ctx, ptr, col, cols)
    assert n == len(offs) and np.array_equal(got_offs, offs)
    pitch = 3008
    d_dat, miss = gpu_to_dia(ctx, ptr, col, val, offs, pitch)
    assert miss == 0
    want = oracle.csr_spmv(ptr, col, val, x[1:])
    d_x = dev(x)
    d_y = dev_empty(rows + 1, dtype)
    ctx.call("cme213_spmv_dia_" + sfx(dtype), rows, cols, pitch, len(offs), dev(offs), d_dat,
             d_x.data_ptr() + x.dtype.itemsize, d_y.data_ptr() + x.dtype.itemsize)
    assert np.array_equal(host(d_y)[1:], want)


def test_dia_errors_and_empty(ctx, cme):
    ctx.call("cme213_spmv_dia_f32", 0, 0, 0, 0, None, None, None, None)
    d = dev(np.zeros(8, np.float32))
    o = dev(np.zeros(1, np.int32))
    with pytest.raises(cme.Cme213Error):
        ctx.call("cme213_spmv_dia_f32", 8, 8, 4, 1, o, d, d, d)          # pitch < rows
    with pytest.raises(cme.Cme213Error):
        ctx.call("cme213_spmv_dia_f32", 8, 8, 8, 1, None, d, d, d)       # null offsets
    ctx.call("cme213_spmv_dia_f32", 8, 8, 8, 0, None, None, None, d)     # no diagonals: y = 0
    assert not host(d).any()


def test_dia_full_size_matches_csr(ctx, cme):
    # config C5's operator at 256^3 (16.8 M rows): DIA and CSR agree bit for bit on the device
    import torch
    gdim = 256
    rows = gdim ** 3
    nnz = C.c_int64()
    d_ptr = torch.empty(rows + 1, dtype=torch.int32, device="cuda")
    ctx.call("cme213_gen_laplace3d_f64", gdim, gdim, gdim, 0, rows, d_ptr, None, None, C.byref(nnz))
    d_col = torch.empty(nnz.value, dtype=torch.int32, device="cuda")
    d_val = torch.empty(nnz.value, dtype=torch.float64, device="cuda")
    ctx.call("cme213_gen_laplace3d_f64", gdim, gdim, gdim, 0, rows, d_ptr, d_col, d_val, C.byref(nnz))
    d_off = torch.empty(16, dtype=torch.int32, device="cuda")
    n = C.c_int()
    ctx.call("cme213_csr_diagonals", rows, rows, d_ptr, d_col, 16, d_off, C.byref(n))
    assert n.value == 7 and d_off[:7].tolist() == [-gdim * gdim, -gdim, -1, 0, 1, gdim, gdim * gdim]
    pitch = int(cme.load().cme213_ell_pitch(rows))
    d_dat = torch.empty(7 * pitch, dtype=torch.float64, device="cuda")
    miss = C.c_int64(-1)
    ctx.call("cme213_csr_to_dia_f64", rows, d_ptr, d_col, d_val, 7, d_off, pitch, d_dat, C.byref(miss))
    assert miss.value == 0
    x = torch.empty(rows, dtype=torch.float64, device="cuda")
    ctx.call("cme213_fill_uniform_f64", x, rows, 4)
    y_dia, y_csr = torch.empty_like(x), torch.empty_like(x)
    ctx.call("cme213_spmv_dia_f64", rows, rows, pitch, 7, d_off, d_dat, x, y_dia)
    ctx.call("cme213_spmv_csr_f64", 0, rows, d_ptr, d_col, d_val, x, y_csr, 0)
    torch.cuda.synchronize()
    assert torch.equal(y_dia, y_csr)


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("shape", [(1, 1, 1), (4, 1, 1), (7, 5, 1), (8, 8, 1), (9, 7, 5), (32, 16, 8), (36, 3, 6), (64, 64, 9), (132, 9, 40),
                                   (256, 20, 4), (4, 4, 70), (512, 8, 3), (1024, 4, 6), (2048, 2, 3), (520, 6, 2), (96, 10, 7)])
def test_matrix_free_stencil_equals_csr_bit_exact(ctx, dtype, shape):
    nx, ny, nz = shape
    rows = nx * ny * nz
    ptr, col, val = g.laplace3d(nx, ny, nz, dtype=dtype)       # centre 6
    x = np.random.default_rng(rows + 1).uniform(-1, 1, rows).astype(dtype)
    d_y = dev_empty(rows, dtype)
    ctx.call("cme213_stencil_laplace_" + sfx(dtype), nx, ny, nz, 6.0, 0, rows, dev(x), d_y)
    assert np.array_equal(host(d_y), oracle.csr_spmv(ptr, col, val, x))
    # a row range (as a shard would call it): untouched outside, identical inside
    lo, hi = (rows // 3) // 4 * 4, max((rows // 3) // 4 * 4, (2 * rows // 3) // 4 * 4)
    d_y2 = dev(np.full(rows, 7, dtype))
    ctx.call("cme213_stencil_laplace_" + sfx(dtype), nx, ny, nz, 6.0, lo, hi, dev(x), d_y2)
    got = host(d_y2)
    assert np.array_equal(got[lo:hi], host(d_y)[lo:hi]) and (got[:lo] == 7).all() and (got[hi:] == 7).all()
    # whole planes (how a z-slab shard calls it) -> the plane-marching kernel when nx % 4 == 0
    if nz >= 6:
        zl, zh = 1, nz - 1
        d_y4 = dev(np.full(rows, 7, dtype))
        ctx.call("cme213_stencil_laplace_" + sfx(dtype), nx, ny, nz, 6.0, zl * nx * ny, zh * nx * ny, dev(x), d_y4)
        got = host(d_y4)
        assert np.array_equal(got[zl * nx * ny:zh * nx * ny], host(d_y)[zl * nx * ny:zh * nx * ny])
        assert (got[:zl * nx * ny] == 7).all() and (got[zh * nx * ny:] == 7).all()
    # odd range -> scalar kernel
    if rows > 3:
        d_y3 = dev(np.full(rows, 7, dtype))
        ctx.call("cme213_stencil_laplace_" + sfx(dtype), nx, ny, nz, 6.0, 1, rows - 1, dev(x), d_y3)
        assert np.array_equal(host(d_y3)[1:-1], host(d_y)[1:-1])


def test_matrix_free_stencil_2d_generator_and_errors(ctx, cme):
    nx, ny = 24, 13
    ptr, col, val = g.laplace2d(nx, ny)                         # cme213_gen_laplace2d_f32's matrix: centre 4
    x = np.random.default_rng(2).uniform(-1, 1, nx * ny).astype(np.float32)
    d_y = dev_empty(nx * ny, np.float32)
    ctx.call("cme213_stencil_laplace_f32", nx, ny, 1, 4.0, 0, nx * ny, dev(x), d_y)
    assert np.array_equal(host(d_y), oracle.csr_spmv(ptr, col, val, x))
    d_x = dev(x)
    for args in ((0, ny, 1, 4.0, 0, 0, d_x, d_y), (nx, ny, 1, 4.0, 0, nx * ny + 1, d_x, d_y), (nx, ny, 1, 4.0, 0, 8, d_x, d_x),
                 (nx, ny, 1, 4.0, 0, 8, None, d_y)):
        with pytest.raises(cme.Cme213Error):
            ctx.call("cme213_stencil_laplace_f32", *args)


def test_matrix_free_stencil_full_size_matches_dia(ctx, cme):
    # 256^3: matrix-free == DIA == CSR on the device (the DIA/CSR pair is checked above)
    import torch
    gdim = 256
    rows = gdim ** 3
    nnz = C.c_int64()
    d_ptr = torch.empty(rows + 1, dtype=torch.int32, device="cuda")
    ctx.call("cme213_gen_laplace3d_f64", gdim, gdim, gdim, 0, rows, d_ptr, None, None, C.byref(nnz))
    d_col = torch.empty(nnz.value, dtype=torch.int32, device="cuda")
    d_val = torch.empty(nnz.value, dtype=torch.float64, device="cuda")
    ctx.call("cme213_gen_laplace3d_f64", gdim, gdim, gdim, 0, rows, d_ptr, d_col, d_val, C.byref(nnz))
    x = torch.empty(rows, dtype=torch.float64, device="cuda")
    ctx.call("cme213_fill_uniform_f64", x, rows, 4)
    y_mf, y_csr = torch.empty_like(x), torch.empty_like(x)
    ctx.call("cme213_stencil_laplace_f64", gdim, gdim, gdim, 6.0, 0, rows, x, y_mf)
    ctx.call("cme213_spmv_csr_f64", 0, rows, d_ptr, d_col, d_val, x, y_csr, 0)
    torch.cuda.synchronize()
    assert torch.equal(y_mf, y_csr)
