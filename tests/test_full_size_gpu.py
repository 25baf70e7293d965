"""GPU parity at the STATED sizes of BASELINE.json's configs C1, C4 and C5 (C2 and C3 have theirs in test_scan_gpu.py
and test_transpose_gpu.py).  Inputs are built by the device generators (the same calls bench.py makes); the oracle
checks what a CPU can check in seconds (the whole of C1, a 1 Mi-row prefix of C4, the first 32 planes of C5) and the
rest is pinned through format-independent identities on the device: ELL == CSR-stream bit for bit (both keep the CPU
loop's order and rounding), HYB within the reduction-order bound, DIA == matrix-free == CSR bit for bit."""
import ctypes as C

import numpy as np
import pytest

import oracle
from conftest import host

pytestmark = pytest.mark.gpu

STREAM = 3


def _free(torch):
    import gc
    gc.collect()
    torch.cuda.empty_cache()


def test_c1_csr_fp32_poisson_1024x1024_bit_exact_and_10_ulp(ctx):
    """Config C1 at its full size: 1,048,576 rows / 5,238,784 nonzeros, every CSR kernel against the CPU loop, plus the
    homework's own acceptance check (<= 10 ULP per element, hw/hw1/programming/pagerank.cu:216-239)."""
    import torch
    n = 1024
    rows = n * n
    nnz = C.c_int64()
    d_ptr = torch.empty(rows + 1, dtype=torch.int32, device="cuda")
    ctx.call("cme213_gen_laplace2d_f32", n, n, d_ptr, None, None, C.byref(nnz))
    assert nnz.value == 5238784
    d_col = torch.empty(nnz.value, dtype=torch.int32, device="cuda")
    d_val = torch.empty(nnz.value, dtype=torch.float32, device="cuda")
    ctx.call("cme213_gen_laplace2d_f32", n, n, d_ptr, d_col, d_val, C.byref(nnz))
    d_x = torch.empty(rows, dtype=torch.float32, device="cuda")
    ctx.call("cme213_fill_uniform_f32", d_x, rows, 2)
    ptr, col, x = host(d_ptr), host(d_col), host(d_x)
    for values in ("stencil", "random"):                      # diag 4 / off-diag -1, then values ~ U(-1, 1)
        if values == "random":
            ctx.call("cme213_fill_uniform_f32", d_val, nnz.value, 1)
        val = host(d_val)
        want = oracle.csr_spmv(ptr, col, val, x)
        for kernel in (0, 1, STREAM, 4):                      # AUTO, scalar, stream, stream-padded: all keep the CPU order
            d_y = torch.full((rows,), -7.0, dtype=torch.float32, device="cuda")
            ctx.call("cme213_spmv_csr_f32", 0, rows, d_ptr, d_col, d_val, d_x, d_y, kernel)
            got = host(d_y)
            assert np.array_equal(got, want), "kernel %d, %s values" % (kernel, values)
            assert oracle.count_ulp_errors(got, want, 10) == 0
        d_y = torch.empty(rows, dtype=torch.float32, device="cuda")
        ctx.call("cme213_spmv_csr_f32", 0, rows, d_ptr, d_col, d_val, d_x, d_y, 2 | (4 << 8))   # CSR-vector, 4 lanes per row
        bound = oracle.csr_abs_bound(ptr, col, val, x)
        assert np.all(np.abs(host(d_y).astype(np.float64) - want) <= 2 * 6 * 2.0 ** -24 * bound)
    _free(torch)


def test_c4_powerlaw_16mi_rows_ell_hyb_vs_csr(ctx, cme):
    """Config C4 at its full size: 16,777,216 rows, ~262 M nonzeros, Pareto row lengths capped at 256.  Pure ELL pads
    to K = 256: 2^32 padded entries, 34 GB -- the 64-bit indexing case."""
    import torch
    rows = 1 << 24
    lib = cme.load()
    d_ptr = torch.empty(rows + 1, dtype=torch.int32, device="cuda")
    nnz = C.c_int64()
    ctx.call("cme213_gen_powerlaw_lengths", rows, 4, 265, 100, 256, 6, d_ptr, C.byref(nnz))
    d_col = torch.empty(nnz.value, dtype=torch.int32, device="cuda")
    d_val = torch.empty(nnz.value, dtype=torch.float32, device="cuda")
    ctx.call("cme213_gen_random_csr_fill_f32", rows, rows, d_ptr, 0, 7, d_col, d_val)
    d_x = torch.empty(rows, dtype=torch.float32, device="cuda")
    ctx.call("cme213_fill_uniform_f32", d_x, rows, 8)
    y_csr = torch.empty(rows, dtype=torch.float32, device="cuda")
    ctx.call("cme213_spmv_csr_f32", 0, rows, d_ptr, d_col, d_val, d_x, y_csr, STREAM)

    # (1) a 1 Mi-row prefix against the CPU loop, bit for bit
    pre = 1 << 20
    ptr = host(d_ptr[:pre + 1])
    m = int(ptr[-1])
    want = oracle.csr_spmv(ptr, host(d_col[:m]), host(d_val[:m]), host(d_x))
    assert np.array_equal(host(y_csr[:pre]), want)

    # (2) pure ELL, K = longest row: bit-identical to CSR-stream over all 16 Mi rows
    K, maxlen = C.c_int(), C.c_int()
    ctx.call("cme213_hyb_pick_k", rows, d_ptr, C.byref(K), C.byref(maxlen))
    assert maxlen.value == 256 and 1 <= K.value < 256
    pitch = int(lib.cme213_ell_pitch(rows))
    assert maxlen.value * pitch >= 1 << 32                    # the reason this test exists
    e_idx = torch.empty(maxlen.value * pitch, dtype=torch.int32, device="cuda")
    e_val = torch.empty(maxlen.value * pitch, dtype=torch.float32, device="cuda")
    ctx.call("cme213_csr_to_ell_f32", rows, rows, d_ptr, d_col, d_val, maxlen.value, pitch, e_idx, e_val)
    y = torch.full((rows,), -7.0, dtype=torch.float32, device="cuda")
    ctx.call("cme213_spmv_ell_f32", rows, pitch, maxlen.value, e_idx, e_val, d_x, y)
    torch.cuda.synchronize()
    assert torch.equal(y, y_csr)
    # the last ELL column (entries beyond 2^32) really holds the longest rows' last entries
    last = e_val[(maxlen.value - 1) * pitch:(maxlen.value - 1) * pitch + rows]
    lens = (d_ptr[1:] - d_ptr[:-1])
    assert torch.equal(last != 0, lens == maxlen.value)
    del e_idx, e_val, last
    _free(torch)

    # (3) HYB with the SC09 split: rows inside the ELL part are bit-identical, the others within gamma_k * sum|a x|
    Kh = K.value
    e_idx = torch.empty(Kh * pitch, dtype=torch.int32, device="cuda")
    e_val = torch.empty(Kh * pitch, dtype=torch.float32, device="cuda")
    ctx.call("cme213_csr_to_ell_f32", rows, rows, d_ptr, d_col, d_val, Kh, pitch, e_idx, e_val)
    tail = C.c_int64()
    ctx.call("cme213_csr_to_coo_tail_f32", rows, d_ptr, d_col, d_val, Kh, C.byref(tail), None, None, None)
    assert tail.value == int(torch.clamp(lens.to(torch.int64) - Kh, min=0).sum())
    t_r = torch.empty(tail.value, dtype=torch.int32, device="cuda")
    t_c = torch.empty(tail.value, dtype=torch.int32, device="cuda")
    t_v = torch.empty(tail.value, dtype=torch.float32, device="cuda")
    ctx.call("cme213_csr_to_coo_tail_f32", rows, d_ptr, d_col, d_val, Kh, None, t_r, t_c, t_v)
    y.fill_(-7.0)
    ctx.call("cme213_spmv_hyb_f32", rows, pitch, Kh, e_idx, e_val, tail.value, t_r, t_c, t_v, d_x, y)
    bound = torch.empty(rows, dtype=torch.float32, device="cuda")      # sum |a_ij x_j| per row, on the device
    ctx.call("cme213_spmv_csr_f32", 0, rows, d_ptr, d_col, d_val.abs(), d_x.abs(), bound, STREAM)
    torch.cuda.synchronize()
    short = lens <= Kh
    assert torch.equal(y[short], y_csr[short])
    tol = 2.0 * (lens.to(torch.float64) + 2) * 2.0 ** -24 * bound.to(torch.float64) * (1 + 2.0 ** -20)
    assert bool(((y.to(torch.float64) - y_csr.to(torch.float64)).abs() <= tol).all())
    _free(torch)


def test_c5_csr_dia_matrix_free_512_cubed(ctx, cme):
    """Config C5 at its full size (134,217,728 rows, 937,951,232 nonzeros, fp64): the CSR kernel the benchmark times
    against the CPU loop on the first 32 planes, and against DIA and the matrix-free stencil over the whole grid."""
    import torch
    g = 512
    rows = g ** 3
    nnz = C.c_int64()
    d_ptr = torch.empty(rows + 1, dtype=torch.int32, device="cuda")
    ctx.call("cme213_gen_laplace3d_f64", g, g, g, 0, rows, d_ptr, None, None, C.byref(nnz))
    assert nnz.value == 937951232
    d_col = torch.empty(nnz.value, dtype=torch.int32, device="cuda")
    d_val = torch.empty(nnz.value, dtype=torch.float64, device="cuda")
    ctx.call("cme213_gen_laplace3d_f64", g, g, g, 0, rows, d_ptr, d_col, d_val, C.byref(nnz))
    d_x = torch.empty(rows, dtype=torch.float64, device="cuda")
    ctx.call("cme213_fill_uniform_f64", d_x, rows, 9)
    y_csr = torch.empty(rows, dtype=torch.float64, device="cuda")
    ctx.call("cme213_spmv_csr_f64", 0, rows, d_ptr, d_col, d_val, d_x, y_csr, 0)          # AUTO: what bench.py runs
    y_mf = torch.empty_like(y_csr)
    ctx.call("cme213_stencil_laplace_f64", g, g, g, 6.0, 0, rows, d_x, y_mf)
    torch.cuda.synchronize()
    assert torch.equal(y_mf, y_csr)

    planes = 32
    pre = planes * g * g
    ptr = host(d_ptr[:pre + 1])
    m = int(ptr[-1])
    want = oracle.csr_spmv(ptr, host(d_col[:m]), host(d_val[:m]), host(d_x[:pre + g * g]), omp=True)
    assert np.array_equal(host(y_csr[:pre]), want)
    # and the LAST planes, where the 32-bit element offsets are largest (nnz ~ 0.94e9 < 2^31, byte offsets > 2^33)
    lo = rows - 4 * g * g
    ptr = host(d_ptr[lo:]).astype(np.int64)
    b0, b1 = int(ptr[0]), int(ptr[-1])
    xs = host(d_x[lo - g * g:])
    want = oracle.csr_spmv((ptr - b0).astype(np.int32), host(d_col[b0:b1]) - (lo - g * g), host(d_val[b0:b1]), xs)
    assert np.array_equal(host(y_csr[lo:]), want)

    # DIA, converted on the device from the same CSR arrays
    d_off = torch.empty(16, dtype=torch.int32, device="cuda")
    nd = C.c_int()
    ctx.call("cme213_csr_diagonals", rows, rows, d_ptr, d_col, 16, d_off, C.byref(nd))
    assert nd.value == 7
    pitch = int(cme.load().cme213_ell_pitch(rows))
    d_dat = torch.empty(7 * pitch, dtype=torch.float64, device="cuda")
    miss = C.c_int64(-1)
    ctx.call("cme213_csr_to_dia_f64", rows, d_ptr, d_col, d_val, 7, d_off, pitch, d_dat, C.byref(miss))
    assert miss.value == 0
    del d_col, d_val
    _free(torch)
    y_dia = torch.empty_like(y_csr)
    ctx.call("cme213_spmv_dia_f64", rows, rows, pitch, 7, d_off, d_dat, d_x, y_dia)
    torch.cuda.synchronize()
    assert torch.equal(y_dia, y_csr)
    del d_dat, y_dia, y_mf, y_csr
    _free(torch)


def test_c4_hyb_with_csr_tail_16mi_rows_bit_identical_to_csr(ctx, cme):
    """Config C4 at full size through the format chooser: cme213_spmv_pick_format picks HYB and its K for the power-law
    matrix, the tail runs as a CSR matrix of its own, and the product equals CSR-stream's bit for bit over all 16 Mi rows."""
    import torch
    rows = 1 << 24
    lib = cme.load()
    d_ptr = torch.empty(rows + 1, dtype=torch.int32, device="cuda")
    nnz = C.c_int64()
    ctx.call("cme213_gen_powerlaw_lengths", rows, 4, 265, 100, 256, 6, d_ptr, C.byref(nnz))
    d_col = torch.empty(nnz.value, dtype=torch.int32, device="cuda")
    d_val = torch.empty(nnz.value, dtype=torch.float32, device="cuda")
    ctx.call("cme213_gen_random_csr_fill_f32", rows, rows, d_ptr, 0, 7, d_col, d_val)
    d_x = torch.empty(rows, dtype=torch.float32, device="cuda")
    ctx.call("cme213_fill_uniform_f32", d_x, rows, 8)
    choice = cme.capi.FormatChoice()
    ctx.call("cme213_spmv_pick_format", rows, d_ptr, C.byref(choice))
    assert choice.format == cme.capi.FORMAT_HYB and 1 <= choice.K <= 32 and choice.cost_vs_csr < 1.0
    K = choice.K
    pitch = int(lib.cme213_ell_pitch(rows))
    e_idx = torch.empty(K * pitch, dtype=torch.int32, device="cuda")
    e_val = torch.empty(K * pitch, dtype=torch.float32, device="cuda")
    ctx.call("cme213_csr_to_ell_f32", rows, rows, d_ptr, d_col, d_val, K, pitch, e_idx, e_val)
    t_p = torch.empty(rows + 1, dtype=torch.int32, device="cuda")
    tail = C.c_int64()
    ctx.call("cme213_csr_tail_row_ptr", rows, d_ptr, K, t_p, C.byref(tail))
    assert abs(tail.value / nnz.value - choice.tail_fraction) < 1e-6
    t_r = torch.empty(tail.value, dtype=torch.int32, device="cuda")
    t_c = torch.empty(tail.value, dtype=torch.int32, device="cuda")
    t_v = torch.empty(tail.value, dtype=torch.float32, device="cuda")
    ctx.call("cme213_csr_to_coo_tail_f32", rows, d_ptr, d_col, d_val, K, None, t_r, t_c, t_v)
    del t_r
    y_csr = torch.empty(rows, dtype=torch.float32, device="cuda")
    ctx.call("cme213_spmv_csr_f32", 0, rows, d_ptr, d_col, d_val, d_x, y_csr, STREAM)
    y = torch.full((rows,), -7.0, dtype=torch.float32, device="cuda")
    ctx.call("cme213_spmv_hyb_csrtail_f32", rows, pitch, K, e_idx, e_val, t_p, t_c, t_v, d_x, y)
    y_auto = torch.empty_like(y)
    ctx.call("cme213_spmv_csr_f32", 0, rows, d_ptr, d_col, d_val, d_x, y_auto, 0)     # the dispatcher's choice for this matrix: an order-preserving kernel
    torch.cuda.synchronize()
    assert torch.equal(y, y_csr)
    assert torch.equal(y_auto, y_csr)
    _free(torch)
