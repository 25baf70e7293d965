"""GPU parity: CSR SpMV kernels and the PageRank propagate step through the C ABI vs the oracle."""
import numpy as np
import pytest

import oracle
from oracle import generators as g
from conftest import dev, dev_empty, host, golden

pytestmark = pytest.mark.gpu

SCALAR, VECTOR, STREAM, AUTO = 1, 2, 3, 0


def gpu_csr(ctx, ptr, col, val, x, kernel, row_begin=0, row_end=None):
    rows = len(ptr) - 1
    row_end = rows if row_end is None else row_end
    fn = "cme213_spmv_csr_f64" if val.dtype == np.float64 else "cme213_spmv_csr_f32"
    d_y = dev(np.full(rows, -7.0, dtype=val.dtype))
    ctx.call(fn, row_begin, row_end, dev(ptr), dev(col if len(col) else np.zeros(4, np.int32)),
             dev(val if len(val) else np.zeros(4, val.dtype)), dev(x), d_y, kernel)
    return host(d_y)


def check_tol(got, ptr, col, val, x, extra=1):
    want = oracle.csr_spmv(ptr, col, val, x).astype(np.float64)
    bound = oracle.csr_abs_bound(ptr, col, val, x)
    k = np.diff(ptr) + extra
    u = np.finfo(val.dtype).eps / 2
    # gamma_k * sum|a_ij x_j|; the oracle itself carries one such error, hence the factor 2
    assert np.all(np.abs(got.astype(np.float64) - want) <= 2 * k * u * bound + 1e-300)


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("kernel", [SCALAR, STREAM, AUTO, 7, 7 | (1 << 8)])
def test_csr_sequential_kernels_bit_exact_on_laplacian(ctx, dtype, kernel):
    # config C1 in miniature: 5-point Laplacian, natural order
    ptr, col, val = g.laplace2d(67, 45, dtype)
    x = np.random.default_rng(2).uniform(-1, 1, 67 * 45).astype(dtype)
    assert np.array_equal(gpu_csr(ctx, ptr, col, val, x, kernel), oracle.csr_spmv(ptr, col, val, x))


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("rows,cols,mean", [(1, 1, 1.0), (255, 300, 3.0), (257, 300, 0.3), (5000, 4000, 9.0), (3000, 20000, 40.0)])
def test_csr_kernels_on_random_matrices(ctx, dtype, rows, cols, mean):
    ptr, col, val = g.random_csr(rows, cols, mean, seed=rows + cols, dtype=dtype)  # includes empty rows
    x = np.random.default_rng(3).uniform(-1, 1, cols).astype(dtype)
    want = oracle.csr_spmv(ptr, col, val, x)
    assert np.array_equal(gpu_csr(ctx, ptr, col, val, x, SCALAR), want)
    assert np.array_equal(gpu_csr(ctx, ptr, col, val, x, STREAM), want)
    assert np.array_equal(gpu_csr(ctx, ptr, col, val, x, 4), want)   # CME213_CSR_STREAM_PADDED
    assert np.array_equal(gpu_csr(ctx, ptr, col, val, x, 7), want)   # CME213_CSR_STREAM_BULK (mean 40: several passes per block)
    assert np.array_equal(gpu_csr(ctx, ptr, col, val, x, 7 | (1 << 8)), want)   # CME213_CSR_STREAM_PIPE (mean 9 / 40: blocks that do not fit a stage)
    assert np.array_equal(gpu_csr(ctx, ptr, col, val, x, AUTO), want)
    for lanes in (2, 4, 8, 16, 32):
        check_tol(gpu_csr(ctx, ptr, col, val, x, VECTOR | (lanes << 8)), ptr, col, val, x)


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_csr_regular_rows_take_the_padded_variant(ctx, dtype):
    # every row has 16 entries: AUTO's inspector picks the padded stream kernel; results stay bit-exact
    rows, cols = 5000, 7000
    rng = np.random.default_rng(11)
    ptr = (np.arange(rows + 1) * 16).astype(np.int32)
    col = np.concatenate([np.sort(rng.choice(cols, 16, replace=False)) for _ in range(rows)]).astype(np.int32)
    val = rng.uniform(-1, 1, rows * 16).astype(dtype)
    x = rng.uniform(-1, 1, cols).astype(dtype)
    want = oracle.csr_spmv(ptr, col, val, x)
    for kernel in (AUTO, STREAM, 4):
        assert np.array_equal(gpu_csr(ctx, ptr, col, val, x, kernel), want)


def test_csr_stream_rows_longer_than_a_tile(ctx):
    # rows with more nonzeros than one 2048-entry shared-memory pass, next to empty and short rows
    rng = np.random.default_rng(4)
    cols = 9000
    length = np.array([0, 5000, 3, 0, 2048, 2049, 1, 7000, 0, 0, 2, 4097] + [3] * 600 + [0, 6000])
    ptr = np.concatenate([[0], np.cumsum(length)]).astype(np.int32)
    col = np.concatenate([np.sort(rng.choice(cols, l, replace=False)) for l in length]).astype(np.int32)
    val = rng.uniform(-1, 1, ptr[-1]).astype(np.float32)
    x = rng.uniform(-1, 1, cols).astype(np.float32)
    want = oracle.csr_spmv(ptr, col, val, x)
    assert np.array_equal(gpu_csr(ctx, ptr, col, val, x, STREAM), want)
    assert np.array_equal(gpu_csr(ctx, ptr, col, val, x, SCALAR), want)
    check_tol(gpu_csr(ctx, ptr, col, val, x, VECTOR | (32 << 8)), ptr, col, val, x)


@pytest.mark.parametrize("kernel", [SCALAR, STREAM, VECTOR | (8 << 8), 7, 7 | (1 << 8)])
def test_csr_row_ranges_leave_other_rows_untouched(ctx, kernel):
    # the multi-GPU path launches boundary rows and interior rows separately
    ptr, col, val = g.laplace3d(9, 8, 7)
    n = 9 * 8 * 7
    x = np.random.default_rng(5).uniform(-1, 1, n)
    want = oracle.csr_spmv(ptr, col, val, x)
    for lo, hi in [(0, 72), (72, n - 72), (n - 72, n), (100, 101), (300, 300)]:
        got = gpu_csr(ctx, ptr, col, val, x, kernel, lo, hi)
        assert np.all(got[:lo] == -7.0) and np.all(got[hi:] == -7.0)
        if kernel == (VECTOR | (8 << 8)):
            assert np.allclose(got[lo:hi], want[lo:hi], rtol=1e-14, atol=1e-14)
        else:
            assert np.array_equal(got[lo:hi], want[lo:hi])


def test_csr_bad_arguments(ctx, cme):
    d = dev(np.zeros(8, np.float32)); di = dev(np.zeros(8, np.int32))
    with pytest.raises(cme.Cme213Error):
        ctx.call("cme213_spmv_csr_f32", 5, 2, di, di, d, d, d, AUTO)
    with pytest.raises(cme.Cme213Error):
        ctx.call("cme213_spmv_csr_f32", 0, 2, None, di, d, d, d, AUTO)
    with pytest.raises(cme.Cme213Error):
        ctx.call("cme213_spmv_csr_f32", 0, 2, di, di, d, d, d, 77)
    with pytest.raises(cme.Cme213Error):
        ctx.call("cme213_spmv_csr_f32", 0, 2, di, di, d, d, d, VECTOR | (3 << 8))


@pytest.mark.parametrize("kernel", [SCALAR, STREAM])
def test_pagerank_propagate_bit_exact(ctx, kernel):
    ind, edges, inv, nodes = g.pagerank_graph(40000, 8, seed=9)
    d_out = dev_empty(len(nodes), np.float32)
    ctx.call("cme213_pagerank_propagate", dev(ind), dev(edges), dev(nodes), d_out, dev(inv), len(nodes), kernel)
    want = oracle.pagerank_propagate(ind, edges, nodes, inv)
    got = host(d_out)
    assert np.array_equal(got, want)
    assert oracle.count_ulp_errors(got, want, 10) == 0  # the homework's own check, pagerank.cu:216-239


def test_pagerank_iterations_match_reference_golden(ctx):
    # golden = the reference's host_graph_iterate on this graph (tests/golden/make_golden.py)
    f = golden("pagerank_ref_n5000.npz")
    n = len(f["nodes"])
    d_ind, d_edges, d_inv = dev(f["indices"]), dev(f["edges"]), dev(f["inv"])
    A, B = dev(f["nodes"]), dev_empty(n, np.float32)
    for it in range(0, 20, 2):
        ctx.call("cme213_pagerank_propagate", d_ind, d_edges, A, B, d_inv, n, AUTO)
        ctx.call("cme213_pagerank_propagate", d_ind, d_edges, B, A, d_inv, n, AUTO)
        if it == 0:
            assert np.array_equal(host(A), f["out2"])
    assert np.array_equal(host(A), f["out20"])


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_csr_ordered_launch_is_bitwise_identical(ctx, cme, dtype):
    # any permutation of the 256-row blocks gives the same bits; the 3-D strip order is one such permutation
    import importlib
    import torch
    sh = importlib.import_module("2012-04_stanford_cme213_b200.sharded")
    nx, ny, nz = 20, 17, 13
    ptr, col, val = g.laplace3d(nx, ny, nz, dtype=dtype)
    n = nx * ny * nz
    x = np.random.default_rng(6).uniform(-1, 1, n).astype(dtype)
    want = oracle.csr_spmv(ptr, col, val, x)
    br = cme.load().cme213_spmv_csr_block_rows()
    fn = "cme213_spmv_csr_f64_ordered" if dtype == np.float64 else "cme213_spmv_csr_f32_ordered"
    for lo, hi in [(0, n), (nx * ny, n - nx * ny), (300, 301)]:
        nblk = -(-(hi - lo) // br)
        for order in (np.random.default_rng(7).permutation(nblk).astype(np.int32),
                      sh.grid3d_block_order(nx, ny, nz, lo, hi, br, 4).cpu().numpy()):
            assert sorted(order.tolist()) == list(range(nblk))
            d_y = dev(np.full(n, -7.0, dtype=dtype))
            ctx.call(fn, lo, hi, dev(ptr), dev(col), dev(val), dev(x), d_y, dev(order), nblk)
            got = host(d_y)
            assert np.array_equal(got[lo:hi], want[lo:hi]) and np.all(got[:lo] == -7.0) and np.all(got[hi:] == -7.0)
    # a subset of the blocks: only those rows are written (the multi-GPU boundary launch)
    nblk = -(-n // br)
    sub = np.array([nblk - 1, 0, 3], np.int32)
    d_y = dev(np.full(n, -7.0, dtype=dtype))
    ctx.call(fn, 0, n, dev(ptr), dev(col), dev(val), dev(x), d_y, dev(sub), len(sub))
    got = host(d_y)
    mask = np.zeros(n, bool)
    for b in sub:
        mask[b * br:(b + 1) * br] = True
    assert np.array_equal(got[mask], want[mask]) and np.all(got[~mask] == -7.0)
