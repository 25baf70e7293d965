"""GPU: device-side input generators vs their numpy restatement (bit-exact), so any input a bench or a
full-size test builds on the device can be rebuilt on the host."""
import ctypes as C

import numpy as np
import pytest

from oracle import generators as g
from conftest import dev, dev_empty, host

pytestmark = pytest.mark.gpu


def test_fills(ctx):
    n = 100003
    d = dev_empty(n, np.float32)
    ctx.call("cme213_fill_uniform_f32", d, n, 42)
    assert np.array_equal(host(d), g.fill_uniform_f32(n, 42))
    d64 = dev_empty(n, np.float64)
    ctx.call("cme213_fill_uniform_f64", d64, n, 43)
    assert np.array_equal(host(d64), g.fill_uniform_f64(n, 43))
    du = dev_empty(n, np.uint32)
    ctx.call("cme213_fill_hash_u32", du, n, 44, 0)
    assert np.array_equal(host(du, np.uint32), g.fill_hash_u32(n, 44))
    ctx.call("cme213_fill_hash_u32", du, n, 45, 1000)
    assert np.array_equal(host(du, np.uint32), g.fill_hash_u32(n, 45, 1000))
    bits = dev_empty((n + 31) // 32, np.uint32)
    ctx.call("cme213_fill_random_flags", bits, n, 46, (1 << 32) // 64)
    assert np.array_equal(host(bits, np.uint32), g.pack_flags(g.random_flags(n, 46, (1 << 32) // 64)))


def test_laplacians(ctx):
    nx, ny = 37, 29
    rows = nx * ny
    ptr, col, val = g.laplace2d(nx, ny)
    d_ptr, d_col, d_val = dev_empty(rows + 1, np.int32), dev_empty(len(col), np.int32), dev_empty(len(col), np.float32)
    nnz = C.c_int64()
    ctx.call("cme213_gen_laplace2d_f32", nx, ny, d_ptr, d_col, d_val, C.byref(nnz))
    assert nnz.value == len(col)
    assert np.array_equal(host(d_ptr), ptr) and np.array_equal(host(d_col), col) and np.array_equal(host(d_val), val)
    nx, ny, nz = 9, 7, 11
    for lo, hi in [(0, nx * ny * nz), (nx * ny * 3, nx * ny * 8), (5, 6)]:
        ptr, col, val = g.laplace3d(nx, ny, nz, lo, hi)
        d_ptr, d_col, d_val = dev_empty(hi - lo + 1, np.int32), dev_empty(len(col), np.int32), dev_empty(len(col), np.float64)
        ctx.call("cme213_gen_laplace3d_f64", nx, ny, nz, lo, hi, d_ptr, d_col, d_val, C.byref(nnz))
        assert nnz.value == len(col)
        assert np.array_equal(host(d_ptr), ptr) and np.array_equal(host(d_col), col) and np.array_equal(host(d_val), val)


@pytest.mark.parametrize("window", [0, 4096])
def test_powerlaw_matrix(ctx, window):
    rows = 50000
    ptr = g.powerlaw_lengths(rows, 4, 265, 100, 256, seed=6)
    d_ptr = dev_empty(rows + 1, np.int32)
    nnz = C.c_int64()
    ctx.call("cme213_gen_powerlaw_lengths", rows, 4, 265, 100, 256, 6, d_ptr, C.byref(nnz))
    assert nnz.value == ptr[-1] and np.array_equal(host(d_ptr), ptr)
    d_col, d_val = dev_empty(nnz.value, np.int32), dev_empty(nnz.value, np.float32)
    ctx.call("cme213_gen_random_csr_fill_f32", rows, rows, d_ptr, window, 7, d_col, d_val)
    col, val = g.random_csr_fill(ptr, rows, window, seed=7)
    assert np.array_equal(host(d_col), col) and np.array_equal(host(d_val), val)


def test_pagerank_graph(ctx):
    n, avg = 30000, 8
    ind, edges, inv, nodes = g.pagerank_graph(n, avg, seed=21)
    d_ind, d_edges = dev_empty(n + 1, np.uint32), dev_empty(len(edges), np.uint32)
    d_inv, d_nodes = dev_empty(n, np.float32), dev_empty(n, np.float32)
    nnz = C.c_int64()
    ctx.call("cme213_gen_pagerank_graph", n, avg, 21, d_ind, d_edges, d_inv, d_nodes, C.byref(nnz))
    assert nnz.value == len(edges)
    assert np.array_equal(host(d_ind, np.uint32), ind) and np.array_equal(host(d_edges, np.uint32), edges)
    assert np.array_equal(host(d_inv), inv) and np.array_equal(host(d_nodes), nodes)


@pytest.mark.parametrize("n,cuts", [(1, 0), (2, 5), (10, 0), (100, 30), (100, 400), (1000003, 15000), (1 << 24, (1 << 24) // 64)])
def test_segment_offsets_like_readmm(ctx, n, cuts):
    # readMM.py:44-49: sorted random cut points framed by 0 and n; built on the device with the radix sort and the scan
    d_s = dev_empty(cuts + 2, np.int32)
    p = C.c_int64()
    ctx.call("cme213_gen_segment_offsets", n, cuts, 77, d_s, C.byref(p))
    want = g.segment_offsets(n, cuts, 77)
    assert p.value == len(want)
    s = host(d_s)[:p.value]
    assert np.array_equal(s, want)
    # the loader's invariants (aux/mp1-util.h:118-147): s[0] = 0, s[p-1] = n, strictly increasing
    assert s[0] == 0 and s[-1] == n and (np.diff(s) > 0).all()


def test_segment_offsets_feed_the_flag_builder(ctx):
    n, cuts = 300000, 5000
    d_s = dev_empty(cuts + 2, np.int32)
    p = C.c_int64()
    ctx.call("cme213_gen_segment_offsets", n, cuts, 5, d_s, C.byref(p))
    bits = dev_empty((n + 31) // 32, np.uint32)
    ctx.call("cme213_segment_flags_from_offsets", d_s, int(p.value), n, bits)
    want = np.zeros(n, bool)
    want[g.segment_offsets(n, cuts, 5)[:-1]] = True
    assert np.array_equal(host(bits, np.uint32), g.pack_flags(want))
