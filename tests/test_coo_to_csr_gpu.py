"""GPU parity: key-value radix sort and the COO -> CSR conversion built on it, through the C ABI vs numpy."""
import numpy as np
import pytest

import oracle
from oracle import generators as g
from conftest import dev, dev_empty, host

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n,end_bit", [(1, 8), (33, 3), (8192, 8), (8193, 13), (100003, 20), (2000003, 32), (70000, 0)])
def test_pairs_sort_is_stable_on_the_key_bits(ctx, n, end_bit):
    rng = np.random.default_rng(n + end_bit)
    keys = rng.integers(0, 2**32, n, dtype=np.uint64).astype(np.uint32)
    vals = np.arange(n, dtype=np.uint32)
    d_k, d_v = dev(keys), dev(vals)
    ctx.call("cme213_radix_sort_pairs_u32", d_k, dev_empty(n, np.uint32), d_v, dev_empty(n, np.uint32), n, end_bit)
    mask = np.uint32((1 << end_bit) - 1) if end_bit < 32 else np.uint32(0xFFFFFFFF)
    order = np.argsort(keys & mask, kind="stable")
    assert np.array_equal(host(d_k, np.uint32), keys[order])
    assert np.array_equal(host(d_v, np.uint32), vals[order])


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("rows,cols,nnz", [(1, 1, 1), (5, 7, 0), (10, 10, 40), (1000, 300, 20000), (70000, 90000, 500000), (3, 200000, 100000)])
def test_coo_to_csr_matches_numpy(ctx, dtype, rows, cols, nnz):
    rng = np.random.default_rng(rows + nnz)
    r = rng.integers(0, rows, nnz).astype(np.int32)
    c = rng.integers(0, cols, nnz).astype(np.int32)
    if nnz > 10:                       # some exact duplicates: they must keep their input order
        r[5:10], c[5:10] = r[0], c[0]
        r[rng.integers(0, nnz, nnz // 7)] = rows - 1          # a heavy last row, and empty rows elsewhere
    v = rng.uniform(-1, 1, nnz).astype(dtype)
    d_ptr, d_col, d_val = dev_empty(rows + 1, np.int32), dev_empty(max(nnz, 1), np.int32), dev_empty(max(nnz, 1), dtype)
    sfx = "f64" if dtype == np.float64 else "f32"
    ctx.call("cme213_coo_to_csr_" + sfx, rows, cols, nnz, dev(r) if nnz else None, dev(c) if nnz else None,
             dev(v) if nnz else None, d_ptr, d_col if nnz else None, d_val if nnz else None)
    ptr, col, val = g.coo_to_csr(rows, r, c, v)
    assert np.array_equal(host(d_ptr), ptr)
    assert np.array_equal(host(d_col)[:nnz], col)
    assert np.array_equal(host(d_val)[:nnz], val)
    if nnz:                            # and the CSR arrays feed the SpMV: same bits as the oracle's CSR loop
        x = rng.uniform(-1, 1, cols).astype(dtype)
        d_y = dev_empty(rows, dtype)
        ctx.call("cme213_spmv_csr_" + sfx, 0, rows, d_ptr, d_col, d_val, dev(x), d_y, 8)   # CME213_CSR_AUTO_ORDERED: CPU order whatever the row lengths
        assert np.array_equal(host(d_y), oracle.csr_spmv(ptr, col, val, x))


def test_coo_to_csr_rejects_entries_outside_the_matrix(ctx, cme):
    r, c = np.array([0, 1, 5], np.int32), np.array([0, 1, 1], np.int32)
    v = np.ones(3, np.float32)
    with pytest.raises(cme.Cme213Error):
        ctx.call("cme213_coo_to_csr_f32", 5, 5, 3, dev(r), dev(c), dev(v), dev_empty(6, np.int32), dev_empty(3, np.int32),
                 dev_empty(3, np.float32))
    with pytest.raises(cme.Cme213Error):
        ctx.call("cme213_coo_to_csr_f32", 5, 5, 3, None, dev(c), dev(v), dev_empty(6, np.int32), dev_empty(3, np.int32),
                 dev_empty(3, np.float32))
