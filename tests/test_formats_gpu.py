"""GPU parity: ELL / COO / HYB SpMV and the device-side CSR -> ELL/HYB conversion vs the oracle."""
import ctypes as C

import numpy as np
import pytest

import oracle
from oracle import generators as g
from conftest import dev, dev_empty, host, golden

pytestmark = pytest.mark.gpu


def sfx(dt):
    return "f64" if dt == np.float64 else "f32"


def gpu_ell(ctx, rows, pitch, K, idx, val, x):
    d_y = dev_empty(rows, val.dtype)
    ctx.call("cme213_spmv_ell_" + sfx(val.dtype), rows, pitch, K, dev(idx if len(idx) else np.zeros(4, np.int32)),
             dev(val if len(val) else np.zeros(4, val.dtype)), dev(x), d_y)
    return host(d_y)


def test_bell_garland_fig2(ctx):
    f = golden("bell_fig2.npz")
    ptr, idx, val = f["csr_ptr"], f["csr_idx"], f["csr_val"]
    x = np.array([1, 2, 3, 4], np.float32)
    y = np.array([15, 28, 50, 28], np.float32)
    pitch = 4
    d_idx, d_val = dev_empty(3 * pitch, np.int32), dev_empty(3 * pitch, np.float32)
    ctx.call("cme213_csr_to_ell_f32", 4, 4, dev(ptr), dev(idx), dev(val), 3, pitch, d_idx, d_val)
    assert np.array_equal(host(d_val), f["ell_val_colmajor"])  # SC09 Fig. 4 linearisation
    assert np.array_equal(gpu_ell(ctx, 4, pitch, 3, host(d_idx), host(d_val), x), y)
    d_y = dev(np.zeros(4, np.float32))
    ctx.call("cme213_spmv_coo_accumulate_f32", 9, dev(f["coo_row"]), dev(idx), dev(val), dev(x), d_y)
    assert np.array_equal(host(d_y), y)


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("rows,cols,mean", [(1, 5, 2.0), (700, 650, 6.0), (4099, 5000, 11.0), (10000, 10000, 1.0)])
def test_csr_to_ell_and_ell_spmv_bit_exact(ctx, cme, dtype, rows, cols, mean):
    ptr, col, val = g.random_csr(rows, cols, mean, seed=rows, dtype=dtype)
    x = np.random.default_rng(1).uniform(-1, 1, cols).astype(dtype)
    K = int(np.diff(ptr).max())
    pitch = int(cme.load().cme213_ell_pitch(rows))
    assert pitch >= rows and pitch % 64 == 0
    d_idx, d_val = dev_empty(max(K * pitch, 4), np.int32), dev_empty(max(K * pitch, 4), dtype)
    ctx.call("cme213_csr_to_ell_" + sfx(dtype), rows, cols, dev(ptr), dev(col if len(col) else np.zeros(4, np.int32)),
             dev(val if len(val) else np.zeros(4, dtype)), K, pitch, d_idx, d_val)
    o_idx, o_val = oracle.csr_to_ell(ptr, col, val, cols, K, pitch)
    got_idx, got_val = host(d_idx)[:K * pitch], host(d_val)[:K * pitch]
    assert np.array_equal(got_val, o_val)
    assert np.array_equal(got_idx[o_val != 0], o_idx[o_val != 0])  # padding columns only need to be valid
    assert got_idx.min(initial=0) >= 0 and got_idx.max(initial=0) < cols
    want = oracle.csr_spmv(ptr, col, val, x)
    assert np.array_equal(gpu_ell(ctx, rows, pitch, K, got_idx, got_val, x), want)           # 4 rows / thread
    assert np.array_equal(oracle.ell_spmv(rows, pitch, K, got_idx, got_val, x), want)


def test_ell_scalar_fallback_for_odd_pitch(ctx):
    ptr, col, val = g.random_csr(333, 400, 5.0, seed=2)
    x = np.random.default_rng(2).uniform(-1, 1, 400).astype(np.float32)
    K = int(np.diff(ptr).max())
    idx, dat = oracle.csr_to_ell(ptr, col, val, 400, K, 333)  # pitch 333: not a multiple of 4
    assert np.array_equal(gpu_ell(ctx, 333, 333, K, idx, dat, x), oracle.csr_spmv(ptr, col, val, x))


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("nnz_rows", [(9, 4), (5000, 700), (300000, 20000), (20000, 20000)])
def test_coo_segmented_reduction(ctx, dtype, nnz_rows):
    nnz, rows = nnz_rows
    rng = np.random.default_rng(nnz)
    r = np.sort(rng.integers(0, rows, nnz)).astype(np.int32)
    if nnz > 9000:
        r[1000:9000] = r[1000]  # one row spanning several 4096-entry tiles
        r = np.sort(r)
    c = rng.integers(0, rows, nnz).astype(np.int32)
    v = rng.uniform(-1, 1, nnz).astype(dtype)
    x = rng.uniform(-1, 1, rows).astype(dtype)
    y0 = rng.uniform(-1, 1, rows).astype(dtype)
    d_y = dev(y0)
    ctx.call("cme213_spmv_coo_accumulate_" + sfx(dtype), nnz, dev(r), dev(c), dev(v), dev(x), d_y)
    got = host(d_y).astype(np.float64)
    prod = v.astype(np.float64) * x[c].astype(np.float64)
    want = y0.astype(np.float64) + np.bincount(r, prod, rows)
    bound = np.abs(y0).astype(np.float64) + np.bincount(r, np.abs(prod), rows)
    k = np.bincount(r, minlength=rows) + 2
    assert np.all(np.abs(got - want) <= k * (np.finfo(dtype).eps / 2) * bound + 1e-300)
    untouched = np.bincount(r, minlength=rows) == 0
    assert np.array_equal(host(d_y)[untouched], y0[untouched])


def test_hyb_split_and_spmv(ctx, cme):
    rows = 30000
    lp = g.powerlaw_lengths(rows, 4, 265, 100, 256, seed=6)
    col, val = g.random_csr_fill(lp, rows, 0, seed=7)
    x = np.random.default_rng(8).uniform(-1, 1, rows).astype(np.float32)
    K, maxlen = C.c_int(), C.c_int()
    ctx.call("cme213_hyb_pick_k", rows, dev(lp), C.byref(K), C.byref(maxlen))
    assert K.value == oracle.hyb_pick_k(lp) and maxlen.value == int(np.diff(lp).max())
    K = K.value
    pitch = int(cme.load().cme213_ell_pitch(rows))
    d_ptr, d_col, d_val = dev(lp), dev(col), dev(val)
    d_eidx, d_eval = dev_empty(K * pitch, np.int32), dev_empty(K * pitch, np.float32)
    ctx.call("cme213_csr_to_ell_f32", rows, rows, d_ptr, d_col, d_val, K, pitch, d_eidx, d_eval)
    tail = C.c_int64()
    ctx.call("cme213_csr_to_coo_tail_f32", rows, d_ptr, d_col, d_val, K, C.byref(tail), None, None, None)
    tr, tc, tv = oracle.csr_to_coo_tail(lp, col, val, K)
    assert tail.value == len(tr) and tail.value > 0
    d_tr, d_tc, d_tv = dev_empty(tail.value, np.int32), dev_empty(tail.value, np.int32), dev_empty(tail.value, np.float32)
    ctx.call("cme213_csr_to_coo_tail_f32", rows, d_ptr, d_col, d_val, K, None, d_tr, d_tc, d_tv)
    assert np.array_equal(host(d_tr), tr) and np.array_equal(host(d_tc), tc) and np.array_equal(host(d_tv), tv)
    d_y = dev_empty(rows, np.float32)
    ctx.call("cme213_spmv_hyb_f32", rows, pitch, K, d_eidx, d_eval, tail.value, d_tr, d_tc, d_tv, dev(x), d_y)
    got = host(d_y).astype(np.float64)
    want = oracle.csr_spmv(lp, col, val, x).astype(np.float64)
    bound = oracle.csr_abs_bound(lp, col, val, x)
    assert np.all(np.abs(got - want) <= 2 * (np.diff(lp) + 2) * 2.0 ** -24 * bound)
    # rows that fit in the ELL part are summed in CSR order: bit-exact
    short = np.diff(lp) <= K
    assert np.array_equal(host(d_y)[short], oracle.csr_spmv(lp, col, val, x)[short])
    # pure ELL (K = max length) and HYB agree with CSR-stream on the same matrix
    d_y2 = dev_empty(rows, np.float32)
    ctx.call("cme213_spmv_csr_f32", 0, rows, d_ptr, d_col, d_val, dev(x), d_y2, 3)
    assert np.array_equal(host(d_y2), oracle.csr_spmv(lp, col, val, x))


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_coo_segmented_reduction_bulk_path(ctx, dtype):
    # >= 2^20 entries: persistent bulk-copy kernel; long rows cross tile boundaries, some rows are empty
    nnz, rows = 3 * (1 << 20) + 7, 90000
    rng = np.random.default_rng(5)
    r = np.sort(rng.integers(0, rows, nnz)).astype(np.int32)
    r[500000:900000] = r[500000]
    r = np.sort(r)
    c = rng.integers(0, rows, nnz).astype(np.int32)
    v = rng.integers(-3, 4, nnz).astype(dtype)          # small integers: sums are exact in any order
    x = rng.integers(-3, 4, rows).astype(dtype)
    y0 = rng.integers(-3, 4, rows).astype(dtype)
    d_y = dev(y0)
    ctx.call("cme213_spmv_coo_accumulate_" + sfx(dtype), nnz, dev(r), dev(c), dev(v), dev(x), d_y)
    want = y0.astype(np.float64) + np.bincount(r, v.astype(np.float64) * x[c], rows)
    assert np.array_equal(host(d_y).astype(np.float64), want)
    assert np.array_equal(host(d_y), oracle.coo_spmv_accumulate(r, c, v, x, y0))


def test_hyb_f64_matches_csr_oracle(ctx, cme):
    rows = 6000
    ptr, col, val = g.random_csr(rows, rows, 7.0, seed=9, dtype=np.float64)
    x = np.random.default_rng(10).uniform(-1, 1, rows)
    K = 6
    pitch = int(cme.load().cme213_ell_pitch(rows))
    d_ptr, d_col, d_val = dev(ptr), dev(col), dev(val)
    d_eidx, d_eval = dev_empty(K * pitch, np.int32), dev_empty(K * pitch, np.float64)
    ctx.call("cme213_csr_to_ell_f64", rows, rows, d_ptr, d_col, d_val, K, pitch, d_eidx, d_eval)
    tail = C.c_int64()
    ctx.call("cme213_csr_to_coo_tail_f64", rows, d_ptr, d_col, d_val, K, C.byref(tail), None, None, None)
    tr, tc, tv = oracle.csr_to_coo_tail(ptr, col, val, K)
    assert tail.value == len(tr) > 0
    d_tr, d_tc, d_tv = dev_empty(tail.value, np.int32), dev_empty(tail.value, np.int32), dev_empty(tail.value, np.float64)
    ctx.call("cme213_csr_to_coo_tail_f64", rows, d_ptr, d_col, d_val, K, None, d_tr, d_tc, d_tv)
    assert np.array_equal(host(d_tr), tr) and np.array_equal(host(d_tc), tc) and np.array_equal(host(d_tv), tv)
    d_y = dev_empty(rows, np.float64)
    ctx.call("cme213_spmv_hyb_f64", rows, pitch, K, d_eidx, d_eval, tail.value, d_tr, d_tc, d_tv, dev(x), d_y)
    want = oracle.csr_spmv(ptr, col, val, x)
    bound = oracle.csr_abs_bound(ptr, col, val, x)
    assert np.all(np.abs(host(d_y) - want) <= 2 * (np.diff(ptr) + 2) * 2.0 ** -53 * bound + 1e-300)
