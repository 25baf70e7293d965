"""GPU: the CSR dispatcher (cme213_csr_plan_*, CME213_CSR_AUTO), the format chooser (cme213_spmv_pick_format) and HYB
with a CSR tail.  Every dispatched path is checked against the CPU loop: bit for bit where the kernel keeps its order
(stream, stream-padded, stream-keep, scalar, the short blocks of a binned launch, ELL + CSR tail), within the
reduction-order bound gamma_k * sum|a_ij x_j| where lanes of a warp share a row (vector, the long blocks)."""
import ctypes as C

import numpy as np
import pytest

import oracle
from oracle import generators as g
from conftest import dev, dev_empty, host

pytestmark = pytest.mark.gpu
SCALAR, VECTOR, STREAM, PADDED, KEEP, BINNED = 1, 2, 3, 4, 5, 6


def make_plan(ctx, cme, ptr, cols, elem_bytes):
    lib = cme.load()
    h = C.c_void_p()
    d_ptr = dev(ptr)
    rc = lib.cme213_csr_plan_create(ctx.h, 0, len(ptr) - 1, cols, C.c_void_p(d_ptr.data_ptr()), elem_bytes, None, C.byref(h))
    assert rc == 0, lib.cme213_last_error()
    return h, d_ptr


def run_plan(ctx, cme, plan, d_ptr, col, val, x):
    lib = cme.load()
    fn = lib.cme213_spmv_csr_plan_f64 if val.dtype == np.float64 else lib.cme213_spmv_csr_plan_f32
    d_y = dev(np.full(len(d_ptr) - 1, -7.0, dtype=val.dtype))
    P = lambda t: C.c_void_p(t.data_ptr())
    d_col, d_val, d_x = dev(col), dev(val), dev(x)
    rc = fn(ctx.h, plan, P(d_ptr), P(d_col), P(d_val), P(d_x), P(d_y), None)
    assert rc == 0, lib.cme213_last_error()
    return host(d_y)


def within_bound(got, ptr, col, val, x):
    want = oracle.csr_spmv(ptr, col, val, x).astype(np.float64)
    bound = oracle.csr_abs_bound(ptr, col, val, x)
    u = np.finfo(val.dtype).eps / 2
    return np.all(np.abs(got.astype(np.float64) - want) <= 2 * (np.diff(ptr) + 1) * u * bound + 1e-300)


def regular_csr(rows, cols, length, seed, dtype=np.float32, band=None):
    rng = np.random.default_rng(seed)
    ptr = (np.arange(rows + 1, dtype=np.int64) * length).astype(np.int32)
    if band is None:
        col = np.sort(rng.integers(0, cols, (rows, length)), axis=1)
    else:
        col = np.sort(np.clip(np.arange(rows)[:, None] + rng.integers(-band, band, (rows, length)), 0, cols - 1), axis=1)
    return ptr, col.reshape(-1).astype(np.int32), rng.uniform(-1, 1, rows * length).astype(dtype)


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_plan_picks_by_row_length_and_every_choice_matches_the_cpu_loop(ctx, cme, dtype):
    lib = cme.load()
    eb = np.dtype(dtype).itemsize
    cases = []
    ptr, col, val = g.laplace2d(200, 150, dtype)                                 # short irregular rows
    BULK = 7                                                                     # fp32 rows of 8 or fewer: bulk-copy-fed stream kernel
    cases.append(("laplace", ptr, col, val, 200 * 150, (BULK,) if eb == 4 else (STREAM,)))
    cases.append(("regular16",) + regular_csr(20000, 30000, 16, 1, dtype) + (30000, (PADDED,)))
    cases.append(("regular7",) + regular_csr(20000, 30000, 7, 2, dtype) + (30000, (BULK,) if eb == 4 else (STREAM,)))
    cases.append(("regular64",) + regular_csr(6000, 50000, 64, 3, dtype) + (50000, (PADDED,)))
    cases.append(("short2",) + regular_csr(50000, 50000, 2, 6, dtype) + (50000, (SCALAR,)))
    cases.append(("long300",) + regular_csr(1500, 50000, 300, 4, dtype) + (50000, (VECTOR,)))
    cases.append(("long100",) + regular_csr(3000, 50000, 100, 5, dtype) + (50000, (VECTOR,)))
    for name, ptr, col, val, cols, expect in cases:
        x = np.random.default_rng(9).uniform(-1, 1, cols).astype(dtype)
        plan, d_ptr = make_plan(ctx, cme, ptr, cols, eb)
        kind = lib.cme213_csr_plan_kernel(plan) & 0xFF
        text = lib.cme213_csr_plan_describe(plan).decode()
        assert kind in expect, (name, text)
        got = run_plan(ctx, cme, plan, d_ptr, col, val, x)
        if kind == VECTOR:
            assert within_bound(got, ptr, col, val, x), name
        else:
            assert np.array_equal(got, oracle.csr_spmv(ptr, col, val, x)), (name, text)
        # AUTO through the ordinary entry point takes the same decision (its own cached plan) and gives the same bits
        ctx.lib.cme213_ctx_clear_plans(ctx.h)       # cached plans are keyed by address: an earlier test's matrix may have lived here
        d_y = dev(np.zeros(len(ptr) - 1, dtype))
        ctx.call("cme213_spmv_csr_" + ("f64" if eb == 8 else "f32"), 0, len(ptr) - 1, d_ptr, dev(col), dev(val), dev(x), d_y, 0)
        assert np.array_equal(host(d_y), got), name
        lib.cme213_csr_plan_destroy(plan)


def test_ordered_plans_stay_bit_identical_on_long_rows(ctx, cme):
    """CME213_PLAN_ORDERED / CME213_CSR_AUTO_ORDERED: the dispatcher may only pick kernels that keep the CPU loop's order."""
    lib = cme.load()
    ptr, col, val = regular_csr(1500, 50000, 300, 4)
    x = np.random.default_rng(1).uniform(-1, 1, 50000).astype(np.float32)
    want = oracle.csr_spmv(ptr, col, val, x)
    plan, d_ptr = make_plan(ctx, cme, ptr, 50000, 4 | cme.capi.PLAN_ORDERED)
    assert lib.cme213_csr_plan_kernel(plan) & 0xFF in (SCALAR, STREAM, PADDED, 7), lib.cme213_csr_plan_describe(plan)
    assert np.array_equal(run_plan(ctx, cme, plan, d_ptr, col, val, x), want)
    lib.cme213_csr_plan_destroy(plan)
    d_y = dev(np.zeros(1500, np.float32))
    ctx.call("cme213_spmv_csr_f32", 0, 1500, d_ptr, dev(col), dev(val), dev(x), d_y, cme.capi.CSR_AUTO_ORDERED)
    assert np.array_equal(host(d_y), want)
    for kernel in (KEEP,):                       # the explicit evict-last variant: same bits
        d_y = dev(np.zeros(1500, np.float32))
        ctx.call("cme213_spmv_csr_f32", 0, 1500, d_ptr, dev(col), dev(val), dev(x), d_y, kernel)
        assert np.array_equal(host(d_y), want), kernel


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_plan_bins_a_skewed_matrix(ctx, cme, dtype):
    """4 % of the rows are 400 long and sit together in runs (dense blocks of a multiphysics matrix), the rest have 3
    entries: the plan splits the 256-row blocks into a stream list and a vector list."""
    lib = cme.load()
    rows, cols = 40000, 60000
    rng = np.random.default_rng(7)
    lens = np.full(rows, 3)
    for start in (1000, 9000, 9256, 25000, 39700):
        lens[start:start + 300] = 400
    lens[5] = 0
    ptr = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
    col = np.concatenate([np.sort(rng.choice(cols, l, replace=False)) for l in lens]).astype(np.int32)
    val = rng.uniform(-1, 1, ptr[-1]).astype(dtype)
    x = rng.uniform(-1, 1, cols).astype(dtype)
    plan, d_ptr = make_plan(ctx, cme, ptr, cols, np.dtype(dtype).itemsize)
    assert lib.cme213_csr_plan_kernel(plan) == BINNED, lib.cme213_csr_plan_describe(plan)
    got = run_plan(ctx, cme, plan, d_ptr, col, val, x)
    assert not np.any(got == -7.0) or np.all(oracle.csr_spmv(ptr, col, val, x)[got == -7.0] == -7.0)   # every row was written
    assert within_bound(got, ptr, col, val, x)
    want = oracle.csr_spmv(ptr, col, val, x)
    blocks_all_short = np.add.reduceat(lens, np.arange(0, rows, 256)) < 256 * 32
    short_rows = np.repeat(blocks_all_short, 256)[:rows]
    assert np.array_equal(got[short_rows], want[short_rows])         # stream kernel rows: bit-exact
    lib.cme213_csr_plan_destroy(plan)


def test_plan_errors_and_empty(ctx, cme):
    lib = cme.load()
    h = C.c_void_p()
    assert lib.cme213_csr_plan_create(ctx.h, 0, 0, 0, None, 4, None, C.byref(h)) == 0
    assert lib.cme213_spmv_csr_plan_f32(ctx.h, h, None, None, None, None, None, None) == 0       # nothing to do
    assert lib.cme213_spmv_csr_plan_f64(ctx.h, h, None, None, None, None, None, None) == cme.capi.ERR_INVALID   # plan made for 4-byte values
    lib.cme213_csr_plan_destroy(h)
    assert lib.cme213_csr_plan_create(ctx.h, 0, 10, 10, None, 4, None, C.byref(h)) == cme.capi.ERR_INVALID
    assert lib.cme213_csr_plan_create(ctx.h, 0, 10, 10, None, 3, None, C.byref(h)) == cme.capi.ERR_INVALID


def pick(ctx, cme, ptr):
    out = cme.capi.FormatChoice()
    ctx.call("cme213_spmv_pick_format", len(ptr) - 1, dev(ptr), C.byref(out))
    return out


def test_pick_format_rules(ctx, cme):
    # regular rows: ELL costs exactly its entries
    ptr = (np.arange(5001) * 16).astype(np.int32)
    c = pick(ctx, cme, ptr)
    assert (c.format, c.K) == (cme.capi.FORMAT_ELL, 16) and abs(c.ell_padding_ratio - 1.0) < 1e-12 and c.tail_fraction == 0
    # the stencil: rows of 5..7, padding 7/6.9: ELL
    p3, _, _ = g.laplace3d(30, 30, 30)
    assert pick(ctx, cme, p3).format == cme.capi.FORMAT_ELL
    # power-law rows (the C4 recipe): ELL would pad 16x; HYB with the cost-minimal K, checked against the restated rule
    lp = g.powerlaw_lengths(300000, 4, 265, 100, 256, seed=6)
    c = pick(ctx, cme, lp)
    lens = np.diff(lp)
    M, nnz = len(lens), lens.sum()
    costs = {K: K * M + 1.6 * np.maximum(lens - K, 0).sum() for K in range(1, lens.max() + 1)}
    bestK = min(costs, key=costs.get)
    assert lens.max() * M > 1.6 * nnz                      # ELL ruled out
    if costs[bestK] < 1.6 * nnz:
        assert (c.format, c.K) == (cme.capi.FORMAT_HYB, bestK)
        assert abs(c.tail_fraction - np.maximum(lens - bestK, 0).sum() / nnz) < 1e-9
        assert abs(c.cost_vs_csr - costs[bestK] / (1.6 * nnz)) < 1e-9
    else:
        assert c.format == cme.capi.FORMAT_CSR
    # a few huge rows among empty ones: nothing beats CSR
    lens = np.zeros(10000, np.int64)
    lens[::500] = 3000
    c = pick(ctx, cme, np.concatenate([[0], np.cumsum(lens)]).astype(np.int32))
    assert c.format == cme.capi.FORMAT_CSR and c.max_row == 3000


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_hyb_with_csr_tail_is_bit_identical_to_the_cpu_loop(ctx, cme, dtype):
    sfx = "f64" if dtype == np.float64 else "f32"
    rows = 50000
    lp = g.powerlaw_lengths(rows, 4, 265, 100, 256, seed=6)
    col, val = g.random_csr_fill(lp, rows, 0, seed=7)
    val = val.astype(dtype)
    x = np.random.default_rng(8).uniform(-1, 1, rows).astype(dtype)
    want = oracle.csr_spmv(lp, col, val, x)
    pitch = int(cme.load().cme213_ell_pitch(rows))
    d_ptr, d_col, d_val, d_x = dev(lp), dev(col), dev(val), dev(x)
    for K in (1, 6, 9, 40, 256):
        d_ei, d_ev = dev_empty(K * pitch, np.int32), dev_empty(K * pitch, dtype)
        ctx.call("cme213_csr_to_ell_" + sfx, rows, rows, d_ptr, d_col, d_val, K, pitch, d_ei, d_ev)
        tail = C.c_int64()
        d_tp = dev_empty(rows + 1, np.int32)
        ctx.call("cme213_csr_tail_row_ptr", rows, d_ptr, K, d_tp, C.byref(tail))
        tr, tc, tv = oracle.csr_to_coo_tail(lp, col, val, K)
        assert tail.value == len(tr)
        tp = host(d_tp)
        assert tp[0] == 0 and tp[-1] == len(tr) and np.array_equal(np.diff(tp), np.maximum(np.diff(lp) - K, 0))
        n = max(tail.value, 4)
        d_tr, d_tc, d_tv = dev_empty(n, np.int32), dev_empty(n, np.int32), dev_empty(n, dtype)
        if tail.value:
            ctx.call("cme213_csr_to_coo_tail_" + sfx, rows, d_ptr, d_col, d_val, K, None, d_tr, d_tc, d_tv)
        d_y = dev(np.full(rows, -7.0, dtype))
        ctx.call("cme213_spmv_hyb_csrtail_" + sfx, rows, pitch, K, d_ei, d_ev, d_tp if tail.value else None, d_tc, d_tv, d_x, d_y)
        assert np.array_equal(host(d_y), want), K          # ELL columns, then the tail in order = the CPU loop's order
