"""GPU: the row-sharded iterate behind the C ABI (csrc/sharded.cu).  One process drives all the GPUs the box has
through cme213_csr_iterate_sharded_host_f64 (the entry point a C++ driver uses; bin/csr_iterate is that driver), so the
multi-GPU cases need no launcher: with one GPU visible they are skipped and the single-GPU path of the same entry
point is what runs.  HALO and GATHER must reproduce the CPU loop bit for bit (row sharding keeps the order in which a
row is summed); GATHER_SPLIT sums the owned and the foreign columns of a row separately and is held to the
reduction-order bound."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

import oracle
from oracle import generators as g
from conftest import ROOT

pytestmark = pytest.mark.gpu
AUTO, HALO, GATHER, SPLIT = 0, 1, 2, 3
BIN = os.path.join(ROOT, "2012-04_stanford_cme213_b200", "bin")


def ngpus():
    import torch
    return torch.cuda.device_count()


def sharded_host(cme, gpus, ptr, col, val, x, iters, mode):
    lib = cme.load()
    y = np.empty_like(x)
    ms = C.c_float()
    P = lambda a: C.c_void_p(a.ctypes.data)
    rc = lib.cme213_csr_iterate_sharded_host_f64(gpus, len(ptr) - 1, len(col), P(ptr), P(col), P(val), P(x), P(y), iters, mode, C.byref(ms))
    assert rc == 0, lib.cme213_last_error()
    return y


def test_sharded_host_twin_on_one_gpu_matches_cpu_loop(cme):
    ptr, col, val = g.laplace3d(20, 17, 13)
    x = np.random.default_rng(1).uniform(-0.05, 0.05, 20 * 17 * 13)
    want = oracle.csr_iterate(ptr, col, val, x, 6)
    assert np.array_equal(sharded_host(cme, 1, ptr, col, val, x, 6, AUTO), want)
    lib = cme.load()
    assert lib.cme213_csr_iterate_sharded_host_f64(1, 10, 0, None, None, None, None, None, 3, 0, None) == cme.capi.ERR_INVALID   # odd
    assert lib.cme213_csr_iterate_sharded_host_f64(99, 10, 0, None, None, None, None, None, 2, 0, None) == cme.capi.ERR_INVALID


@pytest.mark.parametrize("grid,iters", [((16, 16, 24), 4), ((64, 64, 64), 10), ((40, 24, 9), 6), ((128, 128, 128), 20)])
def test_halo_exchange_bitwise_equal_to_cpu_loop(cme, grid, iters):
    """The fused exchange: one launch per pass, halo rows stored into the neighbour's vector by the kernel, passes
    ordered by flags in peer memory.  Many passes back to back with no host synchronisation in between."""
    n = ngpus()
    if n < 2:
        pytest.skip("needs two GPUs")
    ptr, col, val = g.laplace3d(*grid)
    rows = len(ptr) - 1
    x = np.random.default_rng(2).uniform(-0.05, 0.05, rows)
    want = oracle.csr_iterate(ptr, col, val, x, iters, omp=True)
    for gpus in sorted({2, min(n, 4), n}):
        if rows // gpus < 2 * grid[0] * grid[1]:
            continue
        assert np.array_equal(sharded_host(cme, gpus, ptr, col, val, x, iters, HALO), want), "%d GPUs" % gpus
        assert np.array_equal(sharded_host(cme, gpus, ptr, col, val, x, iters, AUTO), want)   # AUTO picks HALO here


@pytest.mark.parametrize("mode", [GATHER, SPLIT, AUTO])
def test_general_matrix_gather_modes(cme, mode):
    """A matrix with columns all over the place: the halo exchange cannot serve it, AUTO falls back to GATHER."""
    n = ngpus()
    if n < 2:
        pytest.skip("needs two GPUs")
    rows = 200000
    ptr, col, val = g.random_csr(rows, rows, 9.0, seed=5, dtype=np.float64)
    val = val * 0.1                                   # keeps |x| bounded over the passes
    x = np.random.default_rng(3).uniform(-1, 1, rows)
    iters = 6
    want = oracle.csr_iterate(ptr, col, val, x, iters)
    for gpus in sorted({2, n}):
        got = sharded_host(cme, gpus, ptr, col, val, x, iters, mode)
        if mode == SPLIT:
            scale = np.abs(want).max()
            assert np.all(np.abs(got - want) <= 1e-12 * scale)      # reduction order differs: ~ iters * nnz/row * 2^-53
        else:
            assert np.array_equal(got, want), "%d GPUs" % gpus
    lib = cme.load()
    y = np.empty_like(x)
    P = lambda a: C.c_void_p(a.ctypes.data)
    # HALO on this matrix: with two ranks every column lies in "the neighbour", so the halo is simply the whole slice and the
    # result must still be the CPU loop's; from three ranks on the matrix reaches beyond the neighbours and HALO is refused
    assert np.array_equal(sharded_host(cme, 2, ptr, col, val, x, iters, HALO), want)
    if n >= 3:
        rc = lib.cme213_csr_iterate_sharded_host_f64(3, rows, len(col), P(ptr), P(col), P(val), P(x), P(y), 2, HALO, None)
        assert rc == cme.capi.ERR_UNSUPPORTED, lib.cme213_last_error()   # refused, not wrong


def test_skewed_rows_are_balanced_by_nonzeros_and_stay_bitwise(cme):
    n = ngpus()
    if n < 2:
        pytest.skip("needs two GPUs")
    rows = 120000
    lens = np.where(np.arange(rows) < rows // 4, 40, 4)           # the first quarter of the rows holds 3/4 of the nonzeros
    rng = np.random.default_rng(6)
    ptr = np.concatenate([[0], np.cumsum(lens)]).astype(np.int32)
    band = 3000
    col = np.concatenate([np.sort(rng.choice(np.arange(max(0, r - band), min(rows, r + band)), l, replace=False)) for r, l in zip(range(rows), lens)]).astype(np.int32)
    val = rng.uniform(-0.03, 0.03, len(col))
    x = rng.uniform(-1, 1, rows)
    want = oracle.csr_iterate(ptr, col, val, x, 4)
    assert np.array_equal(sharded_host(cme, 2, ptr, col, val, x, 4, AUTO), want)
    lib = cme.load()
    b = (C.c_int64 * 3)()
    assert lib.cme213_partition_rows(rows, 2, C.c_void_p(ptr.ctypes.data), 128, b) == 0
    assert abs(int(ptr[b[1]]) - int(ptr[-1]) // 2) < 40 * 128 and b[1] < rows // 4


def test_csr_iterate_driver(cme):
    gpus = min(2, ngpus())
    r = subprocess.run([os.path.join(BIN, "csr_iterate"), "--gpus", str(gpus), "--grid", "96", "--iters", "10", "--check"],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "Test PASSED" in r.stdout and "0 of 884736 elements differ bitwise" in r.stdout
    assert "gpu sharded csr iterate (%d GPUs) took" % gpus in r.stdout
