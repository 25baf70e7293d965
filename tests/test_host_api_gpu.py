"""GPU parity: the host-pointer driver entry points (the drop-in boundary) vs the reference's CPU results."""
import ctypes as C

import numpy as np
import pytest

import oracle
from oracle import generators as g
from conftest import golden

pytestmark = pytest.mark.gpu


def P(a):
    return a.ctypes.data_as(C.c_void_p)


def test_device_graph_iterate_homework_signature(cme, capfd):
    # hw/hw1/programming/pagerank.cu:86-93: host pointers in/out, 20 iterations, <= 10 ULP vs the CPU
    f = golden("pagerank_ref_n5000.npz")
    n = len(f["nodes"])
    ind, edges, inv, nodes = (np.ascontiguousarray(f[k]) for k in ("indices", "edges", "inv", "nodes"))
    result = np.full(n, np.inf, np.float32)
    cme.load().device_graph_iterate(P(ind), P(edges), P(nodes), P(result), P(inv), 20, n, 8)
    assert np.array_equal(result, f["out20"])            # bit-identical to host_graph_iterate
    assert oracle.count_ulp_errors(result, f["out20"], 10) == 0
    assert "gpu graph propagate took" in capfd.readouterr().out


def test_pagerank_iterate_host_larger_graph_and_errors(cme):
    lib = cme.load()
    ind, edges, inv, nodes = g.pagerank_graph(1 << 17, 8, seed=31)
    result = np.zeros_like(nodes)
    ms = C.c_float()
    assert lib.cme213_pagerank_iterate_host(P(ind), P(edges), P(nodes), P(result), P(inv), 6, len(nodes), C.byref(ms)) == 0
    assert np.array_equal(result, oracle.pagerank_iterate(ind, edges, nodes, inv, 6, omp=True)) and ms.value > 0
    assert lib.cme213_pagerank_iterate_host(P(ind), P(edges), P(nodes), P(result), P(inv), 3, len(nodes), None) == cme.capi.ERR_INVALID
    assert b"even" in lib.cme213_last_error()
    assert lib.cme213_pagerank_iterate_host(None, P(edges), P(nodes), P(result), P(inv), 2, len(nodes), None) == cme.capi.ERR_INVALID


@pytest.mark.parametrize("dtype,fn", [(np.float64, "cme213_csr_iterate_host_f64"), (np.float32, "cme213_csr_iterate_host_f32")])
def test_csr_iterate_host(cme, dtype, fn):
    ptr, col, val = g.laplace3d(12, 11, 10, dtype=dtype)
    n = 12 * 11 * 10
    x = np.random.default_rng(1).uniform(0, 1, n).astype(dtype)
    out = np.zeros_like(x)
    assert getattr(cme.load(), fn)(n, len(col), P(ptr), P(col), P(val), P(x), P(out), 4, None) == 0
    assert np.array_equal(out, oracle.csr_iterate(ptr, col, val, x, 4))
    assert getattr(cme.load(), fn)(n, len(col), P(ptr), P(col), P(val), P(x), P(out), 5, None) == cme.capi.ERR_INVALID


def test_fp_iterate_host_fixture_and_golden(cme):
    lib = cme.load()
    f = golden("fp_small.npz")
    a, s, k, x = f["a"].astype(np.float32), f["s"], f["k"], f["x"].astype(np.float32)
    b = np.zeros(4, np.float32)
    assert lib.cme213_fp_iterate_host_f32(4, 3, 3, 1, P(a), P(s), P(k), P(x), P(b), None) == 0
    assert np.allclose(b, f["b"], atol=2e-6, rtol=0)
    f = golden("fp_ref_n6000.npz")
    n, p, q = len(f["a"]), len(f["s"]), len(f["x"])
    b64 = np.zeros(n)
    ms = C.c_float()
    assert lib.cme213_fp_iterate_host_f64(n, p, q, 3, P(f["a"].copy()), P(f["s"].copy()), P(f["k"].copy()), P(f["x"].copy()), P(b64), C.byref(ms)) == 0
    norms = oracle.error_norms(f["b3"], b64)
    assert norms[2] < 1e-13 and norms[3] < 1e-13
    b32 = np.zeros(n, np.float32)
    a32, x32 = f["a"].astype(np.float32), f["x"].astype(np.float32)
    assert lib.cme213_fp_iterate_host_f32(n, p, q, 3, P(a32), P(f["s"].copy()), P(f["k"].copy()), P(x32), P(b32), None) == 0
    norms = oracle.error_norms(f["b3"], b32.astype(np.float64))
    assert norms[2] < 1e-5 and norms[3] < 1e-5   # the report's fp32-vs-double level (Final_1 pp.1-3)


def test_fp_iterate_host_rejects_malformed_input(cme):
    # aux/mp1-util.h:118-147 asserts these; the library returns an error instead
    lib = cme.load()
    a = np.ones(6, np.float32); x = np.ones(2, np.float32); b = np.zeros(6, np.float32)
    k = np.zeros(6, np.int32)
    for s, kk in [([1, 6], k), ([0, 5], k), ([0, 3, 3, 6], k), ([0, 6], np.array([0, 0, 2, 0, 0, 0], np.int32)),
                  ([0, 6], np.array([0, -1, 0, 0, 0, 0], np.int32))]:
        s = np.array(s, np.int32)
        assert lib.cme213_fp_iterate_host_f32(6, len(s), 2, 1, P(a), P(s), P(kk), P(x), P(b), None) == cme.capi.ERR_INVALID


def test_primitive_host_twins(cme):
    lib = cme.load()
    x = np.random.default_rng(2).integers(0, 2**32, 70001, dtype=np.uint64).astype(np.uint32)
    out = np.zeros_like(x)
    assert lib.cme213_exclusive_scan_u32_host(P(x), P(out), len(x)) == 0
    assert np.array_equal(out, oracle.exclusive_scan_u32(x))
    # one LSD radix pass built on the GPU scan sorts like the reference's pass (radixsort.cpp:123-153)
    keys = np.random.default_rng(3).integers(0, 2**32, 40000, dtype=np.uint64).astype(np.uint32)
    k = keys
    for start in range(0, 32, 8):
        hist = np.bincount((k >> start) & 255, minlength=256).astype(np.uint32)
        ex = np.zeros_like(hist)
        assert lib.cme213_exclusive_scan_u32_host(P(hist), P(ex), 256) == 0
        dest = ex[(k >> start) & 255] + (np.argsort(np.argsort((k >> start) & 255, kind="stable"), kind="stable").astype(np.uint32)
                                          - ex[(k >> start) & 255])
        nk = np.empty_like(k); nk[dest] = k
        assert np.array_equal(nk, oracle.radix_pass_u32(k, start, 8))
        k = nk
    assert np.array_equal(k, np.sort(keys))
    for dt, fn in [(np.float32, lib.cme213_transpose_host_b32), (np.float64, lib.cme213_transpose_host_b64)]:
        a = np.random.default_rng(4).standard_normal((333, 517)).astype(dt)
        t = np.zeros((517, 333), dt)
        assert fn(P(a), P(t), 333, 517) == 0
        assert np.array_equal(t, a.T)


@pytest.mark.parametrize("dtype,sfx", [(np.float64, "f64"), (np.float32, "f32")])
def test_csr_matrix_stays_resident_between_iterate_calls(cme, dtype, sfx):
    """Upload once, iterate many times with different x: only the vectors cross the bus on the later calls."""
    lib = cme.load()
    ptr, col, val = g.laplace3d(14, 9, 11, dtype=dtype)
    n = 14 * 9 * 11
    h = C.c_void_p()
    assert getattr(lib, "cme213_csr_upload_host_" + sfx)(n, len(col), P(ptr), P(col), P(val), C.byref(h)) == 0
    iterate = getattr(lib, "cme213_csr_iterate_" + sfx)
    for seed, iters in ((1, 2), (2, 6), (3, 0)):
        x = np.random.default_rng(seed).uniform(0, 0.2, n).astype(dtype)
        out = np.full_like(x, -1)
        ms = C.c_float(-1)
        assert iterate(h, P(x), P(out), iters, C.byref(ms)) == 0, lib.cme213_last_error()
        assert np.array_equal(out, oracle.csr_iterate(ptr, col, val, x, iters) if iters else x) and ms.value >= 0
    assert iterate(h, P(x), P(out), 3, None) == cme.capi.ERR_INVALID                      # odd, as pagerank.cu:127 asserts
    other = getattr(lib, "cme213_csr_iterate_" + ("f32" if sfx == "f64" else "f64"))
    assert other(h, P(x), P(out), 2, None) == cme.capi.ERR_INVALID                       # wrong precision for this matrix
    assert lib.cme213_csr_matrix_destroy(h) == 0
    assert lib.cme213_csr_matrix_destroy(None) == 0
