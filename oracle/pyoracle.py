"""numpy-facing wrappers around liboracle.so (oracle.c) and, when built, oracle/_ref/*.so (the
reference's own CPU code compiled from /root/reference).  TEST INFRASTRUCTURE ONLY."""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None
_REF = {}


def build(quiet=True):
    """make -C oracle: liboracle.so always; _ref/ only where /root/reference exists."""
    r = subprocess.run(["make", "-C", HERE, "-j4"], capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("oracle build failed:\n" + r.stdout + r.stderr)
    if not quiet:
        print(r.stdout)


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(HERE, "liboracle.so")
        if not os.path.exists(path):
            build()
        _LIB = C.CDLL(path)
        _LIB.oracle_hyb_tail_count.restype = C.c_int64
        _LIB.oracle_count_ulp_errors_f32.restype = C.c_int64
    return _LIB


def ref_lib(name):
    """oracle/_ref/libref_<name>.so or None when it was not built (no reference checkout, no prebuilt copy)."""
    if name not in _REF:
        path = os.path.join(HERE, "_ref", "libref_%s.so" % name)
        _REF[name] = C.CDLL(path) if os.path.exists(path) else None
    return _REF[name]


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _c(a, dt):
    return np.ascontiguousarray(a, dtype=dt)


def num_threads():
    return int(lib().oracle_num_threads())


# ---- PageRank ------------------------------------------------------------------------------------
def pagerank_propagate(indices, edges, nodes_in, inv, omp=False):
    indices, edges = _c(indices, np.uint32), _c(edges, np.uint32)
    nodes_in, inv = _c(nodes_in, np.float32), _c(inv, np.float32)
    out = np.empty_like(nodes_in)
    f = lib().oracle_pagerank_propagate_omp if omp else lib().oracle_pagerank_propagate
    f(_p(indices), _p(edges), _p(nodes_in), _p(out), _p(inv), C.c_int(len(nodes_in)))
    return out


def pagerank_iterate(indices, edges, nodes, inv, iters, omp=False):
    indices, edges = _c(indices, np.uint32), _c(edges, np.uint32)
    A, inv = _c(nodes, np.float32).copy(), _c(inv, np.float32)
    B = np.empty_like(A)
    rc = lib().oracle_pagerank_iterate(_p(indices), _p(edges), _p(A), _p(B), _p(inv), C.c_int(iters), C.c_int(len(A)),
                                       C.c_int(1 if omp else 0))
    if rc != 0:
        raise ValueError("nr_iterations must be even")
    return A


def ref_pagerank_iterate(indices, edges, nodes, inv, iters):
    """The reference's own host_graph_iterate (pagerank.cu:59-67) from oracle/_ref."""
    r = ref_lib("pagerank")
    indices, edges = _c(indices, np.uint32), _c(edges, np.uint32)
    A, inv = _c(nodes, np.float32).copy(), _c(inv, np.float32)
    B = np.empty_like(A)
    r.ref_host_graph_iterate(_p(indices), _p(edges), _p(A), _p(B), _p(inv), C.c_int(iters), C.c_int(len(A)))
    return A


def ref_almost_equal(a, b, max_ulps):
    r = ref_lib("pagerank")
    return bool(r.ref_almost_equal_2s_complement(C.c_float(a), C.c_float(b), C.c_int(max_ulps)))


# ---- CSR -----------------------------------------------------------------------------------------
def csr_spmv(ptr, col, val, x, omp=False):
    ptr, col = _c(ptr, np.int32), _c(col, np.int32)
    dt = np.float64 if np.asarray(val).dtype == np.float64 else np.float32
    val, x = _c(val, dt), _c(x, dt)
    y = np.empty(len(ptr) - 1, dtype=dt)
    name = "oracle_csr_spmv_%s%s" % ("f64" if dt == np.float64 else "f32", "_omp" if omp else "")
    getattr(lib(), name)(C.c_int64(len(ptr) - 1), _p(ptr), _p(col), _p(val), _p(x), _p(y))
    return y


def laplace3d_host(nx, ny, nz, row_begin=0, row_end=None):
    """generators.laplace3d built by the C restatement (OpenMP): the same arrays, fast enough for the whole 512^3 grid."""
    n = nx * ny * nz
    row_end = n if row_end is None else row_end
    rows = row_end - row_begin
    L = lib()
    L.oracle_gen_laplace3d_f64.restype = C.c_int64
    ptr = np.empty(rows + 1, np.int32)
    nnz = L.oracle_gen_laplace3d_f64(C.c_int(nx), C.c_int(ny), C.c_int(nz), C.c_int64(row_begin), C.c_int64(row_end), _p(ptr), None, None)
    col, val = np.empty(nnz, np.int32), np.empty(nnz, np.float64)
    L.oracle_gen_laplace3d_f64(C.c_int(nx), C.c_int(ny), C.c_int(nz), C.c_int64(row_begin), C.c_int64(row_end), _p(ptr), _p(col), _p(val))
    return ptr, col, val


def csr_spmv_into(ptr, col, val, x, y, omp=True):
    """csr_spmv on preallocated, already contiguous arrays (no copies: for timing loops over multi-GB matrices)."""
    f64 = val.dtype == np.float64
    name = "oracle_csr_spmv_%s%s" % ("f64" if f64 else "f32", "_omp" if omp else "")
    getattr(lib(), name)(C.c_int64(len(ptr) - 1), _p(ptr), _p(col), _p(val), _p(x), _p(y))
    return y


def csr_abs_bound(ptr, col, val, x):
    ptr, col = _c(ptr, np.int32), _c(col, np.int32)
    dt = np.float64 if np.asarray(val).dtype == np.float64 else np.float32
    val, x = _c(val, dt), _c(x, dt)
    b = np.empty(len(ptr) - 1, dtype=np.float64)
    name = "oracle_csr_abs_bound_%s" % ("f64" if dt == np.float64 else "f32")
    getattr(lib(), name)(C.c_int64(len(ptr) - 1), _p(ptr), _p(col), _p(val), _p(x), _p(b))
    return b


def csr_iterate(ptr, col, val, x, iters, omp=False):
    ptr, col = _c(ptr, np.int32), _c(col, np.int32)
    dt = np.float64 if np.asarray(val).dtype == np.float64 else np.float32
    val, A = _c(val, dt), _c(x, dt).copy()
    B = np.empty_like(A)
    name = "oracle_csr_iterate_%s" % ("f64" if dt == np.float64 else "f32")
    rc = getattr(lib(), name)(C.c_int64(len(ptr) - 1), _p(ptr), _p(col), _p(val), _p(A), _p(B), C.c_int(iters),
                              C.c_int(1 if omp else 0))
    if rc != 0:
        raise ValueError("nr_iterations must be even")
    return A


# ---- final project ---------------------------------------------------------------------------------
def sparse_mv_scan(a, s, k, x, iters=1, dtype=np.float64):
    a, x = _c(a, dtype).copy(), _c(x, dtype)
    s, k = _c(s, np.int32), _c(k, np.int32)
    b = np.zeros_like(a)
    name = "oracle_sparse_mv_scan_iterate_%s" % ("f64" if dtype == np.float64 else "f32")
    getattr(lib(), name)(C.c_int(len(a)), C.c_int(len(s)), _p(a), _p(s), _p(k), _p(x), _p(b), C.c_int(iters))
    return b


def fp_cpu_iterate(a, s, k, x, iters):
    """fp.cu:124-150 as shipped: xx = x[k] hoisted, OpenMP multiply + per-segment running sum, fp32."""
    cur = _c(a, np.float32).copy()
    s = _c(s, np.int32)
    xx = _c(np.asarray(x, dtype=np.float32)[np.asarray(k)], np.float32)
    lib().oracle_fp_cpu_iterate_f32(C.c_int(len(cur)), C.c_int(len(s)), _p(cur), _p(s), _p(xx), C.c_int(iters))
    return cur


def ref_sparse_mv_scan(a, s, k, x, iters=1):
    """The instructors' sparseMV_scan + swap loop (reference_spMVscan-released.cu:38-54,213-217) from oracle/_ref."""
    r = ref_lib("spmvscan")
    a, x = _c(a, np.float64), _c(x, np.float64)
    s, k = _c(s, np.int32), _c(k, np.int32)
    b = np.zeros_like(a)
    r.ref_sparse_mv_scan_iterate(C.c_int(len(a)), C.c_int(len(s)), C.c_int(len(x)), C.c_int(iters), _p(a), _p(s), _p(k),
                                 _p(x), _p(b))
    return b


def segscan_inclusive(vals, s):
    dt = np.float64 if np.asarray(vals).dtype == np.float64 else np.float32
    vals, s = _c(vals, dt), _c(s, np.int32)
    out = np.zeros_like(vals)
    name = "oracle_segscan_inclusive_%s" % ("f64" if dt == np.float64 else "f32")
    getattr(lib(), name)(C.c_int64(len(vals)), C.c_int(len(s)), _p(vals), _p(s), _p(out))
    return out


def segscan_flags_inclusive(vals, head):
    head = _c(head, np.uint8)
    if np.asarray(vals).dtype.kind in "iu":
        vals = _c(vals, np.int32)
        out = np.empty_like(vals)
        lib().oracle_segscan_flags_inclusive_i32(C.c_int64(len(vals)), _p(vals), _p(head), _p(out))
    else:
        vals = _c(vals, np.float32)
        out = np.empty_like(vals)
        lib().oracle_segscan_flags_inclusive_f32(C.c_int64(len(vals)), _p(vals), _p(head), _p(out))
    return out


def error_norms(a, b):
    """[abs L2, abs Linf, rel L2, rel Linf] of b against a (reference_spMVscan-released.cu:65-144)."""
    a, b = _c(a, np.float64), _c(b, np.float64)
    out = np.zeros(4)
    lib().oracle_error_norms(_p(a), _p(b), C.c_int64(len(a)), _p(out))
    return out


def ref_error_norms(a, b):
    r = ref_lib("spmvscan")
    a, b = _c(a, np.float64), _c(b, np.float64)
    out = np.zeros(4)
    r.ref_error_norms(_p(a), _p(b), C.c_int(len(a)), _p(out))
    return out


# ---- scan / transpose ------------------------------------------------------------------------------
def exclusive_scan_u32(x, omp=False):
    x = _c(x, np.uint32)
    out = np.empty_like(x)
    (lib().oracle_exclusive_scan_u32_omp if omp else lib().oracle_exclusive_scan_u32)(_p(x), _p(out), C.c_int64(len(x)))
    return out


def block_offsets_u32(local, global_excl):
    local = _c(local, np.uint32)
    blocks, buckets = local.shape
    global_excl = _c(global_excl, np.uint32)
    out = np.empty_like(local)
    lib().oracle_block_offsets_u32(_p(local), _p(global_excl), _p(out), C.c_int64(blocks), C.c_int64(buckets))
    return out


def radix_pass_u32(keys, start_bit, num_bits):
    keys = _c(keys, np.uint32)
    out = np.empty_like(keys)
    lib().oracle_radix_pass_u32(_p(keys), _p(out), C.c_int64(len(keys)), C.c_uint32(start_bit), C.c_uint32(num_bits))
    return out


def radix_pass_blocked_u32(keys, start_bit, num_bits, block_size=8192):
    """radixSortParallelPass restated (radixsort.cpp:44-109)."""
    keys = _c(keys, np.uint32)
    out = np.empty_like(keys)
    lib().oracle_radix_pass_blocked_u32(_p(keys), _p(out), C.c_int64(len(keys)), C.c_uint32(start_bit),
                                        C.c_uint32(num_bits), C.c_int64(block_size))
    return out


def radix_sort_u32(keys, num_bits=8, block_size=8192):
    """radixSortParallel (block_size > 0) / radixSortSerial (block_size 0) restated (radixsort.cpp:111-121, :155-161)."""
    keys = _c(keys, np.uint32).copy()
    tmp = np.empty_like(keys)
    lib().oracle_radix_sort_u32(_p(keys), _p(tmp), C.c_int64(len(keys)), C.c_uint32(num_bits), C.c_int64(block_size))
    return keys


def ref_radix_pass(keys, start_bit, num_bits, block_size=8192):
    """The reference's own radixSortParallelPass (block_size > 0) or radixSortSerialPass (0) from oracle/_ref."""
    r = ref_lib("radixsort")
    keys = _c(keys, np.uint32)
    out = np.empty_like(keys)
    if block_size:
        r.ref_radix_pass_parallel(_p(keys), _p(out), C.c_long(len(keys)), C.c_uint(num_bits), C.c_uint(start_bit),
                                  C.c_uint(block_size))
    else:
        r.ref_radix_pass_serial(_p(keys), _p(out), C.c_long(len(keys)), C.c_uint(num_bits), C.c_uint(start_bit))
    return out


def ref_radix_sort(keys, num_bits=8):
    """The reference's parallel sort loop (radixsort.cpp:116-120) from oracle/_ref; returns (sorted, seconds)."""
    r = ref_lib("radixsort")
    r.ref_radix_sort_parallel.restype = C.c_double
    keys = _c(keys, np.uint32).copy()
    secs = r.ref_radix_sort_parallel(_p(keys), C.c_long(len(keys)), C.c_uint(num_bits))
    return keys, secs


def transpose(a, omp=False):
    a = np.ascontiguousarray(a)
    M, N = a.shape
    out = np.empty((N, M), dtype=a.dtype)
    bits = {4: "b32", 8: "b64"}[a.dtype.itemsize]
    getattr(lib(), "oracle_transpose_%s%s" % (bits, "_omp" if omp else ""))(_p(a), _p(out), C.c_int64(M), C.c_int64(N))
    return out


# ---- ELL / COO / HYB ---------------------------------------------------------------------------------
def _dt(val):
    return np.float64 if np.asarray(val).dtype == np.float64 else np.float32


def hyb_pick_k(ptr):
    ptr = _c(ptr, np.int32)
    return int(lib().oracle_hyb_pick_k(C.c_int64(len(ptr) - 1), _p(ptr)))


def csr_to_ell(ptr, col, val, num_cols, K, pitch):
    ptr, col = _c(ptr, np.int32), _c(col, np.int32)
    dt = _dt(val)
    val = _c(val, dt)
    idx = np.zeros(K * pitch, dtype=np.int32)
    dat = np.zeros(K * pitch, dtype=dt)
    name = "oracle_csr_to_ell_%s" % ("f64" if dt == np.float64 else "f32")
    getattr(lib(), name)(C.c_int64(len(ptr) - 1), C.c_int64(num_cols), _p(ptr), _p(col), _p(val), C.c_int(K),
                         C.c_int64(pitch), _p(idx), _p(dat))
    return idx, dat


def ell_spmv(num_rows, pitch, K, idx, dat, x):
    dt = _dt(dat)
    idx, dat, x = _c(idx, np.int32), _c(dat, dt), _c(x, dt)
    y = np.empty(num_rows, dtype=dt)
    name = "oracle_ell_spmv_%s" % ("f64" if dt == np.float64 else "f32")
    getattr(lib(), name)(C.c_int64(num_rows), C.c_int64(pitch), C.c_int(K), _p(idx), _p(dat), _p(x), _p(y))
    return y


def csr_diagonals(ptr, col, num_cols, max_out=1 << 20):
    ptr, col = _c(ptr, np.int32), _c(col, np.int32)
    out = np.zeros(max_out, dtype=np.int32)
    lib().oracle_csr_diagonals.restype = C.c_int64
    n = int(lib().oracle_csr_diagonals(C.c_int64(len(ptr) - 1), C.c_int64(num_cols), _p(ptr), _p(col), C.c_int64(max_out), _p(out)))
    return n, out[:min(n, max_out)].copy()


def csr_to_dia(ptr, col, val, offsets, pitch):
    ptr, col, offsets = _c(ptr, np.int32), _c(col, np.int32), _c(offsets, np.int32)
    dt = _dt(val)
    val = _c(val, dt)
    dat = np.zeros(len(offsets) * pitch, dtype=dt)
    fn = getattr(lib(), "oracle_csr_to_dia_%s" % ("f64" if dt == np.float64 else "f32"))
    fn.restype = C.c_int64
    miss = int(fn(C.c_int64(len(ptr) - 1), _p(ptr), _p(col), _p(val), C.c_int(len(offsets)), _p(offsets), C.c_int64(pitch), _p(dat)))
    return dat, miss


def dia_spmv(num_rows, num_cols, pitch, offsets, dat, x):
    dt = _dt(dat)
    offsets, dat, x = _c(offsets, np.int32), _c(dat, dt), _c(x, dt)
    y = np.empty(num_rows, dtype=dt)
    name = "oracle_dia_spmv_%s" % ("f64" if dt == np.float64 else "f32")
    getattr(lib(), name)(C.c_int64(num_rows), C.c_int64(num_cols), C.c_int64(pitch), C.c_int(len(offsets)), _p(offsets), _p(dat), _p(x), _p(y))
    return y


def csr_to_coo_tail(ptr, col, val, K):
    ptr, col = _c(ptr, np.int32), _c(col, np.int32)
    dt = _dt(val)
    val = _c(val, dt)
    n = int(lib().oracle_hyb_tail_count(C.c_int64(len(ptr) - 1), _p(ptr), C.c_int(K)))
    r, c, v = np.zeros(n, np.int32), np.zeros(n, np.int32), np.zeros(n, dt)
    name = "oracle_csr_to_coo_tail_%s" % ("f64" if dt == np.float64 else "f32")
    getattr(lib(), name)(C.c_int64(len(ptr) - 1), _p(ptr), _p(col), _p(val), C.c_int(K), _p(r), _p(c), _p(v))
    return r, c, v


def coo_spmv_accumulate(rows, cols, vals, x, y):
    dt = _dt(vals)
    rows, cols, vals, x = _c(rows, np.int32), _c(cols, np.int32), _c(vals, dt), _c(x, dt)
    y = _c(y, dt).copy()
    name = "oracle_coo_spmv_accumulate_%s" % ("f64" if dt == np.float64 else "f32")
    getattr(lib(), name)(C.c_int64(len(rows)), _p(rows), _p(cols), _p(vals), _p(x), _p(y))
    return y


# ---- comparators -------------------------------------------------------------------------------------
def count_ulp_errors(got, want, max_ulps=10):
    got, want = _c(got, np.float32), _c(want, np.float32)
    return int(lib().oracle_count_ulp_errors_f32(_p(got), _p(want), C.c_int64(len(got)), C.c_int(max_ulps)))


def almost_equal_ulps(a, b, max_ulps):
    return bool(lib().oracle_almost_equal_ulps_f32(C.c_float(a), C.c_float(b), C.c_int(max_ulps)))
