"""ctypes binding of libcme213_b200.so (the C ABI declared in include/cme213_b200.h).

This is the only way Python code in this repository reaches the kernels.  There is no fallback:
if the shared library is missing or a call fails, an exception is raised.
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libcme213_b200.so")

OK = 0
ERR_INVALID, ERR_CUDA, ERR_ALLOC, ERR_UNSUPPORTED, ERR_TIMEOUT = -1, -2, -3, -4, -5
SHARD_AUTO, SHARD_HALO, SHARD_GATHER, SHARD_GATHER_SPLIT = 0, 1, 2, 3
CSR_AUTO, CSR_SCALAR, CSR_VECTOR, CSR_STREAM, CSR_STREAM_PADDED, CSR_STREAM_KEEP, CSR_BINNED, CSR_STREAM_BULK = 0, 1, 2, 3, 4, 5, 6, 7
CSR_STREAM_PIPE = CSR_STREAM_BULK | (1 << 8)
CSR_AUTO_ORDERED = 8
PLAN_ORDERED = 0x100
FORMAT_CSR, FORMAT_ELL, FORMAT_HYB = 0, 1, 2


def csr_vector(lanes):
    """kernel selector for CSR-vector with `lanes` (2..32, power of two) lanes per row."""
    return CSR_VECTOR | (int(lanes) << 8)


class Cme213Error(RuntimeError):
    def __init__(self, code, text):
        super().__init__("cme213 error %d: %s" % (code, text))
        self.code = code


_p, _i, _i64, _u32, _f = C.c_void_p, C.c_int, C.c_int64, C.c_uint32, C.c_float
_pi, _pi64, _pf = C.POINTER(C.c_int), C.POINTER(C.c_int64), C.POINTER(C.c_float)

# name -> (restype, argtypes).  Every symbol include/cme213_b200.h declares is listed here;
# tests/test_abi.py checks the two stay in sync.
PROTOTYPES = {
    "cme213_ctx_create": (_i, [_i, C.POINTER(_p)]),
    "cme213_ctx_destroy": (_i, [_p]),
    "cme213_ctx_device": (_i, [_p]),
    "cme213_ctx_sm_count": (_i, [_p]),
    "cme213_ctx_launch_count": (_i64, [_p]),
    "cme213_last_error": (C.c_char_p, []),
    "cme213_version": (_i, []),
    "cme213_build_id": (C.c_char_p, []),
    "cme213_transpose_b32": (_i, [_p, _p, _p, _i64, _i64, _i64, _i64, _p]),
    "cme213_transpose_b64": (_i, [_p, _p, _p, _i64, _i64, _i64, _i64, _p]),
    "cme213_exclusive_scan_u32": (_i, [_p, _p, _p, _i64, _p]),
    "cme213_exclusive_scan_u32_static_max_n": (_i64, []),
    "cme213_exclusive_scan_u32_static": (_i, [_p, _p, _i64, _p]),
    "cme213_flag_words": (_i64, [_i64]),
    "cme213_segment_flags_from_offsets": (_i, [_p, _p, _i, _i64, _p, _p]),
    "cme213_segscan_inclusive_f32": (_i, [_p, _p, _p, _p, _i64, _p]),
    "cme213_segscan_inclusive_f64": (_i, [_p, _p, _p, _p, _i64, _p]),
    "cme213_segscan_exclusive_f32": (_i, [_p, _p, _p, _p, _i64, _p]),
    "cme213_mul_gather_segscan_f32": (_i, [_p, _p, _p, _p, _p, _p, _i64, _p]),
    "cme213_mul_gather_segscan_f64": (_i, [_p, _p, _p, _p, _p, _p, _i64, _p]),
    "cme213_gather_f32": (_i, [_p, _p, _p, _p, _i64, _p]),
    "cme213_gather_f64": (_i, [_p, _p, _p, _p, _i64, _p]),
    "cme213_mul_segscan_f32": (_i, [_p, _p, _p, _p, _p, _i64, _p]),
    "cme213_mul_segscan_f64": (_i, [_p, _p, _p, _p, _p, _i64, _p]),
    "cme213_spmv_csr_f32": (_i, [_p, _i64, _i64, _p, _p, _p, _p, _p, _i, _p]),
    "cme213_spmv_csr_f64": (_i, [_p, _i64, _i64, _p, _p, _p, _p, _p, _i, _p]),
    "cme213_csr_plan_create": (_i, [_p, _i64, _i64, _i64, _p, _i, _p, C.POINTER(_p)]),
    "cme213_csr_plan_destroy": (_i, [_p]),
    "cme213_ctx_clear_plans": (_i, [_p]),
    "cme213_csr_plan_kernel": (_i, [_p]),
    "cme213_csr_plan_describe": (C.c_char_p, [_p]),
    "cme213_spmv_csr_plan_f32": (_i, [_p, _p, _p, _p, _p, _p, _p, _p]),
    "cme213_spmv_csr_plan_f64": (_i, [_p, _p, _p, _p, _p, _p, _p, _p]),
    "cme213_spmv_pick_format": (_i, [_p, _i64, _p, _p, _p]),
    "cme213_spmv_hyb_csrtail_f32": (_i, [_p, _i64, _i64, _i, _p, _p, _p, _p, _p, _p, _p, _p]),
    "cme213_spmv_hyb_csrtail_f64": (_i, [_p, _i64, _i64, _i, _p, _p, _p, _p, _p, _p, _p, _p]),
    "cme213_csr_tail_row_ptr": (_i, [_p, _i64, _p, _i, _p, _pi64, _p]),
    "cme213_spmv_csr_block_rows": (_i, []),
    "cme213_spmv_csr_f32_ordered": (_i, [_p, _i64, _i64, _p, _p, _p, _p, _p, _p, _i64, _p]),
    "cme213_spmv_csr_f64_ordered": (_i, [_p, _i64, _i64, _p, _p, _p, _p, _p, _p, _i64, _p]),
    "cme213_spmv_csr_f64_halo_push": (_i, [_p, _i64, _i64, _p, _p, _p, _p, _p, _p, _i64, _i64, _p, _p, _p]),
    "cme213_device_malloc": (_i, [C.c_size_t, C.POINTER(_p)]),
    "cme213_device_free": (_i, [_p]),
    "cme213_ipc_handle_bytes": (_i, []),
    "cme213_ipc_export": (_i, [_p, _p]),
    "cme213_ipc_open": (_i, [_p, C.POINTER(_p)]),
    "cme213_ipc_close": (_i, [_p]),
    "cme213_partition_rows": (_i, [_i64, _i, _p, _i64, _p]),
    "cme213_shard_create_f64": (_i, [_p, _i, _i, _i64, _p, _p, _p, _p, _i, C.POINTER(_p)]),
    "cme213_shard_destroy": (_i, [_p]),
    "cme213_shard_handle_bytes": (_i, []),
    "cme213_shard_export": (_i, [_p, _p]),
    "cme213_shard_connect": (_i, [_p, _p]),
    "cme213_shard_connect_local": (_i, [C.POINTER(_p), _i]),
    "cme213_shard_mode": (_i, [_p]),
    "cme213_shard_rows": (_i, [_p, _pi64, _pi64]),
    "cme213_shard_comm_bytes_per_step": (_i64, [_p]),
    "cme213_shard_vector": (_p, [_p, _i]),
    "cme213_shard_step": (_i, [_p, _i, _p]),
    "cme213_shard_iterate": (_i, [_p, _i, _p]),
    "cme213_shard_finish": (_i, [_p, _p]),
    "cme213_csr_iterate_sharded_host_f64": (_i, [_i, _i64, _i64, _p, _p, _p, _p, _p, _i, _i, _pf]),
    "cme213_pagerank_propagate": (_i, [_p, _p, _p, _p, _p, _p, _i, _i, _p]),
    "cme213_spmv_ell_f32": (_i, [_p, _i64, _i64, _i, _p, _p, _p, _p, _p]),
    "cme213_spmv_ell_f64": (_i, [_p, _i64, _i64, _i, _p, _p, _p, _p, _p]),
    "cme213_spmv_coo_accumulate_f32": (_i, [_p, _i64, _p, _p, _p, _p, _p, _p]),
    "cme213_spmv_coo_accumulate_f64": (_i, [_p, _i64, _p, _p, _p, _p, _p, _p]),
    "cme213_spmv_hyb_f32": (_i, [_p, _i64, _i64, _i, _p, _p, _i64, _p, _p, _p, _p, _p, _p]),
    "cme213_spmv_hyb_f64": (_i, [_p, _i64, _i64, _i, _p, _p, _i64, _p, _p, _p, _p, _p, _p]),
    "cme213_hyb_pick_k": (_i, [_p, _i64, _p, _pi, _pi, _p]),
    "cme213_ell_pitch": (_i64, [_i64]),
    "cme213_csr_to_ell_f32": (_i, [_p, _i64, _i64, _p, _p, _p, _i, _i64, _p, _p, _p]),
    "cme213_csr_to_ell_f64": (_i, [_p, _i64, _i64, _p, _p, _p, _i, _i64, _p, _p, _p]),
    "cme213_csr_to_coo_tail_f32": (_i, [_p, _i64, _p, _p, _p, _i, _pi64, _p, _p, _p, _p]),
    "cme213_csr_to_coo_tail_f64": (_i, [_p, _i64, _p, _p, _p, _i, _pi64, _p, _p, _p, _p]),
    "cme213_spmv_dia_f32": (_i, [_p, _i64, _i64, _i64, _i, _p, _p, _p, _p, _p]),
    "cme213_spmv_dia_f64": (_i, [_p, _i64, _i64, _i64, _i, _p, _p, _p, _p, _p]),
    "cme213_csr_diagonals": (_i, [_p, _i64, _i64, _p, _p, _i, _p, _pi, _p]),
    "cme213_csr_to_dia_f32": (_i, [_p, _i64, _p, _p, _p, _i, _p, _i64, _p, _pi64, _p]),
    "cme213_csr_to_dia_f64": (_i, [_p, _i64, _p, _p, _p, _i, _p, _i64, _p, _pi64, _p]),
    "cme213_stencil_laplace_f32": (_i, [_p, _i, _i, _i, C.c_float, _i64, _i64, _p, _p, _p]),
    "cme213_stencil_laplace_f64": (_i, [_p, _i, _i, _i, C.c_double, _i64, _i64, _p, _p, _p]),
    "cme213_block_offsets_u32": (_i, [_p, _p, _p, _p, _i64, _i64, _p]),
    "cme213_radix_sort_pass_u32": (_i, [_p, _p, _p, _i64, _i, _i, _p]),
    "cme213_radix_sort_u32": (_i, [_p, _p, _p, _i64, _i, _p]),
    "cme213_radix_sort_pairs_u32": (_i, [_p, _p, _p, _p, _p, _i64, _i, _p]),
    "cme213_coo_to_csr_f32": (_i, [_p, _i64, _i64, _i64, _p, _p, _p, _p, _p, _p, _p]),
    "cme213_coo_to_csr_f64": (_i, [_p, _i64, _i64, _i64, _p, _p, _p, _p, _p, _p, _p]),
    "cme213_gen_laplace2d_f32": (_i, [_p, _i, _i, _p, _p, _p, _pi64, _p]),
    "cme213_gen_laplace3d_f64": (_i, [_p, _i, _i, _i, _i64, _i64, _p, _p, _p, _pi64, _p]),
    "cme213_gen_powerlaw_lengths": (_i, [_p, _i64, _i, _u32, _u32, _i, _u32, _p, _pi64, _p]),
    "cme213_gen_random_csr_fill_f32": (_i, [_p, _i64, _i64, _p, _i64, _u32, _p, _p, _p]),
    "cme213_gen_pagerank_graph": (_i, [_p, _i, _i, _u32, _p, _p, _p, _p, _pi64, _p]),
    "cme213_gen_segment_offsets": (_i, [_p, _i64, _i64, _u32, _p, _pi64, _p]),
    "cme213_fill_uniform_f32": (_i, [_p, _p, _i64, _u32, _p]),
    "cme213_fill_uniform_f64": (_i, [_p, _p, _i64, _u32, _p]),
    "cme213_fill_hash_u32": (_i, [_p, _p, _i64, _u32, _u32, _p]),
    "cme213_fill_random_flags": (_i, [_p, _p, _i64, _u32, _u32, _p]),
    "cme213_probe_gather_f32": (_i, [_p, _p, _i64, _p, _i64, _u32, _p]),
    "device_graph_iterate": (None, [_p, _p, _p, _p, _p, _i, _i, _i]),
    "cme213_pagerank_iterate_host": (_i, [_p, _p, _p, _p, _p, _i, _i, _pf]),
    "cme213_csr_upload_host_f64": (_i, [_i64, _i64, _p, _p, _p, C.POINTER(_p)]),
    "cme213_csr_upload_host_f32": (_i, [_i64, _i64, _p, _p, _p, C.POINTER(_p)]),
    "cme213_csr_iterate_f64": (_i, [_p, _p, _p, _i, _pf]),
    "cme213_csr_iterate_f32": (_i, [_p, _p, _p, _i, _pf]),
    "cme213_csr_matrix_destroy": (_i, [_p]),
    "cme213_csr_iterate_host_f64": (_i, [_i64, _i64, _p, _p, _p, _p, _p, _i, _pf]),
    "cme213_csr_iterate_host_f32": (_i, [_i64, _i64, _p, _p, _p, _p, _p, _i, _pf]),
    "cme213_fp_iterate_host_f32": (_i, [_i, _i, _i, _i, _p, _p, _p, _p, _p, _pf]),
    "cme213_fp_iterate_host_f64": (_i, [_i, _i, _i, _i, _p, _p, _p, _p, _p, _pf]),
    "cme213_exclusive_scan_u32_host": (_i, [_p, _p, _i64]),
    "cme213_radix_sort_u32_host": (_i, [_p, _i64, _i, _pf]),
    "cme213_radix_sort_pass_u32_host": (_i, [_p, _p, _i64, C.c_uint, C.c_uint, C.c_uint]),
    "cme213_transpose_host_b32": (_i, [_p, _p, _i64, _i64]),
    "cme213_transpose_host_b64": (_i, [_p, _p, _i64, _i64]),
}

_lib = None


def load():
    """Loads the shared library (once).  Raises if it has not been built: there is no CPU path."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            "libcme213_b200.so is missing (%s). Build it with `python 2012-04_stanford_cme213_b200/build.py` "
            "or __graft_entry__.build(); this package has no CPU fallback." % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in PROTOTYPES.items():
        fn = getattr(lib, name)  # AttributeError here means header and library disagree
        fn.restype = res
        fn.argtypes = args
    # the .so is a git-ignored build artefact: refuse one that was built from other sources than those beside it
    try:
        import importlib.util
        spec = importlib.util.spec_from_file_location("cme213_build", os.path.join(HERE, "build.py"))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        want = mod.source_hash()
    except Exception:  # sources not shipped beside the library: nothing to compare with
        want = None
    have = lib.cme213_build_id().decode()
    if want is not None and have.replace("-tuning", "") != want:
        raise RuntimeError("libcme213_b200.so was built from other sources (library %s, sources %s): rebuild with "
                           "`python 2012-04_stanford_cme213_b200/build.py`" % (have, want))
    _lib = lib
    return lib


def check(rc):
    if rc != OK:
        raise Cme213Error(rc, load().cme213_last_error().decode("utf-8", "replace"))


def ptr(t):
    """Device or host address of a torch tensor / numpy array / int / None as c_void_p."""
    if t is None:
        return None
    if isinstance(t, int):
        return C.c_void_p(t)
    if hasattr(t, "data_ptr"):
        return C.c_void_p(t.data_ptr())
    if hasattr(t, "ctypes"):
        return C.c_void_p(t.ctypes.data)
    raise TypeError("cannot take the address of %r" % type(t))


class FormatChoice(C.Structure):
    _fields_ = [("format", C.c_int), ("K", C.c_int), ("max_row", C.c_int), ("mean_row", C.c_double),
                ("ell_padding_ratio", C.c_double), ("tail_fraction", C.c_double), ("cost_vs_csr", C.c_double)]


class Context:
    """cme213_ctx wrapper.  `stream` arguments are torch.cuda.Stream objects or None (default stream)."""

    def __init__(self, device=0):
        self.lib = load()
        h = C.c_void_p()
        check(self.lib.cme213_ctx_create(int(device), C.byref(h)))
        self.h = h
        self.device = int(device)

    def close(self):
        if getattr(self, "h", None):
            self.lib.cme213_ctx_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def launches(self):
        return int(self.lib.cme213_ctx_launch_count(self.h))

    @property
    def sm_count(self):
        return int(self.lib.cme213_ctx_sm_count(self.h))

    def call(self, name, *args, stream=None):
        """Calls lib.<name>(ctx, *args, stream) and raises on a non-zero status."""
        s = None if stream is None else C.c_void_p(stream.cuda_stream if hasattr(stream, "cuda_stream") else int(stream))
        conv = [ptr(a) if (a is None or hasattr(a, "data_ptr") or hasattr(a, "ctypes")) else a for a in args]
        check(getattr(self.lib, name)(self.h, *conv, s))
