// spmv_ell.cu -- ELL and HYB sparse matrix-vector multiply.
//
// Format and kernel semantics: refs/Bell SC 2009.pdf p.3 Fig. 2/4 and section 3.2; Bell TR 2008
// Fig. 17 (spmv_ell_kernel: column-major data[pitch*n + row], skip val == 0);
// slides/Lecture22.pdf slides 26-28.  HYB = ELL(K) + COO tail, SC09 section 3.5; the tail runs
// through the segmented-reduction scan in scan.cu.
//
// B200 shape: a thread owns FOUR consecutive rows, so each of its loads from the column-major
// arrays is a 128-bit LDG (int4 of column indices, float4 / 2 x double2 of values) and a warp
// covers 512 contiguous bytes per array per ELL column.  The column loop is unrolled so several
// independent index/value loads are in flight before the dependent x gathers issue.  Padded
// entries (value 0) skip the gather.  Sums are sequential in ELL-column order with rounded
// multiply and add, i.e. the same order as the CSR row loop: bit-identical to the CPU loop.
#include "internal.cuh"

namespace {
using namespace cme213;

__device__ __forceinline__ float mul_rn(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ double mul_rn(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ float add_rn(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ double add_rn(double a, double b) { return __dadd_rn(a, b); }

__device__ __forceinline__ void load4(const float *p, float (&v)[4]) {
    float4 t = ld_stream(reinterpret_cast<const float4 *>(p));
    v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
}
__device__ __forceinline__ void load4(const double *p, double (&v)[4]) {
    double2 t0 = ld_stream(reinterpret_cast<const double2 *>(p));
    double2 t1 = ld_stream(reinterpret_cast<const double2 *>(p) + 1);
    v[0] = t0.x; v[1] = t0.y; v[2] = t1.x; v[3] = t1.y;
}

__device__ __forceinline__ void store4(float *p, const float (&v)[4]) { *reinterpret_cast<float4 *>(p) = make_float4(v[0], v[1], v[2], v[3]); }
__device__ __forceinline__ void store4(double *p, const double (&v)[4]) {
    reinterpret_cast<double2 *>(p)[0] = make_double2(v[0], v[1]);
    reinterpret_cast<double2 *>(p)[1] = make_double2(v[2], v[3]);
}

template <typename T, int UNROLL>
__global__ void __launch_bounds__(256)
ell_v4_kernel(int64_t num_rows, int64_t pitch, int K, const int32_t *__restrict__ idx, const T *__restrict__ val,
              const T *__restrict__ x, T *__restrict__ y) {
    const int64_t r4 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (r4 >= num_rows) return;
    T s[4] = {T(0), T(0), T(0), T(0)};
    // the matrix streams by once (evict-first loads) while x is gathered at random and re-used: keep x in L2
    const uint64_t keep = l2_policy_evict_last();
    const int32_t *ip = idx + r4;
    const T *vp = val + r4;
    int n = 0;
    for (; n + UNROLL <= K; n += UNROLL) {
        int4 c[UNROLL];
        T v[UNROLL][4];
#pragma unroll
        for (int u = 0; u < UNROLL; u++) {
            c[u] = ld_stream(reinterpret_cast<const int4 *>(ip + (int64_t)(n + u) * pitch));
            load4(vp + (int64_t)(n + u) * pitch, v[u]);
        }
#pragma unroll
        for (int u = 0; u < UNROLL; u++) {
            if (v[u][0] != T(0)) s[0] = add_rn(s[0], mul_rn(v[u][0], ld_keep(x + c[u].x, keep)));
            if (v[u][1] != T(0)) s[1] = add_rn(s[1], mul_rn(v[u][1], ld_keep(x + c[u].y, keep)));
            if (v[u][2] != T(0)) s[2] = add_rn(s[2], mul_rn(v[u][2], ld_keep(x + c[u].z, keep)));
            if (v[u][3] != T(0)) s[3] = add_rn(s[3], mul_rn(v[u][3], ld_keep(x + c[u].w, keep)));
        }
    }
    for (; n < K; n++) {
        const int4 c = ld_stream(reinterpret_cast<const int4 *>(ip + (int64_t)n * pitch));
        T v[4];
        load4(vp + (int64_t)n * pitch, v);
        if (v[0] != T(0)) s[0] = add_rn(s[0], mul_rn(v[0], ld_keep(x + c.x, keep)));
        if (v[1] != T(0)) s[1] = add_rn(s[1], mul_rn(v[1], ld_keep(x + c.y, keep)));
        if (v[2] != T(0)) s[2] = add_rn(s[2], mul_rn(v[2], ld_keep(x + c.z, keep)));
        if (v[3] != T(0)) s[3] = add_rn(s[3], mul_rn(v[3], ld_keep(x + c.w, keep)));
    }
    if (r4 + 4 <= num_rows && (reinterpret_cast<uintptr_t>(y + r4) & 15u) == 0) {
        store4(y + r4, s);   // one 128-bit (fp32) / two 128-bit (fp64) stores instead of four scalar ones
    } else {
#pragma unroll
        for (int t = 0; t < 4; t++)
            if (r4 + t < num_rows) y[r4 + t] = s[t];
    }
}

// Any pitch / alignment: one row per thread (Bell TR Fig. 17 as written).
template <typename T>
__global__ void __launch_bounds__(256)
ell_scalar_kernel(int64_t num_rows, int64_t pitch, int K, const int32_t *__restrict__ idx, const T *__restrict__ val,
                  const T *__restrict__ x, T *__restrict__ y) {
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= num_rows) return;
    T s = T(0);
    for (int n = 0; n < K; n++) {
        const T v = val[(int64_t)n * pitch + r];
        if (v != T(0)) s = add_rn(s, mul_rn(v, __ldg(x + idx[(int64_t)n * pitch + r])));
    }
    y[r] = s;
}

template <typename T>
int ell_impl(cme213_ctx *ctx, int64_t num_rows, int64_t pitch, int K, const int32_t *d_idx, const T *d_val,
             const T *d_x, T *d_y, void *stream) {
    CME213_CHECK_CTX(ctx);
    if (num_rows < 0 || K < 0 || pitch < num_rows) { set_error("spmv_ell: bad shape"); return CME213_ERR_INVALID; }
    if (num_rows == 0) return CME213_OK;
    if (!d_x || !d_y || (K > 0 && (!d_idx || !d_val))) { set_error("spmv_ell: null pointer"); return CME213_ERR_INVALID; }
    DeviceGuard g(ctx->device);
    cudaStream_t st = (cudaStream_t)stream;
    const bool vec_ok = pitch % 4 == 0 && pitch >= ((num_rows + 3) & ~(int64_t)3) && aligned16(d_idx) && aligned16(d_val);
    if (vec_ok) {
        const int64_t blocks = div_up(div_up(num_rows, 4), 256);
        ell_v4_kernel<T, 4><<<(unsigned)blocks, 256, 0, st>>>(num_rows, pitch, K, d_idx, d_val, d_x, d_y);
        CME213_LAUNCHED(ctx, "ell_v4_kernel");
    } else {
        ell_scalar_kernel<T><<<(unsigned)div_up(num_rows, 256), 256, 0, st>>>(num_rows, pitch, K, d_idx, d_val, d_x, d_y);
        CME213_LAUNCHED(ctx, "ell_scalar_kernel");
    }
    return CME213_OK;
}
}  // namespace

CME213_API int cme213_spmv_ell_f32(cme213_ctx *ctx, int64_t num_rows, int64_t pitch, int num_cols_per_row,
                                   const int32_t *d_ell_idx, const float *d_ell_val, const float *d_x, float *d_y,
                                   void *stream) {
    return ell_impl<float>(ctx, num_rows, pitch, num_cols_per_row, d_ell_idx, d_ell_val, d_x, d_y, stream);
}
CME213_API int cme213_spmv_ell_f64(cme213_ctx *ctx, int64_t num_rows, int64_t pitch, int num_cols_per_row,
                                   const int32_t *d_ell_idx, const double *d_ell_val, const double *d_x, double *d_y,
                                   void *stream) {
    return ell_impl<double>(ctx, num_rows, pitch, num_cols_per_row, d_ell_idx, d_ell_val, d_x, d_y, stream);
}

CME213_API int cme213_spmv_hyb_f32(cme213_ctx *ctx, int64_t num_rows, int64_t pitch, int K, const int32_t *d_ell_idx,
                                   const float *d_ell_val, int64_t coo_nnz, const int32_t *d_coo_row,
                                   const int32_t *d_coo_col, const float *d_coo_val, const float *d_x, float *d_y,
                                   void *stream) {
    int rc = cme213_spmv_ell_f32(ctx, num_rows, pitch, K, d_ell_idx, d_ell_val, d_x, d_y, stream);
    if (rc) return rc;
    return cme213_spmv_coo_accumulate_f32(ctx, coo_nnz, d_coo_row, d_coo_col, d_coo_val, d_x, d_y, stream);
}

CME213_API int cme213_spmv_hyb_f64(cme213_ctx *ctx, int64_t num_rows, int64_t pitch, int K, const int32_t *d_ell_idx,
                                   const double *d_ell_val, int64_t coo_nnz, const int32_t *d_coo_row,
                                   const int32_t *d_coo_col, const double *d_coo_val, const double *d_x, double *d_y,
                                   void *stream) {
    int rc = cme213_spmv_ell_f64(ctx, num_rows, pitch, K, d_ell_idx, d_ell_val, d_x, d_y, stream);
    if (rc) return rc;
    return cme213_spmv_coo_accumulate_f64(ctx, coo_nnz, d_coo_row, d_coo_col, d_coo_val, d_x, d_y, stream);
}

// HYB with the tail kept as a CSR matrix of its own (tail_ptr from cme213_csr_tail_row_ptr; columns / values are the
// COO tail's arrays).  The tail goes through the CSR stream kernel in accumulate mode: 8 bytes per entry instead of
// COO's 12, coalesced read-modify-write of y for the rows that have a tail, and each tail row is summed left to right
// starting from its ELL sum -- the same bits as the CSR loop.  (A load-balanced tail that reduces each tile by a segmented
// scan was measured slower, 0.840 vs 0.808 ms on the banded power-law matrix, and dropped.)
CME213_API int cme213_spmv_hyb_csrtail_f32(cme213_ctx *ctx, int64_t num_rows, int64_t pitch, int K, const int32_t *d_ell_idx,
                                           const float *d_ell_val, const int32_t *d_tail_ptr, const int32_t *d_tail_col,
                                           const float *d_tail_val, const float *d_x, float *d_y, void *stream) {
    int rc = cme213_spmv_ell_f32(ctx, num_rows, pitch, K, d_ell_idx, d_ell_val, d_x, d_y, stream);
    if (rc || !d_tail_ptr) return rc;
    if (!d_tail_col || !d_tail_val) { cme213::set_error("spmv_hyb_csrtail: null tail arrays"); return CME213_ERR_INVALID; }
    return cme213_internal_spmv_csr_f32_acc_keep(ctx, num_rows, d_tail_ptr, d_tail_col, d_tail_val, d_x, d_y, (cudaStream_t)stream);
}
CME213_API int cme213_spmv_hyb_csrtail_f64(cme213_ctx *ctx, int64_t num_rows, int64_t pitch, int K, const int32_t *d_ell_idx,
                                           const double *d_ell_val, const int32_t *d_tail_ptr, const int32_t *d_tail_col,
                                           const double *d_tail_val, const double *d_x, double *d_y, void *stream) {
    int rc = cme213_spmv_ell_f64(ctx, num_rows, pitch, K, d_ell_idx, d_ell_val, d_x, d_y, stream);
    if (rc || !d_tail_ptr) return rc;
    if (!d_tail_col || !d_tail_val) { cme213::set_error("spmv_hyb_csrtail: null tail arrays"); return CME213_ERR_INVALID; }
    return cme213_internal_spmv_csr_f64_acc_keep(ctx, num_rows, d_tail_ptr, d_tail_col, d_tail_val, d_x, d_y, (cudaStream_t)stream);
}
