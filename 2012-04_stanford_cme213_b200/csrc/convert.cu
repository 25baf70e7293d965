// convert.cu -- CSR -> ELL / HYB conversion on the device.
//
// reference: refs/Bell SC 2009.pdf p.3 Fig. 2 (array contents), Fig. 4 (column-major
// linearisation), p.5 section 3.5 (HYB split rule).  The conversion is where the two primitives
// meet the formats: the column-major ELL arrays are produced by the tiled transpose from a
// padded row-major staging block, and the COO tail offsets are an exclusive scan of the
// per-row overflow counts (stream compaction, slides/Lecture16.pdf slides 3-13).
#include "internal.cuh"
#include <vector>

namespace {
using namespace cme213;

constexpr int HIST_BINS = CME213_HIST_BINS;

__global__ void row_length_hist_kernel(int64_t num_rows, const int32_t *__restrict__ row_ptr,
                                       unsigned long long *hist, int *max_len) {
    __shared__ unsigned int sh[HIST_BINS];
    for (int i = threadIdx.x; i < HIST_BINS; i += blockDim.x) sh[i] = 0;
    __syncthreads();
    int local_max = 0;
    for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < num_rows; r += (int64_t)gridDim.x * blockDim.x) {
        const int len = row_ptr[r + 1] - row_ptr[r];
        local_max = len > local_max ? len : local_max;
        atomicAdd(&sh[len < HIST_BINS - 1 ? len : HIST_BINS - 1], 1u);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < HIST_BINS; i += blockDim.x)
        if (sh[i]) atomicAdd(&hist[i], (unsigned long long)sh[i]);
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        const int o = __shfl_down_sync(0xFFFFFFFFu, local_max, d);
        local_max = o > local_max ? o : local_max;
    }
    if ((threadIdx.x & 31) == 0 && local_max > 0) atomicMax(max_len, local_max);
}

// Row-major padded staging block: stage[(r - r_begin) * Kp + n] for rows [r_begin, r_begin + chunk).
template <typename T>
__global__ void ell_stage_fill_kernel(int64_t num_rows, int64_t num_cols, const int32_t *__restrict__ row_ptr,
                                      const int32_t *__restrict__ col, const T *__restrict__ val, int K, int Kp,
                                      int64_t r_begin, int64_t chunk, int32_t *stage_idx, T *stage_val) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= chunk * Kp) return;
    const int64_t lr = t / Kp;
    const int n = (int)(t % Kp);
    const int64_t r = r_begin + lr;
    int32_t ci = 0;
    T v = T(0);
    if (r < num_rows && n < K) {
        const int s = row_ptr[r], len = row_ptr[r + 1] - s;
        if (n < len) {
            ci = col[s + n];
            v = val[s + n];
        } else {
            ci = (int32_t)(r < num_cols ? r : num_cols - 1);  // any valid column; value 0 marks padding
        }
    }
    stage_idx[t] = ci;
    stage_val[t] = v;
}

__global__ void tail_count_kernel(int64_t num_rows, const int32_t *__restrict__ row_ptr, int K, uint32_t *count) {
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r > num_rows) return;
    uint32_t c = 0;
    if (r < num_rows) {
        const int len = row_ptr[r + 1] - row_ptr[r];
        c = len > K ? (uint32_t)(len - K) : 0u;
    }
    count[r] = c;  // count[num_rows] = 0 so the scan's last element is the total
}

template <typename T>
__global__ void tail_fill_kernel(int64_t num_rows, const int32_t *__restrict__ row_ptr, const int32_t *__restrict__ col,
                                 const T *__restrict__ val, int K, const uint32_t *__restrict__ offs, int32_t *coo_row,
                                 int32_t *coo_col, T *coo_val) {
    // one warp per row: tails are long exactly where rows are long
    const int64_t r = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (r >= num_rows) return;
    const int s = row_ptr[r] + K, e = row_ptr[r + 1];
    const int64_t o = offs[r];
    for (int j = s + lane; j < e; j += 32) {
        coo_row[o + (j - s)] = (int32_t)r;
        coo_col[o + (j - s)] = col[j];
        coo_val[o + (j - s)] = val[j];
    }
}

__global__ void row_head_flags_kernel(int64_t num_rows, int64_t row_len, uint32_t *bits) {
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r < num_rows) {
        const int64_t pos = r * row_len;
        atomicOr(bits + (pos >> 5), 1u << (unsigned)(pos & 31));
    }
}
__global__ void add_row_vector_kernel(uint32_t *m, const uint32_t *__restrict__ v, int64_t total, int64_t cols) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x)
        m[i] += __ldg(v + i % cols);
}

template <typename T>
int csr_to_ell_impl(cme213_ctx *ctx, int64_t num_rows, int64_t num_cols, const int32_t *d_row_ptr,
                    const int32_t *d_col, const T *d_val, int K, int64_t pitch, int32_t *d_ell_idx, T *d_ell_val,
                    void *stream) {
    CME213_CHECK_CTX(ctx);
    if (num_rows < 0 || num_cols <= 0 || K < 0 || pitch < num_rows || pitch % 4 != 0) {
        set_error("csr_to_ell: bad shape (pitch must be >= rows and a multiple of 4)");
        return CME213_ERR_INVALID;
    }
    if (num_rows == 0 || K == 0) return CME213_OK;
    if (!d_row_ptr || !d_col || !d_val || !d_ell_idx || !d_ell_val) { set_error("csr_to_ell: null pointer"); return CME213_ERR_INVALID; }
    DeviceGuard g(ctx->device);
    cudaStream_t st = (cudaStream_t)stream;
    const int Kp = (K + 3) & ~3;  // staging pitch: whole 16-byte vectors for the transpose
    // chunk of rows: a multiple of 64, staging kept under ~256 MiB
    int64_t chunk = (int64_t)(256u << 20) / ((int64_t)Kp * (4 + sizeof(T)));
    chunk = chunk < 64 ? 64 : (chunk & ~(int64_t)63);
    if (chunk > pitch) chunk = pitch;
    void *ws = nullptr;
    const size_t idx_bytes = ((size_t)chunk * Kp * 4 + 255) & ~(size_t)255;
    int rc = workspace(ctx, idx_bytes + (size_t)chunk * Kp * sizeof(T), &ws, st);
    if (rc) return rc;
    int32_t *stage_idx = (int32_t *)ws;
    T *stage_val = (T *)((char *)ws + idx_bytes);
    for (int64_t r0 = 0; r0 < pitch; r0 += chunk) {
        const int64_t rows_here = pitch - r0 < chunk ? pitch - r0 : chunk;  // multiple of 4 (pitch % 4 == 0)
        const int64_t total = rows_here * Kp;
        ell_stage_fill_kernel<T><<<(unsigned)div_up(total, 256), 256, 0, st>>>(num_rows, num_cols, d_row_ptr, d_col, d_val, K, Kp,
                                                                              r0, rows_here, stage_idx, stage_val);
        CME213_LAUNCHED(ctx, "ell_stage_fill_kernel");
        // [rows_here x K] row-major (ld Kp)  ->  [K x rows_here] inside the column-major ELL arrays (ld pitch)
        rc = cme213_transpose_b32(ctx, stage_idx, d_ell_idx + r0, rows_here, K, Kp, pitch, stream);
        if (rc) return rc;
        if (sizeof(T) == 4) rc = cme213_transpose_b32(ctx, stage_val, d_ell_val + r0, rows_here, K, Kp, pitch, stream);
        else rc = cme213_transpose_b64(ctx, stage_val, d_ell_val + r0, rows_here, K, Kp, pitch, stream);
        if (rc) return rc;
    }
    return CME213_OK;
}

}  // namespace

CME213_API int64_t cme213_ell_pitch(int64_t num_rows) { return (num_rows + 63) & ~(int64_t)63; }

// Row-length histogram of rows [0, num_rows) of d_row_ptr (bin HIST_BINS-1 collects everything longer) and the longest
// row; synchronises the stream.  Shared by the HYB split rule, the CSR dispatcher and the format chooser (csr_plan.cu).
int cme213_internal_row_length_hist(cme213_ctx *ctx, int64_t num_rows, const int32_t *d_row_ptr, unsigned long long *h_hist,
                                    int *max_len_out, cudaStream_t st) {
    void *ws = nullptr;
    const size_t bytes = HIST_BINS * 8 + 256;
    int rc = workspace(ctx, bytes, &ws, st);
    if (rc) return rc;
    CME213_CUDA(cudaMemsetAsync(ws, 0, bytes, st));
    unsigned long long *d_hist = (unsigned long long *)ws;
    int *d_max = (int *)((char *)ws + HIST_BINS * 8);
    int blocks = (int)(div_up(num_rows, 256) < (int64_t)ctx->sm_count * 8 ? div_up(num_rows, 256) : (int64_t)ctx->sm_count * 8);
    row_length_hist_kernel<<<blocks, 256, 0, st>>>(num_rows, d_row_ptr, d_hist, d_max);
    CME213_LAUNCHED(ctx, "row_length_hist_kernel");
    CME213_CUDA(cudaMemcpyAsync(h_hist, d_hist, HIST_BINS * 8, cudaMemcpyDeviceToHost, st));
    CME213_CUDA(cudaMemcpyAsync(max_len_out, d_max, sizeof(int), cudaMemcpyDeviceToHost, st));
    CME213_CUDA(cudaStreamSynchronize(st));
    return CME213_OK;
}

CME213_API int cme213_hyb_pick_k(cme213_ctx *ctx, int64_t num_rows, const int32_t *d_row_ptr, int *K_out,
                                 int *max_len_out, void *stream) {
    CME213_CHECK_CTX(ctx);
    if (num_rows < 0 || !K_out || (num_rows > 0 && !d_row_ptr)) { set_error("hyb_pick_k: bad arguments"); return CME213_ERR_INVALID; }
    *K_out = 0;
    if (max_len_out) *max_len_out = 0;
    if (num_rows == 0) return CME213_OK;
    DeviceGuard g(ctx->device);
    cudaStream_t st = (cudaStream_t)stream;
    std::vector<unsigned long long> hist(HIST_BINS);
    int max_len = 0;
    int rc = cme213_internal_row_length_hist(ctx, num_rows, d_row_ptr, hist.data(), &max_len, st);
    if (rc) return rc;
    // SC09 section 3.5: add the K-th ELL column while at least max(4096, M/3) rows have >= K entries.
    const unsigned long long thresh = (unsigned long long)(num_rows / 3 > 4096 ? num_rows / 3 : 4096);
    unsigned long long ge = 0;  // rows with length >= l, walking l downward
    int K = 0;
    for (int l = HIST_BINS - 1; l >= 1; l--) {
        ge += hist[l];
        if (ge >= thresh) { K = l; break; }
    }
    *K_out = K;
    if (max_len_out) *max_len_out = max_len;
    return CME213_OK;
}

CME213_API int cme213_csr_to_ell_f32(cme213_ctx *ctx, int64_t num_rows, int64_t num_cols, const int32_t *d_row_ptr,
                                     const int32_t *d_col, const float *d_val, int K, int64_t pitch, int32_t *d_ell_idx,
                                     float *d_ell_val, void *stream) {
    return csr_to_ell_impl<float>(ctx, num_rows, num_cols, d_row_ptr, d_col, d_val, K, pitch, d_ell_idx, d_ell_val, stream);
}
CME213_API int cme213_csr_to_ell_f64(cme213_ctx *ctx, int64_t num_rows, int64_t num_cols, const int32_t *d_row_ptr,
                                     const int32_t *d_col, const double *d_val, int K, int64_t pitch, int32_t *d_ell_idx,
                                     double *d_ell_val, void *stream) {
    return csr_to_ell_impl<double>(ctx, num_rows, num_cols, d_row_ptr, d_col, d_val, K, pitch, d_ell_idx, d_ell_val, stream);
}

template <typename T>
static int csr_to_coo_tail_impl(cme213_ctx *ctx, int64_t num_rows, const int32_t *d_row_ptr,
                                const int32_t *d_col, const T *d_val, int K, int64_t *tail_nnz_out,
                                int32_t *d_coo_row, int32_t *d_coo_col, T *d_coo_val, void *stream) {
    CME213_CHECK_CTX(ctx);
    if (num_rows < 0 || K < 0 || (num_rows > 0 && !d_row_ptr)) { set_error("csr_to_coo_tail: bad arguments"); return CME213_ERR_INVALID; }
    if (tail_nnz_out) *tail_nnz_out = 0;
    if (num_rows == 0) return CME213_OK;
    DeviceGuard g(ctx->device);
    cudaStream_t st = (cudaStream_t)stream;
    const int64_t n = num_rows + 1;
    const size_t count_bytes = ((size_t)n * 4 + 255) & ~(size_t)255;
    void *ws = nullptr;
    int rc = workspace(ctx, count_bytes + cme213_internal_scan_u32_scratch_bytes(n), &ws, st);
    if (rc) return rc;
    uint32_t *d_count = (uint32_t *)ws;
    tail_count_kernel<<<(unsigned)div_up(n, 256), 256, 0, st>>>(num_rows, d_row_ptr, K, d_count);
    CME213_LAUNCHED(ctx, "tail_count_kernel");
    rc = cme213_internal_exclusive_scan_u32(ctx, d_count, d_count, n, (char *)ws + count_bytes, stream);
    if (rc) return rc;
    const bool fill = d_coo_row && d_coo_col && d_coo_val;
    if (tail_nnz_out || !fill) {
        uint32_t total = 0;
        CME213_CUDA(cudaMemcpyAsync(&total, d_count + num_rows, 4, cudaMemcpyDeviceToHost, st));
        CME213_CUDA(cudaStreamSynchronize(st));
        if (tail_nnz_out) *tail_nnz_out = (int64_t)total;
    }
    if (fill) {
        if (!d_col || !d_val) { set_error("csr_to_coo_tail: null pointer"); return CME213_ERR_INVALID; }
        tail_fill_kernel<T><<<(unsigned)div_up(num_rows * 32, 256), 256, 0, st>>>(num_rows, d_row_ptr, d_col, d_val, K, d_count,
                                                                                     d_coo_row, d_coo_col, d_coo_val);
        CME213_LAUNCHED(ctx, "tail_fill_kernel");
    }
    return CME213_OK;
}

// Row pointers of the tail (entries at in-row position >= K) as a CSR matrix of its own: tail_ptr = exclusive scan of
// max(len - K, 0).  Its column / value arrays are the d_coo_col / d_coo_val of cme213_csr_to_coo_tail_* (same order).
CME213_API int cme213_csr_tail_row_ptr(cme213_ctx *ctx, int64_t num_rows, const int32_t *d_row_ptr, int K,
                                       int32_t *d_tail_ptr, int64_t *tail_nnz_out, void *stream) {
    CME213_CHECK_CTX(ctx);
    if (num_rows < 0 || K < 0 || !d_tail_ptr || (num_rows > 0 && !d_row_ptr)) { set_error("csr_tail_row_ptr: bad arguments"); return CME213_ERR_INVALID; }
    if (tail_nnz_out) *tail_nnz_out = 0;
    DeviceGuard g(ctx->device);
    cudaStream_t st = (cudaStream_t)stream;
    const int64_t n = num_rows + 1;
    void *ws = nullptr;
    int rc = workspace(ctx, cme213_internal_scan_u32_scratch_bytes(n), &ws, st);
    if (rc) return rc;
    tail_count_kernel<<<(unsigned)div_up(n, 256), 256, 0, st>>>(num_rows, d_row_ptr, K, (uint32_t *)d_tail_ptr);
    CME213_LAUNCHED(ctx, "tail_count_kernel");
    rc = cme213_internal_exclusive_scan_u32(ctx, (uint32_t *)d_tail_ptr, (uint32_t *)d_tail_ptr, n, ws, stream);
    if (rc) return rc;
    if (tail_nnz_out) {
        uint32_t total = 0;
        CME213_CUDA(cudaMemcpyAsync(&total, d_tail_ptr + num_rows, 4, cudaMemcpyDeviceToHost, st));
        CME213_CUDA(cudaStreamSynchronize(st));
        *tail_nnz_out = (int64_t)total;
    }
    return CME213_OK;
}

CME213_API int cme213_csr_to_coo_tail_f32(cme213_ctx *ctx, int64_t num_rows, const int32_t *d_row_ptr,
                                          const int32_t *d_col, const float *d_val, int K, int64_t *tail_nnz_out,
                                          int32_t *d_coo_row, int32_t *d_coo_col, float *d_coo_val, void *stream) {
    return csr_to_coo_tail_impl<float>(ctx, num_rows, d_row_ptr, d_col, d_val, K, tail_nnz_out, d_coo_row, d_coo_col, d_coo_val, stream);
}
CME213_API int cme213_csr_to_coo_tail_f64(cme213_ctx *ctx, int64_t num_rows, const int32_t *d_row_ptr,
                                          const int32_t *d_col, const double *d_val, int K, int64_t *tail_nnz_out,
                                          int32_t *d_coo_row, int32_t *d_coo_col, double *d_coo_val, void *stream) {
    return csr_to_coo_tail_impl<double>(ctx, num_rows, d_row_ptr, d_col, d_val, K, tail_nnz_out, d_coo_row, d_coo_col, d_coo_val, stream);
}

// Scatter offsets of a blocked LSD radix pass: for every bucket an exclusive scan down the blocks, started at
// the bucket's global offset.  Built from the two primitives exactly as the reference's loop nest suggests:
// transpose [blocks x buckets] -> [buckets x blocks], ONE segmented exclusive scan whose segments are the rows,
// transpose back, add the global offsets.   reference: hw/hw4/programming/radixsort.cpp:85-97.
CME213_API int cme213_block_offsets_u32(cme213_ctx *ctx, const uint32_t *d_local_hist, const uint32_t *d_global_excl,
                                        uint32_t *d_out, int64_t num_blocks, int64_t num_buckets, void *stream) {
    CME213_CHECK_CTX(ctx);
    if (num_blocks < 0 || num_buckets < 0) { set_error("block_offsets_u32: negative size"); return CME213_ERR_INVALID; }
    const int64_t total = num_blocks * num_buckets;
    if (total == 0) return CME213_OK;
    if (!d_local_hist || !d_global_excl || !d_out || d_local_hist == d_out) { set_error("block_offsets_u32: bad pointers"); return CME213_ERR_INVALID; }
    DeviceGuard g(ctx->device);
    cudaStream_t st = (cudaStream_t)stream;
    const size_t t_bytes = ((size_t)total * 4 + 255) & ~(size_t)255;
    const size_t f_bytes = ((size_t)cme213_flag_words(total) * 4 + 255) & ~(size_t)255;
    void *ws = nullptr;
    int rc = workspace(ctx, t_bytes + f_bytes + cme213_internal_segscan_u32_scratch_bytes(total), &ws, st);
    if (rc) return rc;
    uint32_t *d_t = (uint32_t *)ws;
    uint32_t *d_flags = (uint32_t *)((char *)ws + t_bytes);
    void *scratch = (char *)ws + t_bytes + f_bytes;
    rc = cme213_transpose_b32(ctx, d_local_hist, d_t, num_blocks, num_buckets, num_buckets, num_blocks, stream);
    if (rc) return rc;
    CME213_CUDA(cudaMemsetAsync(d_flags, 0, f_bytes, st));
    row_head_flags_kernel<<<(unsigned)div_up(num_buckets, 256), 256, 0, st>>>(num_buckets, num_blocks, d_flags);
    CME213_LAUNCHED(ctx, "row_head_flags_kernel");
    rc = cme213_internal_segscan_exclusive_u32(ctx, d_t, d_flags, d_t, total, scratch, stream);
    if (rc) return rc;
    rc = cme213_transpose_b32(ctx, d_t, d_out, num_buckets, num_blocks, num_blocks, num_buckets, stream);
    if (rc) return rc;
    const int64_t want = div_up(total, 256), cap = (int64_t)ctx->sm_count * 16;
    add_row_vector_kernel<<<(unsigned)(want < cap ? want : cap), 256, 0, st>>>(d_out, d_global_excl, total, num_buckets);
    CME213_LAUNCHED(ctx, "add_row_vector_kernel");
    return CME213_OK;
}
