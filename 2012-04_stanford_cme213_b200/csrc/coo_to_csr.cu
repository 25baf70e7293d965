// coo_to_csr.cu -- unsorted COO triplets -> CSR with ascending columns, on the device.
//
// This is the input side of the path: the final project's inputs come from MatrixMarket files, which scipy's
// mmread hands to hw/hw_final/programming/aux/readMM.py:18-24 as COO triplets (a.row, a.col, a.data).  The
// conversion is two stable key-value radix sorts of a permutation (by column, then by row: LSD order, so equal
// (row, column) pairs keep their input order) followed by one gather and a run-boundary pass for row_ptr.
#include "internal.cuh"

namespace {
using namespace cme213;

__global__ void coo_init_kernel(int64_t nnz, const int32_t *__restrict__ row, const int32_t *__restrict__ col,
                                int64_t num_rows, int64_t num_cols, uint32_t *keys, uint32_t *perm,
                                unsigned long long *bad) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nnz) return;
    const int32_t r = row[i], c = col[i];
    if (r < 0 || r >= num_rows || c < 0 || c >= num_cols) atomicAdd(bad, 1ull);
    keys[i] = (uint32_t)c;
    perm[i] = (uint32_t)i;
}
__global__ void coo_row_keys_kernel(int64_t nnz, const int32_t *__restrict__ row, const uint32_t *__restrict__ perm,
                                    uint32_t *keys) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < nnz) keys[i] = (uint32_t)row[perm[i]];
}
template <typename T>
__global__ void coo_gather_kernel(int64_t nnz, const uint32_t *__restrict__ perm, const int32_t *__restrict__ col,
                                  const T *__restrict__ val, int32_t *col_out, T *val_out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nnz) return;
    const uint32_t p = perm[i];
    col_out[i] = col[p];
    val_out[i] = val[p];
}
// row_ptr[r] = first position whose (sorted) row is >= r: position i fills the rows in (row[i-1], row[i]].
__global__ void coo_row_ptr_kernel(int64_t nnz, int64_t num_rows, const uint32_t *__restrict__ sorted_rows,
                                   int32_t *row_ptr) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i > nnz) return;
    const int64_t prev = i == 0 ? -1 : (int64_t)sorted_rows[i - 1];
    const int64_t cur = i == nnz ? num_rows : (int64_t)sorted_rows[i];
    for (int64_t r = prev + 1; r <= cur; ++r) row_ptr[r] = (int32_t)i;
}

int bits_for(int64_t count) {   // bits needed for indices 0 .. count-1
    int b = 0;
    while (b < 32 && ((int64_t)1 << b) < count) ++b;
    return b;
}

template <typename T>
int coo_to_csr_impl(cme213_ctx *ctx, int64_t num_rows, int64_t num_cols, int64_t nnz, const int32_t *d_coo_row,
                    const int32_t *d_coo_col, const T *d_coo_val, int32_t *d_row_ptr, int32_t *d_col, T *d_val,
                    void *stream) {
    CME213_CHECK_CTX(ctx);
    if (num_rows < 0 || num_cols < 0 || nnz < 0 || num_rows > 0x7FFFFFFELL || num_cols > 0x7FFFFFFFLL || nnz > 0x7FFFFFFFLL) {
        set_error("coo_to_csr: bad sizes");
        return CME213_ERR_INVALID;
    }
    if (!d_row_ptr || (nnz > 0 && (!d_coo_row || !d_coo_col || !d_coo_val || !d_col || !d_val))) {
        set_error("coo_to_csr: null pointer");
        return CME213_ERR_INVALID;
    }
    DeviceGuard g(ctx->device);
    cudaStream_t st = (cudaStream_t)stream;
    // set-up code: the four permutation / key buffers come from one cudaMalloc (the sorts use the context workspace)
    const size_t words = (size_t)(nnz > 0 ? nnz : 1);
    uint32_t *tmp = nullptr;
    CME213_CUDA(cudaMalloc(&tmp, (4 * words + 2) * sizeof(uint32_t)));
    uint32_t *keys = tmp, *keys2 = tmp + words, *perm = tmp + 2 * words, *perm2 = tmp + 3 * words;
    unsigned long long *d_bad = reinterpret_cast<unsigned long long *>(tmp + 4 * words);   // 8-byte aligned: words*16
    int rc = CME213_OK;
    unsigned long long bad = 0;
    const unsigned grid = (unsigned)div_up(nnz + 1, 256);
    cudaError_t e = cudaMemsetAsync(d_bad, 0, 8, st);
    if (e == cudaSuccess && nnz > 0) {
        coo_init_kernel<<<grid, 256, 0, st>>>(nnz, d_coo_row, d_coo_col, num_rows, num_cols, keys, perm, d_bad);
        ctx->launches++;
        e = cudaMemcpyAsync(&bad, d_bad, 8, cudaMemcpyDeviceToHost, st);
        if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    }
    if (e != cudaSuccess) rc = cuda_fail(e, "coo_to_csr");
    if (rc == CME213_OK && bad) {
        set_error("coo_to_csr: %llu entries outside the %lld x %lld matrix", bad, (long long)num_rows, (long long)num_cols);
        rc = CME213_ERR_INVALID;
    }
    if (rc == CME213_OK && nnz > 0) rc = cme213_radix_sort_pairs_u32(ctx, keys, keys2, perm, perm2, nnz, bits_for(num_cols), stream);
    if (rc == CME213_OK && nnz > 0) {
        coo_row_keys_kernel<<<grid, 256, 0, st>>>(nnz, d_coo_row, perm, keys);
        ctx->launches++;
        rc = cme213_radix_sort_pairs_u32(ctx, keys, keys2, perm, perm2, nnz, bits_for(num_rows), stream);
    }
    if (rc == CME213_OK) {
        if (nnz > 0) {
            coo_gather_kernel<T><<<grid, 256, 0, st>>>(nnz, perm, d_coo_col, d_coo_val, d_col, d_val);
            ctx->launches++;
        }
        // rows are filled by the positions around them; when a run of empty rows is long one thread writes them all
        coo_row_ptr_kernel<<<grid, 256, 0, st>>>(nnz, num_rows, keys, d_row_ptr);
        ctx->launches++;
        e = cudaGetLastError();
        if (e != cudaSuccess) rc = cuda_fail(e, "coo_to_csr");
    }
    cudaStreamSynchronize(st);   // the temporaries go away below
    cudaFree(tmp);
    return rc;
}
}  // namespace

CME213_API int cme213_coo_to_csr_f32(cme213_ctx *ctx, int64_t num_rows, int64_t num_cols, int64_t nnz,
                                     const int32_t *d_coo_row, const int32_t *d_coo_col, const float *d_coo_val,
                                     int32_t *d_row_ptr, int32_t *d_col, float *d_val, void *stream) {
    return coo_to_csr_impl<float>(ctx, num_rows, num_cols, nnz, d_coo_row, d_coo_col, d_coo_val, d_row_ptr, d_col, d_val, stream);
}
CME213_API int cme213_coo_to_csr_f64(cme213_ctx *ctx, int64_t num_rows, int64_t num_cols, int64_t nnz,
                                     const int32_t *d_coo_row, const int32_t *d_coo_col, const double *d_coo_val,
                                     int32_t *d_row_ptr, int32_t *d_col, double *d_val, void *stream) {
    return coo_to_csr_impl<double>(ctx, num_rows, num_cols, nnz, d_coo_row, d_coo_col, d_coo_val, d_row_ptr, d_col, d_val, stream);
}
