// generate.cu -- deterministic device-side input generators.
//
// The homework drivers build their inputs inside main() (pagerank.cu:185-204, readMM.py:16-61);
// the large configurations (16 Mi rows / 256 Mi nonzeros, 512^3 grid / 938 M nonzeros) cannot be
// produced on the host in reasonable time, so the same recipes run here, driven by a counter hash
// so that tests can rebuild any input bit for bit in numpy (oracle/generators.py).  Row pointers
// come from cme213_exclusive_scan_u32 over the row lengths.
#include "internal.cuh"

namespace {
using namespace cme213;

// ---- scan of lengths held in d_row_ptr[0..rows] (last slot 0) -> CSR row pointers -----------------
int lengths_to_row_ptr(cme213_ctx *ctx, int32_t *d_row_ptr, int64_t rows, int64_t *nnz_out, cudaStream_t st) {
    void *ws = nullptr;
    int rc = workspace(ctx, cme213_internal_scan_u32_scratch_bytes(rows + 1), &ws, st);
    if (rc) return rc;
    rc = cme213_internal_exclusive_scan_u32(ctx, (const uint32_t *)d_row_ptr, (uint32_t *)d_row_ptr, rows + 1, ws, st);
    if (rc) return rc;
    if (nnz_out) {
        uint32_t total = 0;
        CME213_CUDA(cudaMemcpyAsync(&total, d_row_ptr + rows, 4, cudaMemcpyDeviceToHost, st));
        CME213_CUDA(cudaStreamSynchronize(st));
        *nnz_out = (int64_t)total;
        if (total > 0x7FFFFFFFu) { set_error("generator: nnz %u exceeds 32-bit signed row pointers", total); return CME213_ERR_UNSUPPORTED; }
    }
    return CME213_OK;
}

// ---- Laplacians ---------------------------------------------------------------------------------------
__global__ void lap2d_len_kernel(int nx, int ny, int32_t *len) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, rows = (int64_t)nx * ny;
    if (i > rows) return;
    int l = 0;
    if (i < rows) {
        const int ix = (int)(i % nx), iy = (int)(i / nx);
        l = 1 + (ix > 0) + (ix < nx - 1) + (iy > 0) + (iy < ny - 1);
    }
    len[i] = l;
}
__global__ void lap2d_fill_kernel(int nx, int ny, const int32_t *__restrict__ row_ptr, int32_t *col, float *val) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, rows = (int64_t)nx * ny;
    if (i >= rows) return;
    const int ix = (int)(i % nx), iy = (int)(i / nx);
    int o = row_ptr[i];
    if (iy > 0) { col[o] = (int32_t)(i - nx); val[o++] = -1.f; }
    if (ix > 0) { col[o] = (int32_t)(i - 1); val[o++] = -1.f; }
    col[o] = (int32_t)i; val[o++] = 4.f;
    if (ix < nx - 1) { col[o] = (int32_t)(i + 1); val[o++] = -1.f; }
    if (iy < ny - 1) { col[o] = (int32_t)(i + nx); val[o++] = -1.f; }
}
__global__ void lap3d_len_kernel(int nx, int ny, int nz, int64_t row_begin, int64_t rows_local, int32_t *len) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t > rows_local) return;
    int l = 0;
    if (t < rows_local) {
        const int64_t i = row_begin + t;
        const int ix = (int)(i % nx), iy = (int)((i / nx) % ny), iz = (int)(i / ((int64_t)nx * ny));
        l = 1 + (ix > 0) + (ix < nx - 1) + (iy > 0) + (iy < ny - 1) + (iz > 0) + (iz < nz - 1);
    }
    len[t] = l;
}
__global__ void lap3d_fill_kernel(int nx, int ny, int nz, int64_t row_begin, int64_t rows_local,
                                  const int32_t *__restrict__ row_ptr, int32_t *col, double *val) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= rows_local) return;
    const int64_t i = row_begin + t, plane = (int64_t)nx * ny;
    const int ix = (int)(i % nx), iy = (int)((i / nx) % ny), iz = (int)(i / plane);
    int64_t o = row_ptr[t];
    if (iz > 0) { col[o] = (int32_t)(i - plane); val[o++] = -1.0; }
    if (iy > 0) { col[o] = (int32_t)(i - nx); val[o++] = -1.0; }
    if (ix > 0) { col[o] = (int32_t)(i - 1); val[o++] = -1.0; }
    col[o] = (int32_t)i; val[o++] = 6.0;
    if (ix < nx - 1) { col[o] = (int32_t)(i + 1); val[o++] = -1.0; }
    if (iy < ny - 1) { col[o] = (int32_t)(i + nx); val[o++] = -1.0; }
    if (iz < nz - 1) { col[o] = (int32_t)(i + plane); val[o++] = -1.0; }
}

// ---- power-law rows ---------------------------------------------------------------------------------
__global__ void powerlaw_len_kernel(int64_t rows, int r0, uint32_t num, uint32_t den, int cap, uint32_t seed, int32_t *len) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i > rows) return;
    int l = 0;
    if (i < rows) {
        const uint64_t h = hash32(seed, (uint64_t)i);
        // Pareto(shape 1) tail in integers: delta = floor(num * (2^32 - 1 - h) / (den * (h + 1)))
        const uint64_t delta = ((uint64_t)num * (0xFFFFFFFFull - h)) / ((uint64_t)den * (h + 1ull));
        const uint64_t t = (uint64_t)r0 + delta;
        l = (int)(t < (uint64_t)cap ? t : (uint64_t)cap);
    }
    len[i] = l;
}
__global__ void random_fill_kernel(int64_t rows, int64_t cols, const int32_t *__restrict__ row_ptr, int64_t nnz,
                                   int64_t window, uint32_t seed, int32_t *col, float *val) {
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= nnz) return;
    // row of nonzero j: last r with row_ptr[r] <= j
    int64_t lo = 0, hi = rows;
    while (hi - lo > 1) {
        const int64_t mid = (lo + hi) >> 1;
        if ((int64_t)row_ptr[mid] <= j) lo = mid; else hi = mid;
    }
    const int64_t r = lo;
    const int s = row_ptr[r], len = row_ptr[r + 1] - s;
    const int64_t pos = j - s;
    const int64_t W = (window > 0 && window < cols) ? window : cols;
    int64_t base = r - W / 2;
    if (base < 0) base = 0;
    if (base > cols - W) base = cols - W;
    const uint64_t stride = (uint64_t)(W / len);  // >= 1 when W >= len
    const uint64_t h = hash32(seed, (uint64_t)j);
    const uint64_t c = (uint64_t)base + (uint64_t)pos * stride + ((h * stride) >> 32);
    col[j] = (int32_t)(c < (uint64_t)cols ? c : (uint64_t)cols - 1);
    val[j] = uniform_pm1_f32(hash32(seed ^ 0x5bd1e995u, (uint64_t)j));
}

// ---- PageRank graph ---------------------------------------------------------------------------------
__global__ void pagerank_len_kernel(int n, int avg_edges, uint32_t *indices, float *inv, float *nodes) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i > n) return;
    uint32_t l = 0;
    if (i < n) {
        l = (uint32_t)(i % (2 * avg_edges - 1) + 1);  // pagerank.cu:189
        inv[i] = 1.f / (float)l;                       // pagerank.cu:190
        nodes[i] = 1.f / (float)n;                     // pagerank.cu:201
    }
    indices[i] = l;
}
__global__ void pagerank_edges_kernel(int64_t nnz, int n, uint32_t seed, uint32_t *edges) {
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j < nnz) edges[j] = hash32(seed, (uint64_t)j) % (uint32_t)n;  // rand() % num_elements, pagerank.cu:198
}

// ---- fills ------------------------------------------------------------------------------------------
__global__ void fill_uniform_f32_kernel(float *out, int64_t n, uint32_t seed) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        out[i] = uniform_pm1_f32(hash32(seed, (uint64_t)i));
}
__global__ void fill_uniform_f64_kernel(double *out, int64_t n, uint32_t seed) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        out[i] = uniform_pm1_f64(hash32(seed, (uint64_t)i));
}
__global__ void fill_hash_u32_kernel(uint32_t *out, int64_t n, uint32_t seed, uint32_t modulus) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const uint32_t h = hash32(seed, (uint64_t)i);
        out[i] = modulus ? h % modulus : h;
    }
}
__global__ void fill_flags_kernel(uint32_t *bits, int64_t n, uint32_t seed, uint32_t prob_num) {
    const int64_t words = (n + 31) / 32;
    for (int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; w < words; w += (int64_t)gridDim.x * blockDim.x) {
        uint32_t m = 0;
        for (int b = 0; b < 32; b++) {
            const int64_t i = w * 32 + b;
            if (i < n && hash32(seed, (uint64_t)i) < prob_num) m |= 1u << b;
        }
        if (w == 0) m |= 1u;
        bits[w] = m;
    }
}

int grid_for(cme213_ctx *ctx, int64_t n) {
    const int64_t want = div_up(n, 256), cap = (int64_t)ctx->sm_count * 16;
    return (int)(want < cap ? (want > 0 ? want : 1) : cap);
}
}  // namespace

CME213_API int cme213_gen_laplace2d_f32(cme213_ctx *ctx, int nx, int ny, int32_t *d_row_ptr, int32_t *d_col,
                                        float *d_val, int64_t *nnz_out, void *stream) {
    CME213_CHECK_CTX(ctx);
    if (nx <= 0 || ny <= 0 || !d_row_ptr) { set_error("gen_laplace2d: bad arguments"); return CME213_ERR_INVALID; }
    DeviceGuard g(ctx->device);
    cudaStream_t st = (cudaStream_t)stream;
    const int64_t rows = (int64_t)nx * ny;
    lap2d_len_kernel<<<(unsigned)div_up(rows + 1, 256), 256, 0, st>>>(nx, ny, d_row_ptr);
    CME213_LAUNCHED(ctx, "lap2d_len_kernel");
    int rc = lengths_to_row_ptr(ctx, d_row_ptr, rows, nnz_out, st);
    if (rc) return rc;
    if (d_col && d_val) {
        lap2d_fill_kernel<<<(unsigned)div_up(rows, 256), 256, 0, st>>>(nx, ny, d_row_ptr, d_col, d_val);
        CME213_LAUNCHED(ctx, "lap2d_fill_kernel");
    }
    return CME213_OK;
}

CME213_API int cme213_gen_laplace3d_f64(cme213_ctx *ctx, int nx, int ny, int nz, int64_t row_begin, int64_t row_end,
                                        int32_t *d_row_ptr, int32_t *d_col, double *d_val, int64_t *nnz_out,
                                        void *stream) {
    CME213_CHECK_CTX(ctx);
    const int64_t rows_total = (int64_t)nx * ny * nz;
    if (nx <= 0 || ny <= 0 || nz <= 0 || row_begin < 0 || row_end < row_begin || row_end > rows_total || !d_row_ptr ||
        rows_total > 0x7FFFFFFFLL) {
        set_error("gen_laplace3d: bad arguments");
        return CME213_ERR_INVALID;
    }
    DeviceGuard g(ctx->device);
    cudaStream_t st = (cudaStream_t)stream;
    const int64_t rows = row_end - row_begin;
    lap3d_len_kernel<<<(unsigned)div_up(rows + 1, 256), 256, 0, st>>>(nx, ny, nz, row_begin, rows, d_row_ptr);
    CME213_LAUNCHED(ctx, "lap3d_len_kernel");
    int rc = lengths_to_row_ptr(ctx, d_row_ptr, rows, nnz_out, st);
    if (rc) return rc;
    if (d_col && d_val && rows > 0) {
        lap3d_fill_kernel<<<(unsigned)div_up(rows, 256), 256, 0, st>>>(nx, ny, nz, row_begin, rows, d_row_ptr, d_col, d_val);
        CME213_LAUNCHED(ctx, "lap3d_fill_kernel");
    }
    return CME213_OK;
}

CME213_API int cme213_gen_powerlaw_lengths(cme213_ctx *ctx, int64_t num_rows, int r0, uint32_t num, uint32_t den,
                                           int cap, uint32_t seed, int32_t *d_row_ptr, int64_t *nnz_out, void *stream) {
    CME213_CHECK_CTX(ctx);
    if (num_rows <= 0 || r0 < 0 || den == 0 || cap < 1 || !d_row_ptr) { set_error("gen_powerlaw_lengths: bad arguments"); return CME213_ERR_INVALID; }
    DeviceGuard g(ctx->device);
    cudaStream_t st = (cudaStream_t)stream;
    powerlaw_len_kernel<<<(unsigned)div_up(num_rows + 1, 256), 256, 0, st>>>(num_rows, r0, num, den, cap, seed, d_row_ptr);
    CME213_LAUNCHED(ctx, "powerlaw_len_kernel");
    return lengths_to_row_ptr(ctx, d_row_ptr, num_rows, nnz_out, st);
}

CME213_API int cme213_gen_random_csr_fill_f32(cme213_ctx *ctx, int64_t num_rows, int64_t num_cols,
                                              const int32_t *d_row_ptr, int64_t window, uint32_t seed, int32_t *d_col,
                                              float *d_val, void *stream) {
    CME213_CHECK_CTX(ctx);
    if (num_rows <= 0 || num_cols <= 0 || num_cols > 0x7FFFFFFFLL || !d_row_ptr || !d_col || !d_val) {
        set_error("gen_random_csr_fill: bad arguments");
        return CME213_ERR_INVALID;
    }
    DeviceGuard g(ctx->device);
    cudaStream_t st = (cudaStream_t)stream;
    int32_t nnz32 = 0;
    CME213_CUDA(cudaMemcpyAsync(&nnz32, d_row_ptr + num_rows, 4, cudaMemcpyDeviceToHost, st));
    CME213_CUDA(cudaStreamSynchronize(st));
    const int64_t nnz = nnz32;
    if (nnz <= 0) return CME213_OK;
    random_fill_kernel<<<(unsigned)div_up(nnz, 256), 256, 0, st>>>(num_rows, num_cols, d_row_ptr, nnz, window, seed, d_col, d_val);
    CME213_LAUNCHED(ctx, "random_fill_kernel");
    return CME213_OK;
}

CME213_API int cme213_gen_pagerank_graph(cme213_ctx *ctx, int num_elements, int avg_edges, uint32_t seed,
                                         uint32_t *d_graph_indices, uint32_t *d_graph_edges,
                                         float *d_inv_edges_per_node, float *d_nodes, int64_t *nnz_out, void *stream) {
    CME213_CHECK_CTX(ctx);
    if (num_elements <= 0 || avg_edges <= 0 || !d_graph_indices || !d_inv_edges_per_node || !d_nodes) {
        set_error("gen_pagerank_graph: bad arguments");
        return CME213_ERR_INVALID;
    }
    DeviceGuard g(ctx->device);
    cudaStream_t st = (cudaStream_t)stream;
    pagerank_len_kernel<<<(unsigned)div_up((int64_t)num_elements + 1, 256), 256, 0, st>>>(num_elements, avg_edges, d_graph_indices,
                                                                                         d_inv_edges_per_node, d_nodes);
    CME213_LAUNCHED(ctx, "pagerank_len_kernel");
    int64_t nnz = 0;
    int rc = lengths_to_row_ptr(ctx, (int32_t *)d_graph_indices, num_elements, &nnz, st);
    if (rc) return rc;
    if (nnz_out) *nnz_out = nnz;
    if (d_graph_edges && nnz > 0) {
        pagerank_edges_kernel<<<(unsigned)div_up(nnz, 256), 256, 0, st>>>(nnz, num_elements, seed, d_graph_edges);
        CME213_LAUNCHED(ctx, "pagerank_edges_kernel");
    }
    return CME213_OK;
}

#define FILL_ENTRY(NAME, TYPE, KERNEL, ...)                                                        \
    CME213_CHECK_CTX(ctx);                                                                         \
    if (n < 0 || (n > 0 && !d_out)) { set_error(NAME ": bad arguments"); return CME213_ERR_INVALID; } \
    if (n == 0) return CME213_OK;                                                                  \
    DeviceGuard g(ctx->device);                                                                    \
    KERNEL<<<grid_for(ctx, n), 256, 0, (cudaStream_t)stream>>>(__VA_ARGS__);                        \
    CME213_LAUNCHED(ctx, NAME);                                                                    \
    return CME213_OK;

CME213_API int cme213_fill_uniform_f32(cme213_ctx *ctx, float *d_out, int64_t n, uint32_t seed, void *stream) {
    FILL_ENTRY("fill_uniform_f32", float, fill_uniform_f32_kernel, d_out, n, seed)
}
CME213_API int cme213_fill_uniform_f64(cme213_ctx *ctx, double *d_out, int64_t n, uint32_t seed, void *stream) {
    FILL_ENTRY("fill_uniform_f64", double, fill_uniform_f64_kernel, d_out, n, seed)
}
CME213_API int cme213_fill_hash_u32(cme213_ctx *ctx, uint32_t *d_out, int64_t n, uint32_t seed, uint32_t modulus,
                                    void *stream) {
    FILL_ENTRY("fill_hash_u32", uint32_t, fill_hash_u32_kernel, d_out, n, seed, modulus)
}
CME213_API int cme213_fill_random_flags(cme213_ctx *ctx, uint32_t *d_flag_bits, int64_t n, uint32_t seed,
                                        uint32_t prob_num, void *stream) {
    uint32_t *d_out = d_flag_bits;
    FILL_ENTRY("fill_random_flags", uint32_t, fill_flags_kernel, d_out, n, seed, prob_num)
}

// ---- segment start offsets the way the final project's input script makes them -----------------------------
// hw/hw_final/programming/aux/readMM.py:44-49: s = random.sample(range(n), p); s.sort(); s[0] = 0; s[p-1] = n.
// Here: num_cuts hashed positions in [1, n), sorted with the library's radix sort, duplicates removed by
// mark + exclusive scan + compaction (slides/Lecture16.pdf slides 3-13), framed by 0 and n.
namespace {
__global__ void cut_points_kernel(uint32_t *keys, int64_t num_cuts, uint32_t seed, uint32_t n) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < num_cuts) keys[i] = 1u + hash32(seed, (uint64_t)i) % (n - 1u);
}
__global__ void mark_first_of_run_kernel(const uint32_t *keys, int64_t num_cuts, uint32_t *mark) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i <= num_cuts) mark[i] = (i < num_cuts && (i == 0 || keys[i] != keys[i - 1])) ? 1u : 0u;
}
__global__ void compact_offsets_kernel(const uint32_t *keys, const uint32_t *mark, const uint32_t *excl, int64_t num_cuts,
                                       uint32_t n, int32_t *s_out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < num_cuts && mark[i]) s_out[1 + excl[i]] = (int32_t)keys[i];
    if (i == num_cuts) { s_out[0] = 0; s_out[1 + excl[num_cuts]] = (int32_t)n; }
}
}  // namespace

CME213_API int cme213_gen_segment_offsets(cme213_ctx *ctx, int64_t n, int64_t num_cuts, uint32_t seed, int32_t *d_s_out,
                                          int64_t *p_out, void *stream) {
    CME213_CHECK_CTX(ctx);
    if (p_out) *p_out = 0;
    if (n < 1 || n > 0x7FFFFFFFLL || num_cuts < 0 || num_cuts > 0x7FFFFFFELL || !d_s_out || !p_out || (num_cuts > 0 && n < 2)) {
        set_error("gen_segment_offsets: bad arguments");
        return CME213_ERR_INVALID;
    }
    DeviceGuard g(ctx->device);
    cudaStream_t st = (cudaStream_t)stream;
    // set-up code, not a timed path: its four temporaries come from one cudaMalloc (the sort and the scan use the
    // context workspace themselves)
    const size_t words = (size_t)num_cuts + 1;
    uint32_t *tmp = nullptr;
    CME213_CUDA(cudaMalloc(&tmp, 4 * words * sizeof(uint32_t)));
    uint32_t *keys = tmp, *keys2 = tmp + words, *mark = tmp + 2 * words, *excl = tmp + 3 * words;
    int rc = CME213_OK;
    const unsigned grid = (unsigned)div_up((int64_t)words, 256);
    if (num_cuts > 0) {
        cut_points_kernel<<<grid, 256, 0, st>>>(keys, num_cuts, seed, (uint32_t)n);
        ctx->launches++;
        rc = cme213_radix_sort_u32(ctx, keys, keys2, num_cuts, 8, stream);
    }
    if (rc == CME213_OK) {
        mark_first_of_run_kernel<<<grid, 256, 0, st>>>(keys, num_cuts, mark);
        ctx->launches++;
        rc = cme213_exclusive_scan_u32(ctx, mark, excl, (int64_t)words, stream);
    }
    uint32_t distinct = 0;
    if (rc == CME213_OK) {
        compact_offsets_kernel<<<grid, 256, 0, st>>>(keys, mark, excl, num_cuts, (uint32_t)n, d_s_out);
        ctx->launches++;
        cudaError_t e = cudaMemcpyAsync(&distinct, excl + num_cuts, 4, cudaMemcpyDeviceToHost, st);
        if (e == cudaSuccess) e = cudaStreamSynchronize(st);
        if (e == cudaSuccess) e = cudaGetLastError();
        if (e != cudaSuccess) rc = cuda_fail(e, "gen_segment_offsets");
    } else {
        cudaStreamSynchronize(st);
    }
    cudaFree(tmp);
    if (rc == CME213_OK) *p_out = (int64_t)distinct + 2;
    return rc;
}


// ---- measurement aid: how fast can this GPU gather? --------------------------------------------------------------------
// Every SpMV format here issues one data-dependent 4- or 8-byte load of x per nonzero.  When the columns are spread out,
// each such load costs a whole 32-byte sector from L2 (or from HBM once x is larger than L2), so the rate of random
// gathers -- not the HBM stream -- bounds the kernel.  This probe measures that ceiling with nothing else in the way:
// 16 independent gathers per thread from hashed positions of x[0, q), one result stored per thread.
namespace {
__global__ void __launch_bounds__(256)
probe_gather_kernel(const float *__restrict__ x, uint32_t q, float *__restrict__ out, int64_t threads_total, uint32_t seed) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= threads_total) return;
    const uint64_t pol = l2_policy_evict_last();
    float v[16];
#pragma unroll
    for (int j = 0; j < 16; j++) {
        const uint32_t h = hash32(seed, (uint64_t)t * 16 + j);
        v[j] = ld_keep(x + (uint32_t)(((uint64_t)h * q) >> 32), pol);
    }
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < 16; j++) s += v[j];
    out[t] = s;
}
}  // namespace

CME213_API int cme213_probe_gather_f32(cme213_ctx *ctx, const float *d_x, int64_t q, float *d_out, int64_t num_gathers,
                                       uint32_t seed, void *stream) {
    CME213_CHECK_CTX(ctx);
    if (q < 1 || q > 0xFFFFFFFFLL || num_gathers < 0 || num_gathers % 16 || !d_x || (num_gathers > 0 && !d_out)) {
        set_error("probe_gather_f32: bad arguments (num_gathers must be a multiple of 16)");
        return CME213_ERR_INVALID;
    }
    if (num_gathers == 0) return CME213_OK;
    DeviceGuard g(ctx->device);
    const int64_t threads = num_gathers / 16;
    probe_gather_kernel<<<(unsigned)div_up(threads, 256), 256, 0, (cudaStream_t)stream>>>(d_x, (uint32_t)q, d_out, threads, seed);
    CME213_LAUNCHED(ctx, "probe_gather_kernel");
    return CME213_OK;
}
