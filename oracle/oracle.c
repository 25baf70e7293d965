/*
 * oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * A plain-C, CPU restatement of the algorithms on the hot path of the
 * CME213 (Spring 2012) coursework: CSR / ELL / COO / HYB sparse matrix-vector
 * multiply, the PageRank propagate step, the final project's product +
 * inclusive segmented scan, the exclusive scan from the radix sort, and the
 * out-of-place transpose.  Only tests/, __graft_entry__.smoke() and the
 * cpu_baseline / --impl reference legs of bench.py may load this library, and
 * only as the checker or as the timed CPU baseline.  Nothing under
 * 2012-04_stanford_cme213_b200/ may call into it.
 *
 * Every function cites the reference file:line (paths relative to the
 * reference checkout) or the slide / paper listing it follows.
 *
 * Parity status (see DESIGN.md "Oracle"):
 *   pinned    : oracle_sparse_mv_scan_* (fixture small_{a,x,b}.txt and the
 *               reference's own sparseMV_scan compiled into oracle/_ref),
 *               oracle_pagerank_* (reference's host_graph_propagate compiled
 *               into oracle/_ref), oracle_exclusive_scan_u32 (worked example
 *               of my-refs/scan.pdf p.4; radix pass == sort)
 *   unpinned  : transpose, ELL, COO, HYB (the reference ships slides and
 *               papers only; pinned here to the Bell & Garland Fig. 2 arrays
 *               and cross-checked against the CSR oracle on the same matrix).
 *
 * Build: gcc -O2 -fopenmp -ffp-contract=off -fPIC -shared (see Makefile).
 * -ffp-contract=off keeps a*b+c as a rounded multiply followed by a rounded
 * add, which is what the reference's host code does on x86-64 without -mfma.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define ORACLE_API __attribute__((visibility("default")))

ORACLE_API int oracle_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* ------------------------------------------------------------------------
 * PageRank propagate = CSR gather with per-column weight and affine epilogue.
 * Follows hw/hw1/programming/pagerank.cu:45-56 (host_graph_propagate):
 *   out[i] = 0.5f/N + 0.5f * sum_{j in [ind[i], ind[i+1])} in[e[j]]*inv[e[j]]
 * summed left to right in fp32.
 * ---------------------------------------------------------------------- */
ORACLE_API void oracle_pagerank_propagate(const uint32_t *graph_indices, const uint32_t *graph_edges,
                                          const float *nodes_in, float *nodes_out,
                                          const float *inv_edges_per_node, int array_length) {
    for (int i = 0; i < array_length; i++) {
        float sum = 0.f;
        for (uint32_t j = graph_indices[i]; j < graph_indices[i + 1]; j++) {
            uint32_t e = graph_edges[j];
            sum += nodes_in[e] * inv_edges_per_node[e];
        }
        nodes_out[i] = 0.5f / (float)array_length + 0.5f * sum;
    }
}

/* Same loop, rows split over OpenMP threads: the "all host threads" CPU
 * baseline.  Per-row arithmetic and order are unchanged, so the result is
 * bit-identical to the serial version. */
ORACLE_API void oracle_pagerank_propagate_omp(const uint32_t *graph_indices, const uint32_t *graph_edges,
                                              const float *nodes_in, float *nodes_out,
                                              const float *inv_edges_per_node, int array_length) {
#pragma omp parallel for schedule(static)
    for (int i = 0; i < array_length; i++) {
        float sum = 0.f;
        for (uint32_t j = graph_indices[i]; j < graph_indices[i + 1]; j++) {
            uint32_t e = graph_edges[j];
            sum += nodes_in[e] * inv_edges_per_node[e];
        }
        nodes_out[i] = 0.5f / (float)array_length + 0.5f * sum;
    }
}

/* hw/hw1/programming/pagerank.cu:59-67 (host_graph_iterate): ping-pong A->B,
 * B->A; nr_iterations must be even; the result ends in A.  Returns -1 for an
 * odd count where the reference asserts. */
ORACLE_API int oracle_pagerank_iterate(const uint32_t *graph_indices, const uint32_t *graph_edges,
                                       float *nodes_A, float *nodes_B, const float *inv_edges_per_node,
                                       int nr_iterations, int array_length, int use_omp) {
    if (nr_iterations % 2 != 0) return -1;
    for (int iter = 0; iter < nr_iterations; iter += 2) {
        if (use_omp) {
            oracle_pagerank_propagate_omp(graph_indices, graph_edges, nodes_A, nodes_B, inv_edges_per_node, array_length);
            oracle_pagerank_propagate_omp(graph_indices, graph_edges, nodes_B, nodes_A, inv_edges_per_node, array_length);
        } else {
            oracle_pagerank_propagate(graph_indices, graph_edges, nodes_A, nodes_B, inv_edges_per_node, array_length);
            oracle_pagerank_propagate(graph_indices, graph_edges, nodes_B, nodes_A, inv_edges_per_node, array_length);
        }
    }
    return 0;
}

/* ------------------------------------------------------------------------
 * CSR SpMV, y = A x.  Follows the serial csr_spmv listing of
 * slides/Lecture22.pdf slide 20 (and Bell TR 2008 Fig. 20's per-row loop):
 *   for row: dot = 0; for jj in [ptr[row], ptr[row+1]): dot += data[jj]*x[indices[jj]]; y[row] = dot
 * ---------------------------------------------------------------------- */
#define DEFINE_CSR_SPMV(NAME, T, PRAGMA)                                                             \
    ORACLE_API void NAME(int64_t num_rows, const int32_t *ptr, const int32_t *indices,              \
                         const T *data, const T *x, T *y) {                                          \
        PRAGMA                                                                                       \
        for (int64_t row = 0; row < num_rows; row++) {                                               \
            T dot = 0;                                                                               \
            for (int32_t jj = ptr[row]; jj < ptr[row + 1]; jj++) dot += data[jj] * x[indices[jj]];   \
            y[row] = dot;                                                                            \
        }                                                                                            \
    }
DEFINE_CSR_SPMV(oracle_csr_spmv_f32, float, )
DEFINE_CSR_SPMV(oracle_csr_spmv_f64, double, )
DEFINE_CSR_SPMV(oracle_csr_spmv_f32_omp, float, _Pragma("omp parallel for schedule(static)"))
DEFINE_CSR_SPMV(oracle_csr_spmv_f64_omp, double, _Pragma("omp parallel for schedule(static)"))

/* Per-row error bound helper for the fp tolerance in SURVEY 8(c):
 *   bound[row] = sum_j |a_ij * x_j|  computed in double. */
#define DEFINE_CSR_ABS(NAME, T)                                                                       \
    ORACLE_API void NAME(int64_t num_rows, const int32_t *ptr, const int32_t *indices,               \
                         const T *data, const T *x, double *bound) {                                  \
        for (int64_t row = 0; row < num_rows; row++) {                                                \
            double s = 0;                                                                             \
            for (int32_t jj = ptr[row]; jj < ptr[row + 1]; jj++)                                      \
                s += fabs((double)data[jj] * (double)x[indices[jj]]);                                 \
            bound[row] = s;                                                                           \
        }                                                                                             \
    }
DEFINE_CSR_ABS(oracle_csr_abs_bound_f32, float)
DEFINE_CSR_ABS(oracle_csr_abs_bound_f64, double)

/* Iterated y = A x with ping-pong buffers, the skeleton of
 * hw/hw1/programming/pagerank.cu:59-67 applied to a general CSR matrix
 * (config C5).  nr_iterations even; result ends in A. */
ORACLE_API int oracle_csr_iterate_f64(int64_t num_rows, const int32_t *ptr, const int32_t *indices,
                                      const double *data, double *A, double *B, int nr_iterations, int use_omp) {
    if (nr_iterations % 2 != 0) return -1;
    for (int it = 0; it < nr_iterations; it += 2) {
        if (use_omp) {
            oracle_csr_spmv_f64_omp(num_rows, ptr, indices, data, A, B);
            oracle_csr_spmv_f64_omp(num_rows, ptr, indices, data, B, A);
        } else {
            oracle_csr_spmv_f64(num_rows, ptr, indices, data, A, B);
            oracle_csr_spmv_f64(num_rows, ptr, indices, data, B, A);
        }
    }
    return 0;
}

ORACLE_API int oracle_csr_iterate_f32(int64_t num_rows, const int32_t *ptr, const int32_t *indices,
                                      const float *data, float *A, float *B, int nr_iterations, int use_omp) {
    if (nr_iterations % 2 != 0) return -1;
    for (int it = 0; it < nr_iterations; it += 2) {
        if (use_omp) {
            oracle_csr_spmv_f32_omp(num_rows, ptr, indices, data, A, B);
            oracle_csr_spmv_f32_omp(num_rows, ptr, indices, data, B, A);
        } else {
            oracle_csr_spmv_f32(num_rows, ptr, indices, data, A, B);
            oracle_csr_spmv_f32(num_rows, ptr, indices, data, B, A);
        }
    }
    return 0;
}

/* ------------------------------------------------------------------------
 * Final project: product + inclusive segmented sum scan.
 * Follows hw/hw_final/programming/aux/reference_spMVscan-released.cu:38-54
 * (sparseMV_scan): s holds p segment start offsets, s[0]=0, s[p-1]=n;
 *   b[s_k] = a[s_k]*x[k[s_k]];  b[j] = b[j-1] + a[j]*x[k[j]]  inside a segment.
 * The instructor's checker computes in double; fp.cu:139-149 computes the same
 * thing in float (multiply pass, then per-segment running sum).
 * ---------------------------------------------------------------------- */
#define DEFINE_SPMV_SCAN(NAME, T)                                                                     \
    ORACLE_API void NAME(int n, int p, const T *a, const int32_t *s, const int32_t *k, const T *x,   \
                         T *b) {                                                                      \
        (void)n;                                                                                      \
        for (int seg = 0; seg < p - 1; seg++) {                                                       \
            const int j0 = s[seg];                                                                    \
            b[j0] = a[j0] * x[k[j0]];                                                                 \
            for (int j = s[seg] + 1; j < s[seg + 1]; j++) b[j] = b[j - 1] + a[j] * x[k[j]];           \
        }                                                                                             \
    }
DEFINE_SPMV_SCAN(oracle_sparse_mv_scan_f64, double)
DEFINE_SPMV_SCAN(oracle_sparse_mv_scan_f32, float)

/* N iterations with the caller-side swap of
 * reference_spMVscan-released.cu:213-217: after each pass but the last, a <- b.
 * a is clobbered (it holds the input of the last pass on return). */
#define DEFINE_SPMV_SCAN_ITER(NAME, STEP, T)                                                          \
    ORACLE_API void NAME(int n, int p, T *a, const int32_t *s, const int32_t *k, const T *x, T *b,    \
                         int iters) {                                                                 \
        for (int i = 0; i < iters; i++) {                                                             \
            STEP(n, p, a, s, k, x, b);                                                                \
            if (i != iters - 1) memcpy(a, b, (size_t)n * sizeof(T));                                  \
        }                                                                                             \
    }
DEFINE_SPMV_SCAN_ITER(oracle_sparse_mv_scan_iterate_f64, oracle_sparse_mv_scan_f64, double)
DEFINE_SPMV_SCAN_ITER(oracle_sparse_mv_scan_iterate_f32, oracle_sparse_mv_scan_f32, float)

/* The CPU path of fp.cu as shipped (hw/hw_final/programming/fp.cu:124-150):
 * xx[i] = x[k[i]] hoisted, then per iteration an OpenMP multiply loop and an
 * OpenMP loop over segments doing the in-place running sum.  In place on cur. */
ORACLE_API void oracle_fp_cpu_iterate_f32(int n, int p, float *cur, const int32_t *s, const float *xx, int iters) {
    for (int iter = 0; iter < iters; ++iter) {
#pragma omp parallel for
        for (int i = 0; i < n; ++i) cur[i] *= xx[i];
#pragma omp parallel for
        for (int i = 1; i < p; ++i)
            for (int j = s[i - 1] + 1; j < s[i]; ++j) cur[j] += cur[j - 1];
    }
}

/* Inclusive segmented sum scan alone (fp.cu:142-149), out of place. */
#define DEFINE_SEGSCAN(NAME, T)                                                                       \
    ORACLE_API void NAME(int64_t n, int p, const T *in, const int32_t *s, T *out) {                  \
        (void)n;                                                                                      \
        for (int i = 1; i < p; ++i) {                                                                 \
            if (s[i - 1] < s[i]) out[s[i - 1]] = in[s[i - 1]];                                        \
            for (int j = s[i - 1] + 1; j < s[i]; ++j) out[j] = out[j - 1] + in[j];                    \
        }                                                                                             \
    }
DEFINE_SEGSCAN(oracle_segscan_inclusive_f32, float)
DEFINE_SEGSCAN(oracle_segscan_inclusive_f64, double)

/* Head-flag form of the same scan (slides/Lecture21.pdf "Segmented Scan":
 * flags mark the first element of each segment). */
ORACLE_API void oracle_segscan_flags_inclusive_f32(int64_t n, const float *in, const uint8_t *head, float *out) {
    float run = 0.f;
    for (int64_t i = 0; i < n; i++) {
        run = (head[i] || i == 0) ? in[i] : run + in[i];
        out[i] = run;
    }
}

ORACLE_API void oracle_segscan_flags_inclusive_i32(int64_t n, const int32_t *in, const uint8_t *head, int32_t *out) {
    int32_t run = 0;
    for (int64_t i = 0; i < n; i++) {
        run = (head[i] || i == 0) ? in[i] : run + in[i];
        out[i] = run;
    }
}

/* ------------------------------------------------------------------------
 * Exclusive scan of unsigned 32-bit counts.
 * Follows hw/hw4/programming/radixsort.cpp:76-81 (running sum, in place) and
 * :136-140 (exScan[i] = exScan[i-1] + h[i-1]); sequential scan() of
 * my-refs/scan.pdf p.4.  Arithmetic wraps mod 2^32.
 * ---------------------------------------------------------------------- */
ORACLE_API void oracle_exclusive_scan_u32(const uint32_t *in, uint32_t *out, int64_t n) {
    uint32_t sum = 0;
    for (int64_t i = 0; i < n; ++i) {
        uint32_t x = in[i];
        out[i] = sum;
        sum += x;
    }
}

/* Two-pass blocked OpenMP variant with the structure of
 * radixsort.cpp:53-108 (per-block reduce, serial scan of block totals,
 * per-block down-sweep); used only as the multi-thread CPU baseline. */
ORACLE_API void oracle_exclusive_scan_u32_omp(const uint32_t *in, uint32_t *out, int64_t n) {
    int nt = oracle_num_threads();
    if (nt < 1) nt = 1;
    int64_t nblocks = nt * 8;
    int64_t bs = (n + nblocks - 1) / nblocks;
    if (bs < 1) bs = 1;
    nblocks = (n + bs - 1) / bs;
    uint32_t *tot = (uint32_t *)malloc(sizeof(uint32_t) * (size_t)(nblocks + 1));
#pragma omp parallel for schedule(static)
    for (int64_t b = 0; b < nblocks; b++) {
        int64_t lo = b * bs, hi = lo + bs < n ? lo + bs : n;
        uint32_t s = 0;
        for (int64_t i = lo; i < hi; i++) s += in[i];
        tot[b] = s;
    }
    uint32_t sum = 0;
    for (int64_t b = 0; b < nblocks; b++) {
        uint32_t x = tot[b];
        tot[b] = sum;
        sum += x;
    }
#pragma omp parallel for schedule(static)
    for (int64_t b = 0; b < nblocks; b++) {
        int64_t lo = b * bs, hi = lo + bs < n ? lo + bs : n;
        uint32_t s = tot[b];
        for (int64_t i = lo; i < hi; i++) {
            uint32_t x = in[i];
            out[i] = s;
            s += x;
        }
    }
    free(tot);
}

/* Per-block scatter offsets of the parallel radix pass, hw/hw4/programming/radixsort.cpp:85-97:
 *   final[0][j] = global_exclusive[j];  final[i][j] = final[i-1][j] + local[i-1][j]
 * i.e. for every bucket j an exclusive scan down the blocks, started at the bucket's global offset. */
ORACLE_API void oracle_block_offsets_u32(const uint32_t *local, const uint32_t *global_excl, uint32_t *final,
                                         int64_t num_blocks, int64_t num_buckets) {
    for (int64_t j = 0; j < num_buckets; ++j) final[j] = global_excl[j];
    for (int64_t i = 1; i < num_blocks; ++i)
        for (int64_t j = 0; j < num_buckets; ++j)
            final[i * num_buckets + j] = final[(i - 1) * num_buckets + j] + local[(i - 1) * num_buckets + j];
}

/* One LSD radix pass built on the scan (radixsort.cpp:123-153,
 * radixSortSerialPass): histogram, exclusive scan, stable scatter. */
ORACLE_API void oracle_radix_pass_u32(const uint32_t *keys, uint32_t *out, int64_t n, uint32_t start_bit, uint32_t num_bits) {
    uint32_t nb = 1u << num_bits, mask = nb - 1;
    uint32_t *hist = (uint32_t *)calloc(nb, sizeof(uint32_t));
    uint32_t *ex = (uint32_t *)calloc(nb, sizeof(uint32_t));
    for (int64_t i = 0; i < n; i++) hist[(keys[i] >> start_bit) & mask]++;
    oracle_exclusive_scan_u32(hist, ex, nb);
    for (int64_t i = 0; i < n; i++) out[ex[(keys[i] >> start_bit) & mask]++] = keys[i];
    free(hist);
    free(ex);
}

/* The blocked pass, hw/hw4/programming/radixsort.cpp:44-109 (radixSortParallelPass): per-block histograms
 * (UPSWEEP 1, :50-57), summed over the blocks (:63-70), scanned (:75-81), pushed back down to per-block
 * offsets (:85-97), then each block scatters with its own offsets (:103-108).  OpenMP over the blocks as
 * in the reference; the result equals the serial pass whatever block_size is. */
ORACLE_API void oracle_radix_pass_blocked_u32(const uint32_t *keys, uint32_t *out, int64_t n, uint32_t start_bit,
                                              uint32_t num_bits, int64_t block_size) {
    const uint32_t nb = 1u << num_bits, mask = nb - 1;
    const int64_t blocks = (n + block_size - 1) / block_size;
    if (blocks == 0) return;
    uint32_t *local = (uint32_t *)calloc((size_t)blocks * nb, sizeof(uint32_t));
    uint32_t *final = (uint32_t *)calloc((size_t)blocks * nb, sizeof(uint32_t));
    uint32_t *glob = (uint32_t *)calloc(nb, sizeof(uint32_t));
#pragma omp parallel for schedule(static)
    for (int64_t b = 0; b < blocks; ++b) {
        const int64_t lo = b * block_size, hi = (b == blocks - 1) ? n : lo + block_size;
        uint32_t *h = local + b * nb;
        for (int64_t i = lo; i < hi; ++i) h[(keys[i] >> start_bit) & mask]++;
    }
    for (int64_t b = 0; b < blocks; ++b)
        for (uint32_t j = 0; j < nb; ++j) glob[j] += local[b * nb + j];
    uint32_t sum = 0;
    for (uint32_t j = 0; j < nb; ++j) { uint32_t x = glob[j]; glob[j] = sum; sum += x; }
    oracle_block_offsets_u32(local, glob, final, blocks, nb);
#pragma omp parallel for schedule(static)
    for (int64_t b = 0; b < blocks; ++b) {
        const int64_t lo = b * block_size, hi = (b == blocks - 1) ? n : lo + block_size;
        uint32_t *o = final + b * nb;
        for (int64_t i = lo; i < hi; ++i) out[o[(keys[i] >> start_bit) & mask]++] = keys[i];
    }
    free(local); free(final); free(glob);
}

/* radixSortParallel / radixSortSerial, radixsort.cpp:111-121 and :155-161: passes in pairs over
 * startBit = 0, numBits, 2 numBits ... while the pair's first start bit is below 32; the result is in keys.
 * block_size 0 = the serial pass, otherwise the blocked pass (the reference uses 8192, :115).
 * A start bit of 32 or more is undefined in the reference (shift by >= width); here such a pass sees digit 0
 * for every key, i.e. it copies. */
ORACLE_API void oracle_radix_sort_u32(uint32_t *keys, uint32_t *tmp, int64_t n, uint32_t num_bits, int64_t block_size) {
    for (uint32_t start = 0; start < 32; start += 2 * num_bits) {
        for (int half = 0; half < 2; ++half) {
            const uint32_t sb = start + (half ? num_bits : 0);
            uint32_t *src = half ? tmp : keys, *dst = half ? keys : tmp;
            if (sb >= 32) { memcpy(dst, src, (size_t)n * sizeof(uint32_t)); continue; }
            if (block_size > 0) oracle_radix_pass_blocked_u32(src, dst, n, sb, num_bits, block_size);
            else oracle_radix_pass_u32(src, dst, n, sb, num_bits);
        }
    }
}

/* ------------------------------------------------------------------------
 * Out-of-place transpose, M rows x N cols row-major -> N x M row-major.
 * Follows slides/Lecture06.pdf slide 9/20 line 10
 *   array_out[n*M + m] = array_in[m*N + n]
 * and computeTransposeGold of my-refs/MatrixTranspose.pdf.  Pure copy, so
 * 32- and 64-bit element versions cover float/int and double bit-exactly.
 * ---------------------------------------------------------------------- */
ORACLE_API void oracle_transpose_b32(const uint32_t *in, uint32_t *out, int64_t M, int64_t N) {
    for (int64_t m = 0; m < M; m++)
        for (int64_t n = 0; n < N; n++) out[n * M + m] = in[m * N + n];
}
ORACLE_API void oracle_transpose_b64(const uint64_t *in, uint64_t *out, int64_t M, int64_t N) {
    for (int64_t m = 0; m < M; m++)
        for (int64_t n = 0; n < N; n++) out[n * M + m] = in[m * N + n];
}
/* OpenMP over 64x64 tiles: the multi-thread CPU baseline. */
#define DEFINE_TRANSPOSE_OMP(NAME, T)                                                                 \
    ORACLE_API void NAME(const T *in, T *out, int64_t M, int64_t N) {                                 \
        const int64_t TB = 64;                                                                        \
        _Pragma("omp parallel for collapse(2) schedule(static)")                                      \
        for (int64_t m0 = 0; m0 < M; m0 += TB)                                                        \
            for (int64_t n0 = 0; n0 < N; n0 += TB) {                                                  \
                int64_t m1 = m0 + TB < M ? m0 + TB : M, n1 = n0 + TB < N ? n0 + TB : N;               \
                for (int64_t n = n0; n < n1; n++)                                                     \
                    for (int64_t m = m0; m < m1; m++) out[n * M + m] = in[m * N + n];                 \
            }                                                                                         \
    }
DEFINE_TRANSPOSE_OMP(oracle_transpose_b32_omp, uint32_t)
DEFINE_TRANSPOSE_OMP(oracle_transpose_b64_omp, uint64_t)

/* ------------------------------------------------------------------------
 * ELL: column-major data[num_rows_padded * n + row], indices likewise, padded
 * entries carry value 0.  Follows spmv_ell_kernel, Bell TR 2008 Fig. 17
 * (skip val == 0), with y = A x as in slides/Lecture22.pdf slide 20.
 * ---------------------------------------------------------------------- */
#define DEFINE_ELL_SPMV(NAME, T)                                                                      \
    ORACLE_API void NAME(int64_t num_rows, int64_t pitch, int num_cols_per_row, const int32_t *indices, \
                         const T *data, const T *x, T *y) {                                           \
        for (int64_t row = 0; row < num_rows; row++) {                                                \
            T dot = 0;                                                                                \
            for (int n = 0; n < num_cols_per_row; n++) {                                              \
                int32_t col = indices[pitch * n + row];                                               \
                T val = data[pitch * n + row];                                                        \
                if (val != 0) dot += val * x[col];                                                    \
            }                                                                                         \
            y[row] = dot;                                                                             \
        }                                                                                             \
    }
DEFINE_ELL_SPMV(oracle_ell_spmv_f32, float)
DEFINE_ELL_SPMV(oracle_ell_spmv_f64, double)

/* ------------------------------------------------------------------------
 * DIA: data[pitch * n + row] holds the entry of `row` on diagonal offsets[n]
 * (column = row + offsets[n]), zero where the matrix has none.  Follows
 * spmv_dia_kernel, Bell TR 2008 Fig. 15 / Bell & Garland SC09 section 3.1,
 * with the same "skip val == 0" as the ELL kernel so padding never touches x.
 * ---------------------------------------------------------------------- */
#define DEFINE_DIA_SPMV(NAME, T)                                                                      \
    ORACLE_API void NAME(int64_t num_rows, int64_t num_cols, int64_t pitch, int num_diags,            \
                         const int32_t *offsets, const T *data, const T *x, T *y) {                   \
        for (int64_t row = 0; row < num_rows; row++) {                                                \
            T dot = 0;                                                                                \
            for (int n = 0; n < num_diags; n++) {                                                     \
                int64_t col = row + offsets[n];                                                       \
                T val = data[pitch * n + row];                                                        \
                if (val != 0 && col >= 0 && col < num_cols) dot += val * x[col];                      \
            }                                                                                         \
            y[row] = dot;                                                                             \
        }                                                                                             \
    }
DEFINE_DIA_SPMV(oracle_dia_spmv_f32, float)
DEFINE_DIA_SPMV(oracle_dia_spmv_f64, double)

/* The occupied diagonals of a CSR matrix in ascending order (SC09 section 3.1: "offsets"); returns how many
 * there are and writes at most max_out of them. */
ORACLE_API int64_t oracle_csr_diagonals(int64_t num_rows, int64_t num_cols, const int32_t *ptr, const int32_t *indices,
                                        int64_t max_out, int32_t *out) {
    int64_t slots = num_rows + num_cols - 1, count = 0;
    if (slots <= 0) return 0;
    unsigned char *mark = (unsigned char *)calloc((size_t)slots, 1);
    for (int64_t r = 0; r < num_rows; r++)
        for (int32_t e = ptr[r]; e < ptr[r + 1]; e++) mark[(int64_t)indices[e] - r + (num_rows - 1)] = 1;
    for (int64_t i = 0; i < slots; i++)
        if (mark[i]) { if (count < max_out) out[count] = (int32_t)(i - (num_rows - 1)); count++; }
    free(mark);
    return count;
}

/* CSR -> DIA on a given ascending offsets list; returns the number of entries on no listed diagonal. */
#define DEFINE_CSR_TO_DIA(NAME, T)                                                                    \
    ORACLE_API int64_t NAME(int64_t num_rows, const int32_t *ptr, const int32_t *indices, const T *vals, \
                            int num_diags, const int32_t *offsets, int64_t pitch, T *data) {          \
        int64_t miss = 0;                                                                             \
        memset(data, 0, sizeof(T) * (size_t)num_diags * (size_t)pitch);                               \
        for (int64_t r = 0; r < num_rows; r++)                                                        \
            for (int32_t e = ptr[r]; e < ptr[r + 1]; e++) {                                           \
                int n = -1;                                                                           \
                for (int k = 0; k < num_diags; k++)                                                   \
                    if ((int64_t)offsets[k] == (int64_t)indices[e] - r) { n = k; break; }             \
                if (n < 0) miss++; else data[pitch * n + r] = vals[e];                                \
            }                                                                                         \
        return miss;                                                                                  \
    }
DEFINE_CSR_TO_DIA(oracle_csr_to_dia_f32, float)
DEFINE_CSR_TO_DIA(oracle_csr_to_dia_f64, double)

/* COO, y += A x, one entry at a time in storage order (Bell TR 2008 Fig. 23's
 * serial semantics; the segmented-reduction kernel of Fig. 24 computes the
 * same sums in a different order). */
#define DEFINE_COO_SPMV(NAME, T)                                                                      \
    ORACLE_API void NAME(int64_t nnz, const int32_t *rows, const int32_t *cols, const T *vals,        \
                         const T *x, T *y) {                                                          \
        for (int64_t i = 0; i < nnz; i++) y[rows[i]] += vals[i] * x[cols[i]];                         \
    }
DEFINE_COO_SPMV(oracle_coo_spmv_accumulate_f32, float)
DEFINE_COO_SPMV(oracle_coo_spmv_accumulate_f64, double)

/* HYB split width, Bell & Garland SC09 p.5 section 3.5: grow the ELL part one
 * column at a time while at least max(4096, M/3) rows have that many
 * entries.  Returns K (possibly 0). */
ORACLE_API int oracle_hyb_pick_k(int64_t num_rows, const int32_t *ptr) {
    int32_t maxlen = 0;
    for (int64_t r = 0; r < num_rows; r++) {
        int32_t l = ptr[r + 1] - ptr[r];
        if (l > maxlen) maxlen = l;
    }
    int64_t *hist = (int64_t *)calloc((size_t)maxlen + 2, sizeof(int64_t));
    for (int64_t r = 0; r < num_rows; r++) hist[ptr[r + 1] - ptr[r]]++;
    int64_t thresh = num_rows / 3 > 4096 ? num_rows / 3 : 4096;
    int64_t rows_ge = 0; /* rows with length >= K */
    int K = 0;
    /* rows with length >= K+1, scanning K upward */
    int64_t *ge = (int64_t *)calloc((size_t)maxlen + 2, sizeof(int64_t));
    for (int32_t l = maxlen; l >= 0; l--) ge[l] = ge[l + 1] + hist[l];
    (void)rows_ge;
    while (K < maxlen && ge[K + 1] >= thresh) K++;
    free(hist);
    free(ge);
    return K;
}

/* CSR -> ELL (column-major, pitch >= num_rows).  Rows longer than K are
 * truncated (their tail belongs to the COO part of HYB).  Padding: value 0,
 * index = the row's own index clamped to a valid column (any valid column is
 * allowed, SC09 Fig. 2 marks them '*'). */
#define DEFINE_CSR_TO_ELL(NAME, T)                                                                    \
    ORACLE_API void NAME(int64_t num_rows, int64_t num_cols, const int32_t *ptr, const int32_t *indices, \
                         const T *data, int K, int64_t pitch, int32_t *ell_idx, T *ell_val) {         \
        for (int64_t r = 0; r < num_rows; r++) {                                                      \
            int32_t len = ptr[r + 1] - ptr[r];                                                        \
            int32_t padcol = (int32_t)(r < num_cols ? r : num_cols - 1);                              \
            for (int n = 0; n < K; n++) {                                                             \
                if (n < len) {                                                                        \
                    ell_idx[pitch * n + r] = indices[ptr[r] + n];                                     \
                    ell_val[pitch * n + r] = data[ptr[r] + n];                                        \
                } else {                                                                              \
                    ell_idx[pitch * n + r] = padcol;                                                  \
                    ell_val[pitch * n + r] = 0;                                                       \
                }                                                                                     \
            }                                                                                         \
        }                                                                                             \
        for (int n = 0; n < K; n++)                                                                   \
            for (int64_t r = num_rows; r < pitch; r++) {                                              \
                ell_idx[pitch * n + r] = 0;                                                           \
                ell_val[pitch * n + r] = 0;                                                           \
            }                                                                                         \
    }
DEFINE_CSR_TO_ELL(oracle_csr_to_ell_f32, float)
DEFINE_CSR_TO_ELL(oracle_csr_to_ell_f64, double)

/* COO tail of the HYB split: entries at in-row position >= K, in row order. */
ORACLE_API int64_t oracle_hyb_tail_count(int64_t num_rows, const int32_t *ptr, int K) {
    int64_t c = 0;
    for (int64_t r = 0; r < num_rows; r++) {
        int32_t len = ptr[r + 1] - ptr[r];
        if (len > K) c += len - K;
    }
    return c;
}
#define DEFINE_CSR_TO_COO_TAIL(NAME, T)                                                               \
    ORACLE_API void NAME(int64_t num_rows, const int32_t *ptr, const int32_t *indices, const T *data, \
                         int K, int32_t *coo_row, int32_t *coo_col, T *coo_val) {                     \
        int64_t o = 0;                                                                                \
        for (int64_t r = 0; r < num_rows; r++)                                                        \
            for (int32_t j = ptr[r] + K; j < ptr[r + 1]; j++) {                                       \
                coo_row[o] = (int32_t)r;                                                              \
                coo_col[o] = indices[j];                                                              \
                coo_val[o] = data[j];                                                                 \
                o++;                                                                                  \
            }                                                                                         \
    }
DEFINE_CSR_TO_COO_TAIL(oracle_csr_to_coo_tail_f32, float)
DEFINE_CSR_TO_COO_TAIL(oracle_csr_to_coo_tail_f64, double)

/* ------------------------------------------------------------------------
 * Comparators.
 * AlmostEqual2sComplement, hw/hw1/programming/mp1-util.h:44-61 (fp32).
 * ---------------------------------------------------------------------- */
ORACLE_API int oracle_almost_equal_ulps_f32(float A, float B, int maxUlps) {
    int32_t aInt, bInt;
    memcpy(&aInt, &A, 4);
    if (aInt < 0) aInt = (int32_t)(0x80000000u - (uint32_t)aInt);
    memcpy(&bInt, &B, 4);
    if (bInt < 0) bInt = (int32_t)(0x80000000u - (uint32_t)bInt);
    int64_t d = (int64_t)aInt - (int64_t)bInt;
    if (d < 0) d = -d;
    return d <= maxUlps;
}

/* Number of elements failing the ULP check (pagerank.cu:216-239). */
ORACLE_API int64_t oracle_count_ulp_errors_f32(const float *got, const float *want, int64_t n, int maxUlps) {
    int64_t bad = 0;
    for (int64_t i = 0; i < n; i++)
        if (!oracle_almost_equal_ulps_f32(got[i], want[i], maxUlps)) bad++;
    return bad;
}

/* Error norms of reference_spMVscan-released.cu:65-144:
 * out[0]=||b-a||_2, out[1]=||b-a||_inf, out[2]=||b-a||_2/||a||_2, out[3]=||b-a||_inf/||a||_inf */
ORACLE_API void oracle_error_norms(const double *a, const double *b, int64_t n, double *out) {
    double d2 = 0, a2 = 0, dinf = 0, ainf = 0;
    for (int64_t i = 0; i < n; i++) {
        double d = a[i] - b[i];
        d2 += d * d;
        a2 += a[i] * a[i];
        if (fabs(d) > dinf) dinf = fabs(d);
        if (fabs(a[i]) > ainf) ainf = fabs(a[i]);
    }
    out[0] = sqrt(d2);
    out[1] = dinf;
    out[2] = a2 > 0 ? sqrt(d2) / sqrt(a2) : 0;
    out[3] = ainf > 0 ? dinf / ainf : 0;
}

/* ---- generator restatement: the 7-point Laplacian rows of cme213_gen_laplace3d_f64 (csrc/generate.cu), built on the host
 * with OpenMP so that bench.py's CPU arm can run config C5 on the WHOLE 512^3 matrix (938 M nonzeros) in seconds instead
 * of the minutes numpy needs.  Same row layout as oracle/generators.py:laplace3d (legs -z,-y,-x,centre,+x,+y,+z; centre 6,
 * neighbours -1; natural order, x fastest).  Two passes: row lengths -> serial prefix (the scan of radixsort.cpp:76-81) -> fill. */
ORACLE_API int64_t oracle_gen_laplace3d_f64(int nx, int ny, int nz, int64_t row_begin, int64_t row_end, int32_t *ptr,
                                            int32_t *indices, double *data) {
    const int64_t plane = (int64_t)nx * ny, rows = row_end - row_begin;
    if (rows <= 0) return 0;
#pragma omp parallel for schedule(static)
    for (int64_t t = 0; t < rows; t++) {
        const int64_t i = row_begin + t;
        const int ix = (int)(i % nx), iy = (int)((i / nx) % ny), iz = (int)(i / plane);
        ptr[t + 1] = 1 + (iz > 0) + (iy > 0) + (ix > 0) + (ix < nx - 1) + (iy < ny - 1) + (iz < nz - 1);
    }
    ptr[0] = 0;
    int64_t sum = 0;
    for (int64_t t = 1; t <= rows; t++) { sum += ptr[t]; ptr[t] = (int32_t)sum; }
    if (!indices || !data) return sum;
#pragma omp parallel for schedule(static)
    for (int64_t t = 0; t < rows; t++) {
        const int64_t i = row_begin + t;
        const int ix = (int)(i % nx), iy = (int)((i / nx) % ny), iz = (int)(i / plane);
        int64_t e = ptr[t];
        if (iz > 0) { indices[e] = (int32_t)(i - plane); data[e++] = -1.0; }
        if (iy > 0) { indices[e] = (int32_t)(i - nx); data[e++] = -1.0; }
        if (ix > 0) { indices[e] = (int32_t)(i - 1); data[e++] = -1.0; }
        indices[e] = (int32_t)i; data[e++] = 6.0;
        if (ix < nx - 1) { indices[e] = (int32_t)(i + 1); data[e++] = -1.0; }
        if (iy < ny - 1) { indices[e] = (int32_t)(i + nx); data[e++] = -1.0; }
        if (iz < nz - 1) { indices[e] = (int32_t)(i + plane); data[e++] = -1.0; }
    }
    return sum;
}
