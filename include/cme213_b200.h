/*
 * cme213_b200.h -- C ABI of libcme213_b200.so
 *
 * B200-native (sm_100a) replacement for the data-parallel hot path of the
 * Stanford CME213 (Spring 2012) coursework: CSR / ELL / HYB sparse
 * matrix-vector multiply, the exclusive / segmented scan and the tiled
 * transpose it rests on, and the host-pointer driver entry points the
 * homework programs call.  Plain pointers and sizes only; no C++ or torch
 * types cross this boundary.
 *
 * Conventions
 *   - every function returns int: 0 = CME213_OK, negative = error
 *     (cme213_last_error() gives the text).  Nothing here calls exit(): the
 *     reference prints and exits (hw/hw1/programming/mp1-util.h:8-18), a
 *     library cannot.
 *   - pointers named d_* are device pointers, h_* are host pointers.
 *   - `stream` is a cudaStream_t passed as void* (NULL = default stream).
 *     Device-pointer functions only enqueue work; they do not synchronise.
 *   - indices are 32-bit as in the reference (unsigned int in hw1, int in
 *     hw_final); element counts and byte offsets are 64-bit.
 *   - "reference:" comments cite the file:line (relative to the reference
 *     checkout) that each entry point replaces or restates.
 */
#ifndef CME213_B200_H
#define CME213_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CME213_OK 0
#define CME213_ERR_INVALID (-1)     /* bad argument (null pointer, negative size, odd iteration count ...) */
#define CME213_ERR_CUDA (-2)        /* a CUDA runtime call or kernel launch failed */
#define CME213_ERR_ALLOC (-3)       /* device or host allocation failed */
#define CME213_ERR_UNSUPPORTED (-4) /* size beyond a documented limit */
#define CME213_ERR_TIMEOUT (-5)     /* multi-GPU: a rank gave up waiting for a neighbour (cme213_shard_finish) */

/* CSR kernel selection (cme213_spmv_csr_*).  reference: slides/Lecture22.pdf slides 20-26,
 * refs/Bell Tech Report 2008.pdf Figs. 20, 22. */
#define CME213_CSR_AUTO 0
#define CME213_CSR_SCALAR 1 /* one thread per row, sequential sum: bit-identical to the CPU loop */
#define CME213_CSR_VECTOR 2 /* 2..32 lanes per row, warp-shuffle reduction */
#define CME213_CSR_STREAM 3 /* block streams its nonzeros through shared memory, then one thread per row sums sequentially: bit-identical to the CPU loop */
#define CME213_CSR_STREAM_KEEP 5   /* CSR_STREAM with the x gathers marked evict-last in L2: x fits L2 but is gathered at random while the matrix streams by */
#define CME213_CSR_STREAM_PADDED 4 /* same, shared-memory products padded against bank conflicts: for matrices whose rows all have 8, 16, 32 ... entries.
                                      AUTO picks between 3 and 4 from the mean row length (one small device read per new matrix). */

#define CME213_CSR_AUTO_ORDERED 8 /* the dispatcher restricted to kernels that keep the CPU loop's summation order (scalar, stream, stream-padded):
                                     bit-identical to the CPU loop on every matrix, at the price of long rows being summed by one thread */
#define CME213_PLAN_ORDERED 0x100  /* OR into cme213_csr_plan_create's elem_bytes for the same restriction */
#define CME213_CSR_STREAM_BULK 7 /* CSR_STREAM with the block's col / val range brought into shared memory by the bulk-copy engine (cp.async.bulk)
                                  * instead of through registers: one round trip fewer per block and every gather of a thread in flight at
                                  * once.  Same order of summation (bit-identical).  The dispatcher takes it for fp32 matrices with a mean
                                  * row of 8 entries or fewer (5-point grids: 0.63 -> 0.78 of the copy peak at 4096^2, C1 0.49 -> 0.57);
                                  * fp64 stays with CSR_STREAM (C5: 2.61 against 2.52 ms).  Needs 16-byte aligned col / val (else CSR_STREAM). */
#define CME213_CSR_STREAM_PIPE (CME213_CSR_STREAM_BULK | (1 << 8)) /* the same as a two-stage pipeline in persistent CTAs: the bulk copies of a CTA's next
                                  * row block are in flight while this one gathers.  Meant for matrices whose 128-row blocks hold at most 1024
                                  * entries (the dispatcher takes it for fp32 when the longest row has 7 or fewer: C1 0.57, 4096^2 0.82 of the
                                  * copy peak); a block that holds more is summed one thread per row (correct, slow).  Bit-identical as well. */
#define CME213_CSR_BINNED 6 /* plans only: 256-row blocks split into a short list (stream kernel) and a long list (vector kernel) */

typedef struct cme213_ctx cme213_ctx;
typedef struct cme213_csr_plan cme213_csr_plan;

/* ---- context ---------------------------------------------------------- */
/* Binds to a CUDA device and owns a scratch workspace that grows on demand
 * (cudaMalloc on first use / growth, never inside a steady-state loop).
 * reference: the callee-owns-device-memory rule of hw/hw1/programming/pagerank.cu:85-93.
 * Sharing rules of that one workspace (scans, radix sort, format conversions, cme213_block_offsets_u32 and
 * the generators that scan all keep their descriptors / tables in it):
 *   - a context is driven by one host thread at a time;
 *   - calls on different streams through the same context are ordered by the library: before a call re-uses the
 *     workspace on another stream than the previous user's, that stream waits (event, no host sync) for what the
 *     previous user enqueued.  This wait is skipped while the stream is being captured into a CUDA graph -- order
 *     the streams yourself there, or give each stream its own context (contexts are cheap);
 *   - growth (first use, or a larger problem than any before) calls cudaDeviceSynchronize + cudaFree + cudaMalloc:
 *     run the largest problem once during set-up, outside any stream capture. */
int cme213_ctx_create(int device, cme213_ctx **out);
int cme213_ctx_destroy(cme213_ctx *ctx);
int cme213_ctx_device(const cme213_ctx *ctx);
int cme213_ctx_sm_count(const cme213_ctx *ctx);
/* Number of kernels this context has launched since creation (bench.py's gpu_launches). */
int64_t cme213_ctx_launch_count(const cme213_ctx *ctx);
const char *cme213_last_error(void);
int cme213_version(void);
/* Hash of the sources (csrc/, include/) this binary was built from; the Python binding refuses a library whose id
 * differs from the sources beside it (the .so is a build artefact and could be stale). */
const char *cme213_build_id(void);

/* ---- transpose --------------------------------------------------------- */
/* out[c*ld_out + r] = in[r*ld_in + c] for r < rows, c < cols (row-major, out of place).
 * ld_in >= cols, ld_out >= rows, in elements.  b32 covers float/int32, b64 covers double.
 * reference: slides/Lecture06.pdf slide 9/20 (array_out[n*M+m] = array_in[m*N+n]);
 * slides/Lecture07.pdf slides 2-4 (fastTranspose); my-refs/MatrixTranspose.pdf. */
int cme213_transpose_b32(cme213_ctx *ctx, const void *d_in, void *d_out, int64_t rows, int64_t cols,
                         int64_t ld_in, int64_t ld_out, void *stream);
int cme213_transpose_b64(cme213_ctx *ctx, const void *d_in, void *d_out, int64_t rows, int64_t cols,
                         int64_t ld_in, int64_t ld_out, void *stream);

/* ---- scans ------------------------------------------------------------- */
/* out[0] = 0, out[i] = in[0] + ... + in[i-1] (mod 2^32).  Single pass, decoupled look-back.
 * In place (d_out == d_in) is allowed.
 * reference: hw/hw4/programming/radixsort.cpp:76-81 and :136-140; my-refs/scan.pdf p.4. */
int cme213_exclusive_scan_u32(cme213_ctx *ctx, const uint32_t *d_in, uint32_t *d_out, int64_t n, void *stream);

/* The same scan for callers that must not allocate device memory when they call it (a host that gave the
 * whole GPU to its own pool, a steady-state loop that cannot afford cudaMalloc's device-wide synchronisation
 * or a late out-of-memory error, a plug-in that only ever receives input and output buffers).  It takes no
 * context: the look-back descriptors live in a fixed array inside the library (about 0.5 MiB of device
 * memory on every GPU that loads the library, until the process exits), so the call performs no cudaMalloc,
 * cudaFree or other allocating call.  Consequences:
 *   - n is limited to cme213_exclusive_scan_u32_static_max_n() = 2^28 elements; larger n returns
 *     CME213_ERR_UNSUPPORTED and touches nothing;
 *   - every GPU has its own copy of the array and the call uses the copy of the CURRENT device: make the
 *     device that holds d_in / d_out current (cudaSetDevice) before calling; buffers of another device are
 *     refused with CME213_ERR_INVALID;
 *   - the first call on each device looks up the array's device address (cudaGetSymbolAddress); make that
 *     first call during set-up, before any stream capture begins;
 *   - all calls on a device share the one array (each call clears the part it uses first), so two calls whose
 *     kernels may overlap on the GPU -- different streams without an ordering between them -- race; keep
 *     calls on one device ordered.
 * Results are bit-identical to cme213_exclusive_scan_u32.  In place (d_out == d_in) is allowed.
 * This entry point has no counterpart in the reference. */
int64_t cme213_exclusive_scan_u32_static_max_n(void);
int cme213_exclusive_scan_u32_static(const uint32_t *d_in, uint32_t *d_out, int64_t n, void *stream);

/* Segment descriptions.  The reference stores p segment START offsets s[0]=0 < ... < s[p-1]=n
 * (hw/hw_final/programming/aux/mp1-util.h:118-139).  The kernels read a head-flag bitmask instead
 * (bit i of word i/32 set = element i starts a segment); it depends on s only, so it is built
 * once outside the timed loop, which hw/hw_final/Final.pdf p.3 allows.
 * d_flag_bits must hold cme213_flag_words(n) uint32 words. */
int64_t cme213_flag_words(int64_t n);
int cme213_segment_flags_from_offsets(cme213_ctx *ctx, const int32_t *d_s, int p, int64_t n,
                                      uint32_t *d_flag_bits, void *stream);

/* Inclusive segmented sum scan.  In place allowed.
 * reference: hw/hw_final/programming/fp.cu:142-149 (CPU), :39-59 (GPU SegmentedScan). */
int cme213_segscan_inclusive_f32(cme213_ctx *ctx, const float *d_in, const uint32_t *d_flag_bits, float *d_out,
                                 int64_t n, void *stream);
int cme213_segscan_inclusive_f64(cme213_ctx *ctx, const double *d_in, const uint32_t *d_flag_bits, double *d_out,
                                 int64_t n, void *stream);
/* Exclusive variant (first element of each segment gets 0). */
int cme213_segscan_exclusive_f32(cme213_ctx *ctx, const float *d_in, const uint32_t *d_flag_bits, float *d_out,
                                 int64_t n, void *stream);

/* One step of the final project, fused: b[j] = inclusive segmented sum of a[i]*x[k[i]].
 * d_b == d_a allowed.  The x gather happens on the device inside the step.
 * reference: hw/hw_final/programming/aux/reference_spMVscan-released.cu:38-54 (sparseMV_scan);
 * hw/hw_final/programming/fp.cu:174-180 (transform + SegmentedScan). */
int cme213_mul_gather_segscan_f32(cme213_ctx *ctx, const float *d_a, const int32_t *d_k, const float *d_x,
                                  const uint32_t *d_flag_bits, float *d_b, int64_t n, void *stream);
int cme213_mul_gather_segscan_f64(cme213_ctx *ctx, const double *d_a, const int32_t *d_k, const double *d_x,
                                  const uint32_t *d_flag_bits, double *d_b, int64_t n, void *stream);

/* The same step with the gather hoisted out of the iteration: xx[i] = x[k[i]] once (cme213_gather_*), then
 * b = inclusive segmented sum of a[i]*xx[i] per iteration -- pure streaming, no random access.  The rounded product and
 * the segmented sum are those of the fused step; only the tiling of the scan differs (same tolerance).  d_b == d_a allowed.
 * reference: hw/hw_final/programming/fp.cu:124-125 (xx = x[k] flattened once), :139-149 / :174-180 (multiply, scan). */
int cme213_gather_f32(cme213_ctx *ctx, const float *d_x, const int32_t *d_k, float *d_xx, int64_t n, void *stream);
int cme213_gather_f64(cme213_ctx *ctx, const double *d_x, const int32_t *d_k, double *d_xx, int64_t n, void *stream);
int cme213_mul_segscan_f32(cme213_ctx *ctx, const float *d_a, const float *d_xx, const uint32_t *d_flag_bits, float *d_b,
                           int64_t n, void *stream);
int cme213_mul_segscan_f64(cme213_ctx *ctx, const double *d_a, const double *d_xx, const uint32_t *d_flag_bits, double *d_b,
                           int64_t n, void *stream);

/* ---- CSR SpMV ---------------------------------------------------------- */
/* y[r] = sum_j val[j] * x[col[j]], j in [row_ptr[r], row_ptr[r+1]), for rows r in [row_begin, row_end).
 * row_ptr, y are indexed by r itself (not r - row_begin) so a caller can split one matrix into
 * several launches (boundary rows first, interior rows later).
 * reference: slides/Lecture22.pdf slide 20 (csr_spmv), slides 21/25 (scalar / vector kernels);
 * hw/hw1/programming/pagerank.cu:70-83 (the CSR-scalar shape). */
int cme213_spmv_csr_f32(cme213_ctx *ctx, int64_t row_begin, int64_t row_end, const int32_t *d_row_ptr,
                        const int32_t *d_col, const float *d_val, const float *d_x, float *d_y, int kernel,
                        void *stream);
int cme213_spmv_csr_f64(cme213_ctx *ctx, int64_t row_begin, int64_t row_end, const int32_t *d_row_ptr,
                        const int32_t *d_col, const double *d_val, const double *d_x, double *d_y, int kernel,
                        void *stream);

/* The dispatcher as an object.  CME213_CSR_AUTO in the calls above builds (first call on a matrix: one small kernel and
 * ONE synchronisation) and caches such a plan per (row_ptr, row range); a caller that must never synchronise -- a timed
 * loop, a stream capture -- creates the plan during set-up and multiplies through it.  The plan reads the row-length
 * histogram and picks (thresholds measured on B200, profiles/r02_csr_kernel_sweep.csv):
 *   scalar         mean row length <= 2.5
 *   stream-padded  every row 8, 16, 24 ... 128 entries long
 *   vector         mean >= 96 (16 lanes below a mean of 128, 32 from there on)
 *   BINNED         skewed matrices whose long rows sit together: 256-row blocks with >= 32 entries per row holding >= 30 % of
 *                  the nonzeros go to the vector kernel, the others to the stream kernel (lists built by mark +
 *                  exclusive scan + compaction; two launches)
 *   stream         everything else
 * scalar / stream / stream-padded keep the CPU loop's summation order (bit-identical to it); vector and the long blocks
 * of BINNED sum in another order (reduction-order bound gamma_k * sum|a_ij x_j|).  CME213_CSR_AUTO_ORDERED / CME213_PLAN_ORDERED
 * restrict the choice to the order-preserving kernels.
 * reference: slides/Lecture22.pdf slides 20, 21, 25; refs/Bell Tech Report 2008.pdf Figs. 20, 22 and section 5;
 * hw/hw_final/programming/paper/Final_Report_DongBang_Tsai.tex:228 (warp-per-segment regresses on short segments). */
int cme213_csr_plan_create(cme213_ctx *ctx, int64_t row_begin, int64_t row_end, int64_t num_cols, const int32_t *d_row_ptr,
                           int elem_bytes, void *stream, cme213_csr_plan **out);
int cme213_csr_plan_destroy(cme213_csr_plan *plan);
/* CME213_CSR_AUTO's cached plans are keyed by the ADDRESS of row_ptr (and the row range): if a matrix is freed and another
 * of the same size lands at the same address, the old plan is applied to it.  Every kernel is correct on every matrix, so
 * results stay right -- only the choice may no longer fit.  Call this when matrices come and go (or use explicit plans). */
int cme213_ctx_clear_plans(cme213_ctx *ctx);
int cme213_csr_plan_kernel(const cme213_csr_plan *plan);           /* CME213_CSR_* (| lanes << 8) */
const char *cme213_csr_plan_describe(const cme213_csr_plan *plan); /* one line: choice and the statistics behind it */
int cme213_spmv_csr_plan_f32(cme213_ctx *ctx, const cme213_csr_plan *plan, const int32_t *d_row_ptr, const int32_t *d_col,
                             const float *d_val, const float *d_x, float *d_y, void *stream);
int cme213_spmv_csr_plan_f64(cme213_ctx *ctx, const cme213_csr_plan *plan, const int32_t *d_row_ptr, const int32_t *d_col,
                             const double *d_val, const double *d_x, double *d_y, void *stream);
/* Which FORMAT to convert a CSR matrix to, from the same histogram (synchronises): ELL when padding every row to the
 * longest costs less than running the matrix as CSR, else HYB with the split width K that minimises
 * K * rows + c * tail(K) when that beats CSR, else CSR.  c = 1.6: the measured ratio between the gather rates of the
 * ELL kernel and of the CSR stream kernel that serves the tail.  reference: refs/Bell SC 2009.pdf section 3.5 (the
 * same rule with c = 3 and a 4096-row floor: cme213_hyb_pick_k), Table 1 and Fig. 9. */
#define CME213_FORMAT_CSR 0
#define CME213_FORMAT_ELL 1
#define CME213_FORMAT_HYB 2
typedef struct cme213_format_choice {
    int format;               /* CME213_FORMAT_* */
    int K;                    /* ELL width (ELL: the longest row; HYB: the split) */
    int max_row;
    double mean_row;
    double ell_padding_ratio; /* max_row * rows / nnz */
    double tail_fraction;     /* share of the nonzeros left to the HYB tail */
    double cost_vs_csr;       /* modelled cost of the choice relative to plain CSR (< 1: the conversion pays) */
} cme213_format_choice;
int cme213_spmv_pick_format(cme213_ctx *ctx, int64_t num_rows, const int32_t *d_row_ptr, cme213_format_choice *out, void *stream);

/* Same result (bit for bit), with the caller choosing WHICH blocks of cme213_spmv_csr_block_rows() consecutive
 * rows one launch computes and in what order: CTA b (b < num_blocks) computes block number d_block_order[b],
 * counted from row_begin.  d_block_order holds distinct block numbers < ceil((row_end-row_begin)/block_rows);
 * rows of blocks that are not listed are left untouched.  Uses: the multi-GPU iterate computes the blocks of
 * its first and last halo rows in one launch before the exchange starts; an order that keeps rows touching
 * the same x entries close in time was tried for the 3-D stencil and measured slower (profiles/r01_notes.md).
 * No reference counterpart: the reference's kernels walk rows in index order (pagerank.cu:72). */
int cme213_spmv_csr_block_rows(void);
int cme213_spmv_csr_f32_ordered(cme213_ctx *ctx, int64_t row_begin, int64_t row_end, const int32_t *d_row_ptr,
                                const int32_t *d_col, const float *d_val, const float *d_x, float *d_y,
                                const int32_t *d_block_order, int64_t num_blocks, void *stream);
int cme213_spmv_csr_f64_ordered(cme213_ctx *ctx, int64_t row_begin, int64_t row_end, const int32_t *d_row_ptr,
                                const int32_t *d_col, const double *d_val, const double *d_x, double *d_y,
                                const int32_t *d_block_order, int64_t num_blocks, void *stream);

/* Multi-GPU boundary launch with the exchange fused into the kernel: like cme213_spmv_csr_f64_ordered, and rows
 * [row_begin, row_begin+halo_rows) are also stored into d_peer_prev_y[r], rows [row_end-halo_rows, row_end) into
 * d_peer_next_y[r] (either may be NULL): the neighbouring GPUs' vectors, mapped with cme213_ipc_open and shifted
 * like d_y, so the halo planes travel over NVLink as the rows are produced.  The caller still has to order the
 * ranks (any collective after the launch) before the neighbours read what was pushed.
 * reference: none (the reference's only exchange is the CPU MPI halo of hw/hw5/programming/2dHeat.cpp:503-547). */
int cme213_spmv_csr_f64_halo_push(cme213_ctx *ctx, int64_t row_begin, int64_t row_end, const int32_t *d_row_ptr,
                                  const int32_t *d_col, const double *d_val, const double *d_x, double *d_y,
                                  const int32_t *d_block_order, int64_t num_blocks, int64_t halo_rows,
                                  double *d_peer_prev_y, double *d_peer_next_y, void *stream);
/* Device vectors another process of the same box can map: one cudaMalloc each (exportable as a whole), an opaque
 * handle of cme213_ipc_handle_bytes() bytes to send to the peer, and open/close on the peer's side. */
int cme213_device_malloc(size_t bytes, void **d_out);
int cme213_device_free(void *d_ptr);
int cme213_ipc_handle_bytes(void);
int cme213_ipc_export(const void *d_ptr, void *handle_out);
int cme213_ipc_open(const void *handle, void **d_out);
int cme213_ipc_close(void *d_ptr);

/* ---- row-sharded iterate across the GPUs of one box (config C5) ------------------------------------------
 * x <- A x repeated, A split into contiguous row blocks, one cme213_shard per GPU -- in one process
 * (cme213_shard_connect_local) or one process per GPU (cme213_shard_export + cme213_shard_connect; the handles
 * travel by any means the host has: MPI, torch.distributed, a pipe).  No NCCL and no host synchronisation on the
 * data path; the ranks only have to run the same number of passes.
 * reference: the 1-D stripe decomposition and the overlapped halo exchange of the reference's only distributed
 * code, hw/hw5/programming/2dHeat.cpp:284-306 (decomposition), :696-723 (asyncComputation), :503-547 (exchange);
 * the iteration itself is the ping-pong loop of hw/hw1/programming/pagerank.cu:59-67,128-132.
 * Exchange modes: */
#define CME213_SHARD_AUTO 0          /* HALO when every rank's columns stay inside its two neighbours' rows, else GATHER */
#define CME213_SHARD_HALO 1          /* ONE kernel launch per pass: boundary rows stored into the neighbours' vectors over NVLink
                                        by the kernel itself, passes ordered by flags in peer memory */
#define CME213_SHARD_GATHER 2        /* any matrix: SpMV, then every rank copies its slice into every peer's vector */
#define CME213_SHARD_GATHER_SPLIT 3  /* same exchange overlapped with the next pass's local-column block; the split
                                        changes the summation order of a row (reduction-order tolerance, not bitwise);
                                        uses a second stream of its own, so its steps cannot be stream-captured */
/* HALO and GATHER leave the order in which a row is summed untouched: results are bitwise equal to one GPU. */
typedef struct cme213_shard cme213_shard;
/* Row block bounds (world + 1 entries, bounds[0] = 0, bounds[world] = num_rows) balanced by nonzeros when the HOST
 * row pointers are given (NULL: ceil(rows/world) rows each as in 2dHeat.cpp:284-306), rounded to `align` rows. */
int cme213_partition_rows(int64_t num_rows, int world, const int32_t *h_row_ptr, int64_t align, int64_t *bounds_out);
/* One rank's part: rows [bounds[rank], bounds[rank+1]) as CSR with row pointers rebased to 0 and GLOBAL column
 * indices (the arrays stay the caller's).  Allocates the two full-length ping-pong vectors and a small flag block
 * (all exportable), and finds the ghost rows the block reads (synchronises). */
int cme213_shard_create_f64(cme213_ctx *ctx, int rank, int world, int64_t num_rows_total, const int64_t *bounds,
                            const int32_t *d_row_ptr_local, const int32_t *d_col, const double *d_val, int mode,
                            cme213_shard **out);
int cme213_shard_destroy(cme213_shard *shard);
/* One process per GPU: every rank exports a handle of cme213_shard_handle_bytes() bytes, gathers everybody's
 * handles in rank order and connects.  One process: pass all shards in rank order (peer access is enabled). */
int cme213_shard_handle_bytes(void);
int cme213_shard_export(cme213_shard *shard, void *handle_out);
int cme213_shard_connect(cme213_shard *shard, const void *handles_of_all_ranks);
int cme213_shard_connect_local(cme213_shard *const *shards_in_rank_order, int world);
int cme213_shard_mode(const cme213_shard *shard);                       /* the mode in use after connecting */
int cme213_shard_rows(const cme213_shard *shard, int64_t *row_begin, int64_t *row_end);
int64_t cme213_shard_comm_bytes_per_step(const cme213_shard *shard);    /* bytes this rank sends per pass */
/* Full-length vector 0 or 1 (device pointer).  Before the first step vector `src` must hold x on every rank --
 * at least the rank's own rows and the ghost rows it reads.  Own rows of the result are [row_begin, row_end). */
double *cme213_shard_vector(cme213_shard *shard, int which);
/* vector[1-src] = A vector[src] for this rank's rows, exchange included; only enqueues work.  Every rank must
 * call it the same number of times with the same src sequence.  The pass number lives in device memory, so a
 * HALO / GATHER step can be captured into a CUDA graph and replayed. */
int cme213_shard_step(cme213_shard *shard, int src, void *stream);
/* nr_iterations (even) steps 0->1, 1->0, ...: the result ends in vector 0. */
int cme213_shard_iterate(cme213_shard *shard, int nr_iterations, void *stream);
/* Synchronises the stream and reports CME213_ERR_TIMEOUT if a wait for a neighbour gave up (~20 s).  Peers may
 * still be writing GHOST rows of this rank's vectors afterwards: put a barrier of your own between finish and
 * cme213_shard_destroy or a read of ghost rows. */
int cme213_shard_finish(cme213_shard *shard, void *stream);
/* Host-pointer twin for a C++ driver (one process, num_gpus devices 0..num_gpus-1): partitions by nonzeros, uploads
 * each row block, runs nr_iterations (even) passes and gathers the result.  kernel_ms_out = the passes alone
 * (max over the GPUs).  The sharded counterpart of cme213_csr_iterate_host_f64 / device_graph_iterate. */
int cme213_csr_iterate_sharded_host_f64(int num_gpus, int64_t num_rows, int64_t nnz, const int32_t *h_row_ptr,
                                        const int32_t *h_col, const double *h_val, const double *h_x_input,
                                        double *h_x_result, int nr_iterations, int mode, float *kernel_ms_out);

/* PageRank propagate: out[i] = 0.5f/N + 0.5f * sum_j in[e[j]] * inv[e[j]].
 * reference: hw/hw1/programming/pagerank.cu:45-56 (CPU), :70-83 (GPU). */
int cme213_pagerank_propagate(cme213_ctx *ctx, const uint32_t *d_graph_indices, const uint32_t *d_graph_edges,
                              const float *d_nodes_in, float *d_nodes_out, const float *d_inv_edges_per_node,
                              int array_length, int kernel, void *stream);

/* ---- ELL / COO / HYB --------------------------------------------------- */
/* ELL, column-major: entry n of row r lives at [n*pitch + r]; padded entries have value 0.
 * y = A x.  reference: refs/Bell Tech Report 2008.pdf Fig. 17 (spmv_ell_kernel);
 * refs/Bell SC 2009.pdf p.3 Fig. 2/4, section 3.2; slides/Lecture22.pdf slides 26-28. */
int cme213_spmv_ell_f32(cme213_ctx *ctx, int64_t num_rows, int64_t pitch, int num_cols_per_row,
                        const int32_t *d_ell_idx, const float *d_ell_val, const float *d_x, float *d_y, void *stream);
int cme213_spmv_ell_f64(cme213_ctx *ctx, int64_t num_rows, int64_t pitch, int num_cols_per_row,
                        const int32_t *d_ell_idx, const double *d_ell_val, const double *d_x, double *d_y, void *stream);
/* COO sorted by row, y[row] += sum of the row's entries (segmented reduction, one writer per row,
 * no atomics).  reference: refs/Bell Tech Report 2008.pdf Fig. 24 (segmented_reduction);
 * slides/Lecture22.pdf slides 29-30. */
int cme213_spmv_coo_accumulate_f32(cme213_ctx *ctx, int64_t nnz, const int32_t *d_row, const int32_t *d_col,
                                   const float *d_val, const float *d_x, float *d_y, void *stream);
int cme213_spmv_coo_accumulate_f64(cme213_ctx *ctx, int64_t nnz, const int32_t *d_row, const int32_t *d_col,
                                   const double *d_val, const double *d_x, double *d_y, void *stream);
/* HYB = ELL(K) followed by the COO tail.  reference: refs/Bell SC 2009.pdf p.5 section 3.5. */
int cme213_spmv_hyb_f32(cme213_ctx *ctx, int64_t num_rows, int64_t pitch, int K, const int32_t *d_ell_idx,
                        const float *d_ell_val, int64_t coo_nnz, const int32_t *d_coo_row,
                        const int32_t *d_coo_col, const float *d_coo_val, const float *d_x, float *d_y, void *stream);
int cme213_spmv_hyb_f64(cme213_ctx *ctx, int64_t num_rows, int64_t pitch, int K, const int32_t *d_ell_idx,
                        const double *d_ell_val, int64_t coo_nnz, const int32_t *d_coo_row,
                        const int32_t *d_coo_col, const double *d_coo_val, const double *d_x, double *d_y, void *stream);

/* HYB with the tail as a CSR matrix of its own instead of COO triplets: d_tail_ptr from cme213_csr_tail_row_ptr, its
 * column / value arrays are the d_coo_col / d_coo_val of cme213_csr_to_coo_tail_* (same order).  The tail runs through
 * the CSR stream kernel in accumulate mode (8 bytes per entry instead of COO's 12, coalesced update of y for the rows
 * that have a tail).  A row is summed ELL columns first, then its tail left to right: bit-identical to the CPU loop.
 * d_tail_ptr == NULL: no tail.  reference: refs/Bell SC 2009.pdf section 3.5 (the split); the tail kernel is the CSR
 * shape of slides/Lecture22.pdf slide 20. */
int cme213_spmv_hyb_csrtail_f32(cme213_ctx *ctx, int64_t num_rows, int64_t pitch, int K, const int32_t *d_ell_idx,
                                const float *d_ell_val, const int32_t *d_tail_ptr, const int32_t *d_tail_col,
                                const float *d_tail_val, const float *d_x, float *d_y, void *stream);
int cme213_spmv_hyb_csrtail_f64(cme213_ctx *ctx, int64_t num_rows, int64_t pitch, int K, const int32_t *d_ell_idx,
                                const double *d_ell_val, const int32_t *d_tail_ptr, const int32_t *d_tail_col,
                                const double *d_tail_val, const double *d_x, double *d_y, void *stream);

/* ---- format conversion on the device ----------------------------------- */
/* Row pointers (num_rows + 1) of the tail left after an ELL(K) split: exclusive scan of max(len - K, 0). */
int cme213_csr_tail_row_ptr(cme213_ctx *ctx, int64_t num_rows, const int32_t *d_row_ptr, int K, int32_t *d_tail_ptr,
                            int64_t *tail_nnz_out, void *stream);
/* HYB split width: largest K such that at least max(4096, num_rows/3) rows have >= K entries.
 * Synchronises `stream`.  reference: refs/Bell SC 2009.pdf p.5 section 3.5. */
int cme213_hyb_pick_k(cme213_ctx *ctx, int64_t num_rows, const int32_t *d_row_ptr, int *K_out, int *max_len_out,
                      void *stream);
/* CSR -> ELL(K), column-major with the given pitch (>= num_rows, multiple of 4 recommended).
 * Rows longer than K are truncated.  Built as a padded row-major fill followed by the tiled
 * transpose above, in row chunks so the staging buffer stays small. */
int64_t cme213_ell_pitch(int64_t num_rows);
int cme213_csr_to_ell_f32(cme213_ctx *ctx, int64_t num_rows, int64_t num_cols, const int32_t *d_row_ptr,
                          const int32_t *d_col, const float *d_val, int K, int64_t pitch, int32_t *d_ell_idx,
                          float *d_ell_val, void *stream);
int cme213_csr_to_ell_f64(cme213_ctx *ctx, int64_t num_rows, int64_t num_cols, const int32_t *d_row_ptr,
                          const int32_t *d_col, const double *d_val, int K, int64_t pitch, int32_t *d_ell_idx,
                          double *d_ell_val, void *stream);
/* Entries at in-row position >= K, in row order.  Call with d_coo_* == NULL to get the count only
 * (synchronises); the offsets are an exclusive scan of max(len-K,0). */
int cme213_csr_to_coo_tail_f32(cme213_ctx *ctx, int64_t num_rows, const int32_t *d_row_ptr, const int32_t *d_col,
                               const float *d_val, int K, int64_t *tail_nnz_out, int32_t *d_coo_row,
                               int32_t *d_coo_col, float *d_coo_val, void *stream);
int cme213_csr_to_coo_tail_f64(cme213_ctx *ctx, int64_t num_rows, const int32_t *d_row_ptr, const int32_t *d_col,
                               const double *d_val, int K, int64_t *tail_nnz_out, int32_t *d_coo_row,
                               int32_t *d_coo_col, double *d_coo_val, void *stream);

/* Key-value form of the sort: the 32-bit payload d_vals travels with d_keys; only the key bits [0, end_bit) are
 * sorted on (8-bit digits), stable; the result ends in d_keys / d_vals.  No counterpart in the reference (its sort is
 * keys only); it exists for cme213_coo_to_csr_*. */
int cme213_radix_sort_pairs_u32(cme213_ctx *ctx, uint32_t *d_keys, uint32_t *d_keys_tmp, uint32_t *d_vals,
                                uint32_t *d_vals_tmp, int64_t n, int end_bit, void *stream);
/* Unsorted COO triplets -> CSR with ascending columns (duplicates kept, in input order): two stable key-value sorts of a
 * permutation (by column, then by row), one gather, row_ptr from the run boundaries.  The input side of the path:
 * MatrixMarket files reach hw/hw_final/programming/aux/readMM.py:18-24 as COO triplets (a.row, a.col, a.data).
 * d_row_ptr needs num_rows + 1 entries, d_col / d_val nnz.  Entries outside the matrix -> CME213_ERR_INVALID.
 * Synchronises; allocates its temporaries (set-up code). */
int cme213_coo_to_csr_f32(cme213_ctx *ctx, int64_t num_rows, int64_t num_cols, int64_t nnz, const int32_t *d_coo_row,
                          const int32_t *d_coo_col, const float *d_coo_val, int32_t *d_row_ptr, int32_t *d_col,
                          float *d_val, void *stream);
int cme213_coo_to_csr_f64(cme213_ctx *ctx, int64_t num_rows, int64_t num_cols, int64_t nnz, const int32_t *d_coo_row,
                          const int32_t *d_coo_col, const double *d_coo_val, int32_t *d_row_ptr, int32_t *d_col,
                          double *d_val, void *stream);

/* ---- DIA (the format bound for the stencil operators of C1 / C5) --------------------------- */
/* y = A x, A in DIA: d_data[n * pitch + r] is the entry of row r on diagonal d_offsets[n] (column r + offset),
 * zero where there is none; offsets ascending.  Zero entries are skipped and the sum runs in diagonal order
 * with rounded multiply and add, so for a CSR matrix with ascending columns the result is bit-identical to
 * cme213_spmv_csr_* (CME213_CSR_STREAM / SCALAR).  pitch >= num_rows; a pitch that is a multiple of 4 with
 * 16-byte aligned d_data takes the 128-bit path (cme213_ell_pitch() gives one).
 * reference: refs/Bell SC 2009.pdf section 3.1, Fig. 1/3; refs/Bell Tech Report 2008.pdf Fig. 15
 * (spmv_dia_kernel); slides/Lecture22.pdf slide 29; the stencil itself: hw/hw2/programming/2dHeat.cu:361-366. */
int cme213_spmv_dia_f32(cme213_ctx *ctx, int64_t num_rows, int64_t num_cols, int64_t pitch, int num_diags,
                        const int32_t *d_offsets, const float *d_data, const float *d_x, float *d_y, void *stream);
int cme213_spmv_dia_f64(cme213_ctx *ctx, int64_t num_rows, int64_t num_cols, int64_t pitch, int num_diags,
                        const int32_t *d_offsets, const double *d_data, const double *d_x, double *d_y, void *stream);
/* The occupied diagonals (col - row) of a CSR matrix, ascending, by mark + exclusive scan + compaction
 * (slides/Lecture16.pdf slides 3-13).  *num_diags_out = how many there are (synchronises); at most max_diags of
 * them are written to d_offsets_out.  Column indices outside [0, num_cols) -> CME213_ERR_INVALID (as in
 * cme213_coo_to_csr_*).  Scratch: 8 bytes per (num_rows + num_cols) in the context workspace. */
int cme213_csr_diagonals(cme213_ctx *ctx, int64_t num_rows, int64_t num_cols, const int32_t *d_row_ptr,
                         const int32_t *d_col, int max_diags, int32_t *d_offsets_out, int *num_diags_out, void *stream);
/* CSR -> DIA on the ascending offsets list d_offsets (num_diags * pitch elements are written, zero-filled).
 * *unmatched_out (may be NULL; non-NULL synchronises) = entries that lie on none of the listed diagonals.
 * Duplicate (row, col) entries are summed (DIA has one slot per position). */
int cme213_csr_to_dia_f32(cme213_ctx *ctx, int64_t num_rows, const int32_t *d_row_ptr, const int32_t *d_col,
                          const float *d_val, int num_diags, const int32_t *d_offsets, int64_t pitch, float *d_data,
                          int64_t *unmatched_out, void *stream);
int cme213_csr_to_dia_f64(cme213_ctx *ctx, int64_t num_rows, const int32_t *d_row_ptr, const int32_t *d_col,
                          const double *d_val, int num_diags, const int32_t *d_offsets, int64_t pitch, double *d_data,
                          int64_t *unmatched_out, void *stream);

/* Matrix-free form of the generators' operator (cme213_gen_laplace2d_f32: nz = 1, centre 4; cme213_gen_laplace3d_f64:
 * centre 6): y[i] = centre * x[i] - sum of the existing grid neighbours of point i, for rows [row_begin, row_end) of
 * the nx * ny * nz grid in natural order (x fastest).  Legs and summation order are those of the CSR rows, so y is
 * bit-identical to the CSR / DIA product; only x and y move (16 / 8 bytes per row): the upper bound for C1 / C5.
 * d_x and d_y are full-length vectors and must not alias.  Fastest when the range is whole grid lines (row_begin and
 * row_end multiples of nx), nx a multiple of 4 (fp32) / 2 (fp64) and the vectors 16-byte aligned; any range works.
 * reference: hw/hw2/programming/2dHeat.cu:361-366 (stencil2) and :431-461 (gpuGlobal). */
int cme213_stencil_laplace_f32(cme213_ctx *ctx, int nx, int ny, int nz, float centre, int64_t row_begin,
                               int64_t row_end, const float *d_x, float *d_y, void *stream);
int cme213_stencil_laplace_f64(cme213_ctx *ctx, int nx, int ny, int nz, double centre, int64_t row_begin,
                               int64_t row_end, const double *d_x, double *d_y, void *stream);

/* Scatter offsets of a blocked LSD radix pass: out[i][j] = global_excl[j] + sum of local[i'][j] over blocks i' < i
 * (arrays [num_blocks x num_buckets], row-major).  Transpose + one segmented exclusive scan + transpose.
 * reference: hw/hw4/programming/radixsort.cpp:85-97 (local_histogram_final). */
int cme213_block_offsets_u32(cme213_ctx *ctx, const uint32_t *d_local_hist, const uint32_t *d_global_excl,
                             uint32_t *d_out, int64_t num_blocks, int64_t num_buckets, void *stream);

/* ---- LSD radix sort on the scan ---------------------------- */
/* One stable counting pass on the digit (key >> start_bit) & ((1 << num_bits) - 1), num_bits 1..8:
 * per-8192-key-block histograms, ONE exclusive scan of the digit-major histogram table, stable scatter.
 * reference: hw/hw4/programming/radixsort.cpp:44-109 (radixSortParallelPass; its blockSize is 8192 too, :115).
 * d_keys_in and d_keys_out must not overlap; n <= 2^31-1. */
int cme213_radix_sort_pass_u32(cme213_ctx *ctx, const uint32_t *d_keys_in, uint32_t *d_keys_out, int64_t n,
                               int start_bit, int num_bits, void *stream);
/* Sorts d_keys ascending in place, using d_keys_tmp (n words) as the other half of the ping-pong.
 * reference: radixsort.cpp:111-121 (radixSortParallel(keys, keys_tmp, numBits)); num_bits 1..16 as there
 * (the assert at :113).  The device digit is min(num_bits, 8) bits wide -- the sorted result does not depend
 * on it -- and the result always ends in d_keys, as the reference's paired passes guarantee. */
int cme213_radix_sort_u32(cme213_ctx *ctx, uint32_t *d_keys, uint32_t *d_keys_tmp, int64_t n, int num_bits,
                          void *stream);

/* ---- input generators (device side, deterministic) ---------------------- */
/* The homework drivers generate their inputs in main(); the large configurations cannot be
 * generated on the host in reasonable time, so the same recipes run on the device.  Row pointers
 * are produced by cme213_exclusive_scan_u32 over the row lengths. */
/* 5-point Laplacian on an nx x ny grid, natural order, sorted columns, diag 4 / off-diag -1.
 * rows = nx*ny.  d_row_ptr[rows+1].  Pass d_col = d_val = NULL to get nnz only.
 * reference: the order-2 stencil of hw/hw2/programming/2dHeat.cu:361-366 as a matrix. */
int cme213_gen_laplace2d_f32(cme213_ctx *ctx, int nx, int ny, int32_t *d_row_ptr, int32_t *d_col, float *d_val,
                             int64_t *nnz_out, void *stream);
/* 7-point Laplacian on nx x ny x nz, rows [row_begin,row_end) only (a z-slab shard); column indices are
 * global; d_row_ptr[(row_end-row_begin)+1] is rebased to 0.  diag 6 / off-diag -1. */
int cme213_gen_laplace3d_f64(cme213_ctx *ctx, int nx, int ny, int nz, int64_t row_begin, int64_t row_end,
                             int32_t *d_row_ptr, int32_t *d_col, double *d_val, int64_t *nnz_out, void *stream);
/* Power-law rows (Pareto tail, shape 1, capped): len = min(cap, r0 + floor(num*(2^32-1-h)/(den*(h+1)))),
 * h = hash32(seed, row); columns stratified-random and sorted inside a window of `window` columns centred
 * on the row (window = 0: the whole matrix); values uniform in (-1,1), never 0.
 * reference: refs/Bell SC 2009.pdf p.6 Fig. 9 recipe (base + Pareto row lengths). */
int cme213_gen_powerlaw_lengths(cme213_ctx *ctx, int64_t num_rows, int r0, uint32_t num, uint32_t den, int cap,
                                uint32_t seed, int32_t *d_row_ptr, int64_t *nnz_out, void *stream);
int cme213_gen_random_csr_fill_f32(cme213_ctx *ctx, int64_t num_rows, int64_t num_cols, const int32_t *d_row_ptr,
                                   int64_t window, uint32_t seed, int32_t *d_col, float *d_val, void *stream);
/* PageRank graph of hw/hw1/programming/pagerank.cu:185-204 with a counter hash in place of rand():
 * node i has (i % (2*avg_edges-1))+1 edges to hash-random nodes, inv = 1/edges, x0 = 1/N. */
int cme213_gen_pagerank_graph(cme213_ctx *ctx, int num_elements, int avg_edges, uint32_t seed,
                              uint32_t *d_graph_indices, uint32_t *d_graph_edges, float *d_inv_edges_per_node,
                              float *d_nodes, int64_t *nnz_out, void *stream);
/* Fills: uniform float/double in (-1,1) (readMM.py:32 distribution), raw 32-bit hash, hash % modulus,
 * and random head flags with probability num/2^32 per element (bit 0 always set). */
int cme213_fill_uniform_f32(cme213_ctx *ctx, float *d_out, int64_t n, uint32_t seed, void *stream);
int cme213_fill_uniform_f64(cme213_ctx *ctx, double *d_out, int64_t n, uint32_t seed, void *stream);
/* Segment start offsets the way the final project's input script draws them
 * (hw/hw_final/programming/aux/readMM.py:44-49: sorted random cut points, s[0] = 0, s[p-1] = n): num_cuts hashed
 * positions in [1, n) -> cme213_radix_sort_u32 -> duplicates dropped by mark + scan + compaction -> framed by 0 and n.
 * d_s_out needs num_cuts + 2 entries; *p_out = entries written (distinct cuts + 2).  Synchronises; allocates its
 * temporaries (set-up code). */
int cme213_gen_segment_offsets(cme213_ctx *ctx, int64_t n, int64_t num_cuts, uint32_t seed, int32_t *d_s_out,
                               int64_t *p_out, void *stream);
int cme213_fill_hash_u32(cme213_ctx *ctx, uint32_t *d_out, int64_t n, uint32_t seed, uint32_t modulus, void *stream);
/* Measurement aid (bench.py --workload gather): d_out[t] = sum of 16 gathers of d_x at hashed positions in [0, q), for
 * t < num_gathers / 16.  Its rate is the ceiling of every format's x gathers on matrices with scattered columns: one
 * 32-byte L2 sector per gather.  No counterpart in the reference (its reports quote GB/s of the arrays only,
 * hw/hw1/programming/PA1_Dong-Bang_Tsai.pdf p.3). */
int cme213_probe_gather_f32(cme213_ctx *ctx, const float *d_x, int64_t q, float *d_out, int64_t num_gathers, uint32_t seed,
                            void *stream);
int cme213_fill_random_flags(cme213_ctx *ctx, uint32_t *d_flag_bits, int64_t n, uint32_t seed, uint32_t prob_num,
                             void *stream);

/* ---- host-pointer driver entry points (drop-in boundary) ---------------- */
/* Exactly the homework's signature: host pointers in, host pointer out, the function owns all device
 * memory, nr_iterations even.  Prints "gpu graph propagate took X ms" like the reference's timer.
 * reference: hw/hw1/programming/pagerank.cu:85-143 (device_graph_iterate). */
void device_graph_iterate(unsigned int *h_graph_indices, unsigned int *h_graph_edges, float *h_graph_nodes_input,
                          float *h_graph_nodes_result, float *h_inv_edges_per_node, int nr_iterations,
                          int num_elements, int avg_edges);
/* Same with a status code, explicit nnz, optional kernel-only time (ms) and no printing. */
int cme213_pagerank_iterate_host(const uint32_t *h_graph_indices, const uint32_t *h_graph_edges,
                                 const float *h_nodes_input, float *h_nodes_result, const float *h_inv_edges_per_node,
                                 int nr_iterations, int num_elements, float *kernel_ms_out);
/* Iterated y = A x on a general CSR matrix with ping-pong buffers (the loop of pagerank.cu:128-132);
 * result after nr_iterations (even) lands in h_x_result. */
int cme213_csr_iterate_host_f64(int64_t num_rows, int64_t nnz, const int32_t *h_row_ptr, const int32_t *h_col,
                                const double *h_val, const double *h_x_input, double *h_x_result,
                                int nr_iterations, float *kernel_ms_out);
int cme213_csr_iterate_host_f32(int64_t num_rows, int64_t nnz, const int32_t *h_row_ptr, const int32_t *h_col,
                                const float *h_val, const float *h_x_input, float *h_x_result, int nr_iterations,
                                float *kernel_ms_out);
/* The same iterate with the matrix kept on the device between calls: upload once (arrays + the dispatcher's plan, so
 * the timed loop never synchronises), then any number of cme213_csr_iterate_* calls that move only x in and out.
 * The one-call forms above are upload + iterate + destroy.  reference: the alloc / copy / loop / copy-back structure of
 * hw/hw1/programming/pagerank.cu:102-142, split at the point where the matrix stops changing. */
typedef struct cme213_csr_matrix cme213_csr_matrix;
int cme213_csr_upload_host_f64(int64_t num_rows, int64_t nnz, const int32_t *h_row_ptr, const int32_t *h_col,
                               const double *h_val, cme213_csr_matrix **out);
int cme213_csr_upload_host_f32(int64_t num_rows, int64_t nnz, const int32_t *h_row_ptr, const int32_t *h_col,
                               const float *h_val, cme213_csr_matrix **out);
int cme213_csr_iterate_f64(cme213_csr_matrix *matrix, const double *h_x_input, double *h_x_result, int nr_iterations,
                           float *kernel_ms_out);
int cme213_csr_iterate_f32(cme213_csr_matrix *matrix, const float *h_x_input, float *h_x_result, int nr_iterations,
                           float *kernel_ms_out);
int cme213_csr_matrix_destroy(cme213_csr_matrix *matrix);
/* The final project's timed loop: N passes of product + inclusive segmented scan with a <- b between
 * passes.  reference: hw/hw_final/programming/fp.cu:154-188 (float);
 * aux/reference_spMVscan-released.cu:213-217 (double, sparseMV_scan + swap). */
int cme213_fp_iterate_host_f32(int n, int p, int q, int iters, const float *h_a, const int32_t *h_s,
                               const int32_t *h_k, const float *h_x, float *h_b, float *kernel_ms_out);
int cme213_fp_iterate_host_f64(int n, int p, int q, int iters, const double *h_a, const int32_t *h_s,
                               const int32_t *h_k, const double *h_x, double *h_b, float *kernel_ms_out);
/* Host-pointer twins of the primitives. */
int cme213_exclusive_scan_u32_host(const uint32_t *h_in, uint32_t *h_out, int64_t n);
/* radixSortParallel on host memory (radixsort.cpp:111-121): h_keys sorted in place; kernel_ms_out (may be NULL)
 * gets the device time of the passes, the figure the reference's main() prints (radixsort.cpp:204-207). */
int cme213_radix_sort_u32_host(uint32_t *h_keys, int64_t n, int num_bits, float *kernel_ms_out);
/* radixSortParallelPass(keys, sorted, numBits, startBit, blockSize) on host memory (radixsort.cpp:44, the scan's only
 * public caller in the reference): one stable counting pass on numBits (1..16) bits from startBit; block_size is accepted
 * and ignored (the result does not depend on it). */
int cme213_radix_sort_pass_u32_host(const uint32_t *h_keys, uint32_t *h_sorted, int64_t n, unsigned num_bits,
                                    unsigned start_bit, unsigned block_size);
int cme213_transpose_host_b32(const void *h_in, void *h_out, int64_t rows, int64_t cols);
int cme213_transpose_host_b64(const void *h_in, void *h_out, int64_t rows, int64_t cols);

#ifdef __cplusplus
}
#endif
#endif /* CME213_B200_H */
