// internal.cuh -- cross-file internal entry points (not exported).
#pragma once
#include "common.cuh"

// scan.cu: exclusive uint32 scan with caller-provided scratch (so a caller that also keeps data in
// the context workspace is not invalidated by a workspace growth).
size_t cme213_internal_scan_u32_scratch_bytes(int64_t n);
int cme213_internal_exclusive_scan_u32(cme213_ctx *ctx, const uint32_t *d_in, uint32_t *d_out, int64_t n,
                                       void *scratch, void *stream);
size_t cme213_internal_segscan_u32_scratch_bytes(int64_t n);
int cme213_internal_segscan_exclusive_u32(cme213_ctx *ctx, const uint32_t *d_in, const uint32_t *d_flag_bits,
                                          uint32_t *d_out, int64_t n, void *scratch, void *stream);
// spmv_csr.cu
int cme213_internal_pagerank_propagate_w(cme213_ctx *ctx, const uint32_t *d_graph_indices,
                                         const uint32_t *d_graph_edges, const float *d_nodes_in, const float *d_w_in,
                                         float *d_nodes_out, float *d_w_out, const float *d_inv, int array_length,
                                         void *stream);

// ---- row-sharded iterate (sharded.cu drives it, the protocol runs inside spmv_csr.cu's stream kernel) --------------
// Every rank keeps a small block of 64-bit words in its own memory: how many of its boundary CTAs have finished (per
// side, counted up for ever: passes completed = count / boundary CTAs per pass), a status word, and one flag per neighbour
// that the NEIGHBOUR writes: "my first `flag` passes are complete, their halo rows are in your vector".
// One pass = one launch: interior row blocks first, boundary row blocks last.  The launch's first CTA raises this rank's flag
// at the neighbours to the number of passes it has completed (the previous kernel has ended, so everything it stored -- into
// the neighbours' vectors too -- has been performed).  A boundary CTA of pass p waits until the neighbour's flag says p-1:
// the neighbour's rows of the previous pass are here, and -- because only boundary rows read ghost entries, and the
// neighbour's pass p-1 is over -- it has also finished reading what this CTA is about to overwrite in its vector.  Then
// the CTA computes and stores its rows locally and into the neighbour's vector.  The wait sits at the end of the grid,
// behind the interior, and nothing on a boundary CTA's path is a fence.
struct ShardSync {
    double *peer_lo, *peer_hi;                 // neighbours' output vectors, shifted like y (null = no neighbour)
    int64_t push_lo_end, push_hi_begin;        // rows < push_lo_end go to peer_lo, rows >= push_hi_begin to peer_hi
    int64_t side_lo_end, side_hi_begin;        // boundary rows of each side (they wait and are counted): the rows pushed or the rows
                                               // that read ghost entries, whichever reaches further
    unsigned long long *cnt_lo, *cnt_hi;       // local: boundary CTAs finished so far, per side
    const unsigned long long *flag_lo, *flag_hi;   // local, written by the neighbours
    unsigned long long *peer_flag_lo, *peer_flag_hi;   // this rank's flag word in each neighbour's memory
    unsigned long long *status;                // local: set to 1 when a wait gave up (a neighbour died or never launched)
    unsigned nb_lo, nb_hi;                     // boundary CTAs per side
    long long timeout_cycles;
};
// x[dst] rows of this shard = A_shard x[src], halo rows pushed to the neighbours and flags raised as described above.
// d_y / the peers are shifted so that index r is the shard's local row r; one kernel launch.
int cme213_internal_spmv_csr_f64_shard(cme213_ctx *ctx, int64_t rows_local, const int32_t *d_row_ptr, const int32_t *d_col,
                                       const double *d_val, const double *d_x, double *d_y, const ShardSync &sync,
                                       cudaStream_t st);
// y[r] += sum_j val[j] x[col[j]] (the remote-column block of a split shard; different rounding order than the unsplit row)
int cme213_internal_spmv_csr_f64_acc(cme213_ctx *ctx, int64_t rows_local, const int32_t *d_row_ptr, const int32_t *d_col,
                                     const double *d_val, const double *d_x, double *d_y, cudaStream_t st);
int cme213_internal_shard_block_rows(void);
// y[r] += (tail row r) . x with the x gathers marked evict-last: the CSR tail of HYB (spmv_ell.cu)
int cme213_internal_spmv_csr_f64_acc_keep(cme213_ctx *ctx, int64_t rows_local, const int32_t *d_row_ptr, const int32_t *d_col,
                                          const double *d_val, const double *d_x, double *d_y, cudaStream_t st);
int cme213_internal_spmv_csr_f32_acc_keep(cme213_ctx *ctx, int64_t rows_local, const int32_t *d_row_ptr, const int32_t *d_col,
                                          const float *d_val, const float *d_x, float *d_y, cudaStream_t st);
// convert.cu: histogram of row lengths (4096 bins, the last collects everything longer) + the longest row; synchronises
constexpr int CME213_HIST_BINS = 4096;
int cme213_internal_row_length_hist(cme213_ctx *ctx, int64_t num_rows, const int32_t *d_row_ptr, unsigned long long *h_hist,
                                    int *max_len_out, cudaStream_t st);

// preload the kernels of the host-pointer loops (spmv_csr.cu, scan.cu) so that lazy module loading stays outside their timers
void cme213_internal_preload_spmv(void);
void cme213_internal_preload_fp(int elem_bytes);
