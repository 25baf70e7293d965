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
