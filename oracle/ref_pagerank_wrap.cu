// TEST INFRASTRUCTURE.  Compiles the reference's own PageRank translation unit
// (hw/hw1/programming/pagerank.cu) where it lies under $(REF) and exports its
// CPU functions with C linkage so the restatement in oracle.c can be checked
// against the real thing.  No reference source is copied: the file is
// #included from the read-only checkout at build time; the output goes to
// oracle/_ref/ (git-ignored).
#define main cme213_ref_pagerank_main
#include "hw/hw1/programming/pagerank.cu"
#undef main

extern "C" {
// hw/hw1/programming/pagerank.cu:45-56
void ref_host_graph_propagate(unsigned int *graph_indices, unsigned int *graph_edges, float *in, float *out,
                              float *inv_edges_per_node, int array_length) {
    host_graph_propagate(graph_indices, graph_edges, in, out, inv_edges_per_node, array_length);
}
// hw/hw1/programming/pagerank.cu:59-67
void ref_host_graph_iterate(unsigned int *graph_indices, unsigned int *graph_edges, float *A, float *B,
                            float *inv_edges_per_node, int nr_iterations, int array_length) {
    host_graph_iterate(graph_indices, graph_edges, A, B, inv_edges_per_node, nr_iterations, array_length);
}
// hw/hw1/programming/mp1-util.h:44-61
int ref_almost_equal_2s_complement(float A, float B, int maxUlps) {
    return AlmostEqual2sComplement(A, B, maxUlps) ? 1 : 0;
}
// hw/hw1/programming/pagerank.cu:86-143 -- the reference's GPU path, usable only on a GPU box.
void ref_device_graph_iterate(unsigned int *ind, unsigned int *edges, float *in, float *out, float *inv,
                              int nr_iterations, int num_elements, int avg_edges) {
    device_graph_iterate(ind, edges, in, out, inv, nr_iterations, num_elements, avg_edges);
}
}
