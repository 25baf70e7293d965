// pagerank -- drop-in for the homework's driver (hw/hw1/programming/pagerank.cu:146-250): builds the
// synthetic link graph, calls device_graph_iterate (now provided by libcme213_b200.so, same signature), runs
// the CPU propagate as the check and compares every element with the homework's 10-ULP comparator.
// Usage: pagerank [num_elements] [avg_edges] [iterations] [seed]   (defaults 2^21, 8, 20, time-independent seed)
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <chrono>
#include "../../include/cme213_b200.h"

static const int maxUlps = 10;  // pagerank.cu:43

// AlmostEqual2sComplement, hw/hw1/programming/mp1-util.h:44-61
static bool almost_equal(float A, float B, int ulps) {
    int a, b;
    memcpy(&a, &A, 4); memcpy(&b, &B, 4);
    if (a < 0) a = (int)(0x80000000u - (unsigned)a);
    if (b < 0) b = (int)(0x80000000u - (unsigned)b);
    long long d = (long long)a - (long long)b;
    return (d < 0 ? -d : d) <= ulps;
}

// the CPU check, pagerank.cu:45-67
static void host_graph_propagate(const unsigned *ind, const unsigned *edges, const float *in, float *out, const float *inv, int n) {
    for (int i = 0; i < n; i++) {
        float sum = 0.f;
        for (unsigned j = ind[i]; j < ind[i + 1]; j++) sum += in[edges[j]] * inv[edges[j]];
        out[i] = 0.5f / (float)n + 0.5f * sum;
    }
}

int main(int argc, char **argv) {
    int num_elements = argc > 1 ? atoi(argv[1]) : 1 << 21;
    int avg_edges = argc > 2 ? atoi(argv[2]) : 8;
    int iterations = argc > 3 ? atoi(argv[3]) : 20;
    unsigned seed = argc > 4 ? (unsigned)atoi(argv[4]) : 12345u;
    if (num_elements <= 0 || avg_edges <= 0 || iterations < 0 || iterations % 2) { printf("bad arguments\n"); return 1; }
    size_t N = (size_t)num_elements;
    unsigned *ind = (unsigned *)malloc((N + 1) * sizeof(unsigned));
    unsigned *edges = (unsigned *)malloc(N * avg_edges * sizeof(unsigned));
    float *inv = (float *)malloc(N * 4), *input = (float *)malloc(N * 4), *result = (float *)malloc(N * 4);
    float *chkA = (float *)malloc(N * 4), *chkB = (float *)malloc(N * 4);
    if (!ind || !edges || !inv || !input || !result || !chkA || !chkB) { printf("couldn't allocate memory\n"); return 1; }
    srand(seed);  // the reference seeds with time(NULL), pagerank.cu:183
    ind[0] = 0;
    for (int i = 0; i < num_elements; i++) {
        int nr_edges = (i % (2 * avg_edges - 1) + 1);
        inv[i] = 1.f / (float)nr_edges;
        ind[i + 1] = ind[i] + nr_edges;
        if (ind[i + 1] >= N * avg_edges) { printf("more edges than we have space for\n"); return 1; }
        for (unsigned j = ind[i]; j < ind[i + 1]; j++) edges[j] = rand() % num_elements;
        input[i] = 1.f / (float)num_elements;
        chkA[i] = input[i];
        result[i] = std::numeric_limits<float>::infinity();
    }
    device_graph_iterate(ind, edges, input, result, inv, iterations, num_elements, avg_edges);
    auto t0 = std::chrono::steady_clock::now();
    for (int it = 0; it < iterations; it += 2) {
        host_graph_propagate(ind, edges, chkA, chkB, inv, num_elements);
        host_graph_propagate(ind, edges, chkB, chkA, inv, num_elements);
    }
    printf("%s took %.2f ms\n", "host graph propagate", std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count());
    int num_errors = 0;
    for (int i = 0; i < num_elements; i++)
        if (!almost_equal(result[i], chkA[i], maxUlps)) {
            if (++num_errors < 10) printf("%d:%.16f::", i, result[i] - chkA[i]);
        }
    if (num_errors > 0) printf("Output of CUDA version and normal version didn't match! \n");
    else printf("Worked! CUDA and reference output match. \n");
    free(ind); free(edges); free(inv); free(input); free(result); free(chkA); free(chkB);
    return num_errors > 0;
}
