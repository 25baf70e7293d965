// radixsort -- drop-in for the homework's driver (hw/hw4/programming/radixsort.cpp:162-214): rand() keys,
// std::sort as the check, then the sort under test -- here cme213_radix_sort_u32_host from libcme213_b200.so in the
// place of radixSortParallel(keys, keys_pingpong, numBits) (:205) -- compared key by key (:209-211).
// Usage: radixsort [n_elements] [numBits]   (defaults 40000000 and 8, radixsort.cpp:165-168)
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "../../include/cme213_b200.h"

typedef unsigned int uint;

static double now() {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

int main(int argc, char **argv) {
    int n_elements = 40000000;
    int numBits = 8;
    if (argc == 2) { n_elements = atoi(argv[1]); }
    if (argc == 3) { numBits = atoi(argv[2]); n_elements = atoi(argv[1]); }
    printf("n_elements: %d\nnumBits: %d\n", n_elements, numBits);
    if (n_elements < 0 || numBits < 1 || numBits > 16) { printf("bad arguments\n"); return 1; }

    std::vector<uint> keys((size_t)n_elements);
    for (size_t i = 0; i < keys.size(); ++i) keys[i] = (uint)rand();

    std::vector<uint> keys_stl = keys;
    double t0 = now();
    std::sort(keys_stl.begin(), keys_stl.end());
    printf("stl: %g\n", now() - t0);

    float kernel_ms = 0.f;
    t0 = now();
    int rc = cme213_radix_sort_u32_host(keys.data(), (int64_t)keys.size(), numBits, &kernel_ms);
    double el = now() - t0;
    if (rc != 0) { printf("cme213_radix_sort_u32_host failed: %s\n", cme213_last_error()); return 1; }
    printf("parallel radix: %g\n", el);                    // the reference's line (:207): host pointers in, host pointers out
    printf("gpu radix passes took %f ms\n", kernel_ms);    // device time of the passes alone

    for (size_t i = 0; i < keys.size(); ++i)
        if (keys_stl[i] != keys[i]) { printf("MISMATCH at %zu: %u != %u\n", i, keys[i], keys_stl[i]); return 1; }
    printf("Worked! GPU radix sort and std::sort output match.\n");
    return 0;
}
