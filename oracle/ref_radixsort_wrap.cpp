// TEST INFRASTRUCTURE.  Compiles the reference's radix sort (hw/hw4/programming/radixsort.cpp) where it lies
// under $(REF) and exports its passes with C linkage.  Nothing is copied; output goes to oracle/_ref/.
// radixSortParallel / radixSortSerial themselves are declared int and fall off their end without a return
// (radixsort.cpp:111-121, :155-161), which is undefined behaviour once optimised, so the drivers below repeat
// their three-line loops around the reference's own (void) pass functions instead of calling them.
#define main cme213_ref_radixsort_main
#include "hw/hw4/programming/radixsort.cpp"
#undef main
#include <cstring>

extern "C" {
// radixsort.cpp:44-109
void ref_radix_pass_parallel(const uint *in, uint *out, long n, uint num_bits, uint start_bit, uint block_size) {
    std::vector<uint> k(in, in + n), s(n);
    radixSortParallelPass(k, s, num_bits, start_bit, block_size);
    std::memcpy(out, s.data(), sizeof(uint) * n);
}
// radixsort.cpp:123-153
void ref_radix_pass_serial(const uint *in, uint *out, long n, uint num_bits, uint start_bit) {
    std::vector<uint> k(in, in + n), s(n);
    radixSortSerialPass(k, s, start_bit, num_bits);
    std::memcpy(out, s.data(), sizeof(uint) * n);
}
// the loop of radixsort.cpp:116-120 (blockSize 8192, :115); returns the seconds spent in the passes
double ref_radix_sort_parallel(uint *keys, long n, uint num_bits) {
    std::vector<uint> k(keys, keys + n), t(n);
    const double t0 = omp_get_wtime();
    for (uint start = 0; start < 32; start += 2 * num_bits) {
        radixSortParallelPass(k, t, num_bits, start, 8192);
        radixSortParallelPass(t, k, num_bits, start + num_bits, 8192);
    }
    const double t1 = omp_get_wtime();
    std::memcpy(keys, k.data(), sizeof(uint) * n);
    return t1 - t0;
}
}
