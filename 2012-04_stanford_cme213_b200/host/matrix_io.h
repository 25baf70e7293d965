// matrix_io.h -- the final project's input container and loader, host side.
// Mirrors `struct matrix` and `load()` of hw/hw_final/programming/aux/mp1-util.h:5-9, 81-169 (same fields,
// same file format: "n p q N", then a[n], s[p], k[n]; x.txt holds x[q]; b.txt optional), but malformed
// input is reported through the return value instead of assert()/exit().
#pragma once
#include <string>
#include <vector>

struct matrix {
    int n, p, q, N;
    std::vector<int> s, k;
    std::vector<double> a, x, b;
};

// Returns nullptr and fills `err` on failure.  b_file may be nullptr or unreadable (b stays zero).
matrix *load(const char *a_file, const char *x_file, const char *b_file, std::string *err);

// b = inclusive segmented scan of a[i]*x[k[i]] on the GPU (double), same contract as
// reference_spMVscan-released.cu:38-54; the caller swaps a and b between iterations (:213-217).
// Returns 0 on success; on failure `err` (if given) holds the library's message.
int sparseMV_scan(matrix *m, std::string *err = nullptr);
