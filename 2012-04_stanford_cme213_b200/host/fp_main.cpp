// fp -- drop-in for the final project's program: `fp a.txt x.txt [anything]`.
// Same contract as hw/hw_final/programming/fp.cu:74-215: reads a.txt / x.txt, runs N iterations of
// product + inclusive segmented scan in float on the GPU, prints the fixed timing line
// (hw/hw_final/Final.pdf p.3), writes b.txt (one line, space separated); a 4th argument also runs the CPU
// loops of fp.cu:130-152 with OpenMP, prints their time, counts elements beyond the reference's tolerance
// (10e-3, fp.cu:193) and writes b_cpu.txt.  All device work goes through libcme213_b200.so.
#include <cmath>
#include <cstdio>
#include <fstream>
#include <iostream>
#include <string>
#include <vector>
#include <chrono>
#include "../../include/cme213_b200.h"
#include "matrix_io.h"

int main(int argc, char **argv) {
    if (argc < 3) {
        printf("Run command: ./fp \"file a.txt\" \"file x.txt\"\n");
        return 0;
    }
    const bool cpu_check = argc == 4;
    std::string err;
    matrix *m = load(argv[1], argv[2], nullptr, &err);
    if (!m) { std::cerr << err << std::endl; return 1; }
    const int n = m->n, p = m->p, q = m->q, iters = m->N;
    std::vector<float> a(m->a.begin(), m->a.end()), x(m->x.begin(), m->x.end()), b(n);
    std::cout << "\nDim of a: " << n << "\nDim of x: " << q << "\nDim of s: " << p << "\n# of iters: " << iters << "\n\n";

    std::vector<float> cpu;
    double cpu_ms = 0;
    if (cpu_check) {
        std::vector<float> xx(n);
        for (int i = 0; i < n; ++i) xx[i] = x[m->k[i]];
        cpu = a;
        auto t0 = std::chrono::steady_clock::now();
        for (int iter = 0; iter < iters; ++iter) {
#pragma omp parallel for
            for (int i = 0; i < n; ++i) cpu[i] *= xx[i];
#pragma omp parallel for
            for (int i = 1; i < p; ++i)
                for (int j = m->s[i - 1] + 1; j < m->s[i]; ++j) cpu[j] += cpu[j - 1];
        }
        cpu_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    }

    float ms = 0.f;
    int rc = cme213_fp_iterate_host_f32(n, p, q, iters, a.data(), m->s.data(), m->k.data(), x.data(), b.data(), &ms);
    if (rc) { std::cerr << "GPU computation failed: " << cme213_last_error() << std::endl; return 1; }
    std::cout << "The running time of my code for " << iters << " iterations is: " << ms << " milliseconds.\n\n";

    if (cpu_check) {
        const double tol = 10e-3;
        long bad = 0;
        std::cout << "The CPU running time of my code for " << iters << " iterations is: " << cpu_ms << " milliseconds.\n\n";
        std::cout << "Checking the correctness by the result from CPU\n\n";
        std::ofstream ofs_cpu("b_cpu.txt");
        for (int i = 0; i < n; ++i) {
            if (std::abs(cpu[i] - b[i]) > tol) bad++;
            ofs_cpu << cpu[i] << " ";
        }
        std::cout << "Elements beyond tolerance " << tol << ": " << bad << "\n";
    }
    std::ofstream ofs_gpu("b.txt");
    for (int i = 0; i < n; ++i) ofs_gpu << b[i] << " ";
    ofs_gpu.close();
    delete m;
    return 0;
}
