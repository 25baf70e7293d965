// check_output -- the instructors' checker (hw/hw_final/programming/aux/CheckOutput/serialMV.cu and
// aux/reference_spMVscan-released.cu:162-237) with the double-precision recomputation done on the GPU:
// `check_output a.txt x.txt b.txt` recomputes b from a.txt/x.txt (N passes of sparseMV_scan with the swap)
// and prints the absolute / relative L2 and Linf distances to the b.txt under test.
#include <cmath>
#include <cstdio>
#include <string>
#include "matrix_io.h"

int main(int argc, char **argv) {
    if (argc != 4) { fprintf(stderr, "./check_output a.txt x.txt b.txt\n"); return 0; }
    std::string err;
    matrix *m = load(argv[1], argv[2], argv[3], &err);
    if (!m) { fprintf(stderr, "%s\n", err.c_str()); return 1; }
    std::vector<double> b_file = m->b;
    for (int i = 0; i < m->N; i++) {
        if (sparseMV_scan(m, &err)) { fprintf(stderr, "%s\n", err.c_str()); return 1; }
        if (i != m->N - 1) m->b.swap(m->a);
    }
    double d2 = 0, a2 = 0, dinf = 0, ainf = 0;
    for (int i = 0; i < m->n; i++) {
        double d = m->b[i] - b_file[i];
        d2 += d * d; a2 += m->b[i] * m->b[i];
        dinf = std::fabs(d) > dinf ? std::fabs(d) : dinf;
        ainf = std::fabs(m->b[i]) > ainf ? std::fabs(m->b[i]) : ainf;
    }
    printf("File:\t absolute l2/linf - relative l2/linf \t%8.1e %8.1e %8.1e %8.1e\n", std::sqrt(d2), dinf,
           a2 > 0 ? std::sqrt(d2) / std::sqrt(a2) : 0.0, ainf > 0 ? dinf / ainf : 0.0);
    delete m;
    return 0;
}
