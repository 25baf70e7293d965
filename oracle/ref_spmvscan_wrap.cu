// TEST INFRASTRUCTURE.  Compiles the instructors' checker
// (hw/hw_final/programming/aux/reference_spMVscan-released.cu, with its own
// aux/mp1-util.h) where it lies under $(REF) and exports sparseMV_scan, the
// iterate-and-swap loop of its main(), its error norms and its file loader
// with C linkage.  Nothing is copied; output goes to oracle/_ref/.
#define main cme213_ref_spmvscan_main
#include "hw/hw_final/programming/aux/reference_spMVscan-released.cu"
#undef main
#include <cstring>

extern "C" {
// reference_spMVscan-released.cu:38-54 driven as in :213-217 (swap between passes).
void ref_sparse_mv_scan_iterate(int n, int p, int q, int N, const double *a, const int *s, const int *k,
                                const double *x, double *b_out) {
    matrix m;
    m.n = n; m.p = p; m.q = q; m.N = N;
    m.a.assign(a, a + n);
    m.b.assign(n, 0.0);
    m.s.assign(s, s + p);
    m.k.assign(k, k + n);
    m.x.assign(x, x + q);
    for (int i = 0; i < N; i++) {
        sparseMV_scan(&m);
        if (i != N - 1) m.b.swap(m.a);
    }
    std::memcpy(b_out, &m.b[0], sizeof(double) * n);
}
// reference_spMVscan-released.cu:65-144
void ref_error_norms(const double *a, const double *b, int n, double *out4) {
    std::vector<double> va(a, a + n), vb(b, b + n);
    out4[0] = L2Distance(va, vb);
    out4[1] = LInfDistance(va, vb);
    out4[2] = relativeL2Error(va, vb);
    out4[3] = relativeLInfError(va, vb);
}
// aux/mp1-util.h:81-169; returns an opaque matrix*, sizes through out params.
void *ref_load(const char *a_file, const char *x_file, const char *b_file, int *n, int *p, int *q, int *N) {
    matrix *m = load(a_file, x_file, b_file);
    *n = m->n; *p = m->p; *q = m->q; *N = m->N;
    return m;
}
void ref_matrix_copy_out(void *vm, double *a, int *s, int *k, double *x, double *b) {
    matrix *m = (matrix *)vm;
    std::memcpy(a, &m->a[0], sizeof(double) * m->n);
    std::memcpy(s, &m->s[0], sizeof(int) * m->p);
    std::memcpy(k, &m->k[0], sizeof(int) * m->n);
    std::memcpy(x, &m->x[0], sizeof(double) * m->q);
    std::memcpy(b, &m->b[0], sizeof(double) * m->n);
}
void ref_matrix_free(void *vm) { delete (matrix *)vm; }
}
