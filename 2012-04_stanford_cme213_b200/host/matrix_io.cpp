#include "matrix_io.h"
#include <fstream>
#include <sstream>
#include "../../include/cme213_b200.h"

static bool fail(std::string *err, const std::string &msg) {
    if (err) *err = msg;
    return false;
}

static bool parse(matrix *m, const char *a_file, const char *x_file, const char *b_file, std::string *err) {
    std::ifstream afs(a_file), xfs(x_file);
    if (!afs.good()) return fail(err, std::string("We could not open file ") + a_file);
    if (!xfs.good()) return fail(err, std::string("We could not open file ") + x_file);
    if (!(afs >> m->n >> m->p >> m->q >> m->N)) return fail(err, "bad header in a file");
    if (m->n <= 0 || m->p <= 0 || m->q <= 0 || m->N <= 0) return fail(err, "n, p, q, N must be positive");
    m->a.resize(m->n); m->b.assign(m->n, 0.0); m->k.resize(m->n); m->s.resize(m->p); m->x.resize(m->q);
    for (int i = 0; i < m->n; ++i) if (!(afs >> m->a[i])) return fail(err, "a file ends inside a");
    for (int i = 0; i < m->p; ++i) {
        if (!(afs >> m->s[i])) return fail(err, "a file ends inside s");
        if (i > 0 && m->s[i] <= m->s[i - 1]) {
            std::ostringstream o;
            o << "Error in input file " << a_file << ": s must increase (index " << i << ")";
            return fail(err, o.str());
        }
    }
    if (m->s[0] != 0 || m->s[m->p - 1] != m->n) return fail(err, "s[0] must be 0 and s[p-1] must be n");
    for (int i = 0; i < m->n; ++i) {
        if (!(afs >> m->k[i])) return fail(err, "a file ends inside k");
        if (m->k[i] < 0 || m->k[i] >= m->q) return fail(err, "k out of range");
    }
    for (int i = 0; i < m->q; ++i) if (!(xfs >> m->x[i])) return fail(err, "x file too short");
    if (b_file) {
        std::ifstream bfs(b_file);
        if (bfs.good()) for (int i = 0; i < m->n; ++i) if (!(bfs >> m->b[i])) break;
    }
    return true;
}

matrix *load(const char *a_file, const char *x_file, const char *b_file, std::string *err) {
    matrix *m = new matrix;
    if (!parse(m, a_file, x_file, b_file, err)) { delete m; return nullptr; }
    return m;
}

int sparseMV_scan(matrix *m, std::string *err) {
    int rc = cme213_fp_iterate_host_f64(m->n, m->p, m->q, 1, m->a.data(), m->s.data(), m->k.data(), m->x.data(),
                                        m->b.data(), nullptr);
    if (rc && err) *err = cme213_last_error();
    return rc;
}
