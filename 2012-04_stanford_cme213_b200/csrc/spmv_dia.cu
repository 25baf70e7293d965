// spmv_dia.cu -- DIA (diagonal) sparse matrix-vector multiply and the CSR -> DIA conversion.
//
// Format and kernel semantics: refs/Bell SC 2009.pdf p.3 Fig. 1/3 and section 3.1; Bell TR 2008 Fig. 15
// (spmv_dia_kernel: data[pitch*n + row], col = row + offsets[n], in-range test); slides/Lecture22.pdf slide 29.
// The operator of configs C1 / C5 is a stencil (hw/hw2/programming/2dHeat.cu:361-366 is its 2-D order-2 form), so
// DIA is the format bound for them: 8 (fp64) or 4 (fp32) bytes per stored entry against 12 / 8 for CSR
// (SC09 Table 1).
//
// B200 shape: as in the ELL kernel a thread owns FOUR consecutive rows, so its loads from the diagonal-major
// array are 128-bit and a warp covers 512 contiguous bytes (fp32) per diagonal; x is read with 128-bit loads
// when the diagonal's offset keeps the four elements 16-byte aligned (every diagonal of a grid whose line length
// is a multiple of 4 except the +-1 ones), scalar loads otherwise.  Zero entries (padding outside the matrix and
// missing stencil legs at the grid boundary) are skipped, and the sum runs in diagonal order with rounded multiply
// and add: for a CSR matrix with ascending columns that is the CSR row loop's order, hence bit-identical results.
#include "internal.cuh"

namespace {
using namespace cme213;

__device__ __forceinline__ float mul_rn(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ double mul_rn(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ float add_rn(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ double add_rn(double a, double b) { return __dadd_rn(a, b); }

__device__ __forceinline__ void load4_stream(const float *p, float (&v)[4]) {
    const float4 t = ld_stream(reinterpret_cast<const float4 *>(p));
    v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
}
__device__ __forceinline__ void load4_stream(const double *p, double (&v)[4]) {
    const double2 t0 = ld_stream(reinterpret_cast<const double2 *>(p));
    const double2 t1 = ld_stream(reinterpret_cast<const double2 *>(p) + 1);
    v[0] = t0.x; v[1] = t0.y; v[2] = t1.x; v[3] = t1.y;
}
__device__ __forceinline__ void load4_cached(const float *p, float (&v)[4]) {
    const float4 t = __ldg(reinterpret_cast<const float4 *>(p));
    v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
}
__device__ __forceinline__ void load4_cached(const double *p, double (&v)[4]) {
    const double2 t0 = __ldg(reinterpret_cast<const double2 *>(p));
    const double2 t1 = __ldg(reinterpret_cast<const double2 *>(p) + 1);
    v[0] = t0.x; v[1] = t0.y; v[2] = t1.x; v[3] = t1.y;
}

// One diagonal's contribution to the four rows r4 .. r4+3.
template <typename T>
__device__ __forceinline__ void dia_accumulate(const T (&v)[4], int64_t c0, int64_t num_cols, const T *__restrict__ x,
                                               bool x_vec_ok, T (&s)[4]) {
    // four elements of 4 bytes (or two pairs of 8 bytes) at a multiple of 4 elements from an aligned base
    constexpr int64_t AL = sizeof(T) == 4 ? 3 : 1;   // 16-byte alignment of x + c0 in elements
    if (x_vec_ok && (c0 & AL) == 0 && c0 >= 0 && c0 + 3 < num_cols) {
        T xv[4];
        load4_cached(x + c0, xv);
#pragma unroll
        for (int t = 0; t < 4; ++t)
            if (v[t] != T(0)) s[t] = add_rn(s[t], mul_rn(v[t], xv[t]));
    } else {
#pragma unroll
        for (int t = 0; t < 4; ++t) {
            const int64_t c = c0 + t;
            if (v[t] != T(0) && c >= 0 && c < num_cols) s[t] = add_rn(s[t], mul_rn(v[t], __ldg(x + c)));
        }
    }
}

template <typename T, int UNROLL>
__global__ void __launch_bounds__(256)
dia_v4_kernel(int64_t num_rows, int64_t num_cols, int64_t pitch, int num_diags, const int32_t *__restrict__ offsets,
              const T *__restrict__ data, const T *__restrict__ x, T *__restrict__ y, int x_vec_ok, int y_vec_ok) {
    const int64_t r4 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (r4 >= num_rows) return;
    T s[4] = {T(0), T(0), T(0), T(0)};
    const T *dp = data + r4;
    int n = 0;
    for (; n + UNROLL <= num_diags; n += UNROLL) {
        T v[UNROLL][4];
        int32_t off[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {
            load4_stream(dp + (int64_t)(n + u) * pitch, v[u]);
            off[u] = __ldg(offsets + n + u);
        }
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) dia_accumulate(v[u], r4 + off[u], num_cols, x, x_vec_ok != 0, s);
    }
    for (; n < num_diags; ++n) {
        T v[4];
        load4_stream(dp + (int64_t)n * pitch, v);
        dia_accumulate(v, r4 + __ldg(offsets + n), num_cols, x, x_vec_ok != 0, s);
    }
    if (r4 + 3 < num_rows && y_vec_ok) {
        if (sizeof(T) == 4) {
            st_stream(reinterpret_cast<float4 *>(y + r4), make_float4((float)s[0], (float)s[1], (float)s[2], (float)s[3]));
        } else {
            st_stream(reinterpret_cast<double2 *>(y + r4), make_double2((double)s[0], (double)s[1]));
            st_stream(reinterpret_cast<double2 *>(y + r4) + 1, make_double2((double)s[2], (double)s[3]));
        }
    } else {
#pragma unroll
        for (int t = 0; t < 4; ++t)
            if (r4 + t < num_rows) y[r4 + t] = s[t];
    }
}

// Any pitch / alignment: one row per thread (Bell TR Fig. 15 as written, plus the zero skip).
template <typename T>
__global__ void __launch_bounds__(256)
dia_scalar_kernel(int64_t num_rows, int64_t num_cols, int64_t pitch, int num_diags, const int32_t *__restrict__ offsets,
                  const T *__restrict__ data, const T *__restrict__ x, T *__restrict__ y) {
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= num_rows) return;
    T s = T(0);
    for (int n = 0; n < num_diags; ++n) {
        const int64_t c = r + offsets[n];
        const T v = data[(int64_t)n * pitch + r];
        if (v != T(0) && c >= 0 && c < num_cols) s = add_rn(s, mul_rn(v, __ldg(x + c)));
    }
    y[r] = s;
}

// ---- CSR -> DIA -----------------------------------------------------------------------------------------
// Position of `off` in the ascending list, or -1.
__device__ __forceinline__ int find_offset(const int32_t *__restrict__ offsets, int num_diags, int32_t off) {
    int lo = 0, hi = num_diags - 1;
    while (lo <= hi) {
        const int mid = (lo + hi) >> 1;
        const int32_t o = __ldg(offsets + mid);
        if (o == off) return mid;
        if (o < off) lo = mid + 1; else hi = mid - 1;
    }
    return -1;
}

template <typename T>
__global__ void __launch_bounds__(256)
csr_to_dia_kernel(int64_t num_rows, const int32_t *__restrict__ row_ptr, const int32_t *__restrict__ col,
                  const T *__restrict__ val, int num_diags, const int32_t *__restrict__ offsets, int64_t pitch,
                  T *__restrict__ data, unsigned long long *unmatched) {
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= num_rows) return;
    unsigned long long miss = 0;
    for (int32_t e = row_ptr[r]; e < row_ptr[r + 1]; ++e) {
        const int n = find_offset(offsets, num_diags, (int32_t)(col[e] - r));
        if (n < 0) ++miss; else atomicAdd(&data[(int64_t)n * pitch + r], val[e]);   // data starts at zero: duplicates of (row, col) are summed
    }
    if (miss) atomicAdd(unmatched, miss);
}

// Which diagonals hold an entry: mark[(col - row) + (num_rows - 1)] = 1.
__global__ void __launch_bounds__(256)
mark_diagonals_kernel(int64_t num_rows, int64_t num_cols, const int32_t *__restrict__ row_ptr,
                      const int32_t *__restrict__ col, uint32_t *__restrict__ mark, unsigned int *__restrict__ bad) {
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= num_rows) return;
    for (int32_t e = row_ptr[r]; e < row_ptr[r + 1]; ++e) {
        const int64_t c = col[e];
        if (c < 0 || c >= num_cols) { atomicAdd(bad, 1u); continue; }   // malformed input: counted, never written
        mark[c - r + (num_rows - 1)] = 1u;
    }
}
// Stream compaction on the scan (slides/Lecture16.pdf slides 3-13): marked slot i goes to out[excl[i]].
__global__ void __launch_bounds__(256)
compact_diagonals_kernel(int64_t slots, int64_t num_rows, const uint32_t *__restrict__ mark,
                         const uint32_t *__restrict__ excl, int32_t *__restrict__ out, int max_out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= slots) return;
    if (mark[i] && excl[i] < (uint32_t)max_out) out[excl[i]] = (int32_t)(i - (num_rows - 1));
}

// ---- matrix-free: the 7-point (nz > 1) / 5-point (nz == 1) Laplacian of the generators, no matrix arrays at all ----
// Same legs, coefficients (centre given by the caller: 6 / 4 in the generators; neighbours -1) and summation order as the CSR rows written by
// cme213_gen_laplace3d_f64 / cme213_gen_laplace2d_f32, so y is bit-identical to the CSR / DIA product.  The
// reference's stencil code is hw/hw2/programming/2dHeat.cu:361-366 (stencil2); 16 (fp64) / 8 (fp32) bytes per row.
template <typename T>
__device__ __forceinline__ T stencil_point(const T *__restrict__ x, int64_t i, int ix, int iy, int iz, int nx, int ny,
                                           int nz, int64_t plane, T centre) {
    T s = T(0);
    if (iz > 0) s = add_rn(s, mul_rn(T(-1), __ldg(x + i - plane)));
    if (iy > 0) s = add_rn(s, mul_rn(T(-1), __ldg(x + i - nx)));
    if (ix > 0) s = add_rn(s, mul_rn(T(-1), __ldg(x + i - 1)));
    s = add_rn(s, mul_rn(centre, __ldg(x + i)));
    if (ix < nx - 1) s = add_rn(s, mul_rn(T(-1), __ldg(x + i + 1)));
    if (iy < ny - 1) s = add_rn(s, mul_rn(T(-1), __ldg(x + i + nx)));
    if (iz < nz - 1) s = add_rn(s, mul_rn(T(-1), __ldg(x + i + plane)));
    return s;
}

// One 16-byte vector of consecutive points per thread (4 floats / 2 doubles), so every warp-wide load is one
// contiguous 512-byte request; the x legs of a thread's edge points come from the neighbouring lanes by shuffle.
// (The first version gave a thread four doubles = two 16-byte loads 32 bytes apart, fetched the x legs with scalar
// loads and divided 64-bit indices: 32 sectors per request, L1 wavefront pipe 98 % busy, 0.48 ms --
// profiles/r01_ncu_stencil_v1_l1bound.txt.  This one: 0.464 ms.)
// Needs nx % NP == 0, 16-byte aligned x and y, row range on multiples of NP.
template <typename T> struct Vec16;
template <> struct Vec16<float> {
    static constexpr int NP = 4;
    static __device__ __forceinline__ void load(const float *p, float (&v)[4]) {
        const float4 t = __ldg(reinterpret_cast<const float4 *>(p));
        v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
    }
    static __device__ __forceinline__ void store(float *p, const float (&v)[4]) {
        st_stream(reinterpret_cast<float4 *>(p), make_float4(v[0], v[1], v[2], v[3]));
    }
};
template <> struct Vec16<double> {
    static constexpr int NP = 2;
    static __device__ __forceinline__ void load(const double *p, double (&v)[2]) {
        const double2 t = __ldg(reinterpret_cast<const double2 *>(p));
        v[0] = t.x; v[1] = t.y;
    }
    static __device__ __forceinline__ void store(double *p, const double (&v)[2]) {
        st_stream(reinterpret_cast<double2 *>(p), make_double2(v[0], v[1]));
    }
};

template <typename T, int GROUPS>
__global__ void __launch_bounds__(256)
stencil_vec_kernel(int nx, int ny, int nz, int64_t row_begin, int64_t row_end, T centre, const T *__restrict__ x,
                   T *__restrict__ y) {
    constexpr int NP = Vec16<T>::NP;
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t plane = (int64_t)nx * ny;
    // a warp owns GROUPS consecutive 512-byte segments, so a CTA of 256 threads covers 2 KiB * GROUPS of x: with
    // lines of that length or less the y legs of one line are the centre loads of its neighbour in the same CTA (L1 hits)
#pragma unroll
    for (int g = 0; g < GROUPS; ++g) {
        const int64_t i = row_begin + ((warp * GROUPS + g) * 32 + lane) * NP;
        const bool alive = i < row_end;      // no early exit: the shuffles below need the whole warp
        T c[NP], s[NP], v[NP];
#pragma unroll
        for (int t = 0; t < NP; ++t) { c[t] = T(0); s[t] = T(0); }
        if (alive) Vec16<T>::load(x + i, c);
        T left = __shfl_up_sync(0xFFFFFFFFu, c[NP - 1], 1), right = __shfl_down_sync(0xFFFFFFFFu, c[0], 1);
        if (!alive) continue;
        // 32-bit index arithmetic (the grid has fewer than 2^31 points): 64-bit divisions cost more than the loads
        const uint32_t iu = (uint32_t)i, line = iu / (uint32_t)nx;
        const int ix = (int)(iu - line * (uint32_t)nx), iz = (int)(line / (uint32_t)ny), iy = (int)(line - (uint32_t)iz * (uint32_t)ny);
        if (lane == 0 && ix > 0) left = __ldg(x + i - 1);
        if ((lane == 31 || i + NP >= row_end) && ix + NP < nx) right = __ldg(x + i + NP);
        if (iz > 0) { Vec16<T>::load(x + i - plane, v);
#pragma unroll
            for (int t = 0; t < NP; ++t) s[t] = add_rn(s[t], mul_rn(T(-1), v[t])); }
        if (iy > 0) { Vec16<T>::load(x + i - nx, v);
#pragma unroll
            for (int t = 0; t < NP; ++t) s[t] = add_rn(s[t], mul_rn(T(-1), v[t])); }
        if (ix > 0) s[0] = add_rn(s[0], mul_rn(T(-1), left));
#pragma unroll
        for (int t = 1; t < NP; ++t) s[t] = add_rn(s[t], mul_rn(T(-1), c[t - 1]));
#pragma unroll
        for (int t = 0; t < NP; ++t) s[t] = add_rn(s[t], mul_rn(centre, c[t]));
#pragma unroll
        for (int t = 0; t < NP - 1; ++t) s[t] = add_rn(s[t], mul_rn(T(-1), c[t + 1]));
        if (ix + NP < nx) s[NP - 1] = add_rn(s[NP - 1], mul_rn(T(-1), right));
        if (iy < ny - 1) { Vec16<T>::load(x + i + nx, v);
#pragma unroll
            for (int t = 0; t < NP; ++t) s[t] = add_rn(s[t], mul_rn(T(-1), v[t])); }
        if (iz < nz - 1) { Vec16<T>::load(x + i + plane, v);
#pragma unroll
            for (int t = 0; t < NP; ++t) s[t] = add_rn(s[t], mul_rn(T(-1), v[t])); }
        Vec16<T>::store(y + i, s);
    }
}

// Whole grid lines: a thread owns one 16-byte vector of G consecutive lines of one plane.  Against the kernel above:
//  * the y legs of the G lines are each other's centre loads: 3G + 2 vector loads for G lines instead of 5G (the capture of
//    the kernel above, profiles/r02_ncu_stencil_vec.txt: DRAM traffic = the algorithmic 2.1 GB, 5.2 GB read from L2, every
//    unit half busy, warps waiting on loads 15 of 22 cycles per instruction and the issue slots 61 % busy);
//  * the plane / line coordinates come from one division per thread (not two per vector), no 64-bit index arithmetic;
//  * a neighbour leg is ONE subtraction: s - v has the bits of s + (-1 * v), the product being exact.
// Same legs in the same order as the CSR rows, so still bit-identical to the CSR / DIA product.
// blockDim = (threads along the line: a multiple of 32, line groups per CTA); needs nx % NP == 0, ny % G == 0, whole lines.
__device__ __forceinline__ float sub_rn(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ double sub_rn(double a, double b) { return __dsub_rn(a, b); }

template <typename T, int G, bool PLANES, int MINB = 0>
__global__ void __launch_bounds__(256, MINB)
stencil_lines_kernel(int nx, int ny, int nz, uint32_t first_line, uint32_t num_groups, T centre, const T *__restrict__ x,
                     T *__restrict__ y) {
    constexpr int NP = Vec16<T>::NP;
    uint32_t iz, iy0;
    if constexpr (PLANES) {     // whole planes: grid = (line groups of a plane / blockDim.y, chunks of a line, planes), no division
        iy0 = (blockIdx.x * blockDim.y + threadIdx.y) * G;
        if (iy0 >= (uint32_t)ny) return;                                // the same for a whole warp
        iz = first_line + blockIdx.z;                                   // first_line is the first PLANE here
    } else {
        const uint32_t group = blockIdx.x * blockDim.y + threadIdx.y;
        if (group >= num_groups) return;
        const uint32_t line0 = first_line + group * G;
        iz = line0 / (uint32_t)ny;
        iy0 = line0 - iz * (uint32_t)ny;
    }
    const int lane = threadIdx.x & 31;
    const int ix = (int)(blockIdx.y * blockDim.x + threadIdx.x) * NP;
    const bool alive = ix < nx;                                         // dead lanes stay for the shuffles
    const int64_t plane = (int64_t)nx * ny, i0 = (int64_t)iz * plane + (int64_t)iy0 * nx + ix;
    const T *p = x + i0;
    const bool has_zm = iz > 0, has_zp = iz + 1 < (uint32_t)nz, has_ym = iy0 > 0, has_yp = iy0 + G < (uint32_t)ny;
    const bool edge_l = lane == 0 && ix > 0, edge_r = lane == 31 && ix + NP < nx;
    // Absent legs stay +0: a - (+0) == a for every a, so the sums below need no boundary tests.
    T c[G + 2][NP], zm[G][NP], zp[G][NP], el[G], er[G];
#pragma unroll
    for (int l = 0; l < G + 2; ++l)
#pragma unroll
        for (int t = 0; t < NP; ++t) c[l][t] = T(0);
#pragma unroll
    for (int l = 0; l < G; ++l) {
        el[l] = T(0); er[l] = T(0);
#pragma unroll
        for (int t = 0; t < NP; ++t) { zm[l][t] = T(0); zp[l][t] = T(0); }
    }
    // every load of the thread before the first use
    if (alive) {
        const T *q = p + plane;
        if (has_zp) {                                                    // the leg that comes from DRAM goes first
#pragma unroll
            for (int l = 0; l < G; ++l, q += nx) Vec16<T>::load(q, zp[l]);
        }
        q = p - nx;
#pragma unroll
        for (int l = 0; l < G + 2; ++l, q += nx)
            if ((l > 0 || has_ym) && (l < G + 1 || has_yp)) Vec16<T>::load(q, c[l]);
        q = p - plane;
        if (has_zm) {
#pragma unroll
            for (int l = 0; l < G; ++l, q += nx) Vec16<T>::load(q, zm[l]);
        }
        if (edge_l) {
            q = p - 1;
#pragma unroll
            for (int l = 0; l < G; ++l, q += nx) el[l] = __ldg(q);
        }
        if (edge_r) {
            q = p + NP;
#pragma unroll
            for (int l = 0; l < G; ++l, q += nx) er[l] = __ldg(q);
        }
    }
    T *out = y + i0;
#pragma unroll
    for (int l = 0; l < G; ++l, out += nx) {
        T left = __shfl_up_sync(0xFFFFFFFFu, c[l + 1][NP - 1], 1), right = __shfl_down_sync(0xFFFFFFFFu, c[l + 1][0], 1);
        if (lane == 0) left = el[l];          // the loaded neighbour, or +0 at the start of the line
        if (lane == 31) right = er[l];        // a dead lane's centre is +0, so the last alive lane of a short line needs no test
        T s[NP];
#pragma unroll
        for (int t = 0; t < NP; ++t) {
            T a = sub_rn(T(0), zm[l][t]);
            a = sub_rn(a, c[l][t]);
            a = sub_rn(a, t > 0 ? c[l + 1][t - 1] : left);
            a = add_rn(a, mul_rn(centre, c[l + 1][t]));
            a = sub_rn(a, t < NP - 1 ? c[l + 1][t + 1] : right);
            a = sub_rn(a, c[l + 2][t]);
            s[t] = sub_rn(a, zp[l][t]);
        }
        if (alive) Vec16<T>::store(out, s);
    }
}

template <typename T, int G, int MINB = 0>
void launch_stencil_lines(int nx, int ny, int nz, int64_t first_line, int64_t lines, T centre, const T *d_x, T *d_y, cudaStream_t st,
                          int cta = 256, int max_tx = 256) {
    constexpr int NP = Vec16<T>::NP;
    const int vecs = nx / NP;
    const int tx = vecs >= max_tx ? max_tx : (vecs + 31) / 32 * 32, ty = cta / tx > 0 ? cta / tx : 1;
    dim3 block((unsigned)tx, (unsigned)ty);
    const unsigned chunks = (unsigned)div_up(vecs, tx);
    if (first_line % ny == 0 && lines % ny == 0 && lines / ny <= 65535 && chunks <= 65535) {
        dim3 grid((unsigned)div_up(ny / G, ty), chunks, (unsigned)(lines / ny));
        stencil_lines_kernel<T, G, true, MINB><<<grid, block, 0, st>>>(nx, ny, nz, (uint32_t)(first_line / ny), 0u, centre, d_x, d_y);
    } else {
        const int64_t groups = lines / G;
        dim3 grid((unsigned)div_up(groups, ty), chunks);
        stencil_lines_kernel<T, G, false, MINB><<<grid, block, 0, st>>>(nx, ny, nz, (uint32_t)first_line, (uint32_t)groups, centre, d_x, d_y);
    }
}

template <typename T>
__global__ void __launch_bounds__(256)
stencil_scalar_kernel(int nx, int ny, int nz, int64_t row_begin, int64_t row_end, T centre, const T *__restrict__ x,
                      T *__restrict__ y) {
    const int64_t i = row_begin + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= row_end) return;
    const int64_t plane = (int64_t)nx * ny;
    const uint32_t iu = (uint32_t)i, line = iu / (uint32_t)nx, iz = line / (uint32_t)ny;
    y[i] = stencil_point(x, i, (int)(iu - line * (uint32_t)nx), (int)(line - iz * (uint32_t)ny), (int)iz, nx, ny, nz, plane, centre);
}

template <typename T>
int stencil_impl(cme213_ctx *ctx, int nx, int ny, int nz, T centre, int64_t row_begin, int64_t row_end, const T *d_x,
                 T *d_y, void *stream) {
    CME213_CHECK_CTX(ctx);
    const int64_t rows_total = (int64_t)nx * ny * nz;
    if (nx <= 0 || ny <= 0 || nz <= 0 || row_begin < 0 || row_end < row_begin || row_end > rows_total ||
        rows_total > 0x7FFFFFFFLL) {   // 32-bit indices, as in the generators
        set_error("stencil_laplace: bad shape");
        return CME213_ERR_INVALID;
    }
    if (row_end == row_begin) return CME213_OK;
    if (!d_x || !d_y || d_x == d_y) { set_error("stencil_laplace: bad pointers"); return CME213_ERR_INVALID; }
    DeviceGuard g(ctx->device);
    cudaStream_t st = (cudaStream_t)stream;
    const int64_t rows = row_end - row_begin;
    constexpr int NP = Vec16<T>::NP;
    if (nx % NP == 0 && row_begin % nx == 0 && rows % nx == 0 && aligned16(d_x) && aligned16(d_y)) {
        const int64_t first_line = row_begin / nx, lines = rows / nx;
        // 512^3 fp64, 256-thread CTAs (profiles/r02_stencil_kernel_sweep.txt): one-vector kernel 0.467 ms, G = 1 0.423, G = 2 0.412,
        // G = 4 0.435 (98 registers)
        int g = (ny % 2 == 0 && first_line % 2 == 0 && lines % 2 == 0) ? 2 : 1;
#ifdef CME213_TUNING   // CME213_STENCIL_CFG = "g,cta threads,max threads along x,min CTAs per SM"; g = 0: the one-vector kernel below
        int forced = -1, cta = 256, mtx = 256, minb = 0;
        if (const char *e = getenv("CME213_STENCIL_CFG")) sscanf(e, "%d,%d,%d,%d", &forced, &cta, &mtx, &minb);
        const bool ok4 = ny % 4 == 0 && first_line % 4 == 0 && lines % 4 == 0;
        bool done = true;
        if (forced == 4 && ok4) launch_stencil_lines<T, 4>(nx, ny, nz, first_line, lines, centre, d_x, d_y, st, cta, mtx);
        else if (forced == 2 && g == 2 && minb == 5) launch_stencil_lines<T, 2, 5>(nx, ny, nz, first_line, lines, centre, d_x, d_y, st, cta, mtx);
        else if (forced == 2 && g == 2 && minb == 6) launch_stencil_lines<T, 2, 6>(nx, ny, nz, first_line, lines, centre, d_x, d_y, st, cta, mtx);
        else if (forced == 2 && g == 2) launch_stencil_lines<T, 2>(nx, ny, nz, first_line, lines, centre, d_x, d_y, st, cta, mtx);
        else if (forced == 1 && minb == 8) launch_stencil_lines<T, 1, 8>(nx, ny, nz, first_line, lines, centre, d_x, d_y, st, cta, mtx);
        else if (forced == 1) launch_stencil_lines<T, 1>(nx, ny, nz, first_line, lines, centre, d_x, d_y, st, cta, mtx);
        else done = false;
        if (done) { CME213_LAUNCHED(ctx, "stencil_lines_kernel(cfg)"); return CME213_OK; }
        if (forced != 0)
#endif
        {
            // small CTAs: a slot is refilled as soon as its two warps have stored (G = 2: 256 threads 0.410, 128 0.391, 64 0.380 ms,
            // 32 0.547 -- 32 CTAs per SM at most; G = 1: 128 threads 0.421, 64 0.547: the CTA launch rate)
            if (g == 2) launch_stencil_lines<T, 2>(nx, ny, nz, first_line, lines, centre, d_x, d_y, st, 64, 64);
            else launch_stencil_lines<T, 1>(nx, ny, nz, first_line, lines, centre, d_x, d_y, st, 128, 128);
            CME213_LAUNCHED(ctx, "stencil_lines_kernel");
            return CME213_OK;
        }
    }
    if (nx % NP == 0 && row_begin % NP == 0 && rows % NP == 0 && aligned16(d_x) && aligned16(d_y)) {
        // 256 threads x one vector each measured best (0.464 ms at 512^3 fp64); 2 or 4 vectors per thread and larger
        // CTAs were slower (0.49-0.85 ms)
        stencil_vec_kernel<T, 1><<<(unsigned)div_up(rows / NP, 256), 256, 0, st>>>(nx, ny, nz, row_begin, row_end, centre, d_x, d_y);
        CME213_LAUNCHED(ctx, "stencil_vec_kernel");
    } else {
        stencil_scalar_kernel<T><<<(unsigned)div_up(rows, 256), 256, 0, st>>>(nx, ny, nz, row_begin, row_end, centre, d_x, d_y);
        CME213_LAUNCHED(ctx, "stencil_scalar_kernel");
    }
    return CME213_OK;
}

template <typename T>
int dia_impl(cme213_ctx *ctx, int64_t num_rows, int64_t num_cols, int64_t pitch, int num_diags,
             const int32_t *d_offsets, const T *d_data, const T *d_x, T *d_y, void *stream) {
    CME213_CHECK_CTX(ctx);
    if (num_rows < 0 || num_cols < 0 || num_diags < 0 || pitch < num_rows) { set_error("spmv_dia: bad shape"); return CME213_ERR_INVALID; }
    if (num_rows == 0) return CME213_OK;
    if (!d_y || (num_diags > 0 && (!d_offsets || !d_data || !d_x))) { set_error("spmv_dia: null pointer"); return CME213_ERR_INVALID; }
    DeviceGuard g(ctx->device);
    cudaStream_t st = (cudaStream_t)stream;
    const bool vec_ok = pitch % 4 == 0 && pitch >= ((num_rows + 3) & ~(int64_t)3) && aligned16(d_data);
    if (vec_ok) {
        const int64_t blocks = div_up(div_up(num_rows, 4), 256);
        dia_v4_kernel<T, 4><<<(unsigned)blocks, 256, 0, st>>>(num_rows, num_cols, pitch, num_diags, d_offsets, d_data, d_x,
                                                             d_y, aligned16(d_x) ? 1 : 0, aligned16(d_y) ? 1 : 0);
        CME213_LAUNCHED(ctx, "dia_v4_kernel");
    } else {
        dia_scalar_kernel<T><<<(unsigned)div_up(num_rows, 256), 256, 0, st>>>(num_rows, num_cols, pitch, num_diags,
                                                                             d_offsets, d_data, d_x, d_y);
        CME213_LAUNCHED(ctx, "dia_scalar_kernel");
    }
    return CME213_OK;
}

template <typename T>
int csr_to_dia_impl(cme213_ctx *ctx, int64_t num_rows, const int32_t *d_row_ptr, const int32_t *d_col, const T *d_val,
                    int num_diags, const int32_t *d_offsets, int64_t pitch, T *d_data, int64_t *unmatched_out,
                    void *stream) {
    CME213_CHECK_CTX(ctx);
    if (unmatched_out) *unmatched_out = 0;
    if (num_rows < 0 || num_diags < 0 || pitch < num_rows) { set_error("csr_to_dia: bad shape"); return CME213_ERR_INVALID; }
    if (num_rows == 0 || num_diags == 0) return CME213_OK;
    if (!d_row_ptr || !d_col || !d_val || !d_offsets || !d_data) { set_error("csr_to_dia: null pointer"); return CME213_ERR_INVALID; }
    DeviceGuard g(ctx->device);
    cudaStream_t st = (cudaStream_t)stream;
    void *ws = nullptr;
    int rc = workspace(ctx, 256, &ws, st);
    if (rc) return rc;
    unsigned long long *d_miss = (unsigned long long *)ws;
    CME213_CUDA(cudaMemsetAsync(d_miss, 0, 8, st));
    CME213_CUDA(cudaMemsetAsync(d_data, 0, (size_t)num_diags * (size_t)pitch * sizeof(T), st));
    csr_to_dia_kernel<T><<<(unsigned)div_up(num_rows, 256), 256, 0, st>>>(num_rows, d_row_ptr, d_col, d_val, num_diags,
                                                                         d_offsets, pitch, d_data, d_miss);
    CME213_LAUNCHED(ctx, "csr_to_dia_kernel");
    if (unmatched_out) {
        unsigned long long miss = 0;
        CME213_CUDA(cudaMemcpyAsync(&miss, d_miss, 8, cudaMemcpyDeviceToHost, st));
        CME213_CUDA(cudaStreamSynchronize(st));
        *unmatched_out = (int64_t)miss;
    }
    return CME213_OK;
}
}  // namespace

CME213_API int cme213_spmv_dia_f32(cme213_ctx *ctx, int64_t num_rows, int64_t num_cols, int64_t pitch, int num_diags,
                                   const int32_t *d_offsets, const float *d_data, const float *d_x, float *d_y,
                                   void *stream) {
    return dia_impl<float>(ctx, num_rows, num_cols, pitch, num_diags, d_offsets, d_data, d_x, d_y, stream);
}
CME213_API int cme213_spmv_dia_f64(cme213_ctx *ctx, int64_t num_rows, int64_t num_cols, int64_t pitch, int num_diags,
                                   const int32_t *d_offsets, const double *d_data, const double *d_x, double *d_y,
                                   void *stream) {
    return dia_impl<double>(ctx, num_rows, num_cols, pitch, num_diags, d_offsets, d_data, d_x, d_y, stream);
}
CME213_API int cme213_csr_to_dia_f32(cme213_ctx *ctx, int64_t num_rows, const int32_t *d_row_ptr, const int32_t *d_col,
                                     const float *d_val, int num_diags, const int32_t *d_offsets, int64_t pitch,
                                     float *d_data, int64_t *unmatched_out, void *stream) {
    return csr_to_dia_impl<float>(ctx, num_rows, d_row_ptr, d_col, d_val, num_diags, d_offsets, pitch, d_data,
                                  unmatched_out, stream);
}
CME213_API int cme213_csr_to_dia_f64(cme213_ctx *ctx, int64_t num_rows, const int32_t *d_row_ptr, const int32_t *d_col,
                                     const double *d_val, int num_diags, const int32_t *d_offsets, int64_t pitch,
                                     double *d_data, int64_t *unmatched_out, void *stream) {
    return csr_to_dia_impl<double>(ctx, num_rows, d_row_ptr, d_col, d_val, num_diags, d_offsets, pitch, d_data,
                                   unmatched_out, stream);
}

CME213_API int cme213_csr_diagonals(cme213_ctx *ctx, int64_t num_rows, int64_t num_cols, const int32_t *d_row_ptr,
                                    const int32_t *d_col, int max_diags, int32_t *d_offsets_out, int *num_diags_out,
                                    void *stream) {
    CME213_CHECK_CTX(ctx);
    if (!num_diags_out) { set_error("csr_diagonals: null count pointer"); return CME213_ERR_INVALID; }
    *num_diags_out = 0;
    if (num_rows < 0 || num_cols < 0 || max_diags < 0) { set_error("csr_diagonals: bad shape"); return CME213_ERR_INVALID; }
    if (num_rows == 0 || num_cols == 0) return CME213_OK;
    if (!d_row_ptr || !d_col || (max_diags > 0 && !d_offsets_out)) { set_error("csr_diagonals: null pointer"); return CME213_ERR_INVALID; }
    DeviceGuard g(ctx->device);
    cudaStream_t st = (cudaStream_t)stream;
    const int64_t slots = num_rows + num_cols - 1;
    const size_t a_bytes = ((size_t)(slots + 1) * 4 + 255) & ~(size_t)255;   // one extra slot: its scan value is the count
    void *ws = nullptr;
    int rc = workspace(ctx, 2 * a_bytes + 256 + cme213_internal_scan_u32_scratch_bytes(slots + 1), &ws, st);
    if (rc) return rc;
    uint32_t *mark = (uint32_t *)ws, *excl = (uint32_t *)((char *)ws + a_bytes);
    unsigned int *bad = (unsigned int *)((char *)ws + 2 * a_bytes);
    void *scratch = (char *)ws + 2 * a_bytes + 256;
    CME213_CUDA(cudaMemsetAsync(mark, 0, a_bytes, st));
    CME213_CUDA(cudaMemsetAsync(bad, 0, 256, st));
    mark_diagonals_kernel<<<(unsigned)div_up(num_rows, 256), 256, 0, st>>>(num_rows, num_cols, d_row_ptr, d_col, mark, bad);
    CME213_LAUNCHED(ctx, "mark_diagonals_kernel");
    rc = cme213_internal_exclusive_scan_u32(ctx, mark, excl, slots + 1, scratch, st);
    if (rc) return rc;
    if (max_diags > 0) {
        compact_diagonals_kernel<<<(unsigned)div_up(slots, 256), 256, 0, st>>>(slots, num_rows, mark, excl, d_offsets_out, max_diags);
        CME213_LAUNCHED(ctx, "compact_diagonals_kernel");
    }
    uint32_t count = 0, nbad = 0;
    CME213_CUDA(cudaMemcpyAsync(&count, excl + slots, 4, cudaMemcpyDeviceToHost, st));
    CME213_CUDA(cudaMemcpyAsync(&nbad, bad, 4, cudaMemcpyDeviceToHost, st));
    CME213_CUDA(cudaStreamSynchronize(st));
    if (nbad) {   // same rule as cme213_coo_to_csr_*: indices outside the matrix are an error, not silently dropped
        set_error("csr_diagonals: %u column indices outside [0, %lld)", nbad, (long long)num_cols);
        return CME213_ERR_INVALID;
    }
    *num_diags_out = (int)count;
    return CME213_OK;
}

CME213_API int cme213_stencil_laplace_f32(cme213_ctx *ctx, int nx, int ny, int nz, float centre, int64_t row_begin,
                                          int64_t row_end, const float *d_x, float *d_y, void *stream) {
    return stencil_impl<float>(ctx, nx, ny, nz, centre, row_begin, row_end, d_x, d_y, stream);
}
CME213_API int cme213_stencil_laplace_f64(cme213_ctx *ctx, int nx, int ny, int nz, double centre, int64_t row_begin,
                                          int64_t row_end, const double *d_x, double *d_y, void *stream) {
    return stencil_impl<double>(ctx, nx, ny, nz, centre, row_begin, row_end, d_x, d_y, stream);
}
