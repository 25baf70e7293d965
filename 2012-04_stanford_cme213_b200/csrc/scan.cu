// scan.cu -- entry points built on the look-back scan: exclusive uint32 scan, (fused) segmented
// scans of the final project, COO segmented reduction.
#include "lookback_scan.cuh"
#include <stdlib.h>
#include <stdio.h>

namespace {
using namespace cme213;

template <typename T> struct V16;
template <> struct V16<uint32_t> { using type = uint4; };
template <> struct V16<float> { using type = float4; };
template <> struct V16<double> { using type = double2; };
template <int IPV> struct IdxVec;
template <> struct IdxVec<4> { using type = int4; };
template <> struct IdxVec<2> { using type = int2; };

struct NoAux {};

// Rounded product, never contracted into an FMA with the additions of the scan: the reference
// rounds a[i]*x[k[i]] to working precision before summing (fp.cu:139-140, thrust multiplies).
__device__ __forceinline__ float mul_rn(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ double mul_rn(double a, double b) { return __dmul_rn(a, b); }

template <int IPV>
__device__ __forceinline__ unsigned head_bits(const uint32_t *bits, int64_t idx) {
    return (bits[idx >> 5] >> (unsigned)(idx & 31)) & ((1u << IPV) - 1u);
}

// ---- plain load / store ------------------------------------------------------------------------
template <typename T_, bool SEG_, bool INCLUSIVE>
struct PlainPolicy {
    using T = T_;
    using Aux = NoAux;
    static constexpr int IPV = 16 / sizeof(T);
    static constexpr bool SEG = SEG_;
    static constexpr bool BULK = true;
    static constexpr bool INCLUSIVE_OUT = INCLUSIVE;
    using V16T = typename V16<T_>::type;
    __host__ __device__ T *out_ptr() const { return out; }
    static constexpr int NIN = SEG_ ? 2 : 1;   // values, and for segmented scans the head-flag words (1 bit per item)
    __host__ __device__ static constexpr int tile_bytes(int s, int tile_items) { return s == 0 ? tile_items * (int)sizeof(T) : tile_items / 8; }
    const T *in;
    T *out;
    const uint32_t *flag_bits;
    __host__ __device__ const void *in_ptr(int s) const { return s == 0 ? (const void *)in : (const void *)flag_bits; }

    __device__ void load_smem(const unsigned char *const (&sin)[NIN], int vec, int, int64_t, int64_t, T (&v)[IPV],
                              unsigned &head, Aux &) const {
        typename V16<T>::type x = *reinterpret_cast<const typename V16<T>::type *>(sin[0] + (size_t)vec * 16);
        const T *e = reinterpret_cast<const T *>(&x);
#pragma unroll
        for (int j = 0; j < IPV; j++) v[j] = e[j];
        head = 0;
        if constexpr (SEG_) head = head_bits<IPV>(reinterpret_cast<const uint32_t *>(sin[NIN - 1]), (int64_t)vec * IPV);
    }
    __device__ void load(int64_t idx, int64_t n, bool fast, T (&v)[IPV], unsigned &head, Aux &) const {
        head = 0;
        if (fast) {
            typename V16<T>::type x = ld_stream(reinterpret_cast<const typename V16<T>::type *>(in + idx));
            const T *e = reinterpret_cast<const T *>(&x);
#pragma unroll
            for (int j = 0; j < IPV; j++) v[j] = e[j];
            if (SEG) head = head_bits<IPV>(flag_bits, idx);
        } else {
#pragma unroll
            for (int j = 0; j < IPV; j++) {
                v[j] = T(0);
                if (idx + j < n) {
                    v[j] = in[idx + j];
                    if (SEG) head |= ((flag_bits[(idx + j) >> 5] >> (unsigned)((idx + j) & 31)) & 1u) << j;
                }
            }
        }
    }
    __device__ void store(int64_t idx, int64_t n, bool fast, const T (&incl)[IPV], const T (&excl)[IPV],
                          const Aux &) const {
        const T(&r)[IPV] = INCLUSIVE ? incl : excl;
        if (fast) {
            typename V16<T>::type x;
            T *e = reinterpret_cast<T *>(&x);
#pragma unroll
            for (int j = 0; j < IPV; j++) e[j] = r[j];
            st_stream(reinterpret_cast<typename V16<T>::type *>(out + idx), x);
        } else {
#pragma unroll
            for (int j = 0; j < IPV; j++)
                if (idx + j < n) out[idx + j] = r[j];
        }
    }
};

// ---- final project step with the gather hoisted: v = a[i] * xx[i], inclusive segmented sum ------------------------
// x[k[i]] is the same in every one of the N iterations (only a changes: a <- b), which the reference exploits by
// flattening xx = x[k] once (fp.cu:124-125, on the host).  Here the flattening runs once on the device inside the timed
// region (cme213_gather_*), and the remaining work per iteration is pure streaming: two value streams and the flag
// words in, one value stream out, through the bulk-copy pipeline -- no random gather, so no L2-sector bound.
template <typename T_>
struct MulPolicy {
    using T = T_;
    using Aux = NoAux;
    static constexpr int IPV = 16 / sizeof(T);
    static constexpr bool SEG = true;
    static constexpr bool BULK = true;
    static constexpr bool INCLUSIVE_OUT = true;
    using V16T = typename V16<T_>::type;
    static constexpr int NIN = 3;   // a, xx, head-flag words
    __host__ __device__ static constexpr int tile_bytes(int s, int tile_items) { return s < 2 ? tile_items * (int)sizeof(T) : tile_items / 8; }
    const T *a;
    const T *xx;
    const uint32_t *flag_bits;
    T *b;
    __host__ __device__ T *out_ptr() const { return b; }
    __host__ __device__ const void *in_ptr(int s) const { return s == 0 ? (const void *)a : s == 1 ? (const void *)xx : (const void *)flag_bits; }

    __device__ void load_smem(const unsigned char *const (&sin)[NIN], int vec, int, int64_t, int64_t, T (&v)[IPV],
                              unsigned &head, Aux &) const {
        typename V16<T>::type av = *reinterpret_cast<const typename V16<T>::type *>(sin[0] + (size_t)vec * 16);
        typename V16<T>::type xv = *reinterpret_cast<const typename V16<T>::type *>(sin[1] + (size_t)vec * 16);
        const T *ae = reinterpret_cast<const T *>(&av), *xe = reinterpret_cast<const T *>(&xv);
#pragma unroll
        for (int j = 0; j < IPV; j++) v[j] = mul_rn(ae[j], xe[j]);
        head = head_bits<IPV>(reinterpret_cast<const uint32_t *>(sin[2]), (int64_t)vec * IPV);
    }
    __device__ void load(int64_t idx, int64_t n, bool fast, T (&v)[IPV], unsigned &head, Aux &) const {
        head = 0;
        if (fast) {
            typename V16<T>::type av = ld_stream(reinterpret_cast<const typename V16<T>::type *>(a + idx));
            typename V16<T>::type xv = ld_stream(reinterpret_cast<const typename V16<T>::type *>(xx + idx));
            const T *ae = reinterpret_cast<const T *>(&av), *xe = reinterpret_cast<const T *>(&xv);
#pragma unroll
            for (int j = 0; j < IPV; j++) v[j] = mul_rn(ae[j], xe[j]);
            head = head_bits<IPV>(flag_bits, idx);
        } else {
#pragma unroll
            for (int j = 0; j < IPV; j++) {
                v[j] = T(0);
                if (idx + j < n) {
                    v[j] = mul_rn(a[idx + j], xx[idx + j]);
                    head |= ((flag_bits[(idx + j) >> 5] >> (unsigned)((idx + j) & 31)) & 1u) << j;
                }
            }
        }
    }
    __device__ void store(int64_t idx, int64_t n, bool fast, const T (&incl)[IPV], const T (&)[IPV], const Aux &) const {
        if (fast) {
            typename V16<T>::type o;
            T *e = reinterpret_cast<T *>(&o);
#pragma unroll
            for (int j = 0; j < IPV; j++) e[j] = incl[j];
            st_stream(reinterpret_cast<typename V16<T>::type *>(b + idx), o);
        } else {
#pragma unroll
            for (int j = 0; j < IPV; j++)
                if (idx + j < n) b[idx + j] = incl[j];
        }
    }
};

// xx[i] = x[k[i]]: four consecutive indices per thread (one 128-bit load of k, one 128-bit store of xx), x gathers evict-last
template <typename T>
__global__ void __launch_bounds__(256) gather_kernel(const T *__restrict__ x, const int32_t *__restrict__ k, T *__restrict__ xx, int64_t n, int vec_ok) {
    const int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (i >= n) return;
    const uint64_t pol = l2_policy_evict_last();
    if (vec_ok && i + 4 <= n) {
        const int4 kv = ld_stream(reinterpret_cast<const int4 *>(k + i));
        const T v0 = ld_keep(x + kv.x, pol), v1 = ld_keep(x + kv.y, pol), v2 = ld_keep(x + kv.z, pol), v3 = ld_keep(x + kv.w, pol);
        if constexpr (sizeof(T) == 4) {
            st_stream(reinterpret_cast<float4 *>(xx + i), make_float4(v0, v1, v2, v3));
        } else {
            st_stream(reinterpret_cast<double2 *>(xx + i), make_double2(v0, v1));
            st_stream(reinterpret_cast<double2 *>(xx + i) + 1, make_double2(v2, v3));
        }
    } else {
        for (int j = 0; j < 4 && i + j < n; j++) xx[i + j] = __ldg(x + k[i + j]);
    }
}

// ---- final project step: v = a[i] * x[k[i]], inclusive segmented sum ------------------------------
template <typename T_>
struct MulGatherPolicy {
    using T = T_;
    using Aux = NoAux;
    static constexpr int IPV = 16 / sizeof(T);
    static constexpr bool SEG = true;
    static constexpr bool BULK = false;  // bound by L2 gather rate, not by the stream: many small CTAs per SM hide it better (profiles/r01_notes.md)
    static constexpr int NIN = 3;   // a, k, head-flag words
    __host__ __device__ static constexpr int tile_bytes(int s, int tile_items) {
        return s == 0 ? tile_items * (int)sizeof(T) : s == 1 ? tile_items * 4 : tile_items / 8;
    }
    const T *a;
    const int32_t *k;
    const T *x;
    const uint32_t *flag_bits;
    T *b;
    __host__ __device__ const void *in_ptr(int s) const {
        return s == 0 ? (const void *)a : s == 1 ? (const void *)k : (const void *)flag_bits;
    }

    __device__ void load_smem(const unsigned char *const (&sin)[NIN], int vec, int, int64_t idx, int64_t, T (&v)[IPV],
                              unsigned &head, Aux &) const {
        typename V16<T>::type av = *reinterpret_cast<const typename V16<T>::type *>(sin[0] + (size_t)vec * 16);
        typename IdxVec<IPV>::type kv = *reinterpret_cast<const typename IdxVec<IPV>::type *>(sin[1] + (size_t)vec * IPV * 4);
        const T *ae = reinterpret_cast<const T *>(&av);
        const int *ke = reinterpret_cast<const int *>(&kv);
        T xv[IPV];
#pragma unroll
        for (int j = 0; j < IPV; j++) xv[j] = __ldg(x + ke[j]);
#pragma unroll
        for (int j = 0; j < IPV; j++) v[j] = mul_rn(ae[j], xv[j]);
        head = head_bits<IPV>(reinterpret_cast<const uint32_t *>(sin[2]), (int64_t)vec * IPV);
    }
    __device__ void load(int64_t idx, int64_t n, bool fast, T (&v)[IPV], unsigned &head, Aux &) const {
        head = 0;
        if (fast) {
            typename V16<T>::type av = ld_stream(reinterpret_cast<const typename V16<T>::type *>(a + idx));
            typename IdxVec<IPV>::type kv = ld_stream(reinterpret_cast<const typename IdxVec<IPV>::type *>(k + idx));
            const T *ae = reinterpret_cast<const T *>(&av);
            const int *ke = reinterpret_cast<const int *>(&kv);
            T xv[IPV];
#pragma unroll
            for (int j = 0; j < IPV; j++) xv[j] = __ldg(x + ke[j]);
#pragma unroll
            for (int j = 0; j < IPV; j++) v[j] = mul_rn(ae[j], xv[j]);
            head = head_bits<IPV>(flag_bits, idx);
        } else {
#pragma unroll
            for (int j = 0; j < IPV; j++) {
                v[j] = T(0);
                if (idx + j < n) {
                    v[j] = mul_rn(a[idx + j], __ldg(x + k[idx + j]));
                    head |= ((flag_bits[(idx + j) >> 5] >> (unsigned)((idx + j) & 31)) & 1u) << j;
                }
            }
        }
    }
    __device__ void store(int64_t idx, int64_t n, bool fast, const T (&incl)[IPV], const T (&)[IPV],
                          const Aux &) const {
        if (fast) {
            typename V16<T>::type o;
            T *e = reinterpret_cast<T *>(&o);
#pragma unroll
            for (int j = 0; j < IPV; j++) e[j] = incl[j];
            st_stream(reinterpret_cast<typename V16<T>::type *>(b + idx), o);
        } else {
#pragma unroll
            for (int j = 0; j < IPV; j++)
                if (idx + j < n) b[idx + j] = incl[j];
        }
    }
};

// ---- COO segmented reduction: heads where the row index changes, row ends add into y ------------
template <typename T_>
struct CooPolicy {
    using T = T_;
    static constexpr int IPV = 16 / sizeof(T);
    static constexpr bool SEG = true;
    struct Aux { int row[IPV]; int next; };
    static constexpr bool BULK = false;  // gather-bound like the fused step
    static constexpr int NIN = 3;
    __host__ __device__ static constexpr int tile_bytes(int s, int tile_items) { return tile_items * (s == 2 ? (int)sizeof(T) : 4); }
    const int32_t *row;
    const int32_t *col;
    const T *val;
    const T *x;
    T *y;
    __host__ __device__ const void *in_ptr(int s) const {
        return s == 0 ? (const void *)row : s == 1 ? (const void *)col : (const void *)val;
    }

    __device__ void load_smem(const unsigned char *const (&sin)[NIN], int vec, int nvec, int64_t idx, int64_t n,
                              T (&v)[IPV], unsigned &head, Aux &aux) const {
        const int *srow = reinterpret_cast<const int *>(sin[0]);
        typename IdxVec<IPV>::type rv = *reinterpret_cast<const typename IdxVec<IPV>::type *>(sin[0] + (size_t)vec * IPV * 4);
        typename IdxVec<IPV>::type cv = *reinterpret_cast<const typename IdxVec<IPV>::type *>(sin[1] + (size_t)vec * IPV * 4);
        typename V16<T>::type vv = *reinterpret_cast<const typename V16<T>::type *>(sin[2] + (size_t)vec * 16);
        const int *re = reinterpret_cast<const int *>(&rv);
        const int *ce = reinterpret_cast<const int *>(&cv);
        const T *ve = reinterpret_cast<const T *>(&vv);
        T xv[IPV];
#pragma unroll
        for (int j = 0; j < IPV; j++) xv[j] = __ldg(x + ce[j]);
        // neighbours across vector boundaries come from the staged tile; only the tile's two ends touch global
        int prev = vec > 0 ? srow[vec * IPV - 1] : (idx > 0 ? __ldg(row + idx - 1) : -1);
        aux.next = vec + 1 < nvec ? srow[(vec + 1) * IPV] : (idx + IPV < n ? __ldg(row + idx + IPV) : -1);
        head = 0;
#pragma unroll
        for (int j = 0; j < IPV; j++) {
            aux.row[j] = re[j];
            v[j] = mul_rn(ve[j], xv[j]);
            if (re[j] != prev) head |= 1u << j;
            prev = re[j];
        }
    }
    __device__ void load(int64_t idx, int64_t n, bool fast, T (&v)[IPV], unsigned &head, Aux &aux) const {
        head = 0;
        int prev = (idx > 0 && idx - 1 < n) ? __ldg(row + idx - 1) : -1;
        if (fast) {
            typename IdxVec<IPV>::type rv = ld_stream(reinterpret_cast<const typename IdxVec<IPV>::type *>(row + idx));
            typename IdxVec<IPV>::type cv = ld_stream(reinterpret_cast<const typename IdxVec<IPV>::type *>(col + idx));
            typename V16<T>::type vv = ld_stream(reinterpret_cast<const typename V16<T>::type *>(val + idx));
            const int *re = reinterpret_cast<const int *>(&rv);
            const int *ce = reinterpret_cast<const int *>(&cv);
            const T *ve = reinterpret_cast<const T *>(&vv);
            T xv[IPV];
#pragma unroll
            for (int j = 0; j < IPV; j++) xv[j] = __ldg(x + ce[j]);
#pragma unroll
            for (int j = 0; j < IPV; j++) {
                aux.row[j] = re[j];
                v[j] = mul_rn(ve[j], xv[j]);
            }
            aux.next = (idx + IPV < n) ? __ldg(row + idx + IPV) : -1;
        } else {
#pragma unroll
            for (int j = 0; j < IPV; j++) {
                aux.row[j] = -1;
                v[j] = T(0);
                if (idx + j < n) {
                    aux.row[j] = row[idx + j];
                    v[j] = mul_rn(val[idx + j], __ldg(x + col[idx + j]));
                }
            }
            aux.next = (idx + IPV < n) ? row[idx + IPV] : -1;
        }
#pragma unroll
        for (int j = 0; j < IPV; j++) {
            if (aux.row[j] != prev) head |= 1u << j;
            prev = aux.row[j];
        }
    }
    __device__ void store(int64_t, int64_t, bool, const T (&incl)[IPV], const T (&)[IPV], const Aux &aux) const {
#pragma unroll
        for (int j = 0; j < IPV; j++) {
            const int nxt = j + 1 < IPV ? aux.row[j + 1] : aux.next;
            if (aux.row[j] >= 0 && aux.row[j] != nxt) y[aux.row[j]] += incl[j];
        }
    }
};

__global__ void flags_from_offsets_kernel(const int32_t *__restrict__ s, int p, int64_t n, uint32_t *bits) {
    // s[0..p-1] are segment starts with s[p-1] == n (aux/mp1-util.h:131-139); the last is not an element.
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < p) {
        const int64_t pos = s[i];
        if (pos >= 0 && pos < n) atomicOr(bits + (pos >> 5), 1u << (unsigned)(pos & 31));
    }
}

template <class P>
int run(cme213_ctx *ctx, const P &p, int64_t n, bool vec_ok, void *stream, const char *name) {
    DeviceGuard g(ctx->device);
#ifdef CME213_TUNING   // tuning hooks (kernel-shape sweeps by environment variable): compiled out of the shipped library
    if constexpr (P::BULK) {
        if (const char *cfg = getenv("CME213_SCAN_CFG")) {  // tuning hook: "threads,vecs,-stages,defer" picks a blocked-kernel shape
            int T_ = 0, V_ = 0, S_ = 0, D_ = 1;
            sscanf(cfg, "%d,%d,%d,%d", &T_, &V_, &S_, &D_);
            cudaStream_t st = (cudaStream_t)stream;
            if (vec_ok && n >= ((int64_t)1 << 20)) {
#define CME213_TRY_BLOCKED(T, V, S, D) if (T_ == T && V_ == V && S_ == -(S) && D_ == D) return scan::launch_bulk<P, T, V, S, 5, true, D>(ctx, p, n, st, name, nullptr, 0);
                if constexpr (P::NIN >= 3) {   // two value streams per tile: smaller tiles
                    CME213_TRY_BLOCKED(640, 3, 3, 1) CME213_TRY_BLOCKED(640, 2, 4, 2) CME213_TRY_BLOCKED(768, 2, 4, 2) CME213_TRY_BLOCKED(512, 3, 4, 2)
                    CME213_TRY_BLOCKED(1024, 2, 3, 1) CME213_TRY_BLOCKED(896, 2, 3, 1) CME213_TRY_BLOCKED(576, 3, 4, 2) CME213_TRY_BLOCKED(448, 5, 3, 1)
                    CME213_TRY_BLOCKED(384, 5, 3, 1) CME213_TRY_BLOCKED(448, 3, 5, 3) CME213_TRY_BLOCKED(384, 3, 6, 4) CME213_TRY_BLOCKED(320, 3, 7, 5)
                } else {
                    CME213_TRY_BLOCKED(1024, 3, 4, 2) CME213_TRY_BLOCKED(896, 5, 3, 1) CME213_TRY_BLOCKED(640, 5, 4, 2)
                    CME213_TRY_BLOCKED(768, 4, 4, 2) CME213_TRY_BLOCKED(512, 6, 4, 2) CME213_TRY_BLOCKED(704, 4, 4, 2) CME213_TRY_BLOCKED(640, 4, 5, 3)
                    CME213_TRY_BLOCKED(576, 5, 4, 2) CME213_TRY_BLOCKED(640, 6, 3, 1) CME213_TRY_BLOCKED(512, 5, 5, 3) CME213_TRY_BLOCKED(832, 4, 4, 2)
                    CME213_TRY_BLOCKED(448, 7, 4, 2) CME213_TRY_BLOCKED(704, 5, 3, 1) CME213_TRY_BLOCKED(768, 5, 3, 1) CME213_TRY_BLOCKED(512, 7, 3, 1)
                    CME213_TRY_BLOCKED(384, 7, 5, 3) CME213_TRY_BLOCKED(320, 7, 6, 4) CME213_TRY_BLOCKED(256, 7, 7, 5) CME213_TRY_BLOCKED(384, 5, 6, 4)
                    CME213_TRY_BLOCKED(448, 5, 5, 3) CME213_TRY_BLOCKED(384, 7, 4, 2)
                }
#undef CME213_TRY_BLOCKED
            }
        }
    }
    if constexpr (!P::BULK) {
        if (const char *cfg = getenv("CME213_DIRECT_CFG")) {  // tuning hook: "threads,vecs,min_blocks" for the one-CTA-per-tile kernel
            int T_ = 0, V_ = 0, B_ = 0;
            sscanf(cfg, "%d,%d,%d", &T_, &V_, &B_);
            cudaStream_t st = (cudaStream_t)stream;
#define CME213_TRY_DIRECT(T, V, B) if (T_ == T && V_ == V && B_ == B) return scan::launch_direct<P, T, V, B>(ctx, p, n, vec_ok, st, name, nullptr);
            CME213_TRY_DIRECT(256, 4, 4) CME213_TRY_DIRECT(128, 8, 8)
            CME213_TRY_DIRECT(256, 4, 6) CME213_TRY_DIRECT(256, 4, 8)
#undef CME213_TRY_DIRECT
        }
    }
#endif
    return scan::launch<P>(ctx, p, n, vec_ok, (cudaStream_t)stream, name);
}

}  // namespace

CME213_API int64_t cme213_flag_words(int64_t n) { return n <= 0 ? 0 : (n + 31) / 32; }

CME213_API int cme213_segment_flags_from_offsets(cme213_ctx *ctx, const int32_t *d_s, int p, int64_t n,
                                                 uint32_t *d_flag_bits, void *stream) {
    CME213_CHECK_CTX(ctx);
    if (p < 0 || n < 0 || (p > 0 && !d_s) || (n > 0 && !d_flag_bits)) {
        set_error("segment_flags_from_offsets: bad arguments");
        return CME213_ERR_INVALID;
    }
    if (n == 0) return CME213_OK;
    DeviceGuard g(ctx->device);
    cudaStream_t st = (cudaStream_t)stream;
    CME213_CUDA(cudaMemsetAsync(d_flag_bits, 0, (size_t)cme213_flag_words(n) * 4, st));
    if (p > 0) {
        flags_from_offsets_kernel<<<(unsigned)div_up(p, 256), 256, 0, st>>>(d_s, p, n, d_flag_bits);
        CME213_LAUNCHED(ctx, "flags_from_offsets_kernel");
    }
    return CME213_OK;
}

CME213_API int cme213_exclusive_scan_u32(cme213_ctx *ctx, const uint32_t *d_in, uint32_t *d_out, int64_t n,
                                         void *stream) {
    CME213_CHECK_CTX(ctx);
    if (n < 0 || (n > 0 && (!d_in || !d_out))) { set_error("exclusive_scan_u32: bad arguments"); return CME213_ERR_INVALID; }
    if (n == 0) return CME213_OK;
    PlainPolicy<uint32_t, false, false> p{d_in, d_out, nullptr};
    const bool ok = aligned16(d_in) && aligned16(d_out);
    return run(ctx, p, n, ok, stream, "exclusive_scan_u32");
}

template <typename T, bool INCLUSIVE>
static int segscan_impl(cme213_ctx *ctx, const T *d_in, const uint32_t *d_flag_bits, T *d_out, int64_t n,
                        void *stream, const char *name) {
    CME213_CHECK_CTX(ctx);
    if (n < 0 || (n > 0 && (!d_in || !d_out || !d_flag_bits))) { set_error("%s: bad arguments", name); return CME213_ERR_INVALID; }
    if (n == 0) return CME213_OK;
    PlainPolicy<T, true, INCLUSIVE> p{d_in, d_out, d_flag_bits};
    return run(ctx, p, n, aligned16(d_in) && aligned16(d_out) && aligned16(d_flag_bits), stream, name);
}
CME213_API int cme213_segscan_inclusive_f32(cme213_ctx *ctx, const float *d_in, const uint32_t *d_flag_bits,
                                            float *d_out, int64_t n, void *stream) {
    return segscan_impl<float, true>(ctx, d_in, d_flag_bits, d_out, n, stream, "segscan_inclusive_f32");
}
CME213_API int cme213_segscan_inclusive_f64(cme213_ctx *ctx, const double *d_in, const uint32_t *d_flag_bits,
                                            double *d_out, int64_t n, void *stream) {
    return segscan_impl<double, true>(ctx, d_in, d_flag_bits, d_out, n, stream, "segscan_inclusive_f64");
}
CME213_API int cme213_segscan_exclusive_f32(cme213_ctx *ctx, const float *d_in, const uint32_t *d_flag_bits,
                                            float *d_out, int64_t n, void *stream) {
    return segscan_impl<float, false>(ctx, d_in, d_flag_bits, d_out, n, stream, "segscan_exclusive_f32");
}

template <typename T>
static int mul_gather_impl(cme213_ctx *ctx, const T *d_a, const int32_t *d_k, const T *d_x,
                           const uint32_t *d_flag_bits, T *d_b, int64_t n, void *stream, const char *name) {
    CME213_CHECK_CTX(ctx);
    if (n < 0 || (n > 0 && (!d_a || !d_k || !d_x || !d_flag_bits || !d_b))) {
        set_error("%s: bad arguments", name);
        return CME213_ERR_INVALID;
    }
    if (n == 0) return CME213_OK;
    MulGatherPolicy<T> p{d_a, d_k, d_x, d_flag_bits, d_b};
    return run(ctx, p, n, aligned16(d_a) && aligned16(d_k) && aligned16(d_b) && aligned16(d_flag_bits), stream, name);
}
CME213_API int cme213_mul_gather_segscan_f32(cme213_ctx *ctx, const float *d_a, const int32_t *d_k, const float *d_x,
                                             const uint32_t *d_flag_bits, float *d_b, int64_t n, void *stream) {
    return mul_gather_impl<float>(ctx, d_a, d_k, d_x, d_flag_bits, d_b, n, stream, "mul_gather_segscan_f32");
}
CME213_API int cme213_mul_gather_segscan_f64(cme213_ctx *ctx, const double *d_a, const int32_t *d_k,
                                             const double *d_x, const uint32_t *d_flag_bits, double *d_b, int64_t n,
                                             void *stream) {
    return mul_gather_impl<double>(ctx, d_a, d_k, d_x, d_flag_bits, d_b, n, stream, "mul_gather_segscan_f64");
}

template <typename T>
static int mul_segscan_impl(cme213_ctx *ctx, const T *d_a, const T *d_xx, const uint32_t *d_flag_bits, T *d_b, int64_t n, void *stream,
                            const char *name) {
    CME213_CHECK_CTX(ctx);
    if (n < 0 || (n > 0 && (!d_a || !d_xx || !d_flag_bits || !d_b))) { set_error("%s: bad arguments", name); return CME213_ERR_INVALID; }
    if (n == 0) return CME213_OK;
    MulPolicy<T> p{d_a, d_xx, d_flag_bits, d_b};
    return run(ctx, p, n, aligned16(d_a) && aligned16(d_xx) && aligned16(d_b) && aligned16(d_flag_bits), stream, name);
}
CME213_API int cme213_mul_segscan_f32(cme213_ctx *ctx, const float *d_a, const float *d_xx, const uint32_t *d_flag_bits, float *d_b,
                                      int64_t n, void *stream) {
    return mul_segscan_impl<float>(ctx, d_a, d_xx, d_flag_bits, d_b, n, stream, "mul_segscan_f32");
}
CME213_API int cme213_mul_segscan_f64(cme213_ctx *ctx, const double *d_a, const double *d_xx, const uint32_t *d_flag_bits, double *d_b,
                                      int64_t n, void *stream) {
    return mul_segscan_impl<double>(ctx, d_a, d_xx, d_flag_bits, d_b, n, stream, "mul_segscan_f64");
}
template <typename T>
static int gather_impl(cme213_ctx *ctx, const T *d_x, const int32_t *d_k, T *d_xx, int64_t n, void *stream) {
    CME213_CHECK_CTX(ctx);
    if (n < 0 || (n > 0 && (!d_x || !d_k || !d_xx))) { set_error("gather: bad arguments"); return CME213_ERR_INVALID; }
    if (n == 0) return CME213_OK;
    DeviceGuard g(ctx->device);
    const int64_t blocks = div_up(div_up(n, 4), 256);
    if (blocks > 0x7FFFFFFFLL) { set_error("gather: n too large"); return CME213_ERR_UNSUPPORTED; }
    gather_kernel<T><<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(d_x, d_k, d_xx, n, (aligned16(d_k) && aligned16(d_xx)) ? 1 : 0);
    CME213_LAUNCHED(ctx, "gather_kernel");
    return CME213_OK;
}
CME213_API int cme213_gather_f32(cme213_ctx *ctx, const float *d_x, const int32_t *d_k, float *d_xx, int64_t n, void *stream) {
    return gather_impl<float>(ctx, d_x, d_k, d_xx, n, stream);
}
CME213_API int cme213_gather_f64(cme213_ctx *ctx, const double *d_x, const int32_t *d_k, double *d_xx, int64_t n, void *stream) {
    return gather_impl<double>(ctx, d_x, d_k, d_xx, n, stream);
}

template <typename T>
static int coo_impl(cme213_ctx *ctx, int64_t nnz, const int32_t *d_row, const int32_t *d_col, const T *d_val,
                    const T *d_x, T *d_y, void *stream, const char *name) {
    CME213_CHECK_CTX(ctx);
    if (nnz < 0 || (nnz > 0 && (!d_row || !d_col || !d_val || !d_x || !d_y))) {
        set_error("%s: bad arguments", name);
        return CME213_ERR_INVALID;
    }
    if (nnz == 0) return CME213_OK;
    CooPolicy<T> p{d_row, d_col, d_val, d_x, d_y};
    return run(ctx, p, nnz, aligned16(d_row) && aligned16(d_col) && aligned16(d_val), stream, name);
}
CME213_API int cme213_spmv_coo_accumulate_f32(cme213_ctx *ctx, int64_t nnz, const int32_t *d_row,
                                              const int32_t *d_col, const float *d_val, const float *d_x, float *d_y,
                                              void *stream) {
    return coo_impl<float>(ctx, nnz, d_row, d_col, d_val, d_x, d_y, stream, "spmv_coo_accumulate_f32");
}
CME213_API int cme213_spmv_coo_accumulate_f64(cme213_ctx *ctx, int64_t nnz, const int32_t *d_row,
                                              const int32_t *d_col, const double *d_val, const double *d_x,
                                              double *d_y, void *stream) {
    return coo_impl<double>(ctx, nnz, d_row, d_col, d_val, d_x, d_y, stream, "spmv_coo_accumulate_f64");
}

// ---- internal: scan with caller-provided scratch (see internal.cuh) ----------------------------------
#include "internal.cuh"
void cme213_internal_preload_fp(int elem_bytes) {
    cudaFuncAttributes at;
    using scan::Cfg;
    if (elem_bytes == 4) {
        using P = MulPolicy<float>;
        using G = MulGatherPolicy<float>;
        cudaFuncGetAttributes(&at, gather_kernel<float>);
        cudaFuncGetAttributes(&at, scan::lookback_scan_kernel<P, Cfg<P>::DIRECT_THREADS, Cfg<P>::DIRECT_VECS, Cfg<P>::DIRECT_MIN_BLOCKS>);
        cudaFuncGetAttributes(&at, scan::scan_bulk_blocked_kernel<P, Cfg<P>::B_THREADS, Cfg<P>::B_VECS, Cfg<P>::B_STAGES, Cfg<P>::DPL, Cfg<P>::B_DEFER>);
        cudaFuncGetAttributes(&at, scan::lookback_scan_kernel<G, Cfg<G>::DIRECT_THREADS, Cfg<G>::DIRECT_VECS, Cfg<G>::DIRECT_MIN_BLOCKS>);
    } else {
        using P = MulPolicy<double>;
        using G = MulGatherPolicy<double>;
        cudaFuncGetAttributes(&at, gather_kernel<double>);
        cudaFuncGetAttributes(&at, scan::lookback_scan_kernel<P, Cfg<P>::DIRECT_THREADS, Cfg<P>::DIRECT_VECS, Cfg<P>::DIRECT_MIN_BLOCKS>);
        cudaFuncGetAttributes(&at, scan::scan_bulk_blocked_kernel<P, Cfg<P>::B_THREADS, Cfg<P>::B_VECS, Cfg<P>::B_STAGES, Cfg<P>::DPL, Cfg<P>::B_DEFER>);
        cudaFuncGetAttributes(&at, scan::lookback_scan_kernel<G, Cfg<G>::DIRECT_THREADS, Cfg<G>::DIRECT_VECS, Cfg<G>::DIRECT_MIN_BLOCKS>);
    }
    cudaGetLastError();
}
size_t cme213_internal_scan_u32_scratch_bytes(int64_t n) {
    return scan::scratch_bytes_for<PlainPolicy<uint32_t, false, false>>(n);
}
int cme213_internal_exclusive_scan_u32(cme213_ctx *ctx, const uint32_t *d_in, uint32_t *d_out, int64_t n,
                                       void *scratch, void *stream) {
    if (n <= 0) return CME213_OK;
    DeviceGuard g(ctx->device);
    PlainPolicy<uint32_t, false, false> p{d_in, d_out, nullptr};
    return scan::launch_with_scratch(ctx, p, n, aligned16(d_in) && aligned16(d_out), (cudaStream_t)stream,
                                     "exclusive_scan_u32", scratch, cme213_internal_scan_u32_scratch_bytes(n));
}

// ---- exclusive scan that allocates nothing -------------------------------------------------------------
// Some hosts of this library may not allocate device memory at the moment they need a scan: a solver that
// handed every byte of the GPU to its own pool at start-up, a service whose steady-state loop must not hit
// cudaMalloc because it synchronises the whole device and can fail late, or a plug-in that is only given
// input and output buffers by its framework.  The regular entry point keeps its look-back descriptors in
// the context workspace, which is cudaMalloc'ed on first use and again whenever a larger input arrives.
// This variant keeps them in a fixed array inside the library image instead, so nothing is ever allocated;
// the price is a hard size limit and ~0.5 MiB of device memory held by every GPU that loads the library.
namespace {
constexpr int64_t STATIC_SCAN_MAX_N = (int64_t)1 << 28;  // elements
using StaticScanPolicy = PlainPolicy<uint32_t, false, false>;
constexpr int STATIC_SCAN_THREADS = 256, STATIC_SCAN_VECS = 4;
constexpr int64_t STATIC_SCAN_TILE = (int64_t)STATIC_SCAN_THREADS * STATIC_SCAN_VECS * StaticScanPolicy::IPV;
constexpr int64_t STATIC_SCAN_MAX_TILES = (STATIC_SCAN_MAX_N + STATIC_SCAN_TILE - 1) / STATIC_SCAN_TILE;
constexpr size_t STATIC_SCAN_BYTES = 256 + (size_t)STATIC_SCAN_MAX_TILES * 8;  // ticket + one 64-bit descriptor per tile

// In host code this name is only a handle for the runtime; the address a kernel may use comes from
// cudaGetSymbolAddress, asked once per device with that device current.
static __device__ __align__(256) unsigned char g_static_scan_scratch[STATIC_SCAN_BYTES];
static void *g_static_scan_ptr[64] = {nullptr};
}  // namespace

CME213_API int64_t cme213_exclusive_scan_u32_static_max_n(void) { return STATIC_SCAN_MAX_N; }

CME213_API int cme213_exclusive_scan_u32_static(const uint32_t *d_in, uint32_t *d_out, int64_t n, void *stream) {
    if (n < 0) { set_error("exclusive_scan_u32_static: negative size"); return CME213_ERR_INVALID; }
    if (n > STATIC_SCAN_MAX_N) {
        set_error("exclusive_scan_u32_static: n = %lld exceeds the built-in scratch (max %lld elements); use "
                  "cme213_exclusive_scan_u32 for larger inputs", (long long)n, (long long)STATIC_SCAN_MAX_N);
        return CME213_ERR_UNSUPPORTED;
    }
    if (n == 0) return CME213_OK;
    if (!d_in || !d_out) { set_error("exclusive_scan_u32_static: null pointer"); return CME213_ERR_INVALID; }
    int dev = -1;
    CME213_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) { set_error("exclusive_scan_u32_static: device index %d not supported", dev); return CME213_ERR_UNSUPPORTED; }
    // The scratch array used is the CURRENT device's copy, so the buffers have to live there too (a query, not an
    // allocation): a buffer of another GPU would otherwise be scanned by a kernel running on the wrong device.
    for (const void *p : {(const void *)d_in, (const void *)d_out}) {
        cudaPointerAttributes at;
        cudaError_t qe = cudaPointerGetAttributes(&at, p);
        if (qe != cudaSuccess) { cudaGetLastError(); return cuda_fail(qe, "cudaPointerGetAttributes"); }
        if (at.type == cudaMemoryTypeDevice && at.device != dev) {
            set_error("exclusive_scan_u32_static: buffer %p lives on device %d but device %d is current; make the "
                      "buffers' device current before the call", p, at.device, dev);
            return CME213_ERR_INVALID;
        }
        if (at.type == cudaMemoryTypeUnregistered) {
            set_error("exclusive_scan_u32_static: %p is not a device pointer", p);
            return CME213_ERR_INVALID;
        }
    }
    if (!g_static_scan_ptr[dev]) {
        // set-up, first call on this device: resolve the device address of the array
        void *ptr = nullptr;
        CME213_CUDA(cudaGetSymbolAddress(&ptr, g_static_scan_scratch));
        g_static_scan_ptr[dev] = ptr;
    }
    cudaStream_t st = (cudaStream_t)stream;
    const int64_t tiles = div_up(n, STATIC_SCAN_TILE);
    // the array still holds the previous call's descriptors and ticket: everything this call reads must start at zero
    CME213_CUDA(cudaMemsetAsync(g_static_scan_ptr[dev], 0, scan::scratch_bytes(tiles, 4), st));
    scan::Scratch s = scan::scratch_carve(g_static_scan_ptr[dev], tiles, 4);
    StaticScanPolicy p{d_in, d_out, nullptr};
    scan::lookback_scan_kernel<StaticScanPolicy, STATIC_SCAN_THREADS, STATIC_SCAN_VECS, 6>
        <<<(unsigned)tiles, STATIC_SCAN_THREADS, 0, st>>>(p, n, s, (aligned16(d_in) && aligned16(d_out)) ? 1 : 0);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_fail(e, "exclusive_scan_u32_static");
    return CME213_OK;
}

// ---- internal: exclusive segmented uint32 scan with caller-provided scratch (convert.cu: radix block offsets) ----
size_t cme213_internal_segscan_u32_scratch_bytes(int64_t n) {
    return scan::scratch_bytes_for<PlainPolicy<uint32_t, true, false>>(n);
}
int cme213_internal_segscan_exclusive_u32(cme213_ctx *ctx, const uint32_t *d_in, const uint32_t *d_flag_bits,
                                          uint32_t *d_out, int64_t n, void *scratch, void *stream) {
    if (n <= 0) return CME213_OK;
    DeviceGuard g(ctx->device);
    PlainPolicy<uint32_t, true, false> p{d_in, d_out, d_flag_bits};
    return scan::launch_with_scratch(ctx, p, n, aligned16(d_in) && aligned16(d_out) && aligned16(d_flag_bits),
                                     (cudaStream_t)stream, "segscan_exclusive_u32", scratch,
                                     cme213_internal_segscan_u32_scratch_bytes(n));
}
