// lookback_scan.cuh -- single-pass (segmented) prefix sum with decoupled look-back.
//
// One kernel serves every scan on the hot path; a policy object supplies how a 16-byte vector of
// items is produced (plain load, a[i]*x[k[i]] gather-multiply, COO product with row-change head
// flags) and consumed (exclusive / inclusive store, COO row-end accumulate into y).
//
// Structure of one tile (256 threads, 4 vectors of 16 bytes per thread):
//   1. the tile index comes from an atomic ticket, so a tile can only wait on tiles that have
//      already started (forward progress does not depend on blockIdx scheduling order);
//   2. sub-tile i (i = 0..3) is the i-th run of 256 consecutive vectors: thread t owns vector
//      i*256 + t of the tile, so every LDG.128/STG.128 of a warp covers 512 contiguous bytes;
//   3. each thread reduces its vectors, each warp scans its 32 per-thread aggregates with
//      shuffles (4 independent chains), and the 8 warps x 4 sub-tiles = 32 warp aggregates are
//      scanned by one more warp-wide shuffle scan (done redundantly by every warp: one barrier);
//   4. warp 0 publishes the tile aggregate, walks back over predecessor descriptors 32 at a time
//      until it meets an inclusive prefix (or, for segmented scans, a tile that contains a
//      segment head), publishes the tile's inclusive prefix and hands the exclusive prefix to
//      the other warps;
//   5. every thread finishes its items in registers and stores them.
// Each element is read once and written once; descriptors add 8-20 bytes per 4096 elements.
//
// Background reading in the reference tree: slides/Lecture16.pdf slides 19-27 (warp scan,
// block scan), slides/Lecture21.pdf (head-flag segmented scan), my-refs/nvr-2008-003.pdf
// Figs. 3-10 (segscan_warp / segscan_block).  The look-back chaining itself is not in the
// reference; its multi-level up-sweep/down-sweep costs a second pass over the data.
#pragma once
#include "common.cuh"

namespace cme213 {
namespace scan {

constexpr int THREADS = 256;
constexpr int WARPS = THREADS / 32;
constexpr int VECS = 4;  // 16-byte vectors per thread; WARPS * VECS must be 32

constexpr unsigned ST_INVALID = 0, ST_AGG = 1, ST_PREFIX = 2;

// ---- scratch layout --------------------------------------------------------------------------
// [0,256)   : ticket counter (first 4 bytes)
// [256, ..) : descriptors.  4-byte value types pack {value, flag, status} into one 64-bit word
//             so a single relaxed store publishes it.  8-byte value types use a status word plus
//             separate aggregate and prefix slots (written once each, published with a release
//             store of the status word, read after an acquire load), so nothing is ever torn.
struct Scratch {
    unsigned *ticket;
    void *desc;       // u64[tiles] (packed) or u32 status[tiles_padded]
    void *agg;        // T[tiles]   (split only)
    void *prefix;     // T[tiles]   (split only)
};

static inline size_t scratch_bytes(int64_t tiles, int value_bytes) {
    size_t t = (size_t)((tiles + 63) & ~(int64_t)63);
    return 256 + (value_bytes == 4 ? t * 8 : t * 4 + t * 8 * 2);
}
static inline Scratch scratch_carve(void *base, int64_t tiles, int value_bytes) {
    size_t t = (size_t)((tiles + 63) & ~(int64_t)63);
    Scratch s;
    char *p = (char *)base;
    s.ticket = (unsigned *)p;
    s.desc = p + 256;
    s.agg = value_bytes == 4 ? nullptr : p + 256 + t * 4;
    s.prefix = value_bytes == 4 ? nullptr : p + 256 + t * 4 + t * 8;
    return s;
}

#ifdef __CUDACC__
__device__ __forceinline__ unsigned long long ld_relaxed_u64(const unsigned long long *p) {
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_u64(unsigned long long *p, unsigned long long v) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned *p) {
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_u32(unsigned *p, unsigned v) {
    asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

template <typename T, int BYTES = sizeof(T)> struct Desc;

// 4-byte values: one 64-bit word = value | flag<<32 | status<<33.
template <typename T> struct Desc<T, 4> {
    __device__ static void publish(const Scratch &s, int64_t tile, unsigned status, T v, unsigned f) {
        unsigned bits;
        memcpy(&bits, &v, 4);
        st_relaxed_u64((unsigned long long *)s.desc + tile,
                       (unsigned long long)bits | ((unsigned long long)(f & 1u) << 32) | ((unsigned long long)status << 33));
    }
    __device__ static void read(const Scratch &s, int64_t tile, unsigned &status, T &v, unsigned &f) {
        unsigned long long d = ld_relaxed_u64((const unsigned long long *)s.desc + tile);
        unsigned bits = (unsigned)d;
        memcpy(&v, &bits, 4);
        f = (unsigned)(d >> 32) & 1u;
        status = (unsigned)(d >> 33);
    }
};
// 8-byte values: status word (status | flag<<2) + write-once value slots.
template <typename T> struct Desc<T, 8> {
    __device__ static void publish(const Scratch &s, int64_t tile, unsigned status, T v, unsigned f) {
        unsigned long long bits;
        memcpy(&bits, &v, 8);
        unsigned long long *slot = (unsigned long long *)(status == ST_AGG ? s.agg : s.prefix) + tile;
        st_relaxed_u64(slot, bits);
        st_release_u32((unsigned *)s.desc + tile, status | ((f & 1u) << 2));
    }
    __device__ static void read(const Scratch &s, int64_t tile, unsigned &status, T &v, unsigned &f) {
        unsigned w = ld_acquire_u32((const unsigned *)s.desc + tile);
        status = w & 3u;
        f = (w >> 2) & 1u;
        v = T(0);
        if (status != ST_INVALID) {
            unsigned long long bits =
                ld_relaxed_u64((const unsigned long long *)(status == ST_AGG ? s.agg : s.prefix) + tile);
            memcpy(&v, &bits, 8);
        }
    }
};

// Inclusive warp scan of one value per lane; for segmented scans `lo` is the nearest lane <= this
// one whose aggregate restarts a segment (-1 if none): contributions from below `lo` are dropped.
template <typename T, bool SEG>
__device__ __forceinline__ T warp_scan_inclusive(T x, int lane, int lo) {
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        T y = __shfl_up_sync(0xFFFFFFFFu, x, d);
        if (lane >= d && (!SEG || lane - d >= lo)) x = y + x;
    }
    return x;
}

// Policy concept:
//   using T; using Aux (per-vector state carried from load to store, may be empty);
//   static constexpr int IPV (items per 16-byte vector); static constexpr bool SEG;
//   __device__ void load(int64_t idx, int64_t n, bool fast, T (&v)[IPV], unsigned &head, Aux &aux) const;
//   __device__ void store(int64_t idx, int64_t n, bool fast, const T (&incl)[IPV], const T (&excl)[IPV],
//                         const Aux &aux) const;
// `idx` is the index of the first item of the vector (a multiple of IPV); `fast` means the whole
// tile is in range and vector accesses are allowed.  Out-of-range items must load as 0 / no flag.
template <class P>
__global__ void __launch_bounds__(THREADS)
lookback_scan_kernel(P p, int64_t n, Scratch scratch, int vec_ok) {
    using T = typename P::T;
    constexpr int IPV = P::IPV;
    constexpr bool SEG = P::SEG;
    constexpr int64_t TILE_ITEMS = (int64_t)THREADS * VECS * IPV;
    static_assert(WARPS * VECS == 32, "warp aggregates must fill exactly one warp");

    __shared__ unsigned s_tile;
    __shared__ T s_wagg_v[32];
    __shared__ unsigned s_wagg_f[32];
    __shared__ T s_carry;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_tile = atomicAdd(scratch.ticket, 1u);
    __syncthreads();
    const int64_t tile = s_tile;
    const int64_t tile_base = tile * TILE_ITEMS;
    const bool fast = vec_ok && (tile_base + TILE_ITEMS <= n);

    // ---- 1. load -------------------------------------------------------------------------
    T v[VECS][IPV];
    unsigned hf[VECS];
    typename P::Aux aux[VECS];
#pragma unroll
    for (int i = 0; i < VECS; i++) {
        const int64_t idx = tile_base + ((int64_t)i * THREADS + tid) * IPV;
        p.load(idx, n, fast, v[i], hf[i], aux[i]);
    }
    // ---- 2. per-thread aggregates, warp scans -------------------------------------------------
    T wv[VECS];          // inclusive scan of thread aggregates within the warp
    unsigned wm[VECS];   // ballot of "this thread's vector contains a head"
#pragma unroll
    for (int i = 0; i < VECS; i++) {
        T a = v[i][0];
#pragma unroll
        for (int j = 1; j < IPV; j++) a = (SEG && ((hf[i] >> j) & 1u)) ? v[i][j] : a + v[i][j];
        const unsigned m = SEG ? __ballot_sync(0xFFFFFFFFu, hf[i] != 0) : 0u;
        const int lo = SEG ? 31 - __clz(m & (0xFFFFFFFFu >> (31 - lane))) : -1;
        wv[i] = warp_scan_inclusive<T, SEG>(a, lane, lo);
        wm[i] = m;
        if (lane == 31) {
            s_wagg_v[i * WARPS + warp] = wv[i];
            s_wagg_f[i * WARPS + warp] = (m != 0);
        }
    }
    __syncthreads();
    // ---- 3. scan of the 32 warp aggregates (every warp, redundantly) --------------------------
    const T gv = s_wagg_v[lane];
    const unsigned gm = SEG ? __ballot_sync(0xFFFFFFFFu, s_wagg_f[lane] != 0) : 0u;
    const int glo = SEG ? 31 - __clz(gm & (0xFFFFFFFFu >> (31 - lane))) : -1;
    const T ginc = warp_scan_inclusive<T, SEG>(gv, lane, glo);
    const T tile_agg = __shfl_sync(0xFFFFFFFFu, ginc, 31);
    const unsigned tile_f = (gm != 0);

    // ---- 4. decoupled look-back (warp 0) ------------------------------------------------------
    if (warp == 0) {
        T carry = T(0);
        if (tile == 0) {
            if (lane == 0) Desc<T>::publish(scratch, 0, ST_PREFIX, tile_agg, tile_f);
        } else {
            if (lane == 0) Desc<T>::publish(scratch, tile, ST_AGG, tile_agg, tile_f);
            int64_t base = tile - 1;
            while (true) {
                const int64_t pred = base - lane;
                unsigned st;
                T dv;
                unsigned df;
                do {
                    if (pred >= 0) {
                        Desc<T>::read(scratch, pred, st, dv, df);
                    } else {
                        st = ST_PREFIX; dv = T(0); df = 0;
                    }
                } while (__any_sync(0xFFFFFFFFu, st == ST_INVALID));
                // a predecessor ends the walk if it carries a full prefix, or (segmented) if it
                // contains a segment head: nothing older can reach this tile.
                const unsigned term = __ballot_sync(0xFFFFFFFFu, st == ST_PREFIX || (SEG && df));
                const int first = term ? __ffs(term) - 1 : 31;
                T c = lane <= first ? dv : T(0);
                // lanes 0..first hold tiles base..base-first (newest first); inside that window
                // only the oldest may contain a head, so a plain tree sum (older on the left) is
                // the segmented sum.
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) {
                    T o = __shfl_down_sync(0xFFFFFFFFu, c, d);
                    if (lane + d < 32) c = o + c;
                }
                carry = c + carry;
                if (term) break;
                base -= 32;
            }
            carry = __shfl_sync(0xFFFFFFFFu, carry, 0);
            if (lane == 0) {
                const T inc = (SEG && tile_f) ? tile_agg : carry + tile_agg;
                Desc<T>::publish(scratch, tile, ST_PREFIX, inc, tile_f);
            }
        }
        if (lane == 0) s_carry = carry;
    }
    __syncthreads();
    const T carry_tile = s_carry;

    // ---- 5. finish in registers and store -----------------------------------------------------
#pragma unroll
    for (int i = 0; i < VECS; i++) {
        const int g = i * WARPS + warp;  // position of this warp's aggregate in the tile order
        // exclusive prefix over the warp aggregates before g, and whether any of them had a head
        T e = __shfl_sync(0xFFFFFFFFu, ginc, g > 0 ? g - 1 : 0);
        if (g == 0) e = T(0);
        const bool head_before_warp = SEG && (gm & ((1u << g) - 1u)) != 0;
        // exclusive prefix over the lanes before this one inside the warp
        T w = __shfl_up_sync(0xFFFFFFFFu, wv[i], 1);
        if (lane == 0) w = T(0);
        const bool head_before_lane = SEG && (wm[i] & ((1u << lane) - 1u)) != 0;
        T c = w;
        if (!head_before_lane) {
            c = e + c;
            if (!head_before_warp) c = carry_tile + c;
        }
        T incl[IPV], excl[IPV];
#pragma unroll
        for (int j = 0; j < IPV; j++) {
            if (SEG && ((hf[i] >> j) & 1u)) c = T(0);
            excl[j] = c;
            c = c + v[i][j];
            incl[j] = c;
        }
        const int64_t idx = tile_base + ((int64_t)i * THREADS + tid) * IPV;
        p.store(idx, n, fast, incl, excl, aux[i]);
    }
}
#endif  // __CUDACC__

template <class P> static inline int64_t tile_items() { return (int64_t)THREADS * VECS * P::IPV; }

// Host launcher: clear + carve the scratch, launch one CTA per tile.  `scratch` must hold
// scratch_bytes(tiles, sizeof(T)) bytes; launch() takes it from the context workspace.
template <class P>
int launch_with_scratch(cme213_ctx *ctx, const P &p, int64_t n, bool vec_ok, cudaStream_t st, const char *name,
                        void *scratch_mem) {
    using T = typename P::T;
    const int64_t tiles = div_up(n, tile_items<P>());
    if (tiles > 0x7FFFFFFFLL) { set_error("%s: n too large", name); return CME213_ERR_UNSUPPORTED; }
    const size_t bytes = scratch_bytes(tiles, sizeof(T));
    CME213_CUDA(cudaMemsetAsync(scratch_mem, 0, bytes, st));
    Scratch s = scratch_carve(scratch_mem, tiles, sizeof(T));
#ifdef __CUDACC__
    lookback_scan_kernel<P><<<(unsigned)tiles, THREADS, 0, st>>>(p, n, s, vec_ok ? 1 : 0);
#endif
    CME213_LAUNCHED(ctx, name);
    return CME213_OK;
}

template <class P>
int launch(cme213_ctx *ctx, const P &p, int64_t n, bool vec_ok, cudaStream_t st, const char *name) {
    const int64_t tiles = div_up(n, tile_items<P>());
    if (tiles > 0x7FFFFFFFLL) { set_error("%s: n too large", name); return CME213_ERR_UNSUPPORTED; }
    void *ws = nullptr;
    int rc = workspace(ctx, scratch_bytes(tiles, sizeof(typename P::T)), &ws);
    if (rc) return rc;
    return launch_with_scratch<P>(ctx, p, n, vec_ok, st, name, ws);
}

}  // namespace scan
}  // namespace cme213
