// lookback_scan.cuh -- single-pass (segmented) prefix sum with decoupled look-back.
//
// One kernel family serves every scan on the hot path; a policy object supplies how a 16-byte
// vector of items is produced (plain load, a[i]*x[k[i]] gather-multiply, COO product with
// row-change head flags) and consumed (exclusive / inclusive store, COO row-end accumulate).
//
// Tile = THREADS threads x VECS vectors of 16 bytes.  Sub-tile i is the i-th run of THREADS
// consecutive vectors: thread t owns vector i*THREADS + t, so a warp's 128-bit accesses cover 512
// contiguous bytes.  Per tile:
//   1. the tile index comes from an atomic ticket, so a tile can only wait on tiles that have
//      already started (forward progress does not depend on blockIdx scheduling order);
//   2. each thread reduces its vectors, each warp scans its 32 thread aggregates with shuffles
//      (VECS independent chains) and parks its VECS warp aggregates in shared memory;
//   3. warp 0 scans the WARPS*VECS warp aggregates, publishes the tile aggregate, walks back over
//      predecessor descriptors 32 at a time until it meets an inclusive prefix (or, for segmented
//      scans, a tile that contains a segment head), publishes the tile's inclusive prefix and hands
//      the exclusive prefix to the other warps;
//   4. every thread finishes its items in registers and stores them with 128-bit stores.
// Each element is read once and written once; descriptors add 8-20 bytes per tile.
//
// Two front ends share that core:
//   * lookback_scan_bulk_kernel (the B200 path): a persistent CTA keeps a ring of STAGES tiles in
//     shared memory, filled by 1-D bulk async copies (cp.async.bulk + mbarrier complete_tx, the TMA
//     engine's linear mode; SASS UBLKCP).  The loads of the next tiles are in flight while warp 0
//     sits in the look-back of the current one, so the bytes in flight per SM no longer depend on
//     the look-back latency -- which is what held the first version of this kernel to half of HBM
//     bandwidth (profiles/r01_notes.md).
//   * lookback_scan_kernel: one CTA per tile, direct global loads.  Any alignment, any size; used
//     for unaligned pointers and small inputs.
//
// Background reading in the reference tree: slides/Lecture16.pdf slides 19-27 (warp scan,
// block scan), slides/Lecture21.pdf (head-flag segmented scan), my-refs/nvr-2008-003.pdf
// Figs. 3-10 (segscan_warp / segscan_block).  The look-back chaining itself is not in the
// reference; its multi-level up-sweep/down-sweep costs a second pass over the data.
#pragma once
#include "common.cuh"
#include <stdlib.h>

namespace cme213 {
namespace scan {

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

// (mbarrier / bulk-copy primitives: common.cuh)

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

// Decoupled look-back of `tile` (> 0) by one warp.  Each lane inspects DPL predecessor descriptors per
// round (a window of 32*DPL tiles, nearest predecessor first: lane l holds tiles tile-1-l*DPL-q), spins
// until all of them are published, and the warp sums everything up to and including the nearest
// terminator: a tile that already carries its inclusive prefix, or (segmented) one that contains a
// segment head -- nothing older can reach this tile.  Returns the exclusive prefix in every lane.
template <typename T, bool SEG, int DPL>
__device__ __forceinline__ T lookback(const Scratch &scratch, int64_t tile, int lane) {
    T carry = T(0);
    int64_t base = tile - 1;
    while (true) {
        unsigned st[DPL], df[DPL];
        T dv[DPL];
        bool invalid;
        do {
            invalid = false;
#pragma unroll
            for (int q = 0; q < DPL; q++) {
                const int64_t pred = base - (int64_t)lane * DPL - q;
                if (pred >= 0) {
                    Desc<T>::read(scratch, pred, st[q], dv[q], df[q]);
                } else {
                    st[q] = ST_PREFIX; dv[q] = T(0); df[q] = 0;
                }
                invalid |= st[q] == ST_INVALID;
            }
        } while (__any_sync(0xFFFFFFFFu, invalid));
        // lane-local: nearest first, older values go on the left of the sum
        T c = T(0);
        bool found = false;
#pragma unroll
        for (int q = 0; q < DPL; q++) {
            if (!found) {
                c = dv[q] + c;
                found = st[q] == ST_PREFIX || (SEG && df[q]);
            }
        }
        const unsigned term = __ballot_sync(0xFFFFFFFFu, found);
        const int first = term ? __ffs(term) - 1 : 31;
        if (lane > first) c = T(0);
        // inside the window only the terminator itself may contain a head, so a plain tree sum with
        // the older lanes on the left is the segmented sum
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            T o = __shfl_down_sync(0xFFFFFFFFu, c, d);
            if (lane + d < 32) c = o + c;
        }
        carry = c + carry;
        if (term) break;
        base -= 32 * DPL;
    }
    return __shfl_sync(0xFFFFFFFFu, carry, 0);
}

// Policy concept:
//   using T; using Aux (per-vector state carried from load to store, may be empty);
//   static constexpr int IPV (items per 16-byte vector); static constexpr bool SEG;
//   static constexpr bool BULK; whether large inputs take the persistent bulk-copy kernel (gather-heavy
//       policies hide their L2 latency better with many small CTAs per SM, i.e. the direct kernel)
//   static constexpr int NIN;  number of input arrays streamed through shared memory by the bulk kernel
//   static constexpr int tile_bytes(int s, int tile_items);  bytes of input stream s per tile (a multiple of 16)
//   const void *in_ptr(int s) const;       its base pointer
//   __device__ void load(int64_t idx, int64_t n, bool fast, T (&v)[IPV], unsigned &head, Aux &aux) const;
//       from global memory; `fast` = whole tile in range and 16-byte vector accesses allowed
//   __device__ void load_smem(const unsigned char *const (&in)[NIN], int vec, int nvec, int64_t idx, int64_t n,
//                             T (&v)[IPV], unsigned &head, Aux &aux) const;
//       from the staged copies: in[s] is stream s of this tile, `vec` the vector's index in the tile
//   __device__ void store(int64_t idx, int64_t n, bool fast, const T (&incl)[IPV], const T (&excl)[IPV],
//                         const Aux &aux) const;
// `idx` is the index of the first item of the vector (a multiple of IPV).  Out-of-range items must
// load as 0 / no flag.

// Shared state of the per-tile core.
template <class P, int THREADS, int VECS> struct CoreSmem {
    static constexpr int NAGG = (THREADS / 32) * VECS;
    typename P::T agg_v[NAGG];
    typename P::T excl_v[NAGG];       // exclusive prefix over aggregates [0, g) since the last head
    typename P::T carry;
    unsigned char agg_f[NAGG];
    unsigned char excl_f[NAGG];       // any head among aggregates [0, g)
};

// One warp scans 32*PL (flag, value) entries held PL per lane in order (lane l owns entries
// l*PL .. l*PL+PL-1).  hb bit k set = entry k starts (or contains) a segment head.  Outputs, per
// entry: the exclusive prefix since the last head, and whether any head precedes it; plus the
// inclusive total since the last head and whether any head exists at all (same in every lane).
template <typename T, bool SEG, int PL>
__device__ __forceinline__ void warp_array_scan(const T (&a)[PL], unsigned hb, int lane, T (&excl)[PL],
                                                unsigned &head_before, T &total, bool &any_head) {
    T inc[PL];
    unsigned fb = 0;  // bit k: a head among this lane's entries [0, k]
    T run = T(0);
#pragma unroll
    for (int k = 0; k < PL; k++) {
        const bool f = SEG && ((hb >> k) & 1u);
        run = (k == 0 || f) ? a[k] : run + a[k];
        if (f) fb |= 0xFFFFFFFFu << k;
        inc[k] = run;
    }
    const unsigned m = SEG ? __ballot_sync(0xFFFFFFFFu, fb != 0) : 0u;
    const int lo = SEG ? 31 - __clz(m & (0xFFFFFFFFu >> (31 - lane))) : -1;
    const T tot = warp_scan_inclusive<T, SEG>(run, lane, lo);
    total = __shfl_sync(0xFFFFFFFFu, tot, 31);
    any_head = m != 0;
    T prev = __shfl_up_sync(0xFFFFFFFFu, tot, 1);
    if (lane == 0) prev = T(0);
    const bool head_prev = SEG && (m & ((1u << lane) - 1u)) != 0;
    head_before = 0;
#pragma unroll
    for (int k = 0; k < PL; k++) {
        const bool head_in_lane = k > 0 && ((fb >> (k - 1)) & 1u);
        T e = prev;
        if (k > 0) e = head_in_lane ? inc[k - 1] : prev + inc[k - 1];
        excl[k] = e;
        if (head_prev || head_in_lane) head_before |= 1u << k;
    }
}

// Per-tile core, split in two so that a front end can put other work between the parts.
//
// tile_local: per-thread aggregates, warp scans (inclusive results in wv, head ballots in wm), first
// barrier, `between()` on every thread (the bulk kernel refills its ring there), then warp 0 scans the
// warp aggregates, leaves their exclusive prefixes in shared memory and calls
// `after_agg(tile_aggregate, tile_has_head)` (all 32 lanes of warp 0): the front ends publish the
// aggregate and obtain tile prefixes there.  No trailing barrier.
template <class P, int THREADS, int VECS, class Between, class AfterAgg>
__device__ __forceinline__ void tile_local(CoreSmem<P, THREADS, VECS> &sm, const typename P::T (&v)[VECS][P::IPV],
                                           const unsigned (&hf)[VECS], typename P::T (&wv)[VECS], unsigned (&wm)[VECS],
                                           Between between, AfterAgg after_agg) {
    using T = typename P::T;
    constexpr int IPV = P::IPV;
    constexpr bool SEG = P::SEG;
    constexpr int WARPS = THREADS / 32;
    constexpr int NAGG = WARPS * VECS;
    constexpr int PL = NAGG / 32;
    static_assert(NAGG % 32 == 0, "warp aggregates must fill whole lanes of one warp");
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
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
            sm.agg_v[i * WARPS + warp] = wv[i];
            sm.agg_f[i * WARPS + warp] = (unsigned char)(m != 0);
        }
    }
    __syncthreads();
    between();
    if (warp == 0) {
        T a[PL], e[PL];
        unsigned hb = 0, before;
#pragma unroll
        for (int k = 0; k < PL; k++) {
            a[k] = sm.agg_v[lane * PL + k];
            if (SEG && sm.agg_f[lane * PL + k]) hb |= 1u << k;
        }
        T tile_agg;
        bool tile_f;
        warp_array_scan<T, SEG, PL>(a, hb, lane, e, before, tile_agg, tile_f);
#pragma unroll
        for (int k = 0; k < PL; k++) {
            sm.excl_v[lane * PL + k] = e[k];
            sm.excl_f[lane * PL + k] = (unsigned char)((before >> k) & 1u);
        }
        after_agg(tile_agg, tile_f);
    }
}

// tile_cst (after a barrier that follows tile_local): cst[i] = what precedes vector i's first item inside
// the tile (since the last head); bit i of `apply` = no head precedes it in the tile, so the tile's
// exclusive prefix still has to be added.
template <class P, int THREADS, int VECS>
__device__ __forceinline__ void tile_cst(const CoreSmem<P, THREADS, VECS> &sm, const typename P::T (&wv)[VECS],
                                         const unsigned (&wm)[VECS], typename P::T (&cst)[VECS], unsigned &apply) {
    using T = typename P::T;
    constexpr bool SEG = P::SEG;
    constexpr int WARPS = THREADS / 32;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    apply = 0;
#pragma unroll
    for (int i = 0; i < VECS; i++) {
        const int g = i * WARPS + warp;
        T w = __shfl_up_sync(0xFFFFFFFFu, wv[i], 1);
        if (lane == 0) w = T(0);
        const bool head_before_lane = SEG && (wm[i] & ((1u << lane) - 1u)) != 0;
        T c = w;
        if (!head_before_lane) {
            c = sm.excl_v[g] + c;
            if (!(SEG && sm.excl_f[g])) apply |= 1u << i;
        }
        cst[i] = c;
    }
}

template <class P, int VECS>
__device__ __forceinline__ void finish_and_store(const P &p, int64_t tile_base, int threads, int64_t n, bool fast,
                                                 const typename P::T (&v)[VECS][P::IPV], const unsigned (&hf)[VECS],
                                                 const typename P::T (&cst)[VECS], unsigned apply, typename P::T carry_tile,
                                                 const typename P::Aux (&aux)[VECS]) {
    using T = typename P::T;
    constexpr int IPV = P::IPV;
#pragma unroll
    for (int i = 0; i < VECS; i++) {
        T c = ((apply >> i) & 1u) ? carry_tile + cst[i] : cst[i];
        T incl[IPV], excl[IPV];
#pragma unroll
        for (int j = 0; j < IPV; j++) {
            if (P::SEG && ((hf[i] >> j) & 1u)) c = T(0);
            excl[j] = c;
            c = c + v[i][j];
            incl[j] = c;
        }
        const int64_t idx = tile_base + ((int64_t)i * threads + threadIdx.x) * IPV;
        p.store(idx, n, fast, incl, excl, aux[i]);
    }
}

// ---- front end 1: one CTA per tile, direct loads, decoupled look-back ----------------------------------
template <class P, int THREADS, int VECS, int MIN_BLOCKS>
__global__ void __launch_bounds__(THREADS, MIN_BLOCKS)
lookback_scan_kernel(P p, int64_t n, Scratch scratch, int vec_ok) {
    using T = typename P::T;
    constexpr int IPV = P::IPV;
    constexpr bool SEG = P::SEG;
    constexpr int64_t TILE_ITEMS = (int64_t)THREADS * VECS * IPV;
    __shared__ unsigned s_tile;
    __shared__ CoreSmem<P, THREADS, VECS> core;
    const int tid = threadIdx.x, lane = tid & 31;
    if (tid == 0) s_tile = atomicAdd(scratch.ticket, 1u);
    __syncthreads();
    const int64_t tile = s_tile;
    const int64_t tile_base = tile * TILE_ITEMS;
    const bool fast = vec_ok && (tile_base + TILE_ITEMS <= n);
    T v[VECS][IPV];
    unsigned hf[VECS];
    typename P::Aux aux[VECS];
#pragma unroll
    for (int i = 0; i < VECS; i++) p.load(tile_base + ((int64_t)i * THREADS + tid) * IPV, n, fast, v[i], hf[i], aux[i]);
    T wv[VECS], cst[VECS];
    unsigned wm[VECS], apply;
    tile_local<P, THREADS, VECS>(core, v, hf, wv, wm, [] {}, [&](T tile_agg, bool tile_f) {
        if (lane == 0) Desc<T>::publish(scratch, tile, tile == 0 ? ST_PREFIX : ST_AGG, tile_agg, tile_f);
        T carry = T(0);
        if (tile != 0) {
            carry = lookback<T, SEG, 1>(scratch, tile, lane);
            if (lane == 0) Desc<T>::publish(scratch, tile, ST_PREFIX, (SEG && tile_f) ? tile_agg : carry + tile_agg, tile_f);
        }
        if (lane == 0) core.carry = carry;
    });
    __syncthreads();
    tile_cst<P, THREADS, VECS>(core, wv, wm, cst, apply);
    finish_and_store<P, VECS>(p, tile_base, THREADS, n, fast, v, hf, cst, apply, core.carry, aux);
}

// ---- front end 2: persistent CTAs, ring of bulk-async-copied tiles, one scanner per wave ----------------
template <class P, int THREADS, int VECS> struct BulkLayout {
    static constexpr int TILE_VECS = THREADS * VECS;
    static constexpr int TILE_ITEMS = TILE_VECS * P::IPV;
    __host__ __device__ static constexpr int stream_off(int s) {
        int o = 0;
        for (int i = 0; i < s; i++) o += P::tile_bytes(i, TILE_ITEMS);
        return o;
    }
    static constexpr int STAGE_BYTES = stream_off(P::NIN);
    __host__ __device__ static constexpr bool aligned() {
        for (int i = 0; i < P::NIN; i++)
            if (P::tile_bytes(i, TILE_ITEMS) % 16 != 0) return false;
        return true;
    }
    static_assert(aligned(), "every staged stream must be a whole number of 16-byte units per tile (bulk copy granularity)");
};

// Cross-CTA state of the wave protocol: slot arrays of {value, flag, ready}.
//   A[tile]  aggregate of the tile (written by its owner)
//   X[tile]  exclusive prefix of the tile (written by the wave's scanner)
//   W[wave]  inclusive prefix at the end of the wave (written by the wave's scanner)
// 4-byte values live in one 64-bit word (value | flag<<32 | ready<<33, one relaxed store publishes
// it); 8-byte values in two such words, one per half.
struct SlotArr {
    unsigned long long *val;
};
struct WaveScratch {
    SlotArr A, X, W;
    unsigned long long *trace;  // debug only (CME213_SCAN_TRACE): per CTA, per wave, 8 globaltimer stamps
};
constexpr int TRACE_MAX_IT = 256;
static inline size_t wave_scratch_bytes(int64_t tiles, int64_t waves, int value_bytes) {
    const size_t t = (size_t)((tiles + 63) & ~(int64_t)63), w = (size_t)((waves + 63) & ~(int64_t)63);
    return (2 * t + w) * (value_bytes == 4 ? 8 : 16);
}
static inline WaveScratch wave_scratch_carve(void *base, int64_t tiles, int64_t waves, int value_bytes) {
    const size_t t = (size_t)((tiles + 63) & ~(int64_t)63);
    const size_t per = value_bytes == 4 ? 1 : 2;
    WaveScratch s;
    s.A.val = (unsigned long long *)base;
    s.X.val = s.A.val + t * per;
    s.W.val = s.X.val + t * per;
    s.trace = nullptr;
    (void)waves;
    return s;
}

template <typename T, int BYTES = sizeof(T)> struct Slot;
template <typename T> struct Slot<T, 4> {
    __device__ static void publish(const SlotArr &s, int64_t i, T v, unsigned f) {
        unsigned bits;
        memcpy(&bits, &v, 4);
        st_relaxed_u64(s.val + i, (unsigned long long)bits | ((unsigned long long)(f & 1u) << 32) | (1ull << 33));
    }
    __device__ static bool read(const SlotArr &s, int64_t i, T &v, unsigned &f) {
        const unsigned long long d = ld_relaxed_u64(s.val + i);
        const unsigned bits = (unsigned)d;
        memcpy(&v, &bits, 4);
        f = (unsigned)(d >> 32) & 1u;
        return (d >> 33) & 1ull;
    }
};
template <typename T> struct Slot<T, 8> {
    // two self-validating words (low half | ready, high half | flag | ready): no ordering between them
    // is needed, so no release store -- a fence here would wait for the CTA's data stores still in flight.
    __device__ static void publish(const SlotArr &s, int64_t i, T v, unsigned f) {
        unsigned long long bits;
        memcpy(&bits, &v, 8);
        st_relaxed_u64(s.val + 2 * i, (bits & 0xFFFFFFFFull) | (1ull << 33));
        st_relaxed_u64(s.val + 2 * i + 1, (bits >> 32) | ((unsigned long long)(f & 1u) << 32) | (1ull << 33));
    }
    __device__ static bool read(const SlotArr &s, int64_t i, T &v, unsigned &f) {
        const unsigned long long lo = ld_relaxed_u64(s.val + 2 * i), hi = ld_relaxed_u64(s.val + 2 * i + 1);
        const unsigned long long bits = (lo & 0xFFFFFFFFull) | (hi << 32);
        memcpy(&v, &bits, 8);
        f = (unsigned)(hi >> 32) & 1u;
        return ((lo >> 33) & (hi >> 33) & 1ull) != 0;
    }
};

// Tiles are dealt round-robin: CTA b handles tiles b, b+G, b+2G, ...  (G = gridDim.x <= 32*DPL), so
// "wave" w = tiles [wG, (w+1)G) is worked on by all CTAs at the same time; they are co-resident
// (cooperative launch).  Every data CTA publishes its tile aggregate; one extra CTA that owns no tiles
// waits for each wave's G aggregates, scans them (one warp, coalesced reads, nobody else reads them),
// adds the running prefix of the previous waves and hands every tile its exclusive prefix.  Each data
// CTA polls only its own slot, one iteration later (the finish of a tile is deferred), so neither the
// scanner's round trips nor a late CTA of the wave stall the others.  All cross-CTA words carry their own ready bit, so no fence or
// atomic is needed (a fence would stall on the CTA's data stores that are still in flight).  Compared with every tile walking back over its
// predecessors' descriptors this keeps the L2 lines holding the descriptors from becoming a
// hot spot: with all tiles of a wave arriving together, 148 warps re-reading the same few lines
// took ~15 us per round trip (profiles/r01_notes.md).
template <class P, int THREADS, int VECS, int STAGES, int DPL>
__global__ void __launch_bounds__(THREADS, (THREADS <= 512 ? 2 : 1))
lookback_scan_bulk_kernel(P p, int64_t n, int64_t num_tiles, WaveScratch ws) {
    using T = typename P::T;
    using L = BulkLayout<P, THREADS, VECS>;
    constexpr int IPV = P::IPV;
    constexpr bool SEG = P::SEG;
    extern __shared__ __align__(128) unsigned char ring[];  // STAGES * STAGE_BYTES
    __shared__ __align__(8) uint64_t full_bar[STAGES];
    __shared__ CoreSmem<P, THREADS, VECS> core;
    const int tid = threadIdx.x, lane = tid & 31;
    const bool filler = tid == THREADS - 32;                 // lane 0 of the last warp feeds the ring
    const int64_t G = gridDim.x - 1;                         // data CTAs; CTA number G only scans

    if ((int64_t)blockIdx.x == G) {
        // The scanner CTA: for every wave, wait for the wave's aggregates (one warp, coalesced reads, nobody
        // else reads them), scan them on top of the running prefix and hand every tile its exclusive prefix.
        // It owns no tiles, so a late wave never delays the data of the next one.
        if (tid >= 32) return;
        T c0 = T(0);  // running prefix (since the last head) at the end of the previous wave
        for (int64_t wave_first = 0; wave_first < num_tiles; wave_first += G) {
            const int64_t wave_cnt = num_tiles - wave_first < G ? num_tiles - wave_first : G;
            T a[DPL], e[DPL];
            unsigned hb = 0, before;
            bool pending;
            do {
                pending = false;
#pragma unroll
                for (int q = 0; q < DPL; q++) {
                    const int64_t j = (int64_t)lane * DPL + q;
                    a[q] = T(0);
                    if (j < wave_cnt) {
                        unsigned f;
                        pending |= !Slot<T>::read(ws.A, wave_first + j, a[q], f);
                        if (SEG && f) hb |= 1u << q; else hb &= ~(1u << q);
                    }
                }
            } while (__any_sync(0xFFFFFFFFu, pending));
            T total;
            bool any;
            warp_array_scan<T, SEG, DPL>(a, hb, lane, e, before, total, any);
#pragma unroll
            for (int q = 0; q < DPL; q++) {
                const int64_t j = (int64_t)lane * DPL + q;
                if (j < wave_cnt) Slot<T>::publish(ws.X, wave_first + j, ((before >> q) & 1u) ? e[q] : c0 + e[q], 0);
            }
            c0 = any ? total : c0 + total;
        }
        return;
    }

    // start the copies of tile t into `stage`
    auto fill = [&](int stage, int64_t t) {
        if (t >= num_tiles) return;
        const int64_t base = t * L::TILE_ITEMS;
        if (base + L::TILE_ITEMS <= n) {
            mbar_expect_tx(&full_bar[stage], (unsigned)L::STAGE_BYTES);
#pragma unroll
            for (int s = 0; s < P::NIN; s++)
                bulk_load(ring + (size_t)stage * L::STAGE_BYTES + L::stream_off(s),
                          (const unsigned char *)p.in_ptr(s) + t * P::tile_bytes(s, L::TILE_ITEMS),
                          (unsigned)P::tile_bytes(s, L::TILE_ITEMS), &full_bar[stage]);
        } else {
            mbar_arrive(&full_bar[stage]);  // ragged last tile: read straight from global with guards
        }
    };

    if (filler) {
        for (int s = 0; s < STAGES; s++) mbar_init(&full_bar[s], 1);
        mbar_fence_init();
        for (int s = 0; s < STAGES; s++) fill(s, (int64_t)s * G + blockIdx.x);
    }
    __syncthreads();

    auto stamp = [&](int it, int k) {
        if (ws.trace && tid == 0 && it < TRACE_MAX_IT) {
            unsigned long long t;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
            ws.trace[((size_t)blockIdx.x * TRACE_MAX_IT + it) * 8 + k] = t;
        }
    };
    // The finish of a tile (add its exclusive prefix, store) is deferred by one iteration: the prefix of
    // tile i is only polled after the local phase of tile i+1, so the wave scanner's round trips overlap
    // with useful work instead of stalling every CTA of the wave.
    T v_p[VECS][IPV], cst_p[VECS];
    unsigned hf_p[VECS], apply_p = 0;
    typename P::Aux aux_p[VECS];
    int64_t tile_p = -1, base_p = 0;
    bool full_p = false;

    int it = 0;
    for (int64_t tile = blockIdx.x; tile < num_tiles; tile += G, it++) {
        const int stage = it % STAGES;
        const int64_t tile_base = tile * L::TILE_ITEMS;
        const bool full = tile_base + L::TILE_ITEMS <= n;
        stamp(it, 0);
        mbar_wait(&full_bar[stage], (unsigned)((it / STAGES) & 1));
        stamp(it, 1);
        T v[VECS][IPV];
        unsigned hf[VECS];
        typename P::Aux aux[VECS];
        if (full) {
            const unsigned char *in[P::NIN];
#pragma unroll
            for (int s = 0; s < P::NIN; s++) in[s] = ring + (size_t)stage * L::STAGE_BYTES + L::stream_off(s);
#pragma unroll
            for (int i = 0; i < VECS; i++) {
                const int vec = i * THREADS + tid;
                p.load_smem(in, vec, L::TILE_VECS, tile_base + (int64_t)vec * IPV, n, v[i], hf[i], aux[i]);
            }
        } else {
#pragma unroll
            for (int i = 0; i < VECS; i++) p.load(tile_base + ((int64_t)i * THREADS + tid) * IPV, n, false, v[i], hf[i], aux[i]);
        }
        T wv[VECS], cst[VECS];
        unsigned wm[VECS], apply;
        tile_local<P, THREADS, VECS>(
            core, v, hf, wv, wm,
            // after the first barrier every thread holds its inputs in registers: the stage is free
            [&] { if (filler) fill(stage, tile + (int64_t)STAGES * G); },
            [&](T tile_agg, bool tile_f) {
                stamp(it, 2);
                if (lane == 0) Slot<T>::publish(ws.A, tile, tile_agg, tile_f);
                stamp(it, 3);
                if (tile_p >= 0) {  // prefix of the PREVIOUS tile: one iteration old, normally there already
                    T carry;
                    unsigned f;
                    while (!Slot<T>::read(ws.X, tile_p, carry, f)) {}
                    if (lane == 0) core.carry = carry;
                }
                stamp(it, 5);
            });
        __syncthreads();
        stamp(it, 6);
        tile_cst<P, THREADS, VECS>(core, wv, wm, cst, apply);
        if (tile_p >= 0) finish_and_store<P, VECS>(p, base_p, THREADS, n, full_p, v_p, hf_p, cst_p, apply_p, core.carry, aux_p);
        stamp(it, 7);
#pragma unroll
        for (int i = 0; i < VECS; i++) {
#pragma unroll
            for (int j = 0; j < IPV; j++) v_p[i][j] = v[i][j];
            hf_p[i] = hf[i]; cst_p[i] = cst[i]; aux_p[i] = aux[i];
        }
        apply_p = apply; tile_p = tile; base_p = tile_base; full_p = full;
    }
    if (tile_p >= 0) {  // drain: the last tile of this CTA
        __syncthreads();  // everybody is done reading core.carry of the previous finish
        if (tid < 32) {
            T carry;
            unsigned f;
            while (!Slot<T>::read(ws.X, tile_p, carry, f)) {}
            if (lane == 0) core.carry = carry;
        }
        __syncthreads();
        finish_and_store<P, VECS>(p, base_p, THREADS, n, full_p, v_p, hf_p, cst_p, apply_p, core.carry, aux_p);
    }
}

// ---- front end 3: as front end 2, but blocked per-thread layout and bulk-copy stores ---------------------
// For policies whose output is one dense array of T (the plain scans).  Differences from front end 2:
//   * a thread owns VECS *consecutive* vectors of the staged tile (odd VECS: its 128-bit shared loads and
//     stores are bank-conflict free), scans them serially and takes part in ONE warp scan per tile instead of
//     one per vector: about a third of the instructions per item;
//   * results go back into the tile's ring slot and leave the SM as one cp.async.bulk store (smem -> global),
//     so neither loads nor stores occupy the LSU or wait in the threads;
//   * ring of STAGES >= 3 slots: [stored last iteration, being refilled] [finishing] [computing] [loading ...].
__device__ __forceinline__ void bulk_store(void *gmem_dst, const void *smem_src, unsigned bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gmem_dst), "r"(smem_u32(smem_src)), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
__device__ __forceinline__ void bulk_store_wait_read_all() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

template <class P, int THREADS, int VECS, int STAGES, int DPL, int DEFER>
__global__ void __launch_bounds__(THREADS, 1)
scan_bulk_blocked_kernel(P p, int64_t n, int64_t num_tiles, WaveScratch ws) {
    using T = typename P::T;
    using L = BulkLayout<P, THREADS, VECS>;
    constexpr int IPV = P::IPV;
    constexpr bool SEG = P::SEG;
    constexpr int WARPS = THREADS / 32;
    constexpr int ITEMS = VECS * IPV;  // per thread, consecutive
    static_assert(WARPS <= 32 && ITEMS <= 32, "one warp scans the warp aggregates; head bits fit a word");
    static_assert(STAGES >= DEFER + 2, "DEFER pending tiles + the one being computed + at least one slot in flight");
    extern __shared__ __align__(128) unsigned char ring[];  // STAGES * STAGE_BYTES
    __shared__ __align__(8) uint64_t full_bar[STAGES];
    __shared__ T s_agg_v[32], s_excl_v[32], s_carry;
    __shared__ unsigned char s_agg_f[32], s_excl_f[32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const bool filler = tid == THREADS - 32;                 // lane 0 of the last warp feeds and drains the ring
    const int64_t G = gridDim.x - 1;                         // data CTAs; CTA number G only scans

    if ((int64_t)blockIdx.x == G) {  // the scanner CTA (see front end 2)
        if (tid >= 32) return;
        T c0 = T(0);
        int wv_i = 0;
        auto sstamp = [&](int k) {
            if (ws.trace && tid == 0 && wv_i < TRACE_MAX_IT) {
                unsigned long long t;
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
                ws.trace[((size_t)G * TRACE_MAX_IT + wv_i) * 8 + k] = t;
            }
        };
        for (int64_t wave_first = 0; wave_first < num_tiles; wave_first += G, wv_i++) {
            const int64_t wave_cnt = num_tiles - wave_first < G ? num_tiles - wave_first : G;
            sstamp(0);
            T a[DPL], e[DPL];
            unsigned hb = 0, before;
            bool pending;
            do {
                pending = false;
#pragma unroll
                for (int q = 0; q < DPL; q++) {
                    const int64_t j = (int64_t)lane * DPL + q;
                    a[q] = T(0);
                    if (j < wave_cnt) {
                        unsigned f;
                        pending |= !Slot<T>::read(ws.A, wave_first + j, a[q], f);
                        if (SEG && f) hb |= 1u << q; else hb &= ~(1u << q);
                    }
                }
            } while (__any_sync(0xFFFFFFFFu, pending));
            sstamp(1);
            T total;
            bool any;
            warp_array_scan<T, SEG, DPL>(a, hb, lane, e, before, total, any);
#pragma unroll
            for (int q = 0; q < DPL; q++) {
                const int64_t j = (int64_t)lane * DPL + q;
                if (j < wave_cnt) Slot<T>::publish(ws.X, wave_first + j, ((before >> q) & 1u) ? e[q] : c0 + e[q], 0);
            }
            c0 = any ? total : c0 + total;
            sstamp(2);
        }
        return;
    }

    auto stage_ptr = [&](int stage) { return ring + (size_t)stage * L::STAGE_BYTES; };
    auto fill = [&](int stage, int64_t t) {
        if (t >= num_tiles) return;
        const int64_t base = t * L::TILE_ITEMS;
        if (base + L::TILE_ITEMS <= n) {
            mbar_expect_tx(&full_bar[stage], (unsigned)L::STAGE_BYTES);
#pragma unroll
            for (int s = 0; s < P::NIN; s++)
                bulk_load(stage_ptr(stage) + L::stream_off(s), (const unsigned char *)p.in_ptr(s) + t * P::tile_bytes(s, L::TILE_ITEMS),
                          (unsigned)P::tile_bytes(s, L::TILE_ITEMS), &full_bar[stage]);
        } else {
            mbar_arrive(&full_bar[stage]);  // ragged last tile: read and written straight from/to global with guards
        }
    };
    auto stamp = [&](int it, int k) {
        if (ws.trace && tid == 0 && it < TRACE_MAX_IT) {
            unsigned long long t;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
            ws.trace[((size_t)blockIdx.x * TRACE_MAX_IT + it) * 8 + k] = t;
        }
    };

    if (filler) {
        for (int s = 0; s < STAGES; s++) mbar_init(&full_bar[s], 1);
        mbar_fence_init();
        for (int s = 0; s < STAGES; s++) fill(s, (int64_t)s * G + blockIdx.x);
    }
    __syncthreads();

    // Tiles whose finish is deferred: the finish of tile i (add its exclusive prefix, write back, store) runs
    // DEFER iterations after its local phase, so the scanner's round trips and a late CTA of the wave are
    // hidden behind useful work.  A pending tile keeps its inputs in its ring slot; only these few values stay
    // in registers.  q[0] is the oldest pending tile.
    struct Pending { T cst; int64_t tile, base; unsigned apply; int stage; };
    Pending q[DEFER];
#pragma unroll
    for (int d = 0; d < DEFER; d++) q[d].tile = -1;

    // re-reads the tile's inputs from its slot, adds the prefix in front of the thread's run, writes the results
    // back to the slot; the caller then hands the slot to the copy engine
    auto finish_slot = [&](const Pending &t, T carry_tile) {
        const unsigned char *in[P::NIN];
#pragma unroll
        for (int s = 0; s < P::NIN; s++) in[s] = stage_ptr(t.stage) + L::stream_off(s);
        T c = t.apply ? carry_tile + t.cst : t.cst;
#pragma unroll
        for (int i = 0; i < VECS; i++) {
            const int vec = tid * VECS + i;
            T v[IPV];
            unsigned h;
            typename P::Aux aux;
            p.load_smem(in, vec, L::TILE_VECS, t.base + (int64_t)vec * IPV, n, v, h, aux);
            typename P::V16T o;
            T *e = reinterpret_cast<T *>(&o);
#pragma unroll
            for (int j = 0; j < IPV; j++) {
                if (SEG && ((h >> j) & 1u)) c = T(0);
                const T ex = c;
                c = c + v[j];
                e[j] = P::INCLUSIVE_OUT ? c : ex;
            }
            *reinterpret_cast<typename P::V16T *>(stage_ptr(t.stage) + (size_t)vec * 16) = o;
        }
        fence_proxy_async_smem();  // make the generic-proxy writes visible to the bulk-copy engine
    };
    auto store_slot = [&](const Pending &t) {
        bulk_store(reinterpret_cast<unsigned char *>(p.out_ptr()) + t.base * (int64_t)sizeof(T), stage_ptr(t.stage),
                   (unsigned)(L::TILE_ITEMS * sizeof(T)));
    };
    auto poll = [&](int64_t t) {  // warp 0: exclusive prefix of tile t into s_carry
        T carry;
        unsigned f;
        while (!Slot<T>::read(ws.X, t, carry, f)) {}
        if (lane == 0) s_carry = carry;
    };

    int it = 0;
    for (int64_t tile = blockIdx.x; tile < num_tiles; tile += G, it++) {
        const int stage = it % STAGES;
        const int64_t tile_base = tile * L::TILE_ITEMS;
        const bool full = tile_base + L::TILE_ITEMS <= n;
        stamp(it, 0);
        // first look at the oldest pending tile's prefix now, so that its L2 round trip overlaps the local phase
        T early_carry = T(0);
        bool early_ok = false;
        if (warp == 0 && q[0].tile >= 0) {
            unsigned f;
            early_ok = Slot<T>::read(ws.X, q[0].tile, early_carry, f);
        }
        mbar_wait(&full_bar[stage], (unsigned)((it / STAGES) & 1));
        stamp(it, 1);
        T v[VECS][IPV];
        unsigned hf = 0;
        if (full) {
            const unsigned char *in[P::NIN];
#pragma unroll
            for (int s = 0; s < P::NIN; s++) in[s] = stage_ptr(stage) + L::stream_off(s);
#pragma unroll
            for (int i = 0; i < VECS; i++) {
                const int vec = tid * VECS + i;
                unsigned h;
                typename P::Aux aux;
                p.load_smem(in, vec, L::TILE_VECS, tile_base + (int64_t)vec * IPV, n, v[i], h, aux);
                hf |= h << (i * IPV);
            }
        } else {
#pragma unroll
            for (int i = 0; i < VECS; i++) {
                unsigned h;
                typename P::Aux aux;
                p.load(tile_base + ((int64_t)tid * VECS + i) * IPV, n, false, v[i], h, aux);
                hf |= h << (i * IPV);
            }
        }
        // thread-serial aggregate of the run, then ONE warp scan
        T a = v[0][0];
#pragma unroll
        for (int k = 1; k < ITEMS; k++) a = (SEG && ((hf >> k) & 1u)) ? v[k / IPV][k % IPV] : a + v[k / IPV][k % IPV];
        const unsigned m = SEG ? __ballot_sync(0xFFFFFFFFu, hf != 0) : 0u;
        const int lo = SEG ? 31 - __clz(m & (0xFFFFFFFFu >> (31 - lane))) : -1;
        const T wv = warp_scan_inclusive<T, SEG>(a, lane, lo);
        if (lane == 31) { s_agg_v[warp] = wv; s_agg_f[warp] = (unsigned char)(m != 0); }
        __syncthreads();  // S1: warp aggregates visible
        if (filler && it >= DEFER + 1) {
            // Slots in use right now: this tile's and the DEFER pending ones.  The tile before those was stored
            // during the previous iteration; once the copy engine has read its slot, the slot takes the tile
            // STAGES-DEFER-1 waves ahead.
            bulk_store_wait_read_all();
            fill((it - DEFER - 1) % STAGES, tile + (int64_t)(STAGES - DEFER - 1) * G);
        }
        if (warp == 0) {
            T ag[1], ex[1];
            ag[0] = lane < WARPS ? s_agg_v[lane] : T(0);
            const unsigned hb = (SEG && lane < WARPS && s_agg_f[lane]) ? 1u : 0u;
            unsigned before;
            T tile_agg;
            bool tile_f;
            warp_array_scan<T, SEG, 1>(ag, hb, lane, ex, before, tile_agg, tile_f);
            stamp(it, 2);
            if (lane == 0) Slot<T>::publish(ws.A, tile, tile_agg, tile_f);
            if (lane < WARPS) { s_excl_v[lane] = ex[0]; s_excl_f[lane] = (unsigned char)(before & 1u); }
            stamp(it, 3);
            if (q[0].tile >= 0) {  // DEFER iterations old: normally the early look already has it
                if (early_ok) { if (lane == 0) s_carry = early_carry; }
                else poll(q[0].tile);
            }
            stamp(it, 5);
        }
        __syncthreads();  // S2
        stamp(it, 6);
        T w = __shfl_up_sync(0xFFFFFFFFu, wv, 1);
        if (lane == 0) w = T(0);
        const bool head_before_lane = SEG && (m & ((1u << lane) - 1u)) != 0;
        Pending cur;
        cur.cst = w; cur.apply = 0; cur.tile = tile; cur.base = tile_base; cur.stage = stage;
        if (!head_before_lane) {
            cur.cst = s_excl_v[warp] + cur.cst;
            cur.apply = !(SEG && s_excl_f[warp]);
        }
        if (!full) {
            // The ragged last tile of the whole input is not in a slot.  It is this CTA's last tile: drain the
            // pending tiles, then finish it straight from registers with guarded stores.
            const T carry_oldest = s_carry;
#pragma unroll
            for (int d = 0; d < DEFER; d++) {
                if (q[d].tile < 0) continue;
                if (d > 0) {
                    __syncthreads();
                    if (warp == 0) poll(q[d].tile);
                    __syncthreads();
                }
                finish_slot(q[d], d == 0 ? carry_oldest : s_carry);
                __syncthreads();
                if (filler) store_slot(q[d]);
                q[d].tile = -1;
            }
            __syncthreads();
            if (warp == 0) poll(tile);
            __syncthreads();
            T c = cur.apply ? s_carry + cur.cst : cur.cst;
#pragma unroll
            for (int i = 0; i < VECS; i++) {
                T incl[IPV], excl[IPV];
#pragma unroll
                for (int j = 0; j < IPV; j++) {
                    if (SEG && ((hf >> (i * IPV + j)) & 1u)) c = T(0);
                    excl[j] = c;
                    c = c + v[i][j];
                    incl[j] = c;
                }
                typename P::Aux aux;
                p.store(tile_base + ((int64_t)tid * VECS + i) * IPV, n, false, incl, excl, aux);
            }
            break;
        }
        if (q[0].tile >= 0) {
            finish_slot(q[0], s_carry);
            __syncthreads();  // S3: the whole tile is back in its slot
            if (filler) store_slot(q[0]);
        }
        stamp(it, 7);
#pragma unroll
        for (int d = 0; d + 1 < DEFER; d++) q[d] = q[d + 1];
        q[DEFER - 1] = cur;
    }
    // drain the pending tiles of this CTA, oldest first
#pragma unroll
    for (int d = 0; d < DEFER; d++) {
        if (q[d].tile < 0) continue;
        __syncthreads();
        if (warp == 0) poll(q[d].tile);
        __syncthreads();
        finish_slot(q[d], s_carry);
        __syncthreads();
        if (filler) store_slot(q[d]);
    }
    if (filler) bulk_store_wait_read_all();  // shared memory must outlive the copy engine's reads
}

#endif  // __CUDACC__

// Tile configuration (tuned on B200, see profiles/r01_notes.md).
template <class P> struct Cfg {
    static constexpr int THREADS = 1024;                       // bulk kernel: one CTA per SM
    static constexpr int VEC_BYTES = BulkLayout<P, 32, 1>::STAGE_BYTES / 32;      // staged bytes per 16-byte value vector
    static constexpr int STAGE_TARGET = (P::SEG ? 64 : 80) * 1024;    // segmented variants carry more state per vector: smaller tiles, no spills
    static constexpr int VECS = STAGE_TARGET / (THREADS * VEC_BYTES) > 0 ? STAGE_TARGET / (THREADS * VEC_BYTES) : 1;
    static constexpr int STAGES = 2;   // large tiles amortise the ~0.7 us prefix round trip better than deep rings (profiles/r01_notes.md)
    static constexpr int DPL = 5;                               // a wave holds at most 32*DPL = 160 tiles
    // blocked bulk kernel (dense-output policies).  Shapes from the sweeps in profiles/r01_notes.md and profiles/r02_scan_shape_sweeps.txt;
    // the number of vectors per thread must be odd (4 or 6 put the threads' runs on a power-of-two shared-memory stride: 0.56-0.76
    // of peak where 5 and 7 reach 0.82-0.84) and threads * vectors a multiple of 64 (flag words are bulk-copied in 16-byte units):
    //   plain scans              704 x 5, 4 ring slots, finish deferred by 2 iterations   (u32 scan 0.96)
    //   segmented scans          384 x 7, 5 slots, deferred by 3                           (fp32 0.838; 640 x 5 x 4 x 2 gave 0.82)
    //   two value streams (the hoisted final-project step: a and xx per tile)   576 x 3, 4 slots, deferred by 2   (7.77 ms per 10
    //                            iterations; 512 x 3: 8.08, 640 x 3 with 3 slots: 9.05)
    static constexpr bool TWO_VALUE_STREAMS = P::BULK && P::NIN >= 3;
    static constexpr int B_THREADS = TWO_VALUE_STREAMS ? 576 : P::SEG ? 384 : 704, B_VECS = TWO_VALUE_STREAMS ? 3 : P::SEG ? 7 : 5,
                         B_STAGES = (P::SEG && !TWO_VALUE_STREAMS) ? 5 : 4, B_DEFER = (P::SEG && !TWO_VALUE_STREAMS) ? 3 : 2;
    static constexpr int DIRECT_THREADS = 256, DIRECT_VECS = 4; // one-tile-per-CTA kernel
    static constexpr int DIRECT_MIN_BLOCKS = 6;   // 40 registers per thread: the gather-heavy policies want occupancy more than registers (fused step 1.68 -> 1.41 ms)
};

// ---- host launchers ----------------------------------------------------------------------------------
// `scratch_mem` must hold scratch_bytes(tiles, sizeof(T)) bytes, or be null to use the context workspace.
template <class P, int THREADS, int VECS, int MIN_BLOCKS>
int launch_direct(cme213_ctx *ctx, const P &p, int64_t n, bool vec_ok, cudaStream_t st, const char *name,
                  void *scratch_mem) {
    using T = typename P::T;
    const int64_t tiles = div_up(n, (int64_t)THREADS * VECS * P::IPV);
    if (tiles > 0x7FFFFFFFLL) { set_error("%s: n too large", name); return CME213_ERR_UNSUPPORTED; }
    if (!scratch_mem) {
        int rc = workspace(ctx, scratch_bytes(tiles, sizeof(T)), &scratch_mem, st);
        if (rc) return rc;
    }
    CME213_CUDA(cudaMemsetAsync(scratch_mem, 0, scratch_bytes(tiles, sizeof(T)), st));
    Scratch s = scratch_carve(scratch_mem, tiles, sizeof(T));
#ifdef __CUDACC__
    lookback_scan_kernel<P, THREADS, VECS, MIN_BLOCKS><<<(unsigned)tiles, THREADS, 0, st>>>(p, n, s, vec_ok ? 1 : 0);
#endif
    CME213_LAUNCHED(ctx, name);
    return CME213_OK;
}

template <class P, int THREADS, int VECS, int STAGES, int DPL, bool BLOCKED = false, int DEFER = 1>
int launch_bulk(cme213_ctx *ctx, const P &p, int64_t n, cudaStream_t st, const char *name, void *scratch_mem,
                size_t scratch_avail) {
    using T = typename P::T;
    using L = BulkLayout<P, THREADS, VECS>;
    int64_t tiles = div_up(n, (int64_t)L::TILE_ITEMS);
    if (tiles > 0x7FFFFFFFLL) { set_error("%s: n too large", name); return CME213_ERR_UNSUPPORTED; }
#ifdef __CUDACC__
    void (*kern)(P, int64_t, int64_t, WaveScratch);
    if constexpr (BLOCKED) kern = scan_bulk_blocked_kernel<P, THREADS, VECS, STAGES, DPL, DEFER>;
    else kern = lookback_scan_bulk_kernel<P, THREADS, VECS, STAGES, DPL>;
    const int smem = STAGES * L::STAGE_BYTES;
    static int per_sm_dev[64] = {0};  // per instantiation and device: resident CTAs per SM
    int &per_sm = per_sm_dev[ctx->device & 63];
    if (per_sm == 0) {
        CME213_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        CME213_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
        CME213_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, THREADS, smem));
        if (per_sm < 1) { set_error("%s: kernel does not fit on an SM", name); per_sm = 0; return CME213_ERR_UNSUPPORTED; }
    }
    int64_t grid = (int64_t)per_sm * ctx->sm_count - 1;   // data CTAs; all co-resident with the scanner CTA (cooperative launch checks it)
    if (grid > 32 * DPL) grid = 32 * DPL;                  // one warp scans a whole wave
    if (grid > tiles) grid = tiles;
    if (grid < 1) grid = 1;
    const int64_t waves = div_up(tiles, grid);
    const size_t bytes = wave_scratch_bytes(tiles, waves, sizeof(T));
    if (!scratch_mem) {
        int rc = workspace(ctx, bytes, &scratch_mem, st);
        if (rc) return rc;
    } else if (bytes > scratch_avail) {
        set_error("%s: scratch too small (%zu > %zu)", name, bytes, scratch_avail);
        return CME213_ERR_INVALID;
    }
    CME213_CUDA(cudaMemsetAsync(scratch_mem, 0, bytes, st));
    WaveScratch ws = wave_scratch_carve(scratch_mem, tiles, waves, sizeof(T));
    static unsigned long long *d_trace = nullptr;
#ifdef CME213_TUNING   // per-phase trace of the persistent scan (scripts/trace_scan.sh): not in the shipped library
    const bool tracing = getenv("CME213_SCAN_TRACE") != nullptr;
#else
    constexpr bool tracing = false;
#endif
    const size_t trace_words = (size_t)(grid + 1) * TRACE_MAX_IT * 8;
    if (tracing) {
        if (!d_trace) CME213_CUDA(cudaMalloc(&d_trace, (size_t)1024 * TRACE_MAX_IT * 8 * 8));
        CME213_CUDA(cudaMemsetAsync(d_trace, 0, trace_words * 8, st));
        ws.trace = d_trace;
    }
    P pp = p;
    void *args[] = {(void *)&pp, (void *)&n, (void *)&tiles, (void *)&ws};
    CME213_CUDA(cudaLaunchCooperativeKernel((const void *)kern, dim3((unsigned)grid + 1), dim3(THREADS), args, (size_t)smem, st));
    if (tracing) {
        // debug aid: average time between the stamps of the kernel, over all CTAs and waves
        unsigned long long *h = (unsigned long long *)malloc(trace_words * 8);
        CME213_CUDA(cudaStreamSynchronize(st));
        CME213_CUDA(cudaMemcpy(h, d_trace, trace_words * 8, cudaMemcpyDeviceToHost));
        double sum[8] = {0}; long cnt[8] = {0}; double iter = 0; long iters = 0;
        long nit = waves < TRACE_MAX_IT ? (long)waves : TRACE_MAX_IT;
        for (long c = 0; c < grid; c++)
            for (long i = 1; i + 1 < nit; i++) {
                unsigned long long *r = h + ((size_t)c * TRACE_MAX_IT + i) * 8, *nx = r + 8;
                if (!r[0] || !r[7] || !nx[0]) continue;
                iter += (double)(nx[0] - r[0]); iters++;
                int prev = 0;
                for (int k = 1; k < 8; k++) { if (!r[k]) continue; sum[k] += (double)(r[k] - r[prev]); cnt[k]++; prev = k; }
            }
        fprintf(stderr, "[scan trace] per_sm %d grid %ld waves %ld  ns/iter %.0f |", per_sm, (long)grid, (long)waves, iters ? iter / iters : 0.0);
        const char *nm[8] = {"", "wait_data", "local+S1", "pub+atomic", "scanner", "poll", "S2+cst", "store"};
        for (int k = 1; k < 8; k++) fprintf(stderr, " %s %.0f(n=%ld)", nm[k], cnt[k] ? sum[k] / cnt[k] : 0.0, cnt[k]);
        fprintf(stderr, "\n");
        if (const char *path = tracing ? getenv("CME213_SCAN_TRACE_FILE") : nullptr) {
            FILE *f = fopen(path, "wb");
            if (f) {
                long hdr[3] = {(long)grid + 1, (long)TRACE_MAX_IT, (long)waves};
                fwrite(hdr, sizeof(long), 3, f);
                fwrite(h, 8, trace_words, f);
                fclose(f);
            }
        }
        free(h);
    }
#endif
    CME213_LAUNCHED(ctx, name);
    return CME213_OK;
}

// Upper bound of the scratch either front end needs for n items (callers that carve it themselves).
template <class P> static inline size_t scratch_bytes_for(int64_t n) {
    using T = typename P::T;
    const int64_t dt = div_up(n > 0 ? n : 1, (int64_t)Cfg<P>::DIRECT_THREADS * Cfg<P>::DIRECT_VECS * P::IPV);
    const size_t direct = scratch_bytes(dt, sizeof(T));
    const size_t wave = wave_scratch_bytes(dt, dt, sizeof(T));  // generous: fewer, larger tiles in reality
    return direct > wave ? direct : wave;
}

// Default dispatch: bulk pipeline when every streamed array is 16-byte aligned and the input is large,
// direct kernel otherwise.
template <class P>
int launch_with_scratch(cme213_ctx *ctx, const P &p, int64_t n, bool vec_ok, cudaStream_t st, const char *name,
                        void *scratch_mem, size_t scratch_avail) {
    if constexpr (P::BULK) {
        // the persistent pipeline pays off from about a million items per stream; with two value streams per tile (a and xx of
        // the hoisted final-project step) its tiles are smaller and the one-CTA-per-tile kernel stays ahead up to ~16 M items
        // (the 15 benchmark shapes of the final project, n = 1-12 M: profiles/r02_reference_gpu_programs.md)
        if (vec_ok && n >= (Cfg<P>::TWO_VALUE_STREAMS ? (int64_t)1 << 24 : (int64_t)1 << 20))
            return launch_bulk<P, Cfg<P>::B_THREADS, Cfg<P>::B_VECS, Cfg<P>::B_STAGES, Cfg<P>::DPL, true, Cfg<P>::B_DEFER>(ctx, p, n, st, name, scratch_mem, scratch_avail);
    }
    return launch_direct<P, Cfg<P>::DIRECT_THREADS, Cfg<P>::DIRECT_VECS, Cfg<P>::DIRECT_MIN_BLOCKS>(ctx, p, n, vec_ok, st, name, scratch_mem);
}

template <class P>
int launch(cme213_ctx *ctx, const P &p, int64_t n, bool vec_ok, cudaStream_t st, const char *name) {
    return launch_with_scratch<P>(ctx, p, n, vec_ok, st, name, nullptr, 0);
}

}  // namespace scan
}  // namespace cme213
