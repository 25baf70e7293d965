// common.cuh -- shared host/device helpers for libcme213_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include "../../include/cme213_b200.h"

#define CME213_API extern "C" __attribute__((visibility("default")))

struct cme213_ctx {
    int device;
    int sm_count;
    void *ws;          // scratch workspace, grown on demand
    size_t ws_bytes;
    cudaStream_t ws_stream;  // stream of the workspace's latest user (see cme213::workspace)
    bool ws_used;
    cudaEvent_t ws_event;
    int64_t launches;  // kernels launched through this context
    // CSR AUTO decisions (spmv_csr.cu), remembered per (row_ptr, row range)
    static constexpr int CSR_CACHE = 16;
    struct CsrChoice { const int32_t *row_ptr; int64_t row_begin, row_end; cme213_csr_plan *plan; };   // plans owned by the context
    CsrChoice csr_cache[CSR_CACHE];
    unsigned csr_cache_next;
};

namespace cme213 {

void set_error(const char *fmt, ...);
int cuda_fail(cudaError_t e, const char *what);
// Returns a workspace of at least `bytes` (256-byte aligned), growing it if needed.  Every call shares the one
// buffer: when the previous user ran on a different stream, `st` is first made to wait (event, no host sync) for
// everything that user enqueued, so calls through one context from several streams do not trample each other's
// descriptors.  Skipped while `st` is being captured (a wait on un-captured work would invalidate the capture):
// there the caller orders its streams.  Growth synchronises the device and must happen outside a capture.
int workspace(cme213_ctx *ctx, size_t bytes, void **out, cudaStream_t st);

struct DeviceGuard {
    int prev;
    bool switched;
    explicit DeviceGuard(int dev) : prev(-1), switched(false) {
        cudaGetDevice(&prev);
        if (prev != dev) { cudaSetDevice(dev); switched = true; }
    }
    ~DeviceGuard() { if (switched) cudaSetDevice(prev); }
};

#define CME213_CHECK_CTX(ctx)                                             \
    do {                                                                  \
        if (!(ctx)) { cme213::set_error("null context"); return CME213_ERR_INVALID; } \
    } while (0)

#define CME213_CUDA(call)                                                 \
    do {                                                                  \
        cudaError_t _e = (call);                                          \
        if (_e != cudaSuccess) return cme213::cuda_fail(_e, #call);       \
    } while (0)

// After a kernel launch: count it and surface launch errors.
#define CME213_LAUNCHED(ctx, name)                                        \
    do {                                                                  \
        (ctx)->launches++;                                                \
        cudaError_t _e = cudaGetLastError();                              \
        if (_e != cudaSuccess) return cme213::cuda_fail(_e, name);        \
    } while (0)

static inline int64_t div_up(int64_t a, int64_t b) { return (a + b - 1) / b; }
static inline bool aligned16(const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// ---- device helpers -------------------------------------------------------
#ifdef __CUDACC__
// Streaming (read-once) loads: keep L1/L2 for the gathered vector x.
template <typename V> __device__ __forceinline__ V ld_stream(const V *p) { return __ldcs(p); }
template <typename V> __device__ __forceinline__ void st_stream(V *p, const V &v) { __stcs(p, v); }

// Gathered vector entries are re-used by neighbouring rows long after the streamed matrix data has gone by:
// keep them in L2 with an evict-last policy while col/val/y use evict-first (streaming) accesses.
__device__ __forceinline__ uint64_t l2_policy_evict_last() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ float ld_keep(const float *p, uint64_t pol) {
    float v;
    asm volatile("ld.global.nc.L2::cache_hint.f32 %0, [%1], %2;" : "=f"(v) : "l"(p), "l"(pol));
    return v;
}
__device__ __forceinline__ double ld_keep(const double *p, uint64_t pol) {
    double v;
    asm volatile("ld.global.nc.L2::cache_hint.f64 %0, [%1], %2;" : "=d"(v) : "l"(p), "l"(pol));
    return v;
}

// Streamed (read-once) data with an explicit evict-first L2 policy instead of the .cs qualifier.
__device__ __forceinline__ uint64_t l2_policy_evict_first() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ int4 ld_first(const int4 *p, uint64_t pol) {
    int4 v;
    asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v4.s32 {%0,%1,%2,%3}, [%4], %5;"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p), "l"(pol));
    return v;
}
__device__ __forceinline__ float4 ld_first(const float4 *p, uint64_t pol) {
    float4 v;
    asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v4.f32 {%0,%1,%2,%3}, [%4], %5;"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p), "l"(pol));
    return v;
}
__device__ __forceinline__ double2 ld_first(const double2 *p, uint64_t pol) {
    double2 v;
    asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v2.f64 {%0,%1}, [%2], %3;" : "=d"(v.x), "=d"(v.y) : "l"(p), "l"(pol));
    return v;
}

__device__ __forceinline__ uint32_t hash32(uint32_t seed, uint64_t idx) {
    // counter-based hash (two rounds of a 32-bit finaliser over a 64-bit counter); restated in
    // numpy by oracle/generators.py so tests can rebuild any generated input bit for bit.
    uint32_t h = seed ^ 0x9E3779B9u;
    h ^= (uint32_t)idx;
    h *= 0x85EBCA6Bu; h ^= h >> 13;
    h ^= (uint32_t)(idx >> 32) * 0xC2B2AE35u;
    h *= 0xC2B2AE35u; h ^= h >> 16;
    h *= 0x27D4EB2Fu; h ^= h >> 15;
    return h;
}
__device__ __forceinline__ float uniform_pm1_f32(uint32_t h) {
    // 24 random bits -> [0,2) in steps of 2^-23, minus 1: exact in fp32, never returns 0.
    float v = (float)(h >> 8) * 1.1920928955078125e-07f - 1.0f;
    return v == 0.0f ? 1.1920928955078125e-07f : v;
}
__device__ __forceinline__ double uniform_pm1_f64(uint32_t h) {
    double v = (double)h * 4.656612873077392578125e-10 - 1.0;  // h * 2^-31 - 1
    return v == 0.0 ? 4.656612873077392578125e-10 : v;
}
// ---- mbarrier / bulk-copy primitives (sm_90+; the linear mode of the TMA engine) ------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, unsigned parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_load(void *smem_dst, const void *gmem_src, unsigned bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(smem_dst)),
                 "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

#endif

}  // namespace cme213
