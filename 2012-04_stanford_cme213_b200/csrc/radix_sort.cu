// radix_sort.cu -- LSD radix sort of uint32 keys, built on the library's exclusive scan.
//
// Reference: hw/hw4/programming/radixsort.cpp:44-109 (radixSortParallelPass: per-block histogram -> histograms
// summed and scanned digit-major -> per-block scatter offsets -> stable scatter) and :111-121 (radixSortParallel:
// passes in pairs so the result lands back in `keys`).  Same pass structure here, one CTA per 8192-key block
// (the reference's blockSize, radixsort.cpp:115):
//   radix_hist_kernel    per-block digit histograms, written digit-major: table[d * pitch + block]
//   exclusive scan       ONE scan over the whole digit-major table replaces UPSWEEP 2 + SCAN + DOWNSWEEP 2
//                        (radixsort.cpp:60-97): entry (d, b) becomes "keys with a smaller digit" + "keys with
//                        digit d in earlier blocks" = the block's scatter offset for digit d
//   radix_scatter_kernel per-warp digit counts, block-wide offsets, then stable ranks item by item (same-digit lanes
//                        found through a shared-memory OR) that put the keys straight into a digit-grouped copy of
//                        the block in shared memory, written out in runs (DOWNSWEEP 1, radixsort.cpp:99-109)
// Algorithmic bytes per pass: 12 n (keys read twice, written once) + 8 * radix * n / 8192 for the table.
#include "internal.cuh"

using namespace cme213;

namespace {

constexpr int RS_THREADS = 512;
constexpr int RS_ITEMS = 16;
constexpr int RS_WARPS = RS_THREADS / 32;
constexpr int RS_TILE = RS_THREADS * RS_ITEMS;   // 8192 keys per block
constexpr int RS_MAX_RADIX = 256;
constexpr int RS_BINS = RS_MAX_RADIX + 1;        // one extra bin for the slots past the end of the last block
constexpr int RS_HIST_TILES = 8;                 // blocks per histogram CTA: 8 table words = one 32-byte sector

// ---------------------------------------------------------------------------------------------------------
// Per-block histograms (radixSortParallelHistoBlock, radixsort.cpp:22-31), digit-major.
__global__ void __launch_bounds__(RS_THREADS)
radix_hist_kernel(const uint32_t *__restrict__ keys, int64_t n, int start_bit, uint32_t mask, int radix,
                  uint32_t *__restrict__ table, int64_t pitch, int vec_ok) {
    __shared__ uint32_t s_warp[RS_WARPS][RS_MAX_RADIX];
    __shared__ uint32_t s_tile[RS_MAX_RADIX][RS_HIST_TILES];
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int t = 0; t < RS_HIST_TILES; ++t) {
        for (int i = tid; i < RS_WARPS * RS_MAX_RADIX; i += RS_THREADS) (&s_warp[0][0])[i] = 0;
        __syncthreads();
        const int64_t base = ((int64_t)blockIdx.x * RS_HIST_TILES + t) * RS_TILE;
        if (base < n) {
            const int64_t left = n - base;
            if (left >= RS_TILE && vec_ok) {
                const uint4 *src = reinterpret_cast<const uint4 *>(keys + base);
#pragma unroll
                for (int it = 0; it < RS_TILE / (RS_THREADS * 4); ++it) {
                    const uint4 v = ld_stream(src + it * RS_THREADS + tid);
                    atomicAdd(&s_warp[warp][(v.x >> start_bit) & mask], 1u);
                    atomicAdd(&s_warp[warp][(v.y >> start_bit) & mask], 1u);
                    atomicAdd(&s_warp[warp][(v.z >> start_bit) & mask], 1u);
                    atomicAdd(&s_warp[warp][(v.w >> start_bit) & mask], 1u);
                }
            } else {
                const int cnt = (int)(left < RS_TILE ? left : RS_TILE);
                for (int j = tid; j < cnt; j += RS_THREADS)
                    atomicAdd(&s_warp[warp][(keys[base + j] >> start_bit) & mask], 1u);
            }
        }
        __syncthreads();
        if (tid < RS_MAX_RADIX) {
            uint32_t sum = 0;
#pragma unroll
            for (int w = 0; w < RS_WARPS; ++w) sum += s_warp[w][tid];
            s_tile[tid][t] = sum;
        }
        __syncthreads();
    }
    if (tid < radix) {
        uint4 *dst = reinterpret_cast<uint4 *>(table + (int64_t)tid * pitch + (int64_t)blockIdx.x * RS_HIST_TILES);
        dst[0] = make_uint4(s_tile[tid][0], s_tile[tid][1], s_tile[tid][2], s_tile[tid][3]);
        dst[1] = make_uint4(s_tile[tid][4], s_tile[tid][5], s_tile[tid][6], s_tile[tid][7]);
    }
}

// ---------------------------------------------------------------------------------------------------------
// Stable scatter of one block (radixSortParallelScatterBlock, radixsort.cpp:33-42).  Keys are held warp-striped
// (warp w owns 512 consecutive keys, item i of lane l is key 32*i + l of them), so "item by item, lane by lane"
// is memory order and the ranks below keep equal digits in input order.
constexpr int RS_START_PAD = 260;
constexpr size_t RS_SCATTER_SMEM = (size_t)(RS_TILE + 2 * RS_WARPS * RS_BINS + RS_START_PAD + RS_MAX_RADIX) * 4;

// Finding the lanes of the warp that hold the same digit.  Measured on B200 (2^28 keys, ms per 8-bit pass of this kernel):
//   __match_any_sync        1.97  MATCH.ANY runs on the address-divergence unit at about one warp instruction per
//                                 64 cycles per SM: 91 % ADU-bound (profiles/r01_ncu_radix_scatter_v1_match.txt)
//   eight ballots + selects 1.27  85 % ALU-bound (..._v2_ballot.txt)
//   shared-memory OR        1.10  every lane ORs its bit into a per-warp, per-digit word and reads the word back:
//                                 89 % shared-memory-bound (..._v3_atomicor.txt)
//   the same + early counts 0.97  (this kernel) one random shared-memory access per key fewer
// Mixing the methods so that the three pipes work side by side was tried (4/6/6 ... 0/8/8 items) and lost to the
// plain shared-memory version once the early counts were in (5.09-5.40 against 4.79 ms per sort).
// PAIRS: a 32-bit payload travels with every key (second shared-memory copy of the block, same slots).
template <bool PAIRS>
__global__ void __launch_bounds__(RS_THREADS, 2)
radix_scatter_kernel(const uint32_t *__restrict__ keys_in, uint32_t *__restrict__ keys_out, int64_t n, int start_bit,
                     uint32_t mask, int radix, const uint32_t *__restrict__ table, int64_t pitch,
                     const uint32_t *__restrict__ vals_in, uint32_t *__restrict__ vals_out) {
    extern __shared__ __align__(16) uint32_t rs_smem[];
    uint32_t *s_keys = rs_smem;                          // [RS_TILE] the block regrouped by digit
    uint32_t *s_wh = s_keys + RS_TILE;                   // [RS_WARPS][RS_BINS] per-warp digit counters, then next free slots
    uint32_t *s_mm = s_wh + RS_WARPS * RS_BINS;          // [RS_WARPS][RS_BINS] match words of the OR method
    uint32_t *s_start = s_mm + RS_WARPS * RS_BINS;       // [RS_BINS] first slot of each digit inside the block
    uint32_t *s_gofs = s_start + RS_START_PAD;           // [radix] global offset minus s_start
    uint32_t *s_vals = s_gofs + RS_MAX_RADIX;            // [RS_TILE] PAIRS only: the payloads, regrouped like the keys
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t base = (int64_t)blockIdx.x * RS_TILE;
    const int cnt = (int)((n - base) < RS_TILE ? (n - base) : RS_TILE);
    const bool full = cnt == RS_TILE;
    const uint32_t tofs = tid < radix ? table[(int64_t)tid * pitch + blockIdx.x] : 0u;

    const int wbase = warp * 32 * RS_ITEMS + lane;
    uint32_t key[RS_ITEMS];
#pragma unroll
    for (int i = 0; i < RS_ITEMS; ++i) {
        const int li = wbase + i * 32;
        key[i] = li < cnt ? ld_stream(keys_in + base + li) : 0xFFFFFFFFu;
    }
    for (int i = tid; i < 2 * RS_WARPS * RS_BINS; i += RS_THREADS) s_wh[i] = 0;   // s_wh and s_mm are adjacent
    __syncthreads();

    // 1. per-warp digit counts first ("early counts"): the block-wide offsets are then known before the ranking,
    //    so the ranking can put every key straight into its slot of the regrouped block.
    uint32_t *wh = s_wh + warp * RS_BINS;
#pragma unroll
    for (int i = 0; i < RS_ITEMS; ++i) {
        const bool valid = full || (wbase + i * 32 < cnt);
        const uint32_t d = valid ? ((key[i] >> start_bit) & mask) : (uint32_t)radix;   // slots past the end: bin `radix`
        atomicAdd(&wh[d], 1u);
    }
    __syncthreads();

    if (tid < RS_BINS) {            // the block's count of each digit
        uint32_t sum = 0;
#pragma unroll
        for (int w = 0; w < RS_WARPS; ++w) sum += s_wh[w * RS_BINS + tid];
        s_start[tid] = sum;
    }
    __syncthreads();
    if (warp == 0) {                // exclusive scan of the RS_BINS counts, 9 per lane
        constexpr int PER = (RS_BINS + 31) / 32;
        uint32_t v[PER], sum = 0;
#pragma unroll
        for (int j = 0; j < PER; ++j) {
            const int idx = lane * PER + j;
            v[j] = idx < RS_BINS ? s_start[idx] : 0u;
            sum += v[j];
        }
        uint32_t incl = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t up = __shfl_up_sync(0xFFFFFFFFu, incl, o);
            if (lane >= o) incl += up;
        }
        uint32_t run = incl - sum;
#pragma unroll
        for (int j = 0; j < PER; ++j) {
            const int idx = lane * PER + j;
            if (idx < RS_BINS) s_start[idx] = run;
            run += v[j];
        }
    }
    __syncthreads();
    if (tid < RS_BINS) {            // counts -> next free slot of (warp, digit) inside the regrouped block
        uint32_t run = s_start[tid];
        if (tid < radix) s_gofs[tid] = tofs - run;
#pragma unroll
        for (int w = 0; w < RS_WARPS; ++w) {
            const uint32_t c = s_wh[w * RS_BINS + tid];
            s_wh[w * RS_BINS + tid] = run;
            run += c;
        }
    }
    __syncthreads();

    // 2. stable ranking, item by item: the highest lane of each same-digit group reserves the group's slots
    uint32_t *mm = s_mm + warp * RS_BINS;
    const uint32_t me = 1u << lane, lt = me - 1u;
#pragma unroll
    for (int i = 0; i < RS_ITEMS; ++i) {
        const bool valid = full || (wbase + i * 32 < cnt);
        const uint32_t d = valid ? ((key[i] >> start_bit) & mask) : (uint32_t)radix;
        atomicOr(&mm[d], me);
        __syncwarp();
        const uint32_t peers = mm[d];
        __syncwarp();
        const int leader = 31 - __clz(peers);
        uint32_t slot = 0;
        if (lane == leader) {
            mm[d] = 0;
            slot = atomicAdd(&wh[d], (uint32_t)__popc(peers));
        }
        __syncwarp();   // orders the leader's reset of mm[d] before the next item's atomicOr (a shuffle alone gives no memory ordering)
        slot = __shfl_sync(0xFFFFFFFFu, slot, leader);
        const uint32_t pos = slot + __popc(peers & lt);
        s_keys[pos] = key[i];
        if (PAIRS) s_vals[pos] = valid ? vals_in[base + wbase + i * 32] : 0u;
    }
    __syncthreads();
#pragma unroll 4
    for (int k = 0; k < RS_ITEMS; ++k) {
        const int j = k * RS_THREADS + tid;
        if (j < cnt) {
            const uint32_t kv = s_keys[j];
            const size_t g = (size_t)(s_gofs[(kv >> start_bit) & mask] + (uint32_t)j);
            keys_out[g] = kv;
            if (PAIRS) vals_out[g] = s_vals[j];
        }
    }
}

struct PassPlan {
    int64_t blocks, pitch, entries;
    size_t table_bytes, total_bytes;
};
PassPlan plan_pass(int64_t n) {
    PassPlan p;
    p.blocks = div_up(n, RS_TILE);
    p.pitch = div_up(p.blocks, RS_HIST_TILES) * RS_HIST_TILES;
    p.entries = p.pitch * RS_MAX_RADIX;          // sized for the widest digit so every pass shares one workspace
    p.table_bytes = ((size_t)p.entries * 4 + 255) & ~(size_t)255;
    p.total_bytes = p.table_bytes + cme213_internal_scan_u32_scratch_bytes(p.entries);
    return p;
}

// One pass with the workspace already in hand.
int run_pass(cme213_ctx *ctx, const uint32_t *d_in, uint32_t *d_out, int64_t n, int start_bit, int num_bits,
             const PassPlan &p, void *ws, cudaStream_t st, const uint32_t *d_vals_in = nullptr,
             uint32_t *d_vals_out = nullptr) {
    const int radix = 1 << num_bits;
    const uint32_t mask = (uint32_t)radix - 1u;
    uint32_t *table = (uint32_t *)ws;
    void *scratch = (char *)ws + p.table_bytes;
    radix_hist_kernel<<<(unsigned)(p.pitch / RS_HIST_TILES), RS_THREADS, 0, st>>>(d_in, n, start_bit, mask, radix, table,
                                                                                 p.pitch, aligned16(d_in) ? 1 : 0);
    CME213_LAUNCHED(ctx, "radix_hist_kernel");
    int rc = cme213_internal_exclusive_scan_u32(ctx, table, table, p.pitch * radix, scratch, st);
    if (rc) return rc;
    // per device and cheap: set on every call rather than tracked
    if (d_vals_in) {
        constexpr size_t smem = RS_SCATTER_SMEM + (size_t)RS_TILE * 4;
        CME213_CUDA(cudaFuncSetAttribute(radix_scatter_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        radix_scatter_kernel<true><<<(unsigned)p.blocks, RS_THREADS, smem, st>>>(d_in, d_out, n, start_bit, mask, radix, table,
                                                                               p.pitch, d_vals_in, d_vals_out);
    } else {
        CME213_CUDA(cudaFuncSetAttribute(radix_scatter_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)RS_SCATTER_SMEM));
        radix_scatter_kernel<false><<<(unsigned)p.blocks, RS_THREADS, RS_SCATTER_SMEM, st>>>(d_in, d_out, n, start_bit, mask,
                                                                                           radix, table, p.pitch, nullptr, nullptr);
    }
    CME213_LAUNCHED(ctx, "radix_scatter_kernel");
    return CME213_OK;
}

int check_sort_args(const char *who, const void *a, const void *b, int64_t n) {
    if (n < 0) { set_error("%s: negative size", who); return CME213_ERR_INVALID; }
    if (n > (int64_t)0x7FFFFFFF) { set_error("%s: more than 2^31-1 keys", who); return CME213_ERR_INVALID; }
    if (n && (!a || !b || a == b)) { set_error("%s: bad pointers", who); return CME213_ERR_INVALID; }
    return CME213_OK;
}

}  // namespace

CME213_API int cme213_radix_sort_pass_u32(cme213_ctx *ctx, const uint32_t *d_keys_in, uint32_t *d_keys_out, int64_t n,
                                          int start_bit, int num_bits, void *stream) {
    CME213_CHECK_CTX(ctx);
    int rc = check_sort_args("radix_sort_pass_u32", d_keys_in, d_keys_out, n);
    if (rc) return rc;
    if (num_bits < 1 || num_bits > 8 || start_bit < 0 || start_bit > 31) {
        set_error("radix_sort_pass_u32: num_bits must be 1..8 and start_bit 0..31");
        return CME213_ERR_INVALID;
    }
    if (n == 0) return CME213_OK;
    DeviceGuard g(ctx->device);
    const PassPlan p = plan_pass(n);
    void *ws = nullptr;
    rc = workspace(ctx, p.total_bytes, &ws, (cudaStream_t)stream);
    if (rc) return rc;
    return run_pass(ctx, d_keys_in, d_keys_out, n, start_bit, num_bits, p, ws, (cudaStream_t)stream);
}

CME213_API int cme213_radix_sort_u32(cme213_ctx *ctx, uint32_t *d_keys, uint32_t *d_keys_tmp, int64_t n, int num_bits,
                                     void *stream) {
    CME213_CHECK_CTX(ctx);
    int rc = check_sort_args("radix_sort_u32", d_keys, d_keys_tmp, n);
    if (rc) return rc;
    if (num_bits < 1 || num_bits > 16) {   // the reference asserts numBits <= 16 (radixsort.cpp:113)
        set_error("radix_sort_u32: num_bits must be 1..16");
        return CME213_ERR_INVALID;
    }
    if (n == 0) return CME213_OK;
    DeviceGuard g(ctx->device);
    const PassPlan p = plan_pass(n);
    void *ws = nullptr;
    rc = workspace(ctx, p.total_bytes, &ws, (cudaStream_t)stream);
    if (rc) return rc;
    // Digits wider than 8 bits are split in two device passes: the sorted result does not depend on the digit width.
    const int bits = num_bits > 8 ? 8 : num_bits;
    uint32_t *src = d_keys, *dst = d_keys_tmp;
    int passes = 0;
    for (int start = 0; start < 32; start += bits) {
        rc = run_pass(ctx, src, dst, n, start, bits, p, ws, (cudaStream_t)stream);
        if (rc) return rc;
        uint32_t *t = src; src = dst; dst = t;
        ++passes;
    }
    if (passes & 1)   // the reference runs its passes in pairs so the result ends in `keys` (radixsort.cpp:117-120)
        CME213_CUDA(cudaMemcpyAsync(d_keys, d_keys_tmp, (size_t)n * 4, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
    return CME213_OK;
}

// Key-value form: d_vals travels with d_keys.  Sorts on the key bits [0, end_bit) only (8-bit digits), which is what
// the COO -> CSR conversion needs (row and column indices use few bits).  Stable; the result ends in d_keys / d_vals.
CME213_API int cme213_radix_sort_pairs_u32(cme213_ctx *ctx, uint32_t *d_keys, uint32_t *d_keys_tmp, uint32_t *d_vals,
                                           uint32_t *d_vals_tmp, int64_t n, int end_bit, void *stream) {
    CME213_CHECK_CTX(ctx);
    int rc = check_sort_args("radix_sort_pairs_u32", d_keys, d_keys_tmp, n);
    if (rc) return rc;
    rc = check_sort_args("radix_sort_pairs_u32", d_vals, d_vals_tmp, n);
    if (rc) return rc;
    if (end_bit < 0 || end_bit > 32) { set_error("radix_sort_pairs_u32: end_bit must be 0..32"); return CME213_ERR_INVALID; }
    if (n == 0 || end_bit == 0) return CME213_OK;
    DeviceGuard g(ctx->device);
    cudaStream_t st = (cudaStream_t)stream;
    const PassPlan p = plan_pass(n);
    void *ws = nullptr;
    rc = workspace(ctx, p.total_bytes, &ws, (cudaStream_t)stream);
    if (rc) return rc;
    uint32_t *ks = d_keys, *kd = d_keys_tmp, *vs = d_vals, *vd = d_vals_tmp;
    int passes = 0;
    for (int start = 0; start < end_bit; start += 8) {
        const int bits = end_bit - start < 8 ? end_bit - start : 8;
        rc = run_pass(ctx, ks, kd, n, start, bits, p, ws, st, vs, vd);
        if (rc) return rc;
        uint32_t *t = ks; ks = kd; kd = t;
        t = vs; vs = vd; vd = t;
        ++passes;
    }
    if (passes & 1) {
        CME213_CUDA(cudaMemcpyAsync(d_keys, d_keys_tmp, (size_t)n * 4, cudaMemcpyDeviceToDevice, st));
        CME213_CUDA(cudaMemcpyAsync(d_vals, d_vals_tmp, (size_t)n * 4, cudaMemcpyDeviceToDevice, st));
    }
    return CME213_OK;
}
