// sharded.cu -- row-sharded iterated y = A x across the GPUs of one box, behind the C ABI (config C5).
//
// Decomposition: contiguous row blocks, the 1-D stripe decomposition of the reference's only distributed code
// (hw/hw5/programming/2dHeat.cpp:284-306), here balanced by nonzeros instead of by rows (cme213_partition_rows).
// Overlap: the reference starts its halo transfers, computes the interior, waits, then computes the borders
// (asyncComputation, 2dHeat.cpp:696-723).  For SpMV the roles turn around: the BORDER rows are what the neighbours
// need, so they are computed first and leave for the neighbours while the interior is still being computed.
//
// One object per GPU (cme213_shard), in one process (cme213_shard_connect_local, peer access) or one process per GPU
// (cme213_shard_export / cme213_shard_connect, CUDA IPC).  No NCCL and no host synchronisation on the data path:
//   HALO          one kernel launch per pass.  Interior row blocks first; the boundary CTAs, last in the grid, wait for the
//                 neighbour's flag ("my previous pass is complete"), store their rows into the local vector AND into the
//                 neighbour's vector over NVLink; the next launch's first CTA raises this rank's flag in the neighbours'
//                 memory (the protocol is described in internal.cuh / spmv_csr.cu).
//   GATHER        general matrices: after the SpMV every rank copies its slice into every peer's vector (peer copies
//                 on the copy engines) and raises its flag at every peer.  Row order untouched: bitwise equal to 1 GPU.
//   GATHER_SPLIT  same exchange on a second stream, overlapped with the next pass's local-column block (the columns a
//                 rank owns need no remote data); the remote-column block follows once the peers' slices are in.
//                 The split changes the order in which a row is summed: equal within the reduction-order bound only.
// Row sharding as such does not change the summation order of a row, so HALO and GATHER results are bitwise identical
// to the single-GPU product (tests/test_sharded_gpu.py, bench.py's parity block).
#include "internal.cuh"
#include <string.h>
#include <vector>

namespace {
using namespace cme213;
typedef unsigned long long u64;

constexpr int LINE = 16;   // u64 words per 128-byte line: every word that another agent polls or writes gets its own line
// cnt_lo / cnt_hi: finished boundary CTAs (HALO); status: a wait gave up; epoch: slices sent (GATHER*); waits: passes that
// have waited for their input (GATHER*); flag[r]: written by rank r
enum { W_CNT_LO = 0, W_CNT_HI = LINE, W_STATUS = 2 * LINE, W_EPOCH = 3 * LINE, W_WAITS = 4 * LINE, W_FLAG0 = 5 * LINE };
size_t sync_bytes(int world) { return (size_t)(5 + world) * LINE * sizeof(u64); }

struct Blob {   // what a rank tells the others (cme213_shard_export)
    cudaIpcMemHandle_t x[2], sync;
    int64_t lo, hi, need_lo, need_hi;   // rows owned; ghost rows read below lo / from hi on
    int32_t rank, world;
};

__global__ void col_range_kernel(int64_t nnz, const int32_t *__restrict__ col, int *mn, int *mx) {
    int lo = 0x7FFFFFFF, hi = -1;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nnz; i += (int64_t)gridDim.x * blockDim.x) {
        const int c = col[i];
        lo = c < lo ? c : lo;
        hi = c > hi ? c : hi;
    }
    for (int d = 16; d > 0; d >>= 1) {
        const int a = __shfl_xor_sync(0xFFFFFFFFu, lo, d), b = __shfl_xor_sync(0xFFFFFFFFu, hi, d);
        lo = a < lo ? a : lo;
        hi = b > hi ? b : hi;
    }
    if ((threadIdx.x & 31) == 0) { atomicMin(mn, lo); atomicMax(mx, hi); }
}

// GATHER modes: wait until every peer's slice of the previous pass has landed here / tell every peer that mine has left
__global__ void gather_wait_kernel(u64 *sync, int world, int rank, long long timeout_cycles) {
    // pass p (p-1 earlier waits) needs every peer's slice of pass p-1.  Not derived from this rank's own sends: in
    // GATHER_SPLIT they run on another stream and may still be under way.
    const int r = threadIdx.x;
    const u64 want = *(volatile u64 *)(sync + W_WAITS);
    __syncthreads();
    if (r == 0) *(volatile u64 *)(sync + W_WAITS) = want + 1;
    if (r >= world || r == rank) return;
    const u64 *flag = sync + W_FLAG0 + (size_t)r * LINE;
    const long long t0 = clock64();
    u64 v;
    do {
        asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(flag) : "memory");
        if (v >= want) break;
        if (clock64() - t0 > timeout_cycles) { atomicExch(sync + W_STATUS, 1ull); break; }
        __nanosleep(64);
    } while (true);
}
struct PeerFlags { u64 *p[64]; };
__global__ void gather_signal_kernel(u64 *sync, PeerFlags peers, int world, int rank) {
    const u64 e = *(volatile u64 *)(sync + W_EPOCH) + 1;
    __syncthreads();
    if (threadIdx.x == 0) *(volatile u64 *)(sync + W_EPOCH) = e;
    const int r = threadIdx.x;
    if (r < world && r != rank && peers.p[r]) {
        __threadfence_system();
        asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(peers.p[r] + W_FLAG0 + (size_t)rank * LINE), "l"(e) : "memory");
    }
}

// GATHER_SPLIT set-up: entries of a row inside / outside the owned column range
__global__ void split_count_kernel(int64_t rows, const int32_t *__restrict__ ptr, const int32_t *__restrict__ col, int lo, int hi,
                                   uint32_t *cnt_loc, uint32_t *cnt_rem) {
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r > rows) return;
    uint32_t a = 0, b = 0;
    if (r < rows)
        for (int e = ptr[r]; e < ptr[r + 1]; e++) { const int c = col[e]; if (c >= lo && c < hi) a++; else b++; }
    cnt_loc[r] = a;   // entry `rows` is 0: its scan value is the total
    cnt_rem[r] = b;
}
__global__ void split_fill_kernel(int64_t rows, const int32_t *__restrict__ ptr, const int32_t *__restrict__ col,
                                  const double *__restrict__ val, int lo, int hi, const int32_t *lp, int32_t *lc, double *lv,
                                  const int32_t *rp, int32_t *rc, double *rv) {
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= rows) return;
    int a = lp[r], b = rp[r];
    for (int e = ptr[r]; e < ptr[r + 1]; e++) {
        const int c = col[e];
        if (c >= lo && c < hi) { lc[a] = c; lv[a++] = val[e]; } else { rc[b] = c; rv[b++] = val[e]; }
    }
}
}  // namespace

struct cme213_shard {
    cme213_ctx *ctx = nullptr;
    int rank = 0, world = 1, mode = CME213_SHARD_HALO, wanted_mode = CME213_SHARD_AUTO;
    int64_t rows_total = 0, lo = 0, hi = 0, nloc = 0, nnz = 0, need_lo = 0, need_hi = 0;
    std::vector<int64_t> bounds;
    const int32_t *row_ptr = nullptr, *col = nullptr;
    const double *val = nullptr;
    double *x[2] = {nullptr, nullptr};
    u64 *sync = nullptr;
    std::vector<double *> px[2];
    std::vector<u64 *> psync;
    bool ipc_opened = false, connected = false;
    int64_t push_lo = 0, push_hi = 0;          // rows the neighbours need from this rank
    long long timeout_cycles = 40000000000LL;  // ~20 s at 1.9 GHz: a peer that never shows up is an error, not a hang
    // GATHER_SPLIT
    cudaStream_t comm = nullptr;
    cudaEvent_t ev_done = nullptr, ev_sent[2] = {nullptr, nullptr};
    bool sent_valid[2] = {false, false};
    int32_t *loc_ptr = nullptr, *loc_col = nullptr, *rem_ptr = nullptr, *rem_col = nullptr;
    double *loc_val = nullptr, *rem_val = nullptr;
};

namespace {
int decide_mode(cme213_shard *s, const std::vector<Blob> &all) {
    // halo exchange works when every rank's ghost rows lie inside its direct neighbours
    bool halo_ok = true;
    for (int r = 0; r < s->world; r++) {
        const int64_t below = r > 0 ? all[r - 1].hi - all[r - 1].lo : 0, above = r + 1 < s->world ? all[r + 1].hi - all[r + 1].lo : 0;
        if (all[r].need_lo > below || all[r].need_hi > above) halo_ok = false;
    }
    int m = s->wanted_mode;
    if (m == CME213_SHARD_AUTO) m = halo_ok ? CME213_SHARD_HALO : CME213_SHARD_GATHER;
    if (m == CME213_SHARD_HALO && !halo_ok) {
        set_error("shard: the matrix reaches beyond the neighbouring row blocks, the halo exchange cannot serve it (use GATHER)");
        return CME213_ERR_UNSUPPORTED;
    }
    s->mode = m;
    s->push_lo = s->rank > 0 ? all[s->rank - 1].need_hi : 0;
    s->push_hi = s->rank + 1 < s->world ? all[s->rank + 1].need_lo : 0;
    return CME213_OK;
}

int build_split(cme213_shard *s) {
    // A_shard = A_loc (owned columns) + A_rem (the rest); row pointers by the exclusive scan, entries in row order
    cme213_ctx *ctx = s->ctx;
    const int64_t n = s->nloc;
    uint32_t *cl = nullptr, *cr = nullptr;
    CME213_CUDA(cudaMalloc(&cl, (size_t)(n + 1) * 4));
    CME213_CUDA(cudaMalloc(&cr, (size_t)(n + 1) * 4));
    CME213_CUDA(cudaMalloc(&s->loc_ptr, (size_t)(n + 1) * 4));
    CME213_CUDA(cudaMalloc(&s->rem_ptr, (size_t)(n + 1) * 4));
    split_count_kernel<<<(unsigned)div_up(n + 1, 256), 256>>>(n, s->row_ptr, s->col, (int)s->lo, (int)s->hi, cl, cr);
    CME213_LAUNCHED(ctx, "split_count_kernel");
    int rc = cme213_exclusive_scan_u32(ctx, cl, (uint32_t *)s->loc_ptr, n + 1, nullptr);
    if (!rc) rc = cme213_exclusive_scan_u32(ctx, cr, (uint32_t *)s->rem_ptr, n + 1, nullptr);
    if (rc) return rc;
    int32_t tl = 0, tr = 0;
    CME213_CUDA(cudaMemcpy(&tl, s->loc_ptr + n, 4, cudaMemcpyDeviceToHost));
    CME213_CUDA(cudaMemcpy(&tr, s->rem_ptr + n, 4, cudaMemcpyDeviceToHost));
    CME213_CUDA(cudaMalloc(&s->loc_col, (size_t)(tl > 0 ? tl : 4) * 4));
    CME213_CUDA(cudaMalloc(&s->loc_val, (size_t)(tl > 0 ? tl : 4) * 8));
    CME213_CUDA(cudaMalloc(&s->rem_col, (size_t)(tr > 0 ? tr : 4) * 4));
    CME213_CUDA(cudaMalloc(&s->rem_val, (size_t)(tr > 0 ? tr : 4) * 8));
    split_fill_kernel<<<(unsigned)div_up(n, 256), 256>>>(n, s->row_ptr, s->col, s->val, (int)s->lo, (int)s->hi, s->loc_ptr, s->loc_col,
                                                         s->loc_val, s->rem_ptr, s->rem_col, s->rem_val);
    CME213_LAUNCHED(ctx, "split_fill_kernel");
    CME213_CUDA(cudaDeviceSynchronize());
    cudaFree(cl);
    cudaFree(cr);
    CME213_CUDA(cudaStreamCreateWithFlags(&s->comm, cudaStreamNonBlocking));
    CME213_CUDA(cudaEventCreateWithFlags(&s->ev_done, cudaEventDisableTiming));
    for (int i = 0; i < 2; i++) CME213_CUDA(cudaEventCreateWithFlags(&s->ev_sent[i], cudaEventDisableTiming));
    return CME213_OK;
}

int finish_connect(cme213_shard *s, const std::vector<Blob> &all) {
    int rc = decide_mode(s, all);
    if (rc) return rc;
    if (s->mode == CME213_SHARD_GATHER_SPLIT && s->world > 1 && s->nloc > 0) {
        DeviceGuard g(s->ctx->device);
        rc = build_split(s);
        if (rc) return rc;
    }
    s->connected = true;
    return CME213_OK;
}

PeerFlags peer_flags(const cme213_shard *s) {
    PeerFlags f;
    for (int r = 0; r < 64; r++) f.p[r] = r < s->world && r != s->rank ? s->psync[r] : nullptr;
    return f;
}

int send_slice(cme213_shard *s, int dst, cudaStream_t st) {
    if (s->nloc > 0)
        for (int k = 1; k < s->world; k++) {
            const int r = (s->rank + k) % s->world;   // every rank starts with a different peer
            CME213_CUDA(cudaMemcpyAsync(s->px[dst][r] + s->lo, s->x[dst] + s->lo, (size_t)s->nloc * 8, cudaMemcpyDeviceToDevice, st));
        }
    gather_signal_kernel<<<1, 64, 0, st>>>(s->sync, peer_flags(s), s->world, s->rank);
    CME213_LAUNCHED(s->ctx, "gather_signal_kernel");
    return CME213_OK;
}
}  // namespace

CME213_API int cme213_partition_rows(int64_t num_rows, int world, const int32_t *h_row_ptr, int64_t align,
                                     int64_t *bounds_out) {
    if (num_rows < 0 || world < 1 || !bounds_out) { set_error("partition_rows: bad arguments"); return CME213_ERR_INVALID; }
    if (align < 1) align = 1;
    bounds_out[0] = 0;
    bounds_out[world] = num_rows;
    const double total = h_row_ptr ? (double)h_row_ptr[num_rows] - (double)h_row_ptr[0] : (double)num_rows;
    for (int r = 1; r < world; r++) {
        int64_t b;
        if (h_row_ptr) {
            // first row whose prefix reaches r/world of the nonzeros (binary search on the row pointers)
            const double want = (double)h_row_ptr[0] + total * r / world;
            int64_t lo = 0, hi = num_rows;
            while (lo < hi) { const int64_t mid = (lo + hi) / 2; if ((double)h_row_ptr[mid] < want) lo = mid + 1; else hi = mid; }
            b = lo;
        } else {
            b = (int64_t)r * ((num_rows + world - 1) / world);   // ceil(rows/P) rows each, the last ranks take what is left (2dHeat.cpp:284-306)
        }
        b = (b + align / 2) / align * align;
        if (b > num_rows) b = num_rows;
        if (b < bounds_out[r - 1]) b = bounds_out[r - 1];
        bounds_out[r] = b;
    }
    return CME213_OK;
}

CME213_API int cme213_shard_create_f64(cme213_ctx *ctx, int rank, int world, int64_t num_rows_total, const int64_t *bounds,
                                       const int32_t *d_row_ptr_local, const int32_t *d_col, const double *d_val,
                                       int mode, cme213_shard **out) {
    CME213_CHECK_CTX(ctx);
    if (!out) { set_error("shard_create: null out pointer"); return CME213_ERR_INVALID; }
    *out = nullptr;
    if (world < 1 || world > 64 || rank < 0 || rank >= world || !bounds || num_rows_total < 0 || mode < 0 || mode > 3) {
        set_error("shard_create: bad arguments (world 1..64)");
        return CME213_ERR_INVALID;
    }
    if (bounds[0] != 0 || bounds[world] != num_rows_total) { set_error("shard_create: bounds must run from 0 to num_rows"); return CME213_ERR_INVALID; }
    for (int r = 0; r < world; r++)
        if (bounds[r + 1] < bounds[r]) { set_error("shard_create: bounds must not decrease"); return CME213_ERR_INVALID; }
    if (num_rows_total > 0x7FFFFFFFLL) { set_error("shard_create: column indices are 32-bit"); return CME213_ERR_UNSUPPORTED; }
    cme213_shard *s = new cme213_shard();
    s->ctx = ctx; s->rank = rank; s->world = world; s->rows_total = num_rows_total; s->wanted_mode = mode;
    s->bounds.assign(bounds, bounds + world + 1);
    s->lo = bounds[rank]; s->hi = bounds[rank + 1]; s->nloc = s->hi - s->lo;
    s->row_ptr = d_row_ptr_local; s->col = d_col; s->val = d_val;
    if (s->nloc > 0 && (!d_row_ptr_local || !d_col || !d_val)) { delete s; set_error("shard_create: null matrix pointer"); return CME213_ERR_INVALID; }
    DeviceGuard g(ctx->device);
    auto fail = [&](int rc) { cme213_shard_destroy(s); return rc; };
    for (int i = 0; i < 2; i++) {
        cudaError_t e = cudaMalloc(&s->x[i], (size_t)(num_rows_total > 0 ? num_rows_total : 1) * 8);
        if (e != cudaSuccess) return fail(cuda_fail(e, "cudaMalloc(shard vector)"));
    }
    cudaError_t e = cudaMalloc(&s->sync, sync_bytes(world));
    if (e != cudaSuccess) return fail(cuda_fail(e, "cudaMalloc(shard sync)"));
    if ((e = cudaMemset(s->sync, 0, sync_bytes(world))) != cudaSuccess) return fail(cuda_fail(e, "cudaMemset"));
    // ghost rows this shard reads: from the smallest / largest column index it holds
    if (s->nloc > 0) {
        int32_t ends[2] = {0, 0};
        if ((e = cudaMemcpy(&ends[0], d_row_ptr_local, 4, cudaMemcpyDeviceToHost)) != cudaSuccess) return fail(cuda_fail(e, "cudaMemcpy"));
        if ((e = cudaMemcpy(&ends[1], d_row_ptr_local + s->nloc, 4, cudaMemcpyDeviceToHost)) != cudaSuccess) return fail(cuda_fail(e, "cudaMemcpy"));
        if (ends[0] != 0) { set_error("shard_create: the shard's row pointers must be rebased to 0"); return fail(CME213_ERR_INVALID); }
        s->nnz = ends[1];
        if (s->nnz > 0) {
            int *mm = (int *)(s->sync + W_STATUS + 2);   // two spare words of the status line
            int init[2] = {0x7FFFFFFF, -1};
            if ((e = cudaMemcpy(mm, init, 8, cudaMemcpyHostToDevice)) != cudaSuccess) return fail(cuda_fail(e, "cudaMemcpy"));
            col_range_kernel<<<ctx->sm_count * 8, 256>>>(s->nnz, d_col, mm, mm + 1);
            ctx->launches++;
            int got[2];
            if ((e = cudaMemcpy(got, mm, 8, cudaMemcpyDeviceToHost)) != cudaSuccess) return fail(cuda_fail(e, "col_range_kernel"));
            if (got[0] < 0 || got[1] >= num_rows_total) { set_error("shard_create: column index outside [0, %lld)", (long long)num_rows_total); return fail(CME213_ERR_INVALID); }
            s->need_lo = got[0] < s->lo ? s->lo - got[0] : 0;
            s->need_hi = got[1] >= s->hi ? got[1] - s->hi + 1 : 0;
            cudaMemset(mm, 0, 8);
        }
    }
    s->px[0].assign(world, nullptr); s->px[1].assign(world, nullptr); s->psync.assign(world, nullptr);
    s->px[0][rank] = s->x[0]; s->px[1][rank] = s->x[1]; s->psync[rank] = s->sync;
    if (world == 1) { s->mode = CME213_SHARD_HALO; s->connected = true; }
    *out = s;
    return CME213_OK;
}

CME213_API int cme213_shard_destroy(cme213_shard *s) {
    if (!s) return CME213_OK;
    DeviceGuard g(s->ctx->device);
    cudaDeviceSynchronize();
    if (s->ipc_opened)
        for (int r = 0; r < s->world; r++) {
            if (r == s->rank) continue;
            if (s->px[0][r]) cudaIpcCloseMemHandle(s->px[0][r]);
            if (s->px[1][r]) cudaIpcCloseMemHandle(s->px[1][r]);
            if (s->psync[r]) cudaIpcCloseMemHandle(s->psync[r]);
        }
    for (void *p : {(void *)s->x[0], (void *)s->x[1], (void *)s->sync, (void *)s->loc_ptr, (void *)s->loc_col, (void *)s->loc_val,
                    (void *)s->rem_ptr, (void *)s->rem_col, (void *)s->rem_val})
        if (p) cudaFree(p);
    if (s->comm) cudaStreamDestroy(s->comm);
    if (s->ev_done) cudaEventDestroy(s->ev_done);
    for (int i = 0; i < 2; i++) if (s->ev_sent[i]) cudaEventDestroy(s->ev_sent[i]);
    cudaGetLastError();
    delete s;
    return CME213_OK;
}

CME213_API int cme213_shard_handle_bytes(void) { return (int)sizeof(Blob); }

CME213_API int cme213_shard_export(cme213_shard *s, void *blob_out) {
    if (!s || !blob_out) { set_error("shard_export: null pointer"); return CME213_ERR_INVALID; }
    DeviceGuard g(s->ctx->device);
    Blob b;
    memset(&b, 0, sizeof(b));
    CME213_CUDA(cudaIpcGetMemHandle(&b.x[0], s->x[0]));
    CME213_CUDA(cudaIpcGetMemHandle(&b.x[1], s->x[1]));
    CME213_CUDA(cudaIpcGetMemHandle(&b.sync, s->sync));
    b.lo = s->lo; b.hi = s->hi; b.need_lo = s->need_lo; b.need_hi = s->need_hi; b.rank = s->rank; b.world = s->world;
    memcpy(blob_out, &b, sizeof(b));
    return CME213_OK;
}

CME213_API int cme213_shard_connect(cme213_shard *s, const void *blobs) {
    if (!s || !blobs) { set_error("shard_connect: null pointer"); return CME213_ERR_INVALID; }
    if (s->connected) return CME213_OK;
    std::vector<Blob> all(s->world);
    memcpy(all.data(), blobs, sizeof(Blob) * (size_t)s->world);
    for (int r = 0; r < s->world; r++)
        if (all[r].rank != r || all[r].world != s->world || all[r].lo != s->bounds[r] || all[r].hi != s->bounds[r + 1]) {
            set_error("shard_connect: handle %d does not belong to rank %d of this partition", r, r);
            return CME213_ERR_INVALID;
        }
    DeviceGuard g(s->ctx->device);
    // decide first (every rank reaches the same decision from the same handles), then map only what the mode needs
    int rc = decide_mode(s, all);
    if (rc) return rc;
    for (int r = 0; r < s->world; r++) {
        if (r == s->rank) continue;
        if (s->mode == CME213_SHARD_HALO && r != s->rank - 1 && r != s->rank + 1) continue;
        CME213_CUDA(cudaIpcOpenMemHandle((void **)&s->px[0][r], all[r].x[0], cudaIpcMemLazyEnablePeerAccess));
        CME213_CUDA(cudaIpcOpenMemHandle((void **)&s->px[1][r], all[r].x[1], cudaIpcMemLazyEnablePeerAccess));
        CME213_CUDA(cudaIpcOpenMemHandle((void **)&s->psync[r], all[r].sync, cudaIpcMemLazyEnablePeerAccess));
        s->ipc_opened = true;
    }
    return finish_connect(s, all);
}

CME213_API int cme213_shard_connect_local(cme213_shard *const *shards, int world) {
    if (!shards || world < 1) { set_error("shard_connect_local: bad arguments"); return CME213_ERR_INVALID; }
    std::vector<Blob> all(world);
    for (int r = 0; r < world; r++) {
        cme213_shard *s = shards[r];
        if (!s || s->rank != r || s->world != world) { set_error("shard_connect_local: shards must be given in rank order"); return CME213_ERR_INVALID; }
        memset(&all[r], 0, sizeof(Blob));
        all[r].lo = s->lo; all[r].hi = s->hi; all[r].need_lo = s->need_lo; all[r].need_hi = s->need_hi; all[r].rank = r; all[r].world = world;
    }
    for (int r = 0; r < world; r++) {
        cme213_shard *s = shards[r];
        if (s->connected) continue;
        DeviceGuard g(s->ctx->device);
        for (int q = 0; q < world; q++) {
            if (q == r) continue;
            const int pd = shards[q]->ctx->device;
            if (pd != s->ctx->device) {
                int can = 0;
                CME213_CUDA(cudaDeviceCanAccessPeer(&can, s->ctx->device, pd));
                if (!can) { set_error("shard_connect_local: device %d cannot access device %d", s->ctx->device, pd); return CME213_ERR_UNSUPPORTED; }
                cudaError_t e = cudaDeviceEnablePeerAccess(pd, 0);
                if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) return cuda_fail(e, "cudaDeviceEnablePeerAccess");
                cudaGetLastError();
            }
            s->px[0][q] = shards[q]->x[0]; s->px[1][q] = shards[q]->x[1]; s->psync[q] = shards[q]->sync;
        }
        int rc = finish_connect(s, all);
        if (rc) return rc;
    }
    return CME213_OK;
}

CME213_API int cme213_shard_mode(const cme213_shard *s) { return s ? s->mode : CME213_ERR_INVALID; }
CME213_API double *cme213_shard_vector(cme213_shard *s, int which) { return (s && (which == 0 || which == 1)) ? s->x[which] : nullptr; }
CME213_API int cme213_shard_rows(const cme213_shard *s, int64_t *row_begin, int64_t *row_end) {
    if (!s) { set_error("shard_rows: null shard"); return CME213_ERR_INVALID; }
    if (row_begin) *row_begin = s->lo;
    if (row_end) *row_end = s->hi;
    return CME213_OK;
}
CME213_API int64_t cme213_shard_comm_bytes_per_step(const cme213_shard *s) {
    if (!s || s->world == 1) return 0;
    if (s->mode == CME213_SHARD_HALO) return (s->push_lo + s->push_hi) * 8;
    return (int64_t)(s->world - 1) * s->nloc * 8;
}

CME213_API int cme213_shard_step(cme213_shard *s, int src, void *stream) {
    if (!s || (src != 0 && src != 1)) { set_error("shard_step: bad arguments"); return CME213_ERR_INVALID; }
    if (!s->connected) { set_error("shard_step: connect the shards first"); return CME213_ERR_INVALID; }
    cme213_ctx *ctx = s->ctx;
    DeviceGuard g(ctx->device);
    cudaStream_t st = (cudaStream_t)stream;
    const int dst = 1 - src;
    double *y = s->x[dst] + s->lo;   // index r = local row r
    if (s->world == 1) return cme213_spmv_csr_f64(ctx, 0, s->nloc, s->row_ptr, s->col, s->val, s->x[src], y, CME213_CSR_STREAM, stream);
    if (s->mode == CME213_SHARD_HALO) {
        const int B = cme213_internal_shard_block_rows();
        const int64_t nblocks = div_up(s->nloc, B);
        ShardSync k;
        memset(&k, 0, sizeof(k));
        const bool has_lo = s->rank > 0, has_hi = s->rank + 1 < s->world;
        // a side's boundary = the rows the neighbour needs from me or that read the neighbour's rows, whichever reaches further
        // (need_* counts rows of the NEIGHBOUR and may exceed this rank's own row count: then every row here is a boundary row)
        int64_t side_lo = has_lo ? (s->push_lo > s->need_lo ? s->push_lo : s->need_lo) : 0;
        int64_t side_hi = has_hi ? (s->push_hi > s->need_hi ? s->push_hi : s->need_hi) : 0;
        if (side_lo > s->nloc) side_lo = s->nloc;
        if (side_hi > s->nloc) side_hi = s->nloc;
        k.peer_lo = has_lo && side_lo > 0 ? s->px[dst][s->rank - 1] + s->lo : nullptr;
        k.peer_hi = has_hi && side_hi > 0 ? s->px[dst][s->rank + 1] + s->lo : nullptr;
        k.push_lo_end = s->push_lo; k.push_hi_begin = s->nloc - s->push_hi;
        k.side_lo_end = side_lo; k.side_hi_begin = s->nloc - side_hi;
        k.nb_lo = k.peer_lo ? (unsigned)div_up(side_lo, B) : 0;
        k.nb_hi = k.peer_hi ? (unsigned)(nblocks - (s->nloc - side_hi) / B) : 0;
        k.cnt_lo = s->sync + W_CNT_LO; k.cnt_hi = s->sync + W_CNT_HI; k.status = s->sync + W_STATUS;
        if (has_lo) { k.flag_lo = s->sync + W_FLAG0 + (size_t)(s->rank - 1) * LINE; k.peer_flag_lo = s->psync[s->rank - 1] + W_FLAG0 + (size_t)s->rank * LINE; }
        if (has_hi) { k.flag_hi = s->sync + W_FLAG0 + (size_t)(s->rank + 1) * LINE; k.peer_flag_hi = s->psync[s->rank + 1] + W_FLAG0 + (size_t)s->rank * LINE; }
        k.timeout_cycles = s->timeout_cycles;
        return cme213_internal_spmv_csr_f64_shard(ctx, s->nloc, s->row_ptr, s->col, s->val, s->x[src], y, k, st);
    }
    if (s->mode == CME213_SHARD_GATHER) {
        gather_wait_kernel<<<1, 64, 0, st>>>(s->sync, s->world, s->rank, s->timeout_cycles);
        CME213_LAUNCHED(ctx, "gather_wait_kernel");
        int rc = cme213_spmv_csr_f64(ctx, 0, s->nloc, s->row_ptr, s->col, s->val, s->x[src], y, CME213_CSR_STREAM, stream);
        if (rc) return rc;
        return send_slice(s, dst, st);
    }
    // GATHER_SPLIT: local-column block first (needs nothing remote), overlapped with the previous pass's copies on s->comm
    if (s->sent_valid[dst]) CME213_CUDA(cudaStreamWaitEvent(st, s->ev_sent[dst], 0));   // x[dst] is still being sent from (two passes ago)
    int rc = cme213_spmv_csr_f64(ctx, 0, s->nloc, s->loc_ptr, s->loc_col, s->loc_val, s->x[src], y, CME213_CSR_STREAM, stream);
    if (rc) return rc;
    gather_wait_kernel<<<1, 64, 0, st>>>(s->sync, s->world, s->rank, s->timeout_cycles);
    CME213_LAUNCHED(ctx, "gather_wait_kernel");
    rc = cme213_internal_spmv_csr_f64_acc(ctx, s->nloc, s->rem_ptr, s->rem_col, s->rem_val, s->x[src], y, st);
    if (rc) return rc;
    CME213_CUDA(cudaEventRecord(s->ev_done, st));
    CME213_CUDA(cudaStreamWaitEvent(s->comm, s->ev_done, 0));
    rc = send_slice(s, dst, s->comm);
    if (rc) return rc;
    CME213_CUDA(cudaEventRecord(s->ev_sent[dst], s->comm));
    s->sent_valid[dst] = true;
    return CME213_OK;
}

CME213_API int cme213_shard_iterate(cme213_shard *s, int nr_iterations, void *stream) {
    if (!s || nr_iterations < 0 || nr_iterations % 2) {
        set_error("shard_iterate: nr_iterations must be even (pagerank.cu:61,127 assert the same)");
        return CME213_ERR_INVALID;
    }
    for (int it = 0; it < nr_iterations; it++) {
        int rc = cme213_shard_step(s, it & 1, stream);
        if (rc) return rc;
    }
    return CME213_OK;
}

CME213_API int cme213_shard_finish(cme213_shard *s, void *stream) {
    if (!s) { set_error("shard_finish: null shard"); return CME213_ERR_INVALID; }
    DeviceGuard g(s->ctx->device);
    CME213_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    if (s->comm) CME213_CUDA(cudaStreamSynchronize(s->comm));
    u64 status = 0;
    CME213_CUDA(cudaMemcpy(&status, s->sync + W_STATUS, 8, cudaMemcpyDeviceToHost));
    if (status) {
        set_error("shard: rank %d gave up waiting for a neighbour (a peer did not run the same number of passes, or died)", s->rank);
        return CME213_ERR_TIMEOUT;
    }
    return CME213_OK;
}
