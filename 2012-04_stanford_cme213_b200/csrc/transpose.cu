// transpose.cu -- out-of-place tiled transpose for 4- and 8-byte elements (bit-exact copy).
//
// Specification followed: slides/Lecture06.pdf slide 9/20 (out[n*M+m] = in[m*N+n]),
// slides/Lecture07.pdf slides 2-4 (shared-memory tile), my-refs/MatrixTranspose.pdf
// (transposeNoBankConflicts).  The B200 design differs from those Fermi listings:
//
//  * every global access is a 128-bit LDG/STG.  A thread loads V = 16/sizeof(T) consecutive
//    elements from each of V consecutive rows (a VxV micro-block), transposes the micro-block in
//    registers, and hands the V resulting 16-byte vectors to shared memory;
//  * shared memory holds 16-byte units, addressed with an XOR swizzle instead of the +1 padding
//    of the slides, so both the write phase (lanes walk along input columns) and the read phase
//    (lanes walk along input rows) hit 8 distinct 16-byte bank groups per quarter-warp: no
//    bank conflicts and no padding, and LDS/STS are 128-bit as well;
//  * tiles are 64 rows x 256 bytes (64x64 fp32, 64x32 fp64: 16 KiB either way), 256 threads, so that
//    8+ CTAs per SM keep >100 KiB of loads in flight per SM, which HBM3e needs (the first version used
//    64x64 fp64 tiles = 32 KiB and was occupancy-limited to 85 % of copy speed, profiles/r01_notes.md).
//
// HBM traffic is the algorithmic minimum: each element is read once and written once, in full
// 32-byte sectors on both sides.  Matrices whose shape or pointers do not allow 16-byte vectors
// take the scalar 32x32 padded-tile kernel (transpose_generic).
#include "common.cuh"

namespace {

constexpr int TILE_R = 64;      // input rows per tile
constexpr int TILE_BYTES_C = 256;  // input row segment per tile: 64 fp32 or 32 fp64 columns (16 KiB tiles both ways)
constexpr int TR_THREADS = 256;

template <typename T> struct Vec16;
template <> struct Vec16<uint32_t> { using type = uint4; };
template <> struct Vec16<uint64_t> { using type = ulonglong2; };

template <typename T>
__global__ void __launch_bounds__(TR_THREADS)
transpose_vec_kernel(const T *__restrict__ in, T *__restrict__ out, int64_t rows, int64_t cols, int64_t ld_in,
                     int64_t ld_out, int tiles_c) {
    constexpr int V = 16 / sizeof(T);             // elements per 16-byte vector: 4 (b32) or 2 (b64)
    constexpr int TILE_C = TILE_BYTES_C / sizeof(T);  // input columns per tile: 64 or 32
    constexpr int NBC = TILE_C / V;               // micro-block columns per tile: 16
    constexpr int NBR = TILE_R / V;               // micro-block rows per tile: 16 or 32
    constexpr int BLOCKS = NBC * NBR;             // micro-blocks per tile: 256 or 512
    constexpr int PER_THREAD = BLOCKS / TR_THREADS;
    using V16 = typename Vec16<T>::type;
    // unit (o, R): o = output row inside the tile (0..TILE_C-1), R = micro-block row index.
    __shared__ V16 smem[TILE_C * NBR];

    // blockIdx.x walks tiles along the input row direction fastest.
    const int64_t tile_id = blockIdx.x;
    const int64_t tr = tile_id / tiles_c, tc = tile_id % tiles_c;
    const int64_t r0 = tr * TILE_R, c0 = tc * TILE_C;
    const int tid = threadIdx.x;

    // ---- load phase: lanes walk along input columns (c fastest) ----
    T blk[PER_THREAD][V][V];
#pragma unroll
    for (int b = 0; b < PER_THREAD; b++) {
        const int id = b * TR_THREADS + tid;
        const int c = id % NBC, R = id / NBC;
        const int64_t gc = c0 + (int64_t)c * V;
#pragma unroll
        for (int i = 0; i < V; i++) {
            const int64_t gr = r0 + (int64_t)R * V + i;
            V16 v = {};
            if (gr < rows && gc < cols) v = cme213::ld_stream(reinterpret_cast<const V16 *>(in + gr * ld_in + gc));
            const T *e = reinterpret_cast<const T *>(&v);
#pragma unroll
            for (int j = 0; j < V; j++) blk[b][i][j] = e[j];
        }
    }
    // ---- register transpose + swizzled 128-bit shared stores ----
#pragma unroll
    for (int b = 0; b < PER_THREAD; b++) {
        const int id = b * TR_THREADS + tid;
        const int c = id % NBC, R = id / NBC;
#pragma unroll
        for (int j = 0; j < V; j++) {
            V16 v;
            T *e = reinterpret_cast<T *>(&v);
#pragma unroll
            for (int i = 0; i < V; i++) e[i] = blk[b][i][j];
            const int o = c * V + j;
            smem[o * NBR + (R ^ (c & 7))] = v;
        }
    }
    __syncthreads();
    // ---- store phase: lanes walk along input rows (R fastest) = along output columns ----
#pragma unroll
    for (int b = 0; b < PER_THREAD; b++) {
        const int id = b * TR_THREADS + tid;
        const int R = id % NBR, c = id / NBR;
        const int64_t orow_base = c0 + (int64_t)c * V;  // output row = input column
        const int64_t ocol = r0 + (int64_t)R * V;       // output column = input row
#pragma unroll
        for (int j = 0; j < V; j++) {
            const int o = c * V + j;
            V16 v = smem[o * NBR + (R ^ (c & 7))];
            const int64_t orow = orow_base + j;
            if (orow < cols && ocol < rows) cme213::st_stream(reinterpret_cast<V16 *>(out + orow * ld_out + ocol), v);
        }
    }
}

// Scalar fallback: any shape, any alignment.  32x32 tile, +1 padding (Lecture07 slide 3).
template <typename T>
__global__ void __launch_bounds__(256)
transpose_generic_kernel(const T *__restrict__ in, T *__restrict__ out, int64_t rows, int64_t cols, int64_t ld_in,
                         int64_t ld_out, int tiles_c) {
    __shared__ T tile[32][33];
    const int64_t tile_id = blockIdx.x;
    const int64_t r0 = (tile_id / tiles_c) * 32, c0 = (tile_id % tiles_c) * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
#pragma unroll
    for (int i = 0; i < 32; i += 8) {
        const int64_t r = r0 + ty + i, c = c0 + tx;
        if (r < rows && c < cols) tile[ty + i][tx] = in[r * ld_in + c];
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 32; i += 8) {
        const int64_t orow = c0 + ty + i, ocol = r0 + tx;
        if (orow < cols && ocol < rows) out[orow * ld_out + ocol] = tile[tx][ty + i];
    }
}

template <typename T>
int transpose_impl(cme213_ctx *ctx, const void *d_in, void *d_out, int64_t rows, int64_t cols, int64_t ld_in,
                   int64_t ld_out, void *stream) {
    CME213_CHECK_CTX(ctx);
    if (rows < 0 || cols < 0 || ld_in < cols || ld_out < rows) {
        cme213::set_error("transpose: bad shape rows=%lld cols=%lld ld_in=%lld ld_out=%lld", (long long)rows,
                          (long long)cols, (long long)ld_in, (long long)ld_out);
        return CME213_ERR_INVALID;
    }
    if (rows == 0 || cols == 0) return CME213_OK;
    if (!d_in || !d_out) { cme213::set_error("transpose: null pointer"); return CME213_ERR_INVALID; }
    if (d_in == d_out) { cme213::set_error("transpose: in-place is not supported"); return CME213_ERR_INVALID; }
    cme213::DeviceGuard g(ctx->device);
    cudaStream_t st = (cudaStream_t)stream;
    constexpr int V = 16 / sizeof(T);
    // 16-byte vectors need aligned rows on both sides and whole vectors per row: a ragged `cols` would make the last
    // load of the last row read past rows*ld_in-(ld_in-cols), which the header does not ask the caller to allocate.
    const bool vec_ok = cme213::aligned16(d_in) && cme213::aligned16(d_out) && rows % V == 0 &&
                        cols % V == 0 && ld_in % V == 0 && ld_out % V == 0;
    if (vec_ok) {
        const int64_t tiles_r = cme213::div_up(rows, TILE_R), tiles_c = cme213::div_up(cols, TILE_BYTES_C / (int)sizeof(T));
        if (tiles_r * tiles_c > 0x7FFFFFFFLL || tiles_c > 0x7FFFFFFFLL) {
            cme213::set_error("transpose: too many tiles");
            return CME213_ERR_UNSUPPORTED;
        }
        transpose_vec_kernel<T><<<(unsigned)(tiles_r * tiles_c), TR_THREADS, 0, st>>>(
            (const T *)d_in, (T *)d_out, rows, cols, ld_in, ld_out, (int)tiles_c);
        CME213_LAUNCHED(ctx, "transpose_vec_kernel");
    } else {
        const int64_t tiles_r = cme213::div_up(rows, 32), tiles_c = cme213::div_up(cols, 32);
        if (tiles_r * tiles_c > 0x7FFFFFFFLL) { cme213::set_error("transpose: too many tiles"); return CME213_ERR_UNSUPPORTED; }
        transpose_generic_kernel<T><<<(unsigned)(tiles_r * tiles_c), 256, 0, st>>>(
            (const T *)d_in, (T *)d_out, rows, cols, ld_in, ld_out, (int)tiles_c);
        CME213_LAUNCHED(ctx, "transpose_generic_kernel");
    }
    return CME213_OK;
}

}  // namespace

CME213_API int cme213_transpose_b32(cme213_ctx *ctx, const void *d_in, void *d_out, int64_t rows, int64_t cols,
                                    int64_t ld_in, int64_t ld_out, void *stream) {
    return transpose_impl<uint32_t>(ctx, d_in, d_out, rows, cols, ld_in, ld_out, stream);
}
CME213_API int cme213_transpose_b64(cme213_ctx *ctx, const void *d_in, void *d_out, int64_t rows, int64_t cols,
                                    int64_t ld_in, int64_t ld_out, void *stream) {
    return transpose_impl<uint64_t>(ctx, d_in, d_out, rows, cols, ld_in, ld_out, stream);
}
