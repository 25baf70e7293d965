// ctx.cu -- context, error text, workspace.
#include "common.cuh"
#include <string.h>
#include <stdlib.h>

namespace cme213 {
static thread_local char g_err[512] = "";

void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int cuda_fail(cudaError_t e, const char *what) {
    set_error("CUDA error %d (%s) at %s", (int)e, cudaGetErrorString(e), what);
    return e == cudaErrorMemoryAllocation ? CME213_ERR_ALLOC : CME213_ERR_CUDA;
}

int workspace(cme213_ctx *ctx, size_t bytes, void **out, cudaStream_t st) {
    if (ctx->ws_used && ctx->ws_stream != st) {
        cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
        if (cudaStreamIsCapturing(st, &cap) != cudaSuccess) { cudaGetLastError(); cap = cudaStreamCaptureStatusNone; }
        if (cap == cudaStreamCaptureStatusNone) {
            if (!ctx->ws_event) CME213_CUDA(cudaEventCreateWithFlags(&ctx->ws_event, cudaEventDisableTiming));
            // everything the previous user enqueued on its stream (calls are serialised on the host, so it is all there)
            if (cudaEventRecord(ctx->ws_event, ctx->ws_stream) == cudaSuccess) {
                CME213_CUDA(cudaStreamWaitEvent(st, ctx->ws_event, 0));
            } else {
                cudaGetLastError();                      // the previous stream is gone (destroyed or mid-capture)
                CME213_CUDA(cudaDeviceSynchronize());
            }
        }
    }
    ctx->ws_stream = st;
    ctx->ws_used = true;
    if (bytes > ctx->ws_bytes) {
        // Growth path: only on first use or when a larger problem arrives.
        if (ctx->ws) {
            CME213_CUDA(cudaDeviceSynchronize());
            CME213_CUDA(cudaFree(ctx->ws));
            ctx->ws = nullptr;
            ctx->ws_bytes = 0;
        }
        size_t want = (bytes + 0xFFFFF) & ~(size_t)0xFFFFF;  // 1 MiB granules
        CME213_CUDA(cudaMalloc(&ctx->ws, want));
        ctx->ws_bytes = want;
    }
    *out = ctx->ws;
    return CME213_OK;
}
}  // namespace cme213

CME213_API const char *cme213_last_error(void) { return cme213::g_err; }
CME213_API int cme213_version(void) { return 200; }
#ifndef CME213_SOURCE_HASH
#define CME213_SOURCE_HASH "unknown"
#endif
CME213_API const char *cme213_build_id(void) { return CME213_SOURCE_HASH; }

CME213_API int cme213_ctx_create(int device, cme213_ctx **out) {
    if (!out) { cme213::set_error("null out pointer"); return CME213_ERR_INVALID; }
    *out = nullptr;
    int count = 0;
    CME213_CUDA(cudaGetDeviceCount(&count));
    if (device < 0 || device >= count) {
        cme213::set_error("device %d out of range (%d visible)", device, count);
        return CME213_ERR_INVALID;
    }
    cme213::DeviceGuard g(device);
    cudaDeviceProp prop;
    CME213_CUDA(cudaGetDeviceProperties(&prop, device));
    cme213_ctx *c = (cme213_ctx *)calloc(1, sizeof(cme213_ctx));
    if (!c) { cme213::set_error("host allocation failed"); return CME213_ERR_ALLOC; }
    c->device = device;
    c->sm_count = prop.multiProcessorCount;
    *out = c;
    return CME213_OK;
}

CME213_API int cme213_ctx_destroy(cme213_ctx *ctx) {
    if (!ctx) return CME213_OK;
    cme213::DeviceGuard g(ctx->device);
    for (int i = 0; i < cme213_ctx::CSR_CACHE; i++) cme213_csr_plan_destroy(ctx->csr_cache[i].plan);
    if (ctx->ws) cudaFree(ctx->ws);
    if (ctx->ws_event) cudaEventDestroy(ctx->ws_event);
    free(ctx);
    return CME213_OK;
}

CME213_API int cme213_ctx_device(const cme213_ctx *ctx) { return ctx ? ctx->device : -1; }
CME213_API int cme213_ctx_sm_count(const cme213_ctx *ctx) { return ctx ? ctx->sm_count : -1; }
CME213_API int64_t cme213_ctx_launch_count(const cme213_ctx *ctx) { return ctx ? ctx->launches : -1; }

// ---- device memory that other processes of the same box can map (multi-GPU halo push) ----------------------
// torch's caching allocator hands out pieces of large blocks, which cannot be exported one by one; vectors that a
// neighbouring rank writes into are therefore allocated here, one cudaMalloc each, and exported whole.
CME213_API int cme213_device_malloc(size_t bytes, void **d_out) {
    if (!d_out) { cme213::set_error("device_malloc: null out pointer"); return CME213_ERR_INVALID; }
    CME213_CUDA(cudaMalloc(d_out, bytes ? bytes : 16));
    return CME213_OK;
}
CME213_API int cme213_device_free(void *d_ptr) {
    if (d_ptr) CME213_CUDA(cudaFree(d_ptr));
    return CME213_OK;
}
CME213_API int cme213_ipc_handle_bytes(void) { return (int)sizeof(cudaIpcMemHandle_t); }
CME213_API int cme213_ipc_export(const void *d_ptr, void *handle_out) {
    if (!d_ptr || !handle_out) { cme213::set_error("ipc_export: null pointer"); return CME213_ERR_INVALID; }
    cudaIpcMemHandle_t h;
    CME213_CUDA(cudaIpcGetMemHandle(&h, const_cast<void *>(d_ptr)));
    memcpy(handle_out, &h, sizeof(h));
    return CME213_OK;
}
CME213_API int cme213_ipc_open(const void *handle, void **d_out) {
    if (!handle || !d_out) { cme213::set_error("ipc_open: null pointer"); return CME213_ERR_INVALID; }
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof(h));
    CME213_CUDA(cudaIpcOpenMemHandle(d_out, h, cudaIpcMemLazyEnablePeerAccess));
    return CME213_OK;
}
CME213_API int cme213_ipc_close(void *d_ptr) {
    if (d_ptr) CME213_CUDA(cudaIpcCloseMemHandle(d_ptr));
    return CME213_OK;
}
