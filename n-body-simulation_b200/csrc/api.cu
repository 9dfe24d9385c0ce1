// api.cu -- the extern "C" surface declared in include/nbody_b200.h (lifetime + direct path).
#include "common.cuh"

char g_nb_create_error[512] = {0};

int nb_direct_launch(nb_context* ctx, const float* pos, int n, int target_begin, int target_count, float* acc,
                     int dim, float softening, int flags, cudaStream_t stream);  // direct.cu
void nb_bh_destroy(nb_context* ctx);                                              // bh.cu
void nb_bh3_destroy(nb_context* ctx);                                             // bh3d.cu

extern "C" {

int nb_create(nb_context** out, int device, int max_bodies) {
    if (!out) return nb_fail(nullptr, NB_ERR_INVALID, "nb_create: out is NULL");
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
        return nb_fail(nullptr, NB_ERR_NO_DEVICE, "nb_create: no CUDA device (this engine has no CPU fallback)", e);
    if (device < 0 || device >= count) return nb_fail(nullptr, NB_ERR_INVALID, "nb_create: device out of range");
    nb_device_guard guard;
    e = guard.enter(device);
    if (e != cudaSuccess) return nb_fail(nullptr, NB_ERR_CUDA, "cudaSetDevice", e);
    cudaDeviceProp prop;
    e = cudaGetDeviceProperties(&prop, device);
    if (e != cudaSuccess) return nb_fail(nullptr, NB_ERR_CUDA, "cudaGetDeviceProperties", e);
    if (prop.major != 10)
        return nb_fail(nullptr, NB_ERR_NO_DEVICE, "nb_create: kernels are built for sm_100a only (need a B200)");
    nb_context* ctx = new nb_context();
    ctx->device = device;
    ctx->sm_count = prop.multiProcessorCount;
    e = cudaStreamCreateWithFlags(&ctx->host_stream, cudaStreamNonBlocking);
    if (e != cudaSuccess) {
        delete ctx;
        return nb_fail(nullptr, NB_ERR_CUDA, "cudaStreamCreate", e);
    }
    ctx->max_bodies = max_bodies > 0 ? max_bodies : 0;
    *out = ctx;
    return NB_OK;
}

int nb_destroy(nb_context* ctx) {
    if (!ctx) return NB_OK;
    nb_device_guard guard;
    guard.enter(ctx->device);
    cudaDeviceSynchronize();
    nb_bh_destroy(ctx);
    nb_bh3_destroy(ctx);
    if (ctx->direct_graph) cudaGraphExecDestroy((cudaGraphExec_t)ctx->direct_graph);
    if (ctx->direct_capture_stream) cudaStreamDestroy((cudaStream_t)ctx->direct_capture_stream);
    if (ctx->partial) cudaFree(ctx->partial);
    if (ctx->tile_flags) cudaFree(ctx->tile_flags);
    if (ctx->stats) cudaFree(ctx->stats);
    if (ctx->dev_pos) cudaFree(ctx->dev_pos);
    if (ctx->dev_acc) cudaFree(ctx->dev_acc);
    if (ctx->host_stream) cudaStreamDestroy(ctx->host_stream);
    delete ctx;
    return NB_OK;
}

const char* nb_last_error(const nb_context* ctx) { return ctx ? ctx->err : g_nb_create_error; }

int nb_device_info(const nb_context* ctx, int* device, int* sm_count, uint64_t* scratch_bytes) {
    if (!ctx) return NB_ERR_INVALID;
    if (device) *device = ctx->device;
    if (sm_count) *sm_count = ctx->sm_count;
    if (scratch_bytes)
        *scratch_bytes = ctx->tile_flags_bytes + ctx->partial_bytes + ctx->stats_bytes + ctx->dev_pos_bytes + ctx->dev_acc_bytes +
                         ctx->bh_bytes + ctx->bh3_bytes;
    return NB_OK;
}

int nb_peer_alloc(nb_context* ctx, uint64_t bytes, void** ptr, unsigned char handle[64]) {
    if (!ctx || !ptr || !handle || bytes == 0) return ctx ? nb_fail(ctx, NB_ERR_INVALID, "nb_peer_alloc: bad argument") : NB_ERR_INVALID;
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    NB_DEVICE(ctx);
    // small cudaMalloc blocks share a 2 MiB page with other allocations and an IPC handle exports whole pages: give
    // every shared buffer pages of its own
    if (bytes < (2ull << 20)) bytes = 2ull << 20;
    NB_CUDA(ctx, cudaMalloc(ptr, bytes));
    cudaIpcMemHandle_t h;
    cudaError_t e = cudaIpcGetMemHandle(&h, *ptr);
    if (e != cudaSuccess) {
        cudaFree(*ptr);
        *ptr = nullptr;
        return nb_fail(ctx, NB_ERR_CUDA, "cudaIpcGetMemHandle", e);
    }
    memcpy(handle, &h, 64);
    return NB_OK;
}

int nb_peer_open(nb_context* ctx, const unsigned char handle[64], void** ptr) {
    if (!ctx || !ptr || !handle) return ctx ? nb_fail(ctx, NB_ERR_INVALID, "nb_peer_open: bad argument") : NB_ERR_INVALID;
    NB_DEVICE(ctx);
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, 64);
    NB_CUDA(ctx, cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
    return NB_OK;
}

int nb_peer_close(nb_context* ctx, void* ptr) {
    if (!ctx) return NB_ERR_INVALID;
    NB_DEVICE(ctx);
    if (ptr) NB_CUDA(ctx, cudaIpcCloseMemHandle(ptr));
    return NB_OK;
}

int nb_peer_free(nb_context* ctx, void* ptr) {
    if (!ctx) return NB_ERR_INVALID;
    NB_DEVICE(ctx);
    if (ptr) NB_CUDA(ctx, cudaFree(ptr));
    return NB_OK;
}

int nb_direct_accel(nb_context* ctx, const float* pos, float* acc, int n, int dim, float softening, int flags,
                    void* stream) {
    return nb_direct_launch(ctx, pos, n, 0, n, acc, dim, softening, flags, (cudaStream_t)stream);
}

int nb_direct_accel_slice(nb_context* ctx, const float* pos, int n, int target_begin, int target_count, float* acc,
                          int dim, float softening, int flags, void* stream) {
    return nb_direct_launch(ctx, pos, n, target_begin, target_count, acc, dim, softening, flags,
                            (cudaStream_t)stream);
}

int nb_direct_accel_host(nb_context* ctx, const float* host_pos, float* host_acc, int n, int dim, float softening,
                         int flags) {
    if (!ctx) return NB_ERR_INVALID;
    if (n < 0 || (n > 0 && (!host_pos || !host_acc))) return nb_fail(ctx, NB_ERR_INVALID, "bad host buffers");
    if (n == 0) return NB_OK;
    NB_DEVICE(ctx);
    size_t bytes = (size_t)n * 16;
    int rc = nb_reserve(ctx, &ctx->dev_pos, &ctx->dev_pos_bytes, bytes);
    if (rc) return rc;
    rc = nb_reserve(ctx, &ctx->dev_acc, &ctx->dev_acc_bytes, bytes);
    if (rc) return rc;
    cudaStream_t s = ctx->host_stream;
    NB_CUDA(ctx, cudaMemcpyAsync(ctx->dev_pos, host_pos, bytes, cudaMemcpyHostToDevice, s));
    rc = nb_direct_launch(ctx, (const float*)ctx->dev_pos, n, 0, n, (float*)ctx->dev_acc, dim, softening, flags, s);
    if (rc) return rc;
    NB_CUDA(ctx, cudaMemcpyAsync(host_acc, ctx->dev_acc, bytes, cudaMemcpyDeviceToHost, s));
    NB_CUDA(ctx, cudaStreamSynchronize(s));
    return NB_OK;
}

int nb_direct_last_launch(const nb_context* ctx, int* jsplit, int* blocks, int* launches) {
    if (!ctx) return NB_ERR_INVALID;
    if (jsplit) *jsplit = ctx->last_jsplit;
    if (blocks) *blocks = ctx->last_blocks;
    if (launches) *launches = ctx->last_launches;
    return NB_OK;
}

}  // extern "C"
