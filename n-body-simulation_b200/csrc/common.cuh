// common.cuh -- engine context, error plumbing and small PTX helpers shared by the sm_100a kernels.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>

#include "../../include/nbody_b200.h"

struct nb_bh_state;   // bh.cu
struct nb_bh3_state;  // bh3d.cu

struct nb_context {
    int device = 0;
    int sm_count = 0;
    char err[512] = {0};
    int max_bodies = 0;

    // direct path scratch
    void* partial = nullptr;      // float4[jsplit][targets] partial sums when the source range is split
    size_t partial_bytes = 0;
    void* tile_flags = nullptr;   // [0..64): m_ref, 1/m_ref; then one byte per source tile (mass hoisting)
    size_t tile_flags_bytes = 0;
    void* stats = nullptr;        // SimulationStats partial sums (stats.cu)
    size_t stats_bytes = 0;
    void* dev_pos = nullptr;      // staging for the *_host entry points
    void* dev_acc = nullptr;
    size_t dev_pos_bytes = 0, dev_acc_bytes = 0;
    int last_jsplit = 0, last_blocks = 0, last_launches = 0;
    // resident CTAs per SM of each direct-kernel variant ON THIS DEVICE, 0 = the shared-memory opt-in has not been
    // made here yet (cudaFuncSetAttribute is per device: a process-wide cache would skip it on a second GPU)
    int direct_occ[2][2][2] = {};

    // nb_direct_step_slice at launch-bound sizes: one captured graph, replayed while the arguments stay the same
    void* direct_graph = nullptr;            // cudaGraphExec_t
    void* direct_capture_stream = nullptr;   // cudaStream_t
    unsigned char direct_graph_key[160] = {0};
    unsigned int scratch_epoch = 0;          // bumped whenever nb_reserve moves a scratch buffer: captured graphs hold the old address

    nb_bh_state* bh = nullptr;
    nb_bh3_state* bh3 = nullptr;
    size_t bh_bytes = 0, bh3_bytes = 0;   // device memory held by the two Barnes-Hut states
    cudaStream_t host_stream = nullptr;  // used by the *_host entry points
};

inline int nb_fail(nb_context* ctx, int code, const char* what, cudaError_t e = cudaSuccess) {
    extern char g_nb_create_error[512];
    char* dst = ctx ? ctx->err : g_nb_create_error;
    if (e != cudaSuccess) snprintf(dst, 512, "%s: %s", what, cudaGetErrorString(e));
    else snprintf(dst, 512, "%s", what);
    return code;
}

#define NB_CUDA(ctx, call)                                                    \
    do {                                                                      \
        cudaError_t e__ = (call);                                             \
        if (e__ != cudaSuccess) return nb_fail((ctx), NB_ERR_CUDA, #call, e__); \
    } while (0)

// Entry points run on the handle's device and leave the caller's current device as they found it (a process that drives
// several GPUs -- or PyTorch, whose allocations follow the current device -- must not find it changed by a library call).
struct nb_device_guard {
    int prev = -1;
    bool changed = false;
    cudaError_t enter(int device) {
        cudaError_t e = cudaGetDevice(&prev);
        if (e != cudaSuccess) return e;
        if (prev == device) return cudaSuccess;
        changed = true;
        return cudaSetDevice(device);
    }
    ~nb_device_guard() {
        if (changed) cudaSetDevice(prev);
    }
};
#define NB_DEVICE(ctx)             \
    nb_device_guard nb_dev_guard__; \
    NB_CUDA((ctx), nb_dev_guard__.enter((ctx)->device))

// Softening: the reference adds eps un-squared to r^2 and includes the self term, so eps = 0 turns every result into
// 0 * inf = NaN (src/DirectComputeKernels.cu:23-30).  A floor of 1e-20 -- far below the FP32
// resolution of any separation that is not itself zero -- keeps the self term at exactly 0 and changes nothing else.
// Negative or NaN softening is an argument error.
inline bool nb_softening_ok(float eps) { return eps >= 0.f; }
inline float nb_effective_softening(float eps) { return eps > 1e-20f ? eps : 1e-20f; }

// Grows a scratch buffer (stream-ordered callers must not have work in flight on the old buffer; the
// engine only grows between evaluations, after a device synchronise).
inline int nb_reserve(nb_context* ctx, void** ptr, size_t* have, size_t want) {
    if (*have >= want) return NB_OK;
    NB_CUDA(ctx, cudaDeviceSynchronize());
    if (*ptr) NB_CUDA(ctx, cudaFree(*ptr));
    *ptr = nullptr;
    *have = 0;
    ctx->scratch_epoch++;
    NB_CUDA(ctx, cudaMalloc(ptr, want));
    *have = want;
    return NB_OK;
}

// ---- packed FP32x2 arithmetic (sm_100: FADD2 / FMUL2 / FFMA2 in SASS) -----------------------------
__device__ __forceinline__ uint64_t pk2(float lo, float hi) {
    uint64_t r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void up2(uint64_t v, float& lo, float& hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ uint64_t fma2(uint64_t a, uint64_t b, uint64_t c) {
    uint64_t d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
__device__ __forceinline__ uint64_t mul2(uint64_t a, uint64_t b) {
    uint64_t d;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
__device__ __forceinline__ uint64_t add2(uint64_t a, uint64_t b) {
    uint64_t d;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
__device__ __forceinline__ uint64_t sub2(uint64_t a, uint64_t b) {
    uint64_t d;
    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
__device__ __forceinline__ float rsqrt_approx(float x) {
    float r;
    asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

// ---- mbarrier + bulk async copy (TMA 1D, UBLKCP in SASS) -------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(smem_dst)),
                 "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
