// stats.cu -- SimulationStats on the engine: the diagnostic sums NbodySystem::simulate prints
// (src/NbodySystem.cpp:128-138) -- "moment" = sum m|v|, kinetic energy, and the FP64 all-pairs potential energy.
// Formulae of src/SimulationStats.cu:29-90 kept as they are: pE = - sum_i sum_j m_i m_j / sqrt(1e-6 + r_ij^2) over ALL
// ordered pairs including i = j (so each pair counts twice and every body contributes -m_i^2 * 1000), a 1/r potential
// in 2D as well, z always included.  What differs: a target slice (for sharding), per-thread FP64 accumulators reduced
// per block and then in block order (deterministic; the reference uses one double atomicAdd per thread), no per-call
// cudaMalloc (the reference leaks three per call, :8-10).
#include "common.cuh"

namespace {

constexpr int ST_THREADS = 128;
constexpr int ST_TILE = 512;

__device__ __forceinline__ double block_sum(double v, double* sh) {
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
    __syncthreads();
    double t = 0.0;
    if (threadIdx.x == 0)
        for (int w = 0; w < ST_THREADS / 32; w++) t += sh[w];
    __syncthreads();
    return t;  // valid in thread 0
}

// partial[blockIdx.x] = potential energy of this block's targets against all n sources
__global__ void __launch_bounds__(ST_THREADS) potential_kernel(const float4* __restrict__ pos, int n, int target_begin,
                                                              int target_count, double* __restrict__ partial) {
    __shared__ float4 tile[ST_TILE];
    __shared__ double sh[ST_THREADS / 32];
    const int t = blockIdx.x * ST_THREADS + threadIdx.x;
    const bool valid = t < target_count;
    const float4 b = pos[target_begin + (valid ? t : 0)];
    const double bx = b.x, by = b.y, bz = b.z, bw = b.w;
    double acc = 0.0;
    for (int base = 0; base < n; base += ST_TILE) {
        const int cnt = min(ST_TILE, n - base);
        for (int j = threadIdx.x; j < cnt; j += ST_THREADS) tile[j] = pos[base + j];
        __syncthreads();
#pragma unroll 4
        for (int j = 0; j < cnt; j++) {
            const float4 s = tile[j];
            const double dx = (double)s.x - bx, dy = (double)s.y - by, dz = (double)s.z - bz;   // :77-79
            const double d2 = 0.000001 + dx * dx + dy * dy + dz * dz;                          // :80
            acc -= ((double)s.w * bw) * rsqrt(d2);                                             // :81-83
        }
        __syncthreads();
    }
    const double tot = block_sum(valid ? acc : 0.0, sh);
    if (threadIdx.x == 0) partial[blockIdx.x] = tot;
}

// moment and kinetic energy of the slice (src/SimulationStats.cu:29-60): v = |vel| (2D: x, y only)
__global__ void __launch_bounds__(ST_THREADS) kinetic_kernel(const float4* __restrict__ pos, const float4* __restrict__ vel,
                                                            int target_begin, int target_count, int dim,
                                                            double* __restrict__ partial_m, double* __restrict__ partial_k) {
    __shared__ double sh[ST_THREADS / 32];
    double lm = 0.0, lk = 0.0;
    for (int i = blockIdx.x * ST_THREADS + threadIdx.x; i < target_count; i += gridDim.x * ST_THREADS) {
        const float4 w = vel[i];
        const double m = pos[target_begin + i].w;
        double v;
        if (dim == 2) v = (double)sqrtf(__fmaf_rn(w.x, w.x, __fmul_rn(w.y, w.y)));   // :34 the argument is a float expression (nvcc contracts it to fma(x, x, y*y)), so `sqrt` there is
                                                                   // the FP32 square root; only its result is promoted
        else v = sqrt((double)w.x * w.x + (double)w.y * w.y + (double)w.z * w.z);   // :51
        lm += v * m;
        lk += 0.5 * v * v * m;
    }
    const double tm = block_sum(lm, sh);
    const double tk = block_sum(lk, sh);
    if (threadIdx.x == 0) {
        partial_m[blockIdx.x] = tm;
        partial_k[blockIdx.x] = tk;
    }
}

__global__ void finish_kernel(const double* __restrict__ pm, const double* __restrict__ pk, int nk, const double* __restrict__ pp,
                              int np, double* __restrict__ out) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    double m = 0.0, k = 0.0, p = 0.0;
    for (int i = 0; i < nk; i++) { m += pm[i]; k += pk[i]; }
    for (int i = 0; i < np; i++) p += pp[i];
    out[0] = m;
    out[1] = k;
    out[2] = p;
}

}  // namespace

extern "C" int nb_stats(nb_context* ctx, const float* pos, const float* vel, int n, int target_begin, int target_count, int dim,
                        double out[3], void* stream_) {
    if (!ctx || !out) return NB_ERR_INVALID;
    if (dim != 2 && dim != 3) return nb_fail(ctx, NB_ERR_INVALID, "dim must be 2 or 3");
    if (n < 0 || target_begin < 0 || target_count < 0 || (long long)target_begin + target_count > n)
        return nb_fail(ctx, NB_ERR_INVALID, "target slice outside [0, n)");
    out[0] = out[1] = out[2] = 0.0;
    if (target_count == 0) return NB_OK;
    if (!pos || !vel) return nb_fail(ctx, NB_ERR_INVALID, "null device pointer");
    NB_DEVICE(ctx);
    cudaStream_t stream = (cudaStream_t)stream_;
    const int pblocks = (target_count + ST_THREADS - 1) / ST_THREADS;
    const int kblocks = 64;   // the reference's launch shape (:13,15)
    int rc = nb_reserve(ctx, &ctx->stats, &ctx->stats_bytes, (size_t)(pblocks + 2 * kblocks + 4) * sizeof(double));
    if (rc) return rc;
    double* d = reinterpret_cast<double*>(ctx->stats);
    double *d_out = d, *pm = d + 4, *pk = pm + kblocks, *pp = pk + kblocks;
    kinetic_kernel<<<kblocks, ST_THREADS, 0, stream>>>((const float4*)pos, (const float4*)vel, target_begin, target_count, dim, pm, pk);
    potential_kernel<<<pblocks, ST_THREADS, 0, stream>>>((const float4*)pos, n, target_begin, target_count, pp);
    finish_kernel<<<1, 32, 0, stream>>>(pm, pk, kblocks, pp, pblocks, d_out);
    NB_CUDA(ctx, cudaGetLastError());
    NB_CUDA(ctx, cudaMemcpyAsync(out, d_out, 3 * sizeof(double), cudaMemcpyDeviceToHost, stream));
    NB_CUDA(ctx, cudaStreamSynchronize(stream));
    return NB_OK;
}
