// integrate.cu -- the kick/drift kernels that bracket a force evaluation in the Leapfrog stepper, per target
// slice (the caller side of the path, needed so that a sharded step never leaves
// the GPU).  Arithmetic mirrors the reference's Verlet_DKD kernels exactly:
//     DKD_Verlet_1  src/IntegrationMethods.cu:72-79   pos += vel * (dt * 0.5)       -- the literal 0.5 is a
//     DKD_Verlet_2  src/IntegrationMethods.cu:81-93   vel += dt * acc (FP32 FMA);      double, so the drift is
//                                                     pos += vel * (dt * 0.5)          evaluated in FP64 and
//                                                                                      rounded once
// Only x, y, z are touched; pos.w (mass) and vel.w stay as they are.
#include "common.cuh"

namespace {

__device__ __forceinline__ float drift1(float p, float v, double hdt) {
    return (float)__fma_rn((double)v, hdt, (double)p);
}

__global__ void verlet_dkd_drift_kernel(float4* __restrict__ pos, const float4* __restrict__ vel, int begin, int count,
                                        float dt) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    const double hdt = (double)dt * 0.5;
    float4 p = pos[begin + i];
    const float4 v = vel[i];
    p.x = drift1(p.x, v.x, hdt);
    p.y = drift1(p.y, v.y, hdt);
    p.z = drift1(p.z, v.z, hdt);
    pos[begin + i] = p;
}

__global__ void verlet_dkd_kick_drift_kernel(float4* __restrict__ pos, float4* __restrict__ vel,
                                             const float4* __restrict__ acc, int begin, int count, float dt) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    const double hdt = (double)dt * 0.5;
    float4 p = pos[begin + i];
    float4 v = vel[i];
    const float4 a = acc[i];
    v.x = fmaf(dt, a.x, v.x);
    v.y = fmaf(dt, a.y, v.y);
    v.z = fmaf(dt, a.z, v.z);
    p.x = drift1(p.x, v.x, hdt);
    p.y = drift1(p.y, v.y, hdt);
    p.z = drift1(p.z, v.z, hdt);
    vel[i] = v;
    pos[begin + i] = p;
}

}  // namespace

extern "C" {

int nb_verlet_dkd_drift(nb_context* ctx, float* pos, const float* vel, int begin, int count, float dt, void* stream) {
    if (!ctx) return NB_ERR_INVALID;
    if (begin < 0 || count < 0) return nb_fail(ctx, NB_ERR_INVALID, "bad slice");
    if (count == 0) return NB_OK;
    verlet_dkd_drift_kernel<<<(count + 255) / 256, 256, 0, (cudaStream_t)stream>>>((float4*)pos, (const float4*)vel,
                                                                                  begin, count, dt);
    NB_CUDA(ctx, cudaGetLastError());
    return NB_OK;
}

int nb_verlet_dkd_kick_drift(nb_context* ctx, float* pos, float* vel, const float* acc, int begin, int count, float dt,
                             void* stream) {
    if (!ctx) return NB_ERR_INVALID;
    if (begin < 0 || count < 0) return nb_fail(ctx, NB_ERR_INVALID, "bad slice");
    if (count == 0) return NB_OK;
    verlet_dkd_kick_drift_kernel<<<(count + 255) / 256, 256, 0, (cudaStream_t)stream>>>(
        (float4*)pos, (float4*)vel, (const float4*)acc, begin, count, dt);
    NB_CUDA(ctx, cudaGetLastError());
    return NB_OK;
}
}
