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
    NB_DEVICE(ctx);
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
    NB_DEVICE(ctx);
    verlet_dkd_kick_drift_kernel<<<(count + 255) / 256, 256, 0, (cudaStream_t)stream>>>(
        (float4*)pos, (float4*)vel, (const float4*)acc, begin, count, dt);
    NB_CUDA(ctx, cudaGetLastError());
    return NB_OK;
}
}

// ------------------------------------------------------------------------------------------------------
// General kick/drift sub-step: every elementwise kernel of src/IntegrationMethods.cu:13-349 is "an optional
// velocity update followed by an optional position update that uses the new velocity", on a target slice.
// The kinds reproduce the reference's mixed FP32/FP64 expression types: a `double`
// literal or coefficient promotes the whole product to FP64 and the result is rounded once on the `+=`.
// ------------------------------------------------------------------------------------------------------
#include "integrate.cuh"

namespace {

__global__ void substep_kernel(float4* __restrict__ pos, float4* __restrict__ vel, const float4* __restrict__ acc,
                               const float4* __restrict__ acc_prev, int begin, int count, SubstepArgs s) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    float4 v = vel[i];
    const float4 a = acc[i];
    float4 ap = make_float4(0.f, 0.f, 0.f, 0.f);
    if (s.kick == NB_KICK_MEAN_F64) ap = acc_prev[i];
    if (s.kick != NB_KICK_NONE) {
        v.x = kick1(s, v.x, a.x, ap.x);
        v.y = kick1(s, v.y, a.y, ap.y);
        v.z = kick1(s, v.z, a.z, ap.z);
        vel[i] = v;
    }
    if (s.drift != NB_DRIFT_NONE) {
        float4 p = pos[begin + i];
        p.x = drift1g(s, p.x, v.x, a.x);
        p.y = drift1g(s, p.y, v.y, a.y);
        p.z = drift1g(s, p.z, v.z, a.z);
        pos[begin + i] = p;
    }
}

}  // namespace

extern "C" int nb_integrate_substep(nb_context* ctx, float* pos, float* vel, const float* acc, const float* acc_prev, int begin,
                                    int count, float dt, int kick_kind, double kick_coef, int drift_kind, double drift_coef,
                                    void* stream) {
    if (!ctx) return NB_ERR_INVALID;
    if (begin < 0 || count < 0) return nb_fail(ctx, NB_ERR_INVALID, "bad slice");
    if (kick_kind < 0 || kick_kind > NB_KICK_MEAN_F64 || drift_kind < 0 || drift_kind > NB_DRIFT_VV_F64)
        return nb_fail(ctx, NB_ERR_INVALID, "unknown kick/drift kind");
    if (kick_kind == NB_KICK_MEAN_F64 && !acc_prev) return nb_fail(ctx, NB_ERR_INVALID, "NB_KICK_MEAN_F64 needs acc_prev");
    if (count == 0) return NB_OK;
    NB_DEVICE(ctx);
    SubstepArgs s{kick_kind, drift_kind, dt, kick_coef, drift_coef};
    substep_kernel<<<(count + 255) / 256, 256, 0, (cudaStream_t)stream>>>((float4*)pos, (float4*)vel, (const float4*)acc,
                                                                         (const float4*)acc_prev, begin, count, s);
    NB_CUDA(ctx, cudaGetLastError());
    return NB_OK;
}
