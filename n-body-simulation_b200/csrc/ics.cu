// ics.cu -- device-side initial conditions, so that large systems are set up where they live.
// Same bodies as n-body-simulation_b200/ics.py (splitmix64 of (seed, stream, index): identical on every host and on the
// device), same shapes as the reference's generators:
//   UniformEllipsoid (src/InitialConditions.cu:111-144): phi ~ U[0, 2 pi), lambda ~ U[0.02, 1),
//       pos = centre + radius * lambda * (cos phi, sin phi), z = 0 (or U[-h, h] for the R3 variant), mass, bulk velocity
//   UniformBox (:78-107): x, y (and z in R3, which the reference draws from the y range, :85) uniform in the box,
//       mass ~ U[massRange], speed ~ U[velRange] in a random in-plane direction plus the bulk velocity
//   DiskModel (:153-196): body `offset` becomes the central mass; after ONE force evaluation every body gets the
//       tangential speed v = sqrt(|a| r) -- nb_ic_disk_velocities, the loop of :169-192 with its FP32 arithmetic.
#include "common.cuh"

namespace {

__device__ __forceinline__ unsigned long long splitmix64(unsigned long long x) {
    x += 0x9E3779B97F4A7C15ull;
    unsigned long long z = x;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
__device__ __forceinline__ double uniform01(unsigned long long seed, unsigned long long stream, unsigned long long idx) {
    const unsigned long long base = splitmix64(seed * 0x100000001B3ull + stream);
    return (double)(splitmix64(idx ^ base) >> 11) * (1.0 / 9007199254740992.0);
}

__global__ void ellipsoid_kernel(float4* __restrict__ pos, float4* __restrict__ vel, int offset, int count, unsigned long long seed,
                                 double cx, double cy, double rx, double ry, double zh, float mass, float central_mass,
                                 float vx, float vy) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    const double phi = uniform01(seed, 0, i) * 6.283185307179586;
    const double lam = 0.02 + 0.98 * uniform01(seed, 1, i);
    float4 p;
    p.x = (float)(cx + rx * lam * cos(phi));
    p.y = (float)(cy + ry * lam * sin(phi));
    p.z = zh > 0.0 ? (float)((2.0 * uniform01(seed, 2, i) - 1.0) * zh) : 0.f;
    p.w = mass;
    if (i == 0 && central_mass >= 0.f) p = make_float4((float)cx, (float)cy, 0.f, central_mass);   // :157
    pos[offset + i] = p;
    if (vel) vel[i] = make_float4(vx, vy, 0.f, 0.f);
}

__global__ void box_kernel(float4* __restrict__ pos, float4* __restrict__ vel, int offset, int count, unsigned long long seed,
                           double cx, double cy, double sx, double sy, int dim, double m0, double m1, double v0, double v1,
                           float bvx, float bvy) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    // explicit _rn operations: no FMA contraction, so the doubles are the host generator's bit for bit
    auto lerp = [](double a, double w, double u) { return __dadd_rn(a, __dmul_rn(w, u)); };
    float4 p;
    p.x = (float)lerp(cx, sx, uniform01(seed, 0, i) - 0.5);
    p.y = (float)lerp(cy, sy, uniform01(seed, 1, i) - 0.5);
    p.z = dim == 3 ? (float)lerp(cy, sy, uniform01(seed, 2, i) - 0.5) : 0.f;   // :85 uses the y range for z
    p.w = (float)lerp(m0, m1 - m0, uniform01(seed, 3, i));
    pos[offset + i] = p;
    if (vel) {
        const double phi = uniform01(seed, 4, i) * 6.283185307179586;
        const double v = lerp(v0, v1 - v0, uniform01(seed, 5, i));
        vel[i] = make_float4((float)(v * cos(phi)) + bvx, (float)(v * sin(phi)) + bvy, 0.f, 0.f);   // :101
    }
}

__global__ void disk_velocity_kernel(const float4* __restrict__ pos, const float4* __restrict__ acc, float4* __restrict__ vel,
                                     int begin, int count, float cx, float cy, int dim, float bvx, float bvy) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    const float4 p = pos[begin + i];
    const float4 a4 = acc[i];
    // src/InitialConditions.cu:169-192; explicit _rn operations: the reference runs this loop on the host without FMA
    const float x = __fsub_rn(p.x, cx), y = __fsub_rn(p.y, cy);
    float z = 0.f;
    float r = __fadd_rn(__fadd_rn(__fmul_rn(x, x), __fmul_rn(y, y)), 0.0001f);
    float a = __fadd_rn(__fmul_rn(a4.x, a4.x), __fmul_rn(a4.y, a4.y));
    if (dim == 3) {
        z = p.z;
        r = __fadd_rn(r, __fmul_rn(z, z));
        a = __fadd_rn(a, __fmul_rn(a4.z, a4.z));
    }
    r = __fsqrt_rn(r);
    a = __fsqrt_rn(a);
    const float v = __fsqrt_rn(__fmul_rn(a, r));
    const float vx = __fdiv_rn(__fmul_rn(v, x), r), vy = __fdiv_rn(__fmul_rn(v, y), r), vz = __fdiv_rn(__fmul_rn(v, z), r);
    vel[i] = make_float4(__fadd_rn(-vy, bvx), __fadd_rn(vx, bvy), vz, 0.f);
}

// Standard (src/InitialConditions.cu:13-76, the reference's "debug version"): theta ~ U[0, 2 pi), r = (Exp(12) + 0.02) * 50000,
// pos = centre + r (cos theta, sin theta), z = sin(r / 2000) * 2000 (FP32, :42), mass 100; body `offset` is the central mass
// 10^6 at the centre.  Velocities after one force evaluation (:49-73): v = sqrt(|F| dist), dist to the central body with the
// 0.0001 softening, direction rotation * (sin theta, -cos theta) with theta from the same generator.
__global__ void standard_kernel(float4* __restrict__ pos, int offset, int count, unsigned long long seed, double cx, double cy) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    const float theta = (float)(uniform01(seed, 0, i) * 6.283185307179586);
    const float r = (float)((-log(1.0 - uniform01(seed, 1, i)) / 12.0 + 0.02) * 50000.0);
    float4 p;
    p.x = (float)(cx + (double)r * cos((double)theta));
    p.y = (float)(cy + (double)r * sin((double)theta));
    p.z = __fmul_rn(sinf(__fdiv_rn(r, 2000.f)), 2000.f);
    p.w = 100.f;
    if (i == 0) p = make_float4((float)cx, (float)cy, 0.f, 1000000.f);
    pos[offset + i] = p;
}

__global__ void standard_velocity_kernel(const float4* __restrict__ pos, const float4* __restrict__ acc, float4* __restrict__ vel,
                                         int offset, int count, unsigned long long seed, float rotation) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    const float4 c = pos[offset], p = pos[offset + i], a = acc[i];
    const float dx = __fsub_rn(c.x, p.x), dy = __fsub_rn(c.y, p.y), dz = __fsub_rn(c.z, p.z);
    const float dist = __fsqrt_rn(__fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz)), 0.0001f));
    const float F = __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(a.x, a.x), __fmul_rn(a.y, a.y)), __fmul_rn(a.z, a.z)));
    const float v = __fsqrt_rn(__fmul_rn(dist, F));
    const double theta = (double)(float)(uniform01(seed, 0, i) * 6.283185307179586);
    float4 w = make_float4(0.f, 0.f, 0.f, 0.f);
    if (i != 0) {
        w.x = __fmul_rn(__fmul_rn(rotation, v), (float)sin(theta));
        w.y = __fmul_rn(__fmul_rn(-rotation, v), (float)cos(theta));
    }
    vel[i] = w;
}

}  // namespace

extern "C" {

int nb_ic_standard(nb_context* ctx, float* pos, int offset, int count, uint64_t seed, double cx, double cy, void* stream) {
    if (!ctx) return NB_ERR_INVALID;
    if (offset < 0 || count < 0 || (count > 0 && !pos)) return nb_fail(ctx, NB_ERR_INVALID, "bad range or null pointer");
    if (count == 0) return NB_OK;
    NB_DEVICE(ctx);
    standard_kernel<<<(count + 255) / 256, 256, 0, (cudaStream_t)stream>>>((float4*)pos, offset, count, seed, cx, cy);
    NB_CUDA(ctx, cudaGetLastError());
    return NB_OK;
}

int nb_ic_standard_velocities(nb_context* ctx, const float* pos, const float* acc, float* vel, int offset, int count, uint64_t seed,
                              float rotation, void* stream) {
    if (!ctx) return NB_ERR_INVALID;
    if (offset < 0 || count < 0 || (count > 0 && (!pos || !acc || !vel))) return nb_fail(ctx, NB_ERR_INVALID, "bad range or null pointer");
    if (count == 0) return NB_OK;
    NB_DEVICE(ctx);
    standard_velocity_kernel<<<(count + 255) / 256, 256, 0, (cudaStream_t)stream>>>((const float4*)pos, (const float4*)acc,
                                                                                   (float4*)vel, offset, count, seed, rotation);
    NB_CUDA(ctx, cudaGetLastError());
    return NB_OK;
}

int nb_ic_uniform_ellipsoid(nb_context* ctx, float* pos, float* vel, int offset, int count, uint64_t seed, double cx, double cy,
                            double rx, double ry, double z_half_height, float mass, float central_mass, float vx, float vy,
                            void* stream) {
    if (!ctx) return NB_ERR_INVALID;
    if (offset < 0 || count < 0 || (count > 0 && !pos)) return nb_fail(ctx, NB_ERR_INVALID, "bad range or null pointer");
    if (count == 0) return NB_OK;
    NB_DEVICE(ctx);
    ellipsoid_kernel<<<(count + 255) / 256, 256, 0, (cudaStream_t)stream>>>((float4*)pos, (float4*)vel, offset, count, seed, cx, cy, rx,
                                                                           ry, z_half_height, mass, central_mass, vx, vy);
    NB_CUDA(ctx, cudaGetLastError());
    return NB_OK;
}

int nb_ic_uniform_box(nb_context* ctx, float* pos, float* vel, int offset, int count, uint64_t seed, double cx, double cy, double sx,
                      double sy, int dim, double mass_lo, double mass_hi, double speed_lo, double speed_hi, float vx, float vy,
                      void* stream) {
    if (!ctx) return NB_ERR_INVALID;
    if (offset < 0 || count < 0 || (count > 0 && !pos)) return nb_fail(ctx, NB_ERR_INVALID, "bad range or null pointer");
    if (dim != 2 && dim != 3) return nb_fail(ctx, NB_ERR_INVALID, "dim must be 2 or 3");
    if (count == 0) return NB_OK;
    NB_DEVICE(ctx);
    box_kernel<<<(count + 255) / 256, 256, 0, (cudaStream_t)stream>>>((float4*)pos, (float4*)vel, offset, count, seed, cx, cy, sx, sy,
                                                                     dim, mass_lo, mass_hi, speed_lo, speed_hi, vx, vy);
    NB_CUDA(ctx, cudaGetLastError());
    return NB_OK;
}

int nb_ic_disk_velocities(nb_context* ctx, const float* pos, const float* acc, float* vel, int begin, int count, float cx, float cy,
                          int dim, float bulk_vx, float bulk_vy, void* stream) {
    if (!ctx) return NB_ERR_INVALID;
    if (begin < 0 || count < 0 || (count > 0 && (!pos || !acc || !vel))) return nb_fail(ctx, NB_ERR_INVALID, "bad range or null pointer");
    if (dim != 2 && dim != 3) return nb_fail(ctx, NB_ERR_INVALID, "dim must be 2 or 3");
    if (count == 0) return NB_OK;
    NB_DEVICE(ctx);
    disk_velocity_kernel<<<(count + 255) / 256, 256, 0, (cudaStream_t)stream>>>((const float4*)pos, (const float4*)acc, (float4*)vel,
                                                                               begin, count, cx, cy, dim, bulk_vx, bulk_vy);
    NB_CUDA(ctx, cudaGetLastError());
    return NB_OK;
}
}
