// ubench.cu -- pipe-level microbenchmarks behind the direct kernel's roofline accounting (profiles/README.md).
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o ubench ubench.cu ; run on a B200.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../../n-body-simulation_b200/csrc/common.cuh"
char g_nb_create_error[512];

#define ITERS 4096
// V0: ffma2 x=x*A+B, A,B invariant     V1: ffma2 x=y*z+x, all varying regs (3 distinct pairs)
// V2: scalar ffma 3 distinct           V3: ffma2 (V1) + 1 MUFU per 6 ffma2
// V4: ffma2 (V1) + 1 LDS.128 per 12    V5: full 3D interaction body, PAIRS=2, sources from registers (rotating)
// V6: body PAIRS=4                      V7: body PAIRS=2 but accumulate as scalar FFMA (mixed)
template <int V>
__global__ void __launch_bounds__(256) k(float* out, float a, float b, int iters) {
    __shared__ float4 sm[64];
    if (threadIdx.x < 64) sm[threadIdx.x] = make_float4(a, b, a, b);
    __syncthreads();
    float s = 0.f;
    if (V == 0 || V == 1 || V == 3 || V == 4) {
        uint64_t x[8], y[8], z[8];
        for (int i = 0; i < 8; i++) { x[i] = pk2(threadIdx.x + i, 1.f); y[i] = pk2(a + i * 1e-3f, a); z[i] = pk2(b, b + i * 1e-3f); }
        const uint64_t A = pk2(a, a), B = pk2(b, b);
        float m0 = a + 1.f, m1 = a + 2.f;
        for (int it = 0; it < iters; it++) {
#pragma unroll
            for (int u = 0; u < 6; u++) {
#pragma unroll
                for (int i = 0; i < 8; i++) {
                    if (V == 0) x[i] = fma2(x[i], A, B);
                    else x[i] = fma2(y[i], z[i], x[i]);
                }
                if (V == 3 && u == 0) {  // 8 MUFU per 48 ffma2
#pragma unroll
                    for (int i = 0; i < 4; i++) { m0 = rsqrt_approx(m0); m1 = rsqrt_approx(m1); }
                }
                if (V == 4 && u == 0) {
#pragma unroll
                    for (int i = 0; i < 4; i++) { float4 v = sm[(it + i + u) & 63]; m0 += v.x; }
                }
            }
        }
        for (int i = 0; i < 8; i++) { float lo, hi; up2(x[i], lo, hi); s += lo + hi; }
        s += m0 + m1;
    } else if (V == 2) {
        float x[16], y[16], z[16];
        for (int i = 0; i < 16; i++) { x[i] = threadIdx.x + i; y[i] = a + i * 1e-3f; z[i] = b + i * 1e-3f; }
        for (int it = 0; it < iters; it++) {
#pragma unroll
            for (int u = 0; u < 6; u++)
#pragma unroll
                for (int i = 0; i < 16; i++) x[i] = fmaf(y[i], z[i], x[i]);
        }
        for (int i = 0; i < 16; i++) s += x[i];
    } else {
        constexpr int PAIRS = (V == 6) ? 4 : 2;
        uint64_t BX[PAIRS], BY[PAIRS], BZ[PAIRS], AX[PAIRS], AY[PAIRS], AZ[PAIRS];
        float fx[2 * PAIRS], fy[2 * PAIRS], fz[2 * PAIRS];
        for (int p = 0; p < PAIRS; p++) {
            BX[p] = pk2(threadIdx.x + p, threadIdx.x - p); BY[p] = pk2(a * p, b * p); BZ[p] = pk2(a + p, b - p);
            AX[p] = AY[p] = AZ[p] = 0;
        }
        for (int i = 0; i < 2 * PAIRS; i++) fx[i] = fy[i] = fz[i] = 0.f;
        const uint64_t E2 = pk2(1.01f, 1.01f);
        const ulonglong2* tp = reinterpret_cast<const ulonglong2*>(sm);
        for (int it = 0; it < iters; it++) {
#pragma unroll 2
            for (int j = 0; j < 24 / PAIRS; j++) {
                const ulonglong2 sa = tp[(2 * j) & 63], sb = tp[(2 * j + 1) & 63];
#pragma unroll
                for (int p = 0; p < PAIRS; p++) {
                    uint64_t dx = sub2(sa.x, BX[p]), dy = sub2(sa.y, BY[p]), dz = sub2(sb.x, BZ[p]);
                    uint64_t r2 = fma2(dx, dx, E2);
                    r2 = fma2(dy, dy, r2);
                    r2 = fma2(dz, dz, r2);
                    float r0, r1;
                    up2(r2, r0, r1);
                    uint64_t inv = pk2(rsqrt_approx(r0), rsqrt_approx(r1));
                    uint64_t sc = mul2(mul2(mul2(inv, inv), inv), sb.y);
                    if (V == 7) {
                        float s0, s1, dx0, dx1, dy0, dy1, dz0, dz1;
                        up2(sc, s0, s1); up2(dx, dx0, dx1); up2(dy, dy0, dy1); up2(dz, dz0, dz1);
                        fx[2 * p] = fmaf(dx0, s0, fx[2 * p]); fx[2 * p + 1] = fmaf(dx1, s1, fx[2 * p + 1]);
                        fy[2 * p] = fmaf(dy0, s0, fy[2 * p]); fy[2 * p + 1] = fmaf(dy1, s1, fy[2 * p + 1]);
                        fz[2 * p] = fmaf(dz0, s0, fz[2 * p]); fz[2 * p + 1] = fmaf(dz1, s1, fz[2 * p + 1]);
                    } else {
                        AX[p] = fma2(dx, sc, AX[p]); AY[p] = fma2(dy, sc, AY[p]); AZ[p] = fma2(dz, sc, AZ[p]);
                    }
                }
            }
        }
        for (int p = 0; p < PAIRS; p++) { float lo, hi; up2(AX[p], lo, hi); s += lo + hi; up2(AY[p], lo, hi); s += lo + hi; up2(AZ[p], lo, hi); s += lo + hi; }
        for (int i = 0; i < 2 * PAIRS; i++) s += fx[i] + fy[i] + fz[i];
    }
    if (s == 123.456f) out[0] = s;
}

template <int V>
void run(const char* name, double lane_ops_per_iter, int blocks_per_sm, float* d_out) {
    cudaDeviceProp prop; cudaGetDeviceProperties(&prop, 0);
    int blocks = prop.multiProcessorCount * blocks_per_sm;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    for (int r = 0; r < 4; r++) {
        cudaEventRecord(e0);
        k<V><<<blocks, 256>>>(d_out, 1.0001f, 0.5f, ITERS);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (r && ms < best) best = ms;
    }
    double ops = (double)blocks * 256 * ITERS * lane_ops_per_iter;   // FP32-pipe lane-ops
    double peak = prop.multiProcessorCount * 128.0 * 1.965e9;
    printf("%-44s blocks/SM=%d  %8.3f ms  %7.2f Tlane-op/s  = %5.1f%% of 128 lanes/clk/SM @1965MHz\n", name, blocks_per_sm, best,
           ops / (best * 1e-3) / 1e12, 100.0 * ops / (best * 1e-3) / peak);
}

int main() {
    float* d; cudaMalloc(&d, 64);
    for (int bps : {2, 4, 8}) {
        run<0>("V0 ffma2 x=x*A+B (invariant A,B)", 6 * 8 * 2, bps, d);
        run<1>("V1 ffma2 x=y*z+x (3 distinct pairs)", 6 * 8 * 2, bps, d);
        run<2>("V2 ffma  x=y*z+x scalar (3 distinct)", 6 * 16, bps, d);
        run<3>("V3 V1 + 8 MUFU per 48 ffma2", 6 * 8 * 2, bps, d);
        run<4>("V4 V1 + 4 LDS.128 per 48 ffma2", 6 * 8 * 2, bps, d);
        run<5>("V5 3D body PAIRS=2 (24 fp32 lane-ops/pair)", 24.0 / 2 * 2 * 24, bps, d);
        run<6>("V6 3D body PAIRS=4", 24.0 / 4 * 4 * 24, bps, d);
        run<7>("V7 3D body PAIRS=2, scalar accumulate", 24.0 / 2 * 2 * 24, bps, d);
    }
    printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
