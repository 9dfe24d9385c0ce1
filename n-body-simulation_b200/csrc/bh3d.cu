// bh3d.cu -- Barnes-Hut in three dimensions (octree) for sm_100a.
//
// NO REFERENCE COUNTERPART: the reference runs Barnes-Hut in R2 only and silently downgrades R3
// (src/ComputeMethods.cpp:35-38); SURVEY.md section 8f-4 lists the octree as the last "next" item.  This file carries
// the 2D pipeline of bh.cu one dimension up with the same rules, so that the two differ only in arity:
//   * padded bounding box with the reference's fmaf padding (src/BarnesHutKernels.cu:190-193), now for x, y, z;
//   * 63-bit tree-path keys: 21 levels of '>=' float bisection (the descent of :266-285 with a third coordinate),
//     digit = (x>=midX) + 2 (y>=midY) + 4 (z>=midZ), first level in bits 62:60;
//   * ONE stable radix sort (bh_shared.cuh); targets are grouped by 32 consecutive bodies of that order (the 2D path
//     needs a second sort only to reproduce the reference's own 32-bit curve);
//   * lock-free bucket-1 region octree from the sorted keys, FP64 prefix-sum monopoles, chains stored once;
//   * the reference's acceptance rule (:398-445): unanimous vote of the group, cell tested against its parent's
//     width with the softening on the distance side, level-1 cells always opened, bodies always summed;
//   * the Newtonian force law of the 3D direct kernel: k = inv^3 m, inv = rsqrt(eps + |d|^2) (src/DirectComputeKernels.cu:64-67).
// Checked against oracle/nbody_oracle.c (oracle_bh3_*) and the FP64 direct sum; parity with the reference is
// undefined for this path.
#include "common.cuh"

#include <algorithm>
#include <cfloat>
#include <vector>

namespace {

constexpr int MAXL3 = 21;
constexpr uint32_t LINK_EMPTY = 0xFFFFFFFFu;
constexpr uint32_t LINK_BODY = 0xFFFFFFFEu;
constexpr uint32_t LINK_BUCKET = 0x80000000u;
constexpr float FAR_AWAY = 3.0e18f;

struct D4 {
    double m, mx, my, mz;
};
__host__ __device__ inline D4 operator+(const D4& a, const D4& b) { return D4{a.m + b.m, a.mx + b.mx, a.my + b.my, a.mz + b.mz}; }

#include "bh_shared.cuh"

template <>
__device__ __forceinline__ D4 shfl_up_t<D4>(D4 v, int o) {
    return D4{__shfl_up_sync(0xffffffffu, v.m, o), __shfl_up_sync(0xffffffffu, v.mx, o), __shfl_up_sync(0xffffffffu, v.my, o),
              __shfl_up_sync(0xffffffffu, v.mz, o)};
}
__device__ __forceinline__ D4 zero_of(D4) { return D4{0.0, 0.0, 0.0, 0.0}; }

}  // namespace

struct nb_bh3_state {
    int cap = 0;
    float4* bounds = nullptr;          // [0] = padded minimum (x, y, z, -), [1] = padded maximum
    float4* bounds_part = nullptr;     // per-block partials, 2 per block
    unsigned int* ticket = nullptr;
    uint64_t* key[2] = {nullptr, nullptr};
    int32_t* idx[2] = {nullptr, nullptr};
    uint32_t* ghist = nullptr;         // [8][256] digit histograms, tickets, look-back status
    uint32_t* os_status = nullptr;
    unsigned int* os_ticket = nullptr;
    size_t os_stride = 0;
    float4* spos = nullptr;            // bodies in tree order (x, y, z, m)
    D4* psum = nullptr;
    D4* psum_blk = nullptr;
    int32_t *ext_a = nullptr, *ext_c = nullptr, *ext_r = nullptr, *flag_scan = nullptr, *flag_blk = nullptr;
    float4* crec = nullptr;            // 8 child records per node: (cx, cy, cz, m)
    float* cthr = nullptr;             // 8 acceptance thresholds per node
    uint32_t* clink = nullptr;         // 8 links per node
    int32_t *node_first = nullptr, *node_last = nullptr;
    int8_t *node_level = nullptr, *node_top = nullptr;
    float4* root_rec = nullptr;        // [0] = (cx, cy, cz, m) of the root cell, [1] = (threshold, link, -, -)
    unsigned long long* counters = nullptr;
    int last_n = 0, sel = 0;
};

namespace {

constexpr int DIGITS3 = 8;

// ---- bounds ---------------------------------------------------------------------------------------------
__global__ void bounds3_kernel(const float4* __restrict__ pos, int n, float4* part, unsigned int* __restrict__ ticket,
                               float4* __restrict__ bounds,
                               unsigned long long* __restrict__ walk_status) {
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float4 p = pos[i];
        lo[0] = fminf(lo[0], p.x); lo[1] = fminf(lo[1], p.y); lo[2] = fminf(lo[2], p.z);
        hi[0] = fmaxf(hi[0], p.x); hi[1] = fmaxf(hi[1], p.y); hi[2] = fmaxf(hi[2], p.z);
    }
    __shared__ float sh[32][6];
    __shared__ bool last;
    auto block_red = [&]() {
#pragma unroll
        for (int k = 0; k < 3; k++)
            for (int o = 16; o > 0; o >>= 1) {
                lo[k] = fminf(lo[k], __shfl_xor_sync(0xffffffffu, lo[k], o));
                hi[k] = fmaxf(hi[k], __shfl_xor_sync(0xffffffffu, hi[k], o));
            }
        if ((threadIdx.x & 31) == 0)
            for (int k = 0; k < 3; k++) { sh[threadIdx.x >> 5][k] = lo[k]; sh[threadIdx.x >> 5][3 + k] = hi[k]; }
        __syncthreads();
        if (threadIdx.x == 0)
            for (int w = 1; w < (int)(blockDim.x >> 5); w++)
                for (int k = 0; k < 3; k++) { lo[k] = fminf(lo[k], sh[w][k]); hi[k] = fmaxf(hi[k], sh[w][3 + k]); }
        __syncthreads();
    };
    block_red();
    if (threadIdx.x == 0) {
        part[2 * blockIdx.x] = make_float4(lo[0], lo[1], lo[2], 0.f);
        part[2 * blockIdx.x + 1] = make_float4(hi[0], hi[1], hi[2], 0.f);
        __threadfence();
        last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
    }
    __syncthreads();
    if (!last) return;
    __threadfence();
    for (int k = 0; k < 3; k++) { lo[k] = FLT_MAX; hi[k] = -FLT_MAX; }
    for (int b = threadIdx.x; b < gridDim.x; b += blockDim.x) {
        const float4 a = __ldcg(&part[2 * b]), c = __ldcg(&part[2 * b + 1]);
        lo[0] = fminf(lo[0], a.x); lo[1] = fminf(lo[1], a.y); lo[2] = fminf(lo[2], a.z);
        hi[0] = fmaxf(hi[0], c.x); hi[1] = fmaxf(hi[1], c.y); hi[2] = fmaxf(hi[2], c.z);
    }
    block_red();
    if (threadIdx.x == 0) {   // the reference's padding, src/BarnesHutKernels.cu:190-193 (one FFMA each)
        bounds[0] = make_float4(__fmaf_rn(fabsf(lo[0]), -0.001f, lo[0]), __fmaf_rn(fabsf(lo[1]), -0.001f, lo[1]),
                                __fmaf_rn(fabsf(lo[2]), -0.001f, lo[2]), 0.f);
        bounds[1] = make_float4(__fmaf_rn(fabsf(hi[0]), 0.001f, hi[0]), __fmaf_rn(fabsf(hi[1]), 0.001f, hi[1]),
                                __fmaf_rn(fabsf(hi[2]), 0.001f, hi[2]), 0.f);
        *ticket = 0u;
        *walk_status = 0ull;
    }
}

// ---- 63-bit tree-path keys + the 8 digit histograms of the sort --------------------------------------------
constexpr int KEYS3_THREADS = 256;
constexpr int KEYS3_PER_BLOCK = 4096;

__global__ void __launch_bounds__(KEYS3_THREADS) keys3_kernel(const float4* __restrict__ pos, int n,
                                                             const float4* __restrict__ bounds, uint64_t* __restrict__ key,
                                                             uint32_t* __restrict__ ghist) {
    __shared__ uint32_t h[DIGITS3][256];
    if (ghist) {
        for (int k = threadIdx.x; k < DIGITS3 * 256; k += KEYS3_THREADS) (&h[0][0])[k] = 0;
        __syncthreads();
    }
    const float4 blo = bounds[0], bhi = bounds[1];
    for (int i = blockIdx.x * KEYS3_PER_BLOCK + threadIdx.x; i < min(n, (blockIdx.x + 1) * KEYS3_PER_BLOCK); i += KEYS3_THREADS) {
        const float4 p = pos[i];
        float lx = blo.x, ly = blo.y, lz = blo.z, hx = bhi.x, hy = bhi.y, hz = bhi.z;
        uint64_t k64 = 0;
#pragma unroll 7
        for (int lev = 0; lev < MAXL3; lev++) {
            unsigned path = 0;
            const float mx = (hx + lx) * 0.5f;
            if (p.x >= mx) { lx = mx; path += 1; } else { hx = mx; }
            const float my = (hy + ly) * 0.5f;
            if (p.y >= my) { ly = my; path += 2; } else { hy = my; }
            const float mz = (hz + lz) * 0.5f;
            if (p.z >= mz) { lz = mz; path += 4; } else { hz = mz; }
            k64 |= (uint64_t)path << (60 - 3 * lev);
        }
        key[i] = k64;
        if (ghist) {
#pragma unroll
            for (int d = 0; d < DIGITS3; d++) atomicAdd(&h[d][(uint32_t)(k64 >> (8 * d)) & 255u], 1u);
        }
    }
    if (ghist) {
        __syncthreads();
        for (int k = threadIdx.x; k < DIGITS3 * 256; k += KEYS3_THREADS) {
            const uint32_t v = (&h[0][0])[k];
            if (v) atomicAdd(&ghist[k], v);
        }
    }
}

// gather bodies into tree order and scan (m, m x, m y, m z) in FP64
__global__ void __launch_bounds__(SCAN_THREADS) gather_scan3_kernel(const float4* __restrict__ pos,
                                                                   const int32_t* __restrict__ order, int n,
                                                                   float4* __restrict__ spos, D4* __restrict__ psum,
                                                                   D4* __restrict__ tot) {
    D4 v[SCAN_ITEMS];
    const int i0 = blockIdx.x * SCAN_CHUNK + threadIdx.x * SCAN_ITEMS;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) {
        const int i = i0 + k;
        if (i < n) {
            const float4 p = pos[order[i]];
            spos[i] = p;
            const double m = p.w;
            v[k] = D4{m, m * (double)p.x, m * (double)p.y, m * (double)p.z};
        } else {
            v[k] = D4{0.0, 0.0, 0.0, 0.0};
        }
    }
    block_scan_chunk<D4>(v, &tot[blockIdx.x]);
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++)
        if (i0 + k < n) psum[i0 + k] = v[k];
}

// ---- octree from the sorted keys (same construction as bh.cu, three bits per level) ----------------------------
__device__ __forceinline__ int cpl3(uint64_t a, uint64_t b) {   // common prefix in levels; bit 63 of a key is always 0
    const uint64_t x = a ^ b;
    return x ? ((__clzll((long long)x) - 1) / 3) : MAXL3;
}
struct Level3 {   // common prefix of two octree keys in levels, for the shared boundary kernel (bh_shared.cuh)
    static __device__ __forceinline__ int of(uint64_t a, uint64_t b) { return cpl3(a, b); }
};

struct TreeView3 {
    const uint64_t* K;
    const float4* spos;
    const D4* psum;
    const D4* psum_blk;
    const int32_t *flag_scan, *flag_blk, *ext_a, *ext_c, *ext_r;
    int n;
};
__device__ __forceinline__ D4 prefix_incl(const TreeView3& t, int i) { return t.psum[i] + t.psum_blk[i >> SCAN_SHIFT]; }
__device__ __forceinline__ int flag_incl(const TreeView3& t, int i) { return t.flag_scan[i] + t.flag_blk[i >> SCAN_SHIFT]; }
__device__ __forceinline__ int level_at(const TreeView3& t, int i) { return (i < 0 || i >= t.n - 1) ? -1 : cpl3(t.K[i], t.K[i + 1]); }
__device__ __forceinline__ int node_of(const TreeView3& t, int j, int L) {
    if (L >= MAXL3) j = t.ext_a[j];                              // bucket: the first boundary leads
    else while (t.ext_c[j] != t.ext_a[j]) j = t.ext_c[j] - 1;     // at most seven hops
    return flag_incl(t, j) - 1;
}
__device__ __forceinline__ float threshold3_of(float s0, int L) {  // (dim/theta)^2 * 0.25^(L-1); never tested below level 2
    if (L < 2) return __int_as_float(0x7f800000);
    return s0 * __int_as_float((127 - 2 * (L - 1)) << 23);
}

__global__ void __launch_bounds__(256) emit3_kernel(TreeView3 t, const float4* __restrict__ bounds, float theta,
                                                   float4* __restrict__ crec, float* __restrict__ cthr,
                                                   uint32_t* __restrict__ clink, int32_t* __restrict__ node_first,
                                                   int32_t* __restrict__ node_last, int8_t* __restrict__ node_level,
                                                   int8_t* __restrict__ node_top, float4* __restrict__ root_rec) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= t.n - 1) return;
    const uint64_t* K = t.K;
    const uint64_t ki = K[i], kn = K[i + 1];
    const int L = cpl3(ki, kn);
    const int a = t.ext_a[i], c = t.ext_c[i], r = t.ext_r[i];
    const int id = node_of(t, i, L);
    const bool first = (c == a);
    const bool last = (r == t.n - 1) || (cpl3(K[r], K[r + 1]) < L);
    if (first) {
        node_first[id] = a;
        node_level[id] = (int8_t)L;
    }
    if (L < MAXL3) {
        const size_t base = 8 * (size_t)id;
        const int ql = (int)((ki >> (60 - 3 * L)) & 7ull), qn = (int)((kn >> (60 - 3 * L)) & 7ull);   // ql < qn
#pragma unroll
        for (int q = 0; q < 8; q++) {
            if ((q > ql && q < qn) || (first && q < ql) || (last && q > qn)) {
                crec[base + q] = make_float4(FAR_AWAY, FAR_AWAY, FAR_AWAY, 0.f);   // massless: adds exactly 0
                cthr[base + q] = 0.f;
                clink[base + q] = LINK_EMPTY;
            }
        }
        if (c == i) {
            crec[base + ql] = t.spos[i];
            cthr[base + ql] = 0.f;
            clink[base + ql] = LINK_BODY;
        }
        if (last && r == i + 1) {
            crec[base + qn] = t.spos[i + 1];
            cthr[base + qn] = 0.f;
            clink[base + qn] = LINK_BODY;
        }
    }
    if (!last) return;
    const int b = r;
    node_last[id] = b;
    if (L >= MAXL3) crec[8 * (size_t)id] = make_float4(__int_as_float(a), __int_as_float(b - a + 1), 0.f, 0.f);
    const float4 blo = bounds[0], bhi = bounds[1];
    const float dim_size = fmaxf(fmaxf(bhi.x - blo.x, bhi.y - blo.y), bhi.z - blo.z);
    float s0 = dim_size / theta;
    s0 *= s0;
    const D4 hi = prefix_incl(t, b);
    D4 lo = D4{0.0, 0.0, 0.0, 0.0};
    if (a > 0) lo = prefix_incl(t, a - 1);
    const double m = hi.m - lo.m;
    float cx = 0.f, cy = 0.f, cz = 0.f;
    if (m > 0.0) { cx = (float)((hi.mx - lo.mx) / m); cy = (float)((hi.my - lo.my) / m); cz = (float)((hi.mz - lo.mz) / m); }
    const float4 rec = make_float4(cx, cy, cz, (float)m);
    const float thr = threshold3_of(s0, L);
    const uint32_t link = (uint32_t)id | (L >= MAXL3 ? LINK_BUCKET : 0u);
    const int dl = level_at(t, a - 1), dr = level_at(t, b);
    if (dl < 0 && dr < 0) {
        node_top[id] = 0;
        root_rec[0] = rec;
        root_rec[1] = make_float4(thr, __uint_as_float(link), 0.f, 0.f);
        return;
    }
    const int Lp = max(dl, dr);
    const int p = (dl >= dr) ? a - 1 : b;
    const int pid = node_of(t, p, Lp);
    const int slot = (int)((K[a] >> (60 - 3 * Lp)) & 7ull);
    node_top[id] = (int8_t)(Lp + 1);
    crec[8 * (size_t)pid + slot] = rec;
    cthr[8 * (size_t)pid + slot] = thr;
    clink[8 * (size_t)pid + slot] = link;
}

// ---- walk (the schedule of walk_lists_kernel in bh.cu: lanes classify 32 different child records at a time) -------
constexpr int W3_WARPS = 4;
constexpr int W3_QCAP = 1024;
constexpr int W3_LCAP = 64;
// as WL_DFS_ROOM in bh.cu: past `qcap` entries a group finishes depth-first, which adds at most 7 entries per
// remaining level (7 * 21 = 147) on top of the batch being pushed (<= 32)
constexpr int W3_DFS_ROOM = 192;
constexpr float FAR_AWAY3 = 3.0e18f;           // coordinate of a padding record: zero mass, contributes exactly 0

__device__ __forceinline__ float rsqrt3_approx(float x) {
    float r;
    asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

template <bool COUNT>
__global__ void __launch_bounds__(W3_WARPS * 32, 6) walk3_kernel(const float4* __restrict__ pos,
                                                                const int32_t* __restrict__ sorted_idx, int n,
                                                                const float4* __restrict__ crec, const float* __restrict__ cthr,
                                                                const uint32_t* __restrict__ clink,
                                                                const float4* __restrict__ spos,
                                                                const float4* __restrict__ root_rec, float eps,
                                                                int group_begin, int group_count, float4* __restrict__ acc,
                                                                unsigned long long* __restrict__ counters, int qcap) {
    __shared__ uint32_t q_s[W3_WARPS][W3_QCAP];
    // pending accepted records, two per slot (FP32x2 evaluation): (cx0, cx1, cy0, cy1) and (cz0, cz1, m0, m1)
    __shared__ float4 lxy_s[W3_WARPS][W3_LCAP / 2];
    __shared__ float4 lzm_s[W3_WARPS][W3_LCAP / 2];
    __shared__ float4 wit_s[W3_WARPS][6];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const unsigned FULLM = 0xffffffffu;
    const int g = blockIdx.x * W3_WARPS + warp;
    if (g >= group_count) return;
    const int bi = (group_begin + g) * 32 + lane;
    const bool valid = bi < n;
    const int body = sorted_idx[valid ? bi : (group_begin + g) * 32];   // lanes past the end shadow the first target
    const float4 me = pos[body];
    const float x = me.x, y = me.y, z = me.z;
    const float xr = __shfl_sync(FULLM, x, 0), yr = __shfl_sync(FULLM, y, 0), zr = __shfl_sync(FULLM, z, 0);
    float bx0 = x, bx1 = x, by0 = y, by1 = y, bz0 = z, bz1 = z;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        bx0 = fminf(bx0, __shfl_xor_sync(FULLM, bx0, o)); bx1 = fmaxf(bx1, __shfl_xor_sync(FULLM, bx1, o));
        by0 = fminf(by0, __shfl_xor_sync(FULLM, by0, o)); by1 = fmaxf(by1, __shfl_xor_sync(FULLM, by1, o));
        bz0 = fminf(bz0, __shfl_xor_sync(FULLM, bz0, o)); bz1 = fmaxf(bz1, __shfl_xor_sync(FULLM, bz1, o));
    }
    // the targets that touch the six faces of the box: witnesses for early rejection
    float4* const wit = wit_s[warp];
    {
        const float ext[6] = {bx0, bx1, by0, by1, bz0, bz1};
        const float mine[6] = {x, x, y, y, z, z};
#pragma unroll
        for (int f = 0; f < 6; f++) {
            const unsigned m = __ballot_sync(FULLM, mine[f] == ext[f]);
            if (lane == __ffs(m) - 1) wit[f] = make_float4(x, y, z, 0.f);
        }
        __syncwarp();
    }
    float ax = 0.f, ay = 0.f, az = 0.f;
    uint64_t AX2 = 0ull, AY2 = 0ull, AZ2 = 0ull;   // packed partial sums of the list evaluation (even / odd records)
    const uint64_t X2 = pk2(x, x), Y2 = pk2(y, y), Z2 = pk2(z, z), E2 = pk2(eps, eps);
    uint32_t* const q = q_s[warp];
    float* const lxy = reinterpret_cast<float*>(lxy_s[warp]);
    float* const lzm = reinterpret_cast<float*>(lzm_s[warp]);
    int qtop = 0, lcount = 0;
    unsigned long long c_open = 0, c_exam = 0, c_inter = 0;
    int c_depth = 0;
    bool depth_first = false;

    auto add = [&](float cx, float cy, float cz, float m) {        // src/DirectComputeKernels.cu:61-67
        const float dx = cx - x, dy = cy - y, dz = cz - z;
        const float r = __fmaf_rn(dz, dz, __fmaf_rn(dy, dy, __fmaf_rn(dx, dx, eps)));
        const float inv = rsqrt3_approx(r);
        const float k = inv * inv * inv * m;
        ax = __fmaf_rn(k, dx, ax);
        ay = __fmaf_rn(k, dy, ay);
        az = __fmaf_rn(k, dz, az);
    };
    auto append = [&](int e, const float4& c) {   // record e of the list
        const int at = (e >> 1) * 4 + (e & 1);
        lxy[at] = c.x;
        lxy[at + 2] = c.y;
        lzm[at] = c.z;
        lzm[at + 2] = c.w;
    };
    auto flush = [&]() {
        if ((lcount & 1) && lane == 0) append(lcount, make_float4(FAR_AWAY3, FAR_AWAY3, FAR_AWAY3, 0.f));   // massless partner
        __syncwarp();
        const int pairs = (lcount + 1) >> 1;
#pragma unroll 2
        for (int k = 0; k < pairs; k++) {
            const ulonglong2 a = reinterpret_cast<const ulonglong2*>(lxy)[k];   // (cx0, cx1) (cy0, cy1)
            const ulonglong2 b = reinterpret_cast<const ulonglong2*>(lzm)[k];   // (cz0, cz1) (m0, m1)
            const uint64_t dx = sub2(a.x, X2), dy = sub2(a.y, Y2), dz = sub2(b.x, Z2);
            const uint64_t r2 = fma2(dz, dz, fma2(dy, dy, fma2(dx, dx, E2)));
            float r0, r1;
            up2(r2, r0, r1);
            const uint64_t inv = pk2(rsqrt3_approx(r0), rsqrt3_approx(r1));
            const uint64_t k2 = mul2(mul2(mul2(inv, inv), inv), b.y);
            AX2 = fma2(k2, dx, AX2);
            AY2 = fma2(k2, dy, AY2);
            AZ2 = fma2(k2, dz, AZ2);
        }
        __syncwarp();
        lcount = 0;
    };

    float4 rec = make_float4(0.f, 0.f, 0.f, 0.f);
    float thr = 0.f;
    uint32_t link = LINK_EMPTY;
    if (lane == 0) {
        rec = root_rec[0];
        thr = root_rec[1].x;
        link = __float_as_uint(root_rec[1].y);
    }
    for (;;) {
        const bool present = link != LINK_EMPTY;
        bool rej = false, amb = false, accp = false;
        if (present) {
            const float dxr = rec.x - xr, dyr = rec.y - yr, dzr = rec.z - zr;
            const float rr = __fmaf_rn(dzr, dzr, __fmaf_rn(dyr, dyr, __fmaf_rn(dxr, dxr, eps)));   // the first target's own vote
            if (!(thr <= rr)) {
                rej = true;
            } else {
                const float ddx = fmaxf(fmaxf(bx0 - rec.x, rec.x - bx1), 0.f);
                const float ddy = fmaxf(fmaxf(by0 - rec.y, rec.y - by1), 0.f);
                const float ddz = fmaxf(fmaxf(bz0 - rec.z, rec.z - bz1), 0.f);
                const float dmin = __fmaf_rn(ddz, ddz, __fmaf_rn(ddy, ddy, __fmaf_rn(ddx, ddx, eps)));
                if (thr <= dmin * 0.999999f) {
                    accp = true;
                } else {   // a second witness: the group's extreme target on the side the record lies on
                    int sel;
                    if (ddx >= ddy && ddx >= ddz) sel = rec.x < bx0 ? 0 : 1;
                    else if (ddy >= ddz) sel = rec.y < by0 ? 2 : 3;
                    else sel = rec.z < bz0 ? 4 : 5;
                    const float4 wt = wit[sel];
                    const float dxw = rec.x - wt.x, dyw = rec.y - wt.y, dzw = rec.z - wt.z;
                    const float rw = __fmaf_rn(dzw, dzw, __fmaf_rn(dyw, dyw, __fmaf_rn(dxw, dxw, eps)));
                    if (!(thr <= rw)) rej = true; else amb = true;
                }
            }
        }
        unsigned ambm = __ballot_sync(FULLM, amb);
        while (ambm) {   // undecided: the real unanimous vote
            const int src = __ffs(ambm) - 1;
            ambm &= ambm - 1;
            const float cx = __shfl_sync(FULLM, rec.x, src), cy = __shfl_sync(FULLM, rec.y, src), cz = __shfl_sync(FULLM, rec.z, src);
            const float cm = __shfl_sync(FULLM, rec.w, src), ct = __shfl_sync(FULLM, thr, src);
            const float dx = cx - x, dy = cy - y, dz = cz - z;
            const float r = __fmaf_rn(dz, dz, __fmaf_rn(dy, dy, __fmaf_rn(dx, dx, eps)));
            if (__all_sync(FULLM, ct <= r)) {
                const float inv = rsqrt3_approx(r);
                const float k = inv * inv * inv * cm;
                ax = __fmaf_rn(k, dx, ax);
                ay = __fmaf_rn(k, dy, ay);
                az = __fmaf_rn(k, dz, az);
                if (COUNT) c_inter++;
            } else if (lane == src) {
                rej = true;
            }
        }
        const unsigned accm = __ballot_sync(FULLM, accp);
        const unsigned rejm = __ballot_sync(FULLM, rej);
        const unsigned below = (1u << lane) - 1u;
        if (COUNT) { c_exam += __popc(__ballot_sync(FULLM, present)); c_inter += __popc(accm); }
        if (accp) append(lcount + __popc(accm & below), rec);
        lcount += __popc(accm);
        const int nrej = __popc(rejm);
        if (qtop + nrej > qcap) depth_first = true;   // room for this batch is reserved (W3_DFS_ROOM)
        if (rej) q[qtop + __popc(rejm & below)] = link;
        qtop += nrej;
        __syncwarp();
        if (COUNT) c_depth = max(c_depth, qtop);
        if (depth_first) {   // the rest of this group's walk one node at a time, every lane voting on the same child
            while (qtop > 0) {
                const uint32_t ent = q[qtop - 1];
                __syncwarp();
                qtop--;
                if (COUNT) c_open++;
                const size_t id = ent & 0x7FFFFFFFu;
                if (ent & LINK_BUCKET) {
                    const float4 h = crec[8 * id];
                    const int first = __float_as_int(h.x), cnt = __float_as_int(h.y);
                    for (int t = 0; t < cnt; t++) {
                        const float4 p = spos[first + t];
                        add(p.x, p.y, p.z, p.w);
                    }
                    if (COUNT) { c_exam += cnt; c_inter += cnt; }
                    continue;
                }
                for (int c = 0; c < 8; c++) {
                    const uint32_t lk = clink[8 * id + c];
                    if (lk == LINK_EMPTY) continue;
                    const float4 r4 = crec[8 * id + c];
                    const float ct = cthr[8 * id + c];
                    const float dx = r4.x - x, dy = r4.y - y, dz = r4.z - z;
                    const float r = __fmaf_rn(dz, dz, __fmaf_rn(dy, dy, __fmaf_rn(dx, dx, eps)));
                    if (COUNT) c_exam++;
                    if (__all_sync(FULLM, ct <= r)) {
                        const float inv = rsqrt3_approx(r);
                        const float k = inv * inv * inv * r4.w;
                        ax = __fmaf_rn(k, dx, ax);
                        ay = __fmaf_rn(k, dy, ay);
                        az = __fmaf_rn(k, dz, az);
                        if (COUNT) c_inter++;
                    } else {
                        if (lane == 0) q[qtop] = lk;
                        qtop++;
                    }
                }
                __syncwarp();
                if (COUNT) c_depth = max(c_depth, qtop);
            }
            break;
        }
        if (qtop == 0) break;
        // pop up to four nodes: lane l takes child (l & 7) of the (l >> 3)-th node from the top
        const int k = min(4, qtop);
        const int slot = lane >> 3;
        uint32_t ent = LINK_EMPTY;
        if (slot < k) ent = q[qtop - 1 - slot];
        __syncwarp();
        qtop -= k;
        if (COUNT) c_open += k;
        link = LINK_EMPTY;
        const bool is_bucket = (ent != LINK_EMPTY) && (ent & LINK_BUCKET);
        if (ent != LINK_EMPTY && !is_bucket) {
            const size_t at = 8 * (size_t)(ent & 0x7FFFFFFFu) + (lane & 7);
            rec = crec[at];
            thr = cthr[at];
            link = clink[at];
        }
        if (lcount >= 32) flush();
        unsigned bm = __ballot_sync(FULLM, is_bucket && (lane & 7) == 0);
        while (bm) {   // depth-cap buckets: every body of the cell
            const int src = __ffs(bm) - 1;
            bm &= bm - 1;
            const uint32_t id = __shfl_sync(FULLM, ent, src) & 0x7FFFFFFFu;
            const float4 h = crec[8 * (size_t)id];
            const int first = __float_as_int(h.x), cnt = __float_as_int(h.y);
            for (int t = 0; t < cnt; t++) {
                const float4 p = spos[first + t];
                add(p.x, p.y, p.z, p.w);
            }
            if (COUNT) { c_exam += cnt; c_inter += cnt; }
        }
    }
    flush();
    {
        float a0, a1;
        up2(AX2, a0, a1); ax += a0 + a1;
        up2(AY2, a0, a1); ay += a0 + a1;
        up2(AZ2, a0, a1); az += a0 + a1;
    }
    if (valid) acc[body] = make_float4(ax, ay, az, 0.f);
    if (depth_first && lane == 0) atomicOr(&counters[4], 1ull);   // informative: a group left the batched schedule
    if (COUNT && lane == 0) {
        atomicAdd(&counters[0], c_open);
        atomicAdd(&counters[1], c_exam);
        atomicMax(&counters[3], (unsigned long long)c_depth);
    }
    if (COUNT && valid) atomicAdd(&counters[2], c_inter);
}

__global__ void zero_acc3_kernel(float4* acc, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) acc[i] = make_float4(0.f, 0.f, 0.f, 0.f);
}

// ---- host side ------------------------------------------------------------------------------------------------
template <typename T>
int alloc3(nb_context* ctx, T** p, size_t count) {
    NB_CUDA(ctx, cudaMalloc((void**)p, count * sizeof(T)));
    ctx->bh3_bytes += count * sizeof(T);
    return NB_OK;
}

void free_state3(nb_bh3_state* st) {
    if (!st) return;
    void* ptrs[] = {st->bounds, st->bounds_part, st->ticket, st->key[0], st->key[1], st->idx[0], st->idx[1], st->ghist, st->spos,
                    st->psum, st->psum_blk, st->ext_a, st->ext_c, st->ext_r, st->flag_scan, st->flag_blk, st->crec, st->cthr,
                    st->clink, st->node_first, st->node_last, st->node_level, st->node_top, st->root_rec, st->counters};
    for (void* p : ptrs)
        if (p) cudaFree(p);
}

int ensure_state3(nb_context* ctx, int n) {
    if (ctx->bh3 && ctx->bh3->cap >= n) return NB_OK;
    NB_CUDA(ctx, cudaDeviceSynchronize());
    if (ctx->bh3) {
        free_state3(ctx->bh3);
        delete ctx->bh3;
        ctx->bh3 = nullptr;
        ctx->bh3_bytes = 0;
    }
    nb_bh3_state* st = new nb_bh3_state();
    ctx->bh3 = st;
    const size_t cap = (size_t)std::max(n, std::max(ctx->max_bodies, 1024));
    const size_t chunks = (cap + SCAN_CHUNK - 1) / SCAN_CHUNK + 1;
    const size_t sblocks = (cap + SORT_TILE - 1) / SORT_TILE;
    int rc = 0;
    rc |= alloc3(ctx, &st->bounds, 2);
    rc |= alloc3(ctx, &st->bounds_part, 2 * 1024);
    rc |= alloc3(ctx, &st->ticket, 1);
    for (int k = 0; k < 2; k++) {
        rc |= alloc3(ctx, &st->key[k], cap);
        rc |= alloc3(ctx, &st->idx[k], cap);
    }
    const size_t os_words = (size_t)DIGITS3 * 256 + 64 + (size_t)DIGITS3 * 256 * sblocks;
    rc |= alloc3(ctx, &st->ghist, os_words);
    st->os_ticket = reinterpret_cast<unsigned int*>(st->ghist + DIGITS3 * 256);
    st->os_status = st->ghist + DIGITS3 * 256 + 64;
    rc |= alloc3(ctx, &st->spos, cap);
    rc |= alloc3(ctx, &st->psum, cap);
    rc |= alloc3(ctx, &st->psum_blk, chunks);
    rc |= alloc3(ctx, &st->ext_a, cap);
    rc |= alloc3(ctx, &st->ext_c, cap);
    rc |= alloc3(ctx, &st->ext_r, cap);
    rc |= alloc3(ctx, &st->flag_scan, cap);
    rc |= alloc3(ctx, &st->flag_blk, chunks);
    rc |= alloc3(ctx, &st->crec, 8 * cap);
    rc |= alloc3(ctx, &st->cthr, 8 * cap);
    rc |= alloc3(ctx, &st->clink, 8 * cap);
    rc |= alloc3(ctx, &st->node_first, cap);
    rc |= alloc3(ctx, &st->node_last, cap);
    rc |= alloc3(ctx, &st->node_level, cap);
    rc |= alloc3(ctx, &st->node_top, cap);
    rc |= alloc3(ctx, &st->root_rec, 2);
    rc |= alloc3(ctx, &st->counters, 8);
    if (rc) return NB_ERR_CUDA;
    NB_CUDA(ctx, cudaMemset(st->ticket, 0, sizeof(unsigned int)));
    NB_CUDA(ctx, cudaMemset(st->counters, 0, 8 * sizeof(unsigned long long)));
    st->cap = (int)cap;
    return NB_OK;
}

// stages: 1 = bounds, 2 = keys, 3 = sort, 4 = tree
int run_build3(nb_context* ctx, const float* pos_, int n, int upto, float theta, cudaStream_t stream) {
    int rc = ensure_state3(ctx, n);
    if (rc) return rc;
    nb_bh3_state* st = ctx->bh3;
    const float4* pos = reinterpret_cast<const float4*>(pos_);
    st->last_n = n;
    const int bblocks = std::min(1024, std::max(1, (n + 1023) / 1024));
    bounds3_kernel<<<bblocks, 256, 0, stream>>>(pos, n, st->bounds_part, st->ticket, st->bounds, st->counters + 4);
    if (upto >= 2) {
        const int tiles = (n + SORT_TILE - 1) / SORT_TILE;
        st->os_stride = (size_t)tiles * 256;
        NB_CUDA(ctx, cudaMemsetAsync(st->ghist, 0, ((size_t)DIGITS3 * 256 + 64 + DIGITS3 * st->os_stride) * sizeof(uint32_t), stream));
        keys3_kernel<<<(n + KEYS3_PER_BLOCK - 1) / KEYS3_PER_BLOCK, KEYS3_THREADS, 0, stream>>>(pos, n, st->bounds, st->key[0], st->ghist);
    }
    if (upto >= 3) {
        rc = radix_sort<uint64_t>(ctx, st->ghist, st->os_status, st->os_stride, st->os_ticket, st->key, st->idx, n, 0, 64, 0, 0,
                                  true, 0, nullptr, &st->sel, stream);
        if (rc) return rc;
    }
    if (upto >= 4 && n >= 2) {
        const int chunks = (n + SCAN_CHUNK - 1) / SCAN_CHUNK;
        const uint64_t* K = st->key[st->sel];
        gather_scan3_kernel<<<chunks, SCAN_THREADS, 0, stream>>>(pos, st->idx[st->sel], n, st->spos, st->psum, st->psum_blk);
        scan_totals_kernel<D4><<<1, 1024, 0, stream>>>(st->psum_blk, chunks);
        const int bchunks = (n - 1 + SCAN_CHUNK - 1) / SCAN_CHUNK;
        boundary_kernel<Level3><<<bchunks, SCAN_THREADS, 0, stream>>>(K, n, st->ext_a, st->ext_c, st->ext_r, st->flag_scan, st->flag_blk);
        scan_totals_kernel<int><<<1, 1024, 0, stream>>>(st->flag_blk, bchunks);
        TreeView3 tv{K, st->spos, st->psum, st->psum_blk, st->flag_scan, st->flag_blk, st->ext_a, st->ext_c, st->ext_r, n};
        emit3_kernel<<<(n - 1 + 255) / 256, 256, 0, stream>>>(tv, st->bounds, theta, st->crec, st->cthr, st->clink, st->node_first,
                                                              st->node_last, st->node_level, st->node_top, st->root_rec);
    }
    NB_CUDA(ctx, cudaGetLastError());
    return NB_OK;
}

}  // namespace

void nb_bh3_destroy(nb_context* ctx) {
    if (ctx->bh3) {
        free_state3(ctx->bh3);
        delete ctx->bh3;
        ctx->bh3 = nullptr;
    }
}

extern "C" {

int nb_bh3d_accel_groups(nb_context* ctx, const float* pos, float* acc, int n, float theta, float softening, int flags,
                         int group_begin, int group_count, void* stream_) {
    if (!ctx) return NB_ERR_INVALID;
    if (n < 0) return nb_fail(ctx, NB_ERR_INVALID, "n must be >= 0");
    if (n > (1 << 28)) return nb_fail(ctx, NB_ERR_INVALID, "n exceeds 2^28 bodies");
    if (n > 0 && (!pos || !acc)) return nb_fail(ctx, NB_ERR_INVALID, "null device pointer");
    if (!(theta > 0.f)) return nb_fail(ctx, NB_ERR_INVALID, "theta must be > 0");
    if (!nb_softening_ok(softening)) return nb_fail(ctx, NB_ERR_INVALID, "softening must be >= 0");
    softening = nb_effective_softening(softening);
    if (n == 0) return NB_OK;
    const int groups = (n + 31) / 32;
    if (group_begin < 0 || group_count < 0 || group_begin + group_count > groups)
        return nb_fail(ctx, NB_ERR_INVALID, "group range outside [0, ceil(n/32))");
    NB_DEVICE(ctx);
    cudaStream_t stream = (cudaStream_t)stream_;
    int rc = run_build3(ctx, pos, n, 4, theta, stream);
    if (rc) return rc;
    if (group_count == 0) return NB_OK;
    nb_bh3_state* st = ctx->bh3;
    if (n == 1) {
        zero_acc3_kernel<<<1, 32, 0, stream>>>((float4*)acc, 1);
        return NB_OK;
    }
    const int blocks = (group_count + W3_WARPS - 1) / W3_WARPS;
    int qcap = W3_QCAP - W3_DFS_ROOM;
    if (const char* v = getenv("NB_BH_WALK_QCAP")) qcap = std::max(0, std::min(qcap, atoi(v)));   // tests: force the depth-first path
    if (flags & NB_BH_COUNT_WALK) {
        cudaMemsetAsync(st->counters, 0, 8 * sizeof(unsigned long long), stream);
        walk3_kernel<true><<<blocks, W3_WARPS * 32, 0, stream>>>((const float4*)pos, st->idx[st->sel], n, st->crec, st->cthr, st->clink,
                                                                st->spos, st->root_rec, softening, group_begin, group_count,
                                                                (float4*)acc, st->counters, qcap);
    } else {
        walk3_kernel<false><<<blocks, W3_WARPS * 32, 0, stream>>>((const float4*)pos, st->idx[st->sel], n, st->crec, st->cthr, st->clink,
                                                                 st->spos, st->root_rec, softening, group_begin, group_count,
                                                                 (float4*)acc, st->counters, qcap);
    }
    NB_CUDA(ctx, cudaGetLastError());
    return NB_OK;
}

int nb_bh3d_accel(nb_context* ctx, const float* pos, float* acc, int n, float theta, float softening, int flags, void* stream) {
    return nb_bh3d_accel_groups(ctx, pos, acc, n, theta, softening, flags, 0, (n + 31) / 32, stream);
}

int nb_bh3d_stages_host(nb_context* ctx, const float* pos, int n, float* bounds6, uint64_t* tree_keys, int32_t* sorted_idx) {
    if (!ctx) return NB_ERR_INVALID;
    if (n <= 0 || !pos) return nb_fail(ctx, NB_ERR_INVALID, "need n > 0 and a device pointer");
    NB_DEVICE(ctx);
    int rc = run_build3(ctx, pos, n, 3, 1.0f, 0);
    if (rc) return rc;
    nb_bh3_state* st = ctx->bh3;
    NB_CUDA(ctx, cudaDeviceSynchronize());
    if (bounds6) {
        float4 b[2];
        NB_CUDA(ctx, cudaMemcpy(b, st->bounds, sizeof(b), cudaMemcpyDeviceToHost));
        bounds6[0] = b[0].x; bounds6[1] = b[0].y; bounds6[2] = b[0].z;
        bounds6[3] = b[1].x; bounds6[4] = b[1].y; bounds6[5] = b[1].z;
    }
    if (sorted_idx) NB_CUDA(ctx, cudaMemcpy(sorted_idx, st->idx[st->sel], (size_t)n * 4, cudaMemcpyDeviceToHost));
    if (tree_keys) {   // keys in body order: recompute into the half the sort no longer needs
        keys3_kernel<<<(n + KEYS3_PER_BLOCK - 1) / KEYS3_PER_BLOCK, KEYS3_THREADS>>>(reinterpret_cast<const float4*>(pos), n, st->bounds,
                                                                                   st->key[st->sel ^ 1], nullptr);
        NB_CUDA(ctx, cudaDeviceSynchronize());
        NB_CUDA(ctx, cudaMemcpy(tree_keys, st->key[st->sel ^ 1], (size_t)n * 8, cudaMemcpyDeviceToHost));
    }
    return NB_OK;
}

int nb_bh3d_export_tree(nb_context* ctx, int* n_cells, uint64_t* prefix, int32_t* level, int32_t* count, double* mass, double* cx,
                        double* cy, double* cz) {
    if (!ctx || !n_cells) return NB_ERR_INVALID;
    nb_bh3_state* st = ctx->bh3;
    if (!st || st->last_n < 1) return nb_fail(ctx, NB_ERR_INVALID, "no octree has been built on this handle");
    NB_DEVICE(ctx);
    NB_CUDA(ctx, cudaDeviceSynchronize());
    const int n = st->last_n;
    if (n < 2) { *n_cells = 0; return NB_OK; }
    const int chunks = (n - 1 + SCAN_CHUNK - 1) / SCAN_CHUNK;
    std::vector<int32_t> fblk(chunks), fs(n - 1);
    NB_CUDA(ctx, cudaMemcpy(fblk.data(), st->flag_blk, chunks * 4, cudaMemcpyDeviceToHost));
    NB_CUDA(ctx, cudaMemcpy(fs.data(), st->flag_scan, (size_t)(n - 1) * 4, cudaMemcpyDeviceToHost));
    const int K = fs[n - 2] + fblk[(n - 2) >> SCAN_SHIFT];
    std::vector<int32_t> first(K), last(K);
    std::vector<int8_t> lev(K), top(K);
    std::vector<uint64_t> keys(n);
    std::vector<D4> ps(n), pb((n + SCAN_CHUNK - 1) / SCAN_CHUNK);
    NB_CUDA(ctx, cudaMemcpy(first.data(), st->node_first, (size_t)K * 4, cudaMemcpyDeviceToHost));
    NB_CUDA(ctx, cudaMemcpy(last.data(), st->node_last, (size_t)K * 4, cudaMemcpyDeviceToHost));
    NB_CUDA(ctx, cudaMemcpy(lev.data(), st->node_level, (size_t)K, cudaMemcpyDeviceToHost));
    NB_CUDA(ctx, cudaMemcpy(top.data(), st->node_top, (size_t)K, cudaMemcpyDeviceToHost));
    NB_CUDA(ctx, cudaMemcpy(keys.data(), st->key[st->sel], (size_t)n * 8, cudaMemcpyDeviceToHost));
    NB_CUDA(ctx, cudaMemcpy(ps.data(), st->psum, (size_t)n * sizeof(D4), cudaMemcpyDeviceToHost));
    NB_CUDA(ctx, cudaMemcpy(pb.data(), st->psum_blk, pb.size() * sizeof(D4), cudaMemcpyDeviceToHost));
    auto pre = [&](int i) { return ps[i] + pb[i >> SCAN_SHIFT]; };
    struct Cell { uint64_t prefix; int level, count; double m, cx, cy, cz; };
    std::vector<Cell> cells;
    for (int id = 0; id < K; id++) {
        const int a = first[id], b = last[id];
        D4 hi = pre(b), lo = a > 0 ? pre(a - 1) : D4{0, 0, 0, 0};
        const double m = hi.m - lo.m;
        const double ccx = m > 0 ? (hi.mx - lo.mx) / m : 0.0, ccy = m > 0 ? (hi.my - lo.my) / m : 0.0,
                     ccz = m > 0 ? (hi.mz - lo.mz) / m : 0.0;
        for (int l = top[id]; l <= lev[id]; l++) {               // expand the compressed chain
            const uint64_t mask = l == 0 ? 0ull : ((~0ull << (63 - 3 * l)) & 0x7FFFFFFFFFFFFFFFull);
            cells.push_back(Cell{keys[a] & mask, l, b - a + 1, m, ccx, ccy, ccz});
        }
    }
    std::sort(cells.begin(), cells.end(), [](const Cell& p, const Cell& q) {
        return p.prefix != q.prefix ? p.prefix < q.prefix : p.level < q.level;
    });
    const int have = *n_cells;
    *n_cells = (int)cells.size();
    if (prefix) {
        if (have < (int)cells.size()) return nb_fail(ctx, NB_ERR_INVALID, "export buffers too small");
        for (size_t k = 0; k < cells.size(); k++) {
            prefix[k] = cells[k].prefix; level[k] = cells[k].level; count[k] = cells[k].count;
            mass[k] = cells[k].m; cx[k] = cells[k].cx; cy[k] = cells[k].cy; cz[k] = cells[k].cz;
        }
    }
    return NB_OK;
}

int nb_bh3d_last_counters(nb_context* ctx, int64_t counters[6]) {
    if (!ctx || !counters) return NB_ERR_INVALID;
    nb_bh3_state* st = ctx->bh3;
    if (!st) return nb_fail(ctx, NB_ERR_INVALID, "no 3D Barnes-Hut evaluation has run on this handle");
    NB_DEVICE(ctx);
    NB_CUDA(ctx, cudaDeviceSynchronize());
    unsigned long long c[8];
    NB_CUDA(ctx, cudaMemcpy(c, st->counters, sizeof(c), cudaMemcpyDeviceToHost));
    int nodes = 0;
    const int n = st->last_n;
    if (n >= 2) {
        int32_t a = 0, b = 0;
        NB_CUDA(ctx, cudaMemcpy(&a, st->flag_scan + (n - 2), 4, cudaMemcpyDeviceToHost));
        NB_CUDA(ctx, cudaMemcpy(&b, st->flag_blk + ((n - 2) >> SCAN_SHIFT), 4, cudaMemcpyDeviceToHost));
        nodes = a + b;
    }
    counters[0] = (int64_t)c[0]; counters[1] = (int64_t)c[1]; counters[2] = (int64_t)c[2]; counters[3] = (int64_t)c[3];
    counters[4] = nodes; counters[5] = (int64_t)c[4];
    return NB_OK;
}

}  // extern "C"
