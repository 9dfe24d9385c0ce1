// direct.cu -- all-pairs softened gravity for sm_100a.
//
// Replaces direct2DComputeKernel / direct3DComputeKernel (reference src/DirectComputeKernels.cu:10-37,46-75).
// Same per-interaction arithmetic as the reference (so that NB_DIRECT_ORDERED reproduces it bit for bit):
//     d   = s_j - b_i                                  (:61-63)
//     r2  = fma(dz,dz, fma(dy,dy, fma(dx,dx, eps)))    (:64, contracted left to right)
//     inv = MUFU.RSQ(r2)                               (:65)
//     s   = ((inv*inv)*inv)*m_j   [2D: (inv*inv)*m_j]  (:66 / :28)
//     a  += d*s  (fma)                                 (:67)
// What is different is everything around it:
//   * FP32x2 packed math (FADD2/FFMA2/FMUL2): one instruction serves two targets of the same thread; the
//     source coordinate enters as a scalar-broadcast operand (SASS `R.F32`), so a source costs one LDS.128.
//     Scalar FFMA with three distinct register operands is register-bank limited to ~64 % of the FP32 lane
//     peak on B200 (profiles/ubench), the packed form reaches 99 %.
//   * each thread owns 2*PAIRS targets in registers; one broadcast LDS.128 per source serves all of them.
//   * source tiles stream from `pos` into shared memory with cp.async.bulk (TMA, UBLKCP) into a 3-stage ring
//     guarded by mbarriers; no __syncthreads in the main loop and no staging copy of the sources.
//   * the source range can be split over blockIdx.y (partials reduced in fixed order -> deterministic) so
//     that small target counts (N = 16,384; an 1/8 slice of 1M on 8 GPUs) still fill 148 SMs.
// Measured dispatch model (profiles/README.md): an SM sub-partition spends 2 cycles per packed FP32x2
// instruction, ~2 per MUFU and 1 per anything else, so the loop is kept free of everything else.
#include "common.cuh"
#include "integrate.cuh"

#include <cstring>

namespace {

// compile-time tuning knobs (profiles/build_variant.sh); the defaults are the measured best
#ifndef NB_DIRECT_TILE
#define NB_DIRECT_TILE 1024
#endif
#ifndef NB_DIRECT_STAGES
#define NB_DIRECT_STAGES 3
#endif
#ifndef NB_DIRECT_UNROLL
#define NB_DIRECT_UNROLL 16
#endif
#ifndef NB_DIRECT_BIG_PAIRS
#define NB_DIRECT_BIG_PAIRS 3
#endif
constexpr int NT = 256;      // threads per CTA
constexpr int TILE = NB_DIRECT_TILE;   // sources per pipeline stage (16 KB)
constexpr int STAGES = NB_DIRECT_STAGES;
constexpr int J_UNROLL = NB_DIRECT_UNROLL;
constexpr int BIG = NB_DIRECT_BIG_PAIRS;   // targets per thread / 2 of the variant used when targets alone fill the GPU
constexpr int REC_BYTES = 16;
constexpr int RING_BYTES = STAGES * TILE * REC_BYTES + 2 * STAGES * 8;
// + the running totals of the two-level sum: 3 components x PAIRS packed pairs per thread
constexpr int smem_bytes(int pairs) { return RING_BYTES + 64 + 3 * pairs * NT * 8; }

// UNIFORM: every source of the tile has the reference mass, the factor m_j is hoisted out of the sum (applied once
// in the epilogue) and an interaction costs 11 instead of 12 FP32-pipe operations.  `mscale` is 1 in the exact
// kernels and 1/m_ref in the hoisting kernels (sources of mixed tiles then carry m_j/m_ref).
template <int DIM, int PAIRS, bool UNIFORM>
__device__ __forceinline__ void interact(const float4 s, const uint64_t (&BX)[PAIRS], const uint64_t (&BY)[PAIRS],
                                         const uint64_t (&BZ)[PAIRS], uint64_t (&AX)[PAIRS], uint64_t (&AY)[PAIRS],
                                         uint64_t (&AZ)[PAIRS], const uint64_t E2, const float mscale, const bool scaled) {
    const float mw = scaled ? s.w * mscale : s.w;
    const uint64_t sx = pk2(s.x, s.x), sy = pk2(s.y, s.y), sz = pk2(s.z, s.z), sm = pk2(mw, mw);
#pragma unroll
    for (int p = 0; p < PAIRS; p++) {
        uint64_t dx = sub2(sx, BX[p]);
        uint64_t dy = sub2(sy, BY[p]);
        uint64_t r2 = fma2(dx, dx, E2);
        r2 = fma2(dy, dy, r2);
        uint64_t dz;
        if (DIM == 3) {
            dz = sub2(sz, BZ[p]);
            r2 = fma2(dz, dz, r2);
        }
        float r0, r1;
        up2(r2, r0, r1);
        uint64_t inv = pk2(rsqrt_approx(r0), rsqrt_approx(r1));
        uint64_t sc = mul2(inv, inv);
        if (DIM == 3) sc = mul2(sc, inv);
        if (!UNIFORM) sc = mul2(sc, sm);
        AX[p] = fma2(dx, sc, AX[p]);
        AY[p] = fma2(dy, sc, AY[p]);
        if (DIM == 3) AZ[p] = fma2(dz, sc, AZ[p]);
    }
}

template <int DIM, int PAIRS, int MINB, bool HOIST>
__global__ void __launch_bounds__(NT, MINB)
direct_kernel(const float4* __restrict__ pos,     // all n bodies: sources, and targets are read from here too
              int n, int target_begin, int target_count, float eps,
              float4* __restrict__ out,           // acc (jsplit == 1) or partial sums [jsplit][target_count]
              int jsplit,
              const unsigned char* __restrict__ tile_uniform,   // HOIST: 1 where all 1024 masses of a tile equal m_ref
              const float* __restrict__ mref,                   // HOIST: [0] = m_ref, [1] = 1/m_ref
              int two_level) {   // 1: tile sums are added to running totals in shared memory (chains of 1024 terms);
                                 // 0: one accumulator per target over the whole source range, the reference's own sum
    float m_ref = 1.f, inv_mref = 1.f;
    bool hoist_on = false;
    if (HOIST) { m_ref = mref[0]; inv_mref = mref[1]; hoist_on = mref[2] != 0.f; }
    extern __shared__ __align__(128) unsigned char smem[];
    const float4* tiles = reinterpret_cast<const float4*>(smem);
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + STAGES * TILE * REC_BYTES);
    uint64_t* empty = full + STAGES;
    uint64_t* tot = reinterpret_cast<uint64_t*>(smem + RING_BYTES + 64) + threadIdx.x;   // [3 * PAIRS][NT], this thread's column
    if (two_level) {
#pragma unroll
        for (int k = 0; k < 3 * PAIRS; k++) tot[k * NT] = 0ull;
    }

    const int tid = threadIdx.x;
    const int js = blockIdx.y;
    const int ntiles = (n + TILE - 1) / TILE;
    const int tile0 = (int)((long long)ntiles * js / jsplit);
    const int tile1 = (int)((long long)ntiles * (js + 1) / jsplit);
    const int my_tiles = tile1 - tile0;

    // targets: thread owns slice-local ids base + k*NT + tid, k < 2*PAIRS (coalesced loads / stores)
    const int base = blockIdx.x * (NT * 2 * PAIRS);
    uint64_t BX[PAIRS], BY[PAIRS], BZ[PAIRS];
    uint64_t AX[PAIRS], AY[PAIRS], AZ[PAIRS];
#pragma unroll
    for (int p = 0; p < PAIRS; p++) {
        int i0 = min(base + (2 * p) * NT + tid, target_count - 1);
        int i1 = min(base + (2 * p + 1) * NT + tid, target_count - 1);
        float4 t0 = pos[target_begin + i0];
        float4 t1 = pos[target_begin + i1];
        BX[p] = pk2(t0.x, t1.x);
        BY[p] = pk2(t0.y, t1.y);
        BZ[p] = pk2(t0.z, t1.z);
        AX[p] = AY[p] = AZ[p] = 0ull;  // (+0.f, +0.f)
    }
    const uint64_t E2 = pk2(eps, eps);

    auto tile_count = [&](int t) { return min(TILE, n - t * TILE); };  // bodies in global tile t
    auto issue = [&](int t, int s) {                                  // one thread: TMA tile t -> stage s
        const uint32_t bytes = (uint32_t)tile_count(t) * REC_BYTES;
        mbar_expect_tx(&full[s], bytes);
        bulk_g2s(smem + s * TILE * REC_BYTES, pos + (size_t)t * TILE, bytes, &full[s]);
    };

    if (tid == 0) {
        for (int s = 0; s < STAGES; s++) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], NT / 32);
        }
        mbar_fence_init();
    }
    __syncthreads();
    if (tid == 0) {
        const int pre = my_tiles < STAGES ? my_tiles : STAGES;
        for (int s = 0; s < pre; s++) issue(tile0 + s, s);
    }

    for (int it = 0; it < my_tiles; it++) {
        const int s = it % STAGES;
        mbar_wait(&full[s], (it / STAGES) & 1);
        const float4* tp = tiles + s * TILE;
        const int cnt = tile_count(tile0 + it);
        if (HOIST && hoist_on && cnt == TILE && tile_uniform[tile0 + it]) {
#pragma unroll J_UNROLL
            for (int j = 0; j < TILE; j++) interact<DIM, PAIRS, true>(tp[j], BX, BY, BZ, AX, AY, AZ, E2, 1.f, false);
        } else if (cnt == TILE) {
#pragma unroll J_UNROLL
            for (int j = 0; j < TILE; j++) interact<DIM, PAIRS, false>(tp[j], BX, BY, BZ, AX, AY, AZ, E2, inv_mref, HOIST);
        } else {  // ragged last tile of the source array
            for (int j = 0; j < cnt; j++) interact<DIM, PAIRS, false>(tp[j], BX, BY, BZ, AX, AY, AZ, E2, inv_mref, HOIST);
        }
        if (two_level) {
            // A target's FP32 sum over 10^5..10^6 sources drifts from the exact sum like a random walk of its rounding
            // errors (measured at N = 1,048,576: 1.6e-5 of the field norm with one accumulator per 262,144 sources, and
            // 4e-3 for the reference kernel's single accumulator).  Adding each tile's 1024-term sum to a running total
            // keeps both chains short; the totals live in shared memory so that no register is held for them.
#pragma unroll
            for (int p = 0; p < PAIRS; p++) {
                tot[(0 * PAIRS + p) * NT] = add2(tot[(0 * PAIRS + p) * NT], AX[p]);
                tot[(1 * PAIRS + p) * NT] = add2(tot[(1 * PAIRS + p) * NT], AY[p]);
                if (DIM == 3) tot[(2 * PAIRS + p) * NT] = add2(tot[(2 * PAIRS + p) * NT], AZ[p]);
                AX[p] = AY[p] = AZ[p] = 0ull;
            }
        }
        __syncwarp();
        if ((tid & 31) == 0) mbar_arrive(&empty[s]);
        // refill the stage consumed one iteration ago (every warp has had a whole tile of time to leave it)
        if (tid == 0 && it >= 1) {
            const int prev = it - 1;
            const int nxt = prev + STAGES;
            if (nxt < my_tiles) {
                const int sp = prev % STAGES;
                mbar_wait(&empty[sp], (prev / STAGES) & 1);
                issue(tile0 + nxt, sp);
            }
        }
    }

    float4* dst = out + (size_t)js * target_count;
#pragma unroll
    for (int p = 0; p < PAIRS; p++) {
        if (two_level) {
            AX[p] = tot[(0 * PAIRS + p) * NT];
            AY[p] = tot[(1 * PAIRS + p) * NT];
            if (DIM == 3) AZ[p] = tot[(2 * PAIRS + p) * NT];
        }
        float x0, x1, y0, y1, z0 = 0.f, z1 = 0.f;
        up2(AX[p], x0, x1);
        up2(AY[p], y0, y1);
        if (DIM == 3) up2(AZ[p], z0, z1);
        if (HOIST) { x0 *= m_ref; x1 *= m_ref; y0 *= m_ref; y1 *= m_ref; z0 *= m_ref; z1 *= m_ref; }
        int i0 = base + (2 * p) * NT + tid;
        int i1 = base + (2 * p + 1) * NT + tid;
        if (i0 < target_count) dst[i0] = make_float4(x0, y0, z0, 0.f);
        if (i1 < target_count) dst[i1] = make_float4(x1, y1, z1, 0.f);
    }
}

// acc[i] = partial[0][i] + partial[1][i] + ... in that order (deterministic).
__global__ void reduce_partials_kernel(const float4* __restrict__ partial, int target_count, int jsplit,
                                       float4* __restrict__ acc) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= target_count) return;
    float4 s = partial[i];
    for (int k = 1; k < jsplit; k++) {
        float4 v = partial[(size_t)k * target_count + i];
        s.x += v.x;
        s.y += v.y;
        s.z += v.z;
    }
    s.w = 0.f;
    acc[i] = s;
}

__global__ void zero_acc_kernel(float4* acc, int n) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) acc[i] = make_float4(0.f, 0.f, 0.f, 0.f);
}

// nb_direct_step_slice: the stepper's kernels around the force evaluation ride on the kernels that run anyway.
//   pre  : drift of the slice's bodies before the force (pos += f(vel)), done by the tile-classification kernel (or a
//          kernel of its own when no classification runs); only meaningful when no exchange of positions has to
//          happen between the drift and the force, i.e. on one GPU
//   post : kick + drift of the slice's bodies, done by the kernel that adds up the source-range partial sums --
//          positions may only move once EVERY block has finished reading them as sources, so never by the force kernel
struct DirectFused {
    float4* pos_rw;      // all n bodies
    float4* vel;         // slice-local
    int begin, count;
    SubstepArgs pre, post;
    float4* acc_out;     // optional, slice-local
    double* stats;       // optional: per 256-target block (sum m|v|, sum m v^2 / 2) of the kicked velocities
    int dim;
};

__global__ void __launch_bounds__(256) pre_drift_kernel(const DirectFused f) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= f.count) return;
    float4 p = f.pos_rw[f.begin + i];
    const float4 v = f.vel[i];
    p.x = drift1g(f.pre, p.x, v.x, 0.f);
    p.y = drift1g(f.pre, p.y, v.y, 0.f);
    p.z = drift1g(f.pre, p.z, v.z, 0.f);
    f.pos_rw[f.begin + i] = p;
}

// acc = partial[0] + partial[1] + ... (fixed order), then the stepper's kick and drift on the same thread's body
__global__ void __launch_bounds__(256) reduce_step_kernel(const float4* __restrict__ partial, int jsplit, const DirectFused f) {
    __shared__ double red[2][8];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const bool valid = i < f.count;
    double lm = 0.0, lk = 0.0;
    if (valid) {
        float4 a = partial[i];
        for (int k = 1; k < jsplit; k++) {
            const float4 w = partial[(size_t)k * f.count + i];
            a.x += w.x;
            a.y += w.y;
            a.z += w.z;
        }
        a.w = 0.f;
        float4 v = f.vel[i];
        float4 p = f.pos_rw[f.begin + i];
        v.x = kick1(f.post, v.x, a.x, 0.f);
        v.y = kick1(f.post, v.y, a.y, 0.f);
        v.z = kick1(f.post, v.z, a.z, 0.f);
        p.x = drift1g(f.post, p.x, v.x, a.x);
        p.y = drift1g(f.post, p.y, v.y, a.y);
        p.z = drift1g(f.post, p.z, v.z, a.z);
        if (f.post.kick != NB_KICK_NONE) f.vel[i] = v;
        if (f.post.drift != NB_DRIFT_NONE) f.pos_rw[f.begin + i] = p;
        if (f.acc_out) f.acc_out[i] = a;
        if (f.stats) {   // src/SimulationStats.cu:29-60
            const double vv = f.dim == 2 ? (double)sqrtf(__fmaf_rn(v.x, v.x, __fmul_rn(v.y, v.y)))
                                         : sqrt((double)v.x * v.x + (double)v.y * v.y + (double)v.z * v.z);
            lm = vv * (double)p.w;
            lk = 0.5 * vv * vv * (double)p.w;
        }
    }
    if (f.stats) {
        for (int o = 16; o > 0; o >>= 1) {
            lm += __shfl_xor_sync(0xffffffffu, lm, o);
            lk += __shfl_xor_sync(0xffffffffu, lk, o);
        }
        if ((threadIdx.x & 31) == 0) { red[0][threadIdx.x >> 5] = lm; red[1][threadIdx.x >> 5] = lk; }
        __syncthreads();
        if (threadIdx.x == 0) {
            double m = 0.0, k = 0.0;
            for (int w = 0; w < 8; w++) { m += red[0][w]; k += red[1][w]; }
            f.stats[2 * blockIdx.x] = m;
            f.stats[2 * blockIdx.x + 1] = k;
        }
    }
}

// Per source tile: are all masses equal to the reference mass (the mass of body n/2)?  The last block to finish then
// decides whether hoisting is safe at all: m_ref must be a normal positive number with a finite reciprocal, and the
// largest |m_j| / m_ref (what a mixed tile carries instead of m_j) must stay finite in FP32.  Otherwise mref[2] = 0 and
// the force kernel takes the exact 12-operation path for every tile with the raw masses.
__global__ void __launch_bounds__(256) classify_tiles_kernel(const float4* __restrict__ pos, int n,
                                                            unsigned char* __restrict__ tile_uniform,
                                                            float* __restrict__ mref,      // [0] m_ref, [1] 1/m_ref, [2] hoisting on
                                                            unsigned int* __restrict__ ticket,
                                                            float* __restrict__ tile_max, const DirectFused fused) {
    __shared__ float red[8];
    if (fused.pre.drift != NB_DRIFT_NONE) {   // the stepper's pre-force drift of this tile's bodies that lie in the slice
        for (int j = threadIdx.x; j < TILE; j += blockDim.x) {
            const int i = blockIdx.x * TILE + j - fused.begin;
            if (i >= 0 && i < fused.count) {
                float4 p = fused.pos_rw[fused.begin + i];
                const float4 v = fused.vel[i];
                p.x = drift1g(fused.pre, p.x, v.x, 0.f);
                p.y = drift1g(fused.pre, p.y, v.y, 0.f);
                p.z = drift1g(fused.pre, p.z, v.z, 0.f);
                fused.pos_rw[fused.begin + i] = p;
            }
        }
        __syncthreads();   // the mass of body n/2 and of this tile are read below; the drift leaves .w alone, but the
                           // writes of this block must not race with its own reads of the same float4
    }
    __shared__ bool last;
    const float m = pos[n / 2].w;
    const bool usable = m >= 1.0e-30f && m <= 1.0e30f;
    const int base = blockIdx.x * TILE;
    int same = 1;
    float mx = 0.f;
    for (int j = threadIdx.x; j < TILE; j += blockDim.x) {
        const bool in = base + j < n;
        const float w = in ? pos[base + j].w : 0.f;
        same &= (in && w == m) ? 1 : 0;
        mx = fmaxf(mx, fabsf(w));
    }
    same = __syncthreads_and(same);
    auto block_max = [&](float v) {
        for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
        if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
        __syncthreads();
        v = red[0];
        for (int w = 1; w < 8; w++) v = fmaxf(v, red[w]);
        __syncthreads();
        return v;
    };
    mx = block_max(mx);
    if (threadIdx.x == 0) {
        tile_uniform[blockIdx.x] = (usable && same) ? 1 : 0;
        tile_max[blockIdx.x] = mx;
        __threadfence();
        last = atomicAdd(ticket, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (!last) return;
    __threadfence();
    float g = 0.f;
    for (int b = threadIdx.x; b < gridDim.x; b += blockDim.x) g = fmaxf(g, __ldcg(&tile_max[b]));
    g = block_max(g);
    if (threadIdx.x == 0) {
        const float inv = usable ? 1.f / m : 1.f;
        const bool ok = usable && g * inv <= 1.0e30f;   // false for inf / NaN as well
        mref[0] = ok ? m : 1.f;
        mref[1] = ok ? inv : 1.f;
        mref[2] = ok ? 1.f : 0.f;
        *ticket = 0u;
    }
}

using kernel_fn = void (*)(const float4*, int, int, int, float, float4*, int, const unsigned char*, const float*, int);

struct Variant {
    kernel_fn fn;
    int pairs;
};

const Variant g_variants[2][2][2] = {   // [hoist][dim - 2][big]; per-device state lives in nb_context::direct_occ
    {{{direct_kernel<2, 2, 3, false>, 2}, {direct_kernel<2, BIG, 2, false>, BIG}},
     {{direct_kernel<3, 2, 3, false>, 2}, {direct_kernel<3, BIG, 2, false>, BIG}}},
    {{{direct_kernel<2, 2, 3, true>, 2}, {direct_kernel<2, BIG, 2, true>, BIG}},
     {{direct_kernel<3, 2, 3, true>, 2}, {direct_kernel<3, BIG, 2, true>, BIG}}},
};

int env_int(const char* name, int dflt) {
    const char* v = getenv(name);
    return v && *v ? atoi(v) : dflt;
}

}  // namespace

static int direct_launch(nb_context* ctx, const float* pos, int n, int target_begin, int target_count, float* acc,
                         int dim, float softening, int flags, cudaStream_t stream, const DirectFused* fused, bool plan_only);

int nb_direct_launch(nb_context* ctx, const float* pos, int n, int target_begin, int target_count, float* acc,
                     int dim, float softening, int flags, cudaStream_t stream) {
    return direct_launch(ctx, pos, n, target_begin, target_count, acc, dim, softening, flags, stream, nullptr, false);
}

// plan_only: make every one-time preparation (kernel attributes, scratch reservations) and launch nothing, so that the
// same call can afterwards run inside a stream capture
static int direct_launch(nb_context* ctx, const float* pos, int n, int target_begin, int target_count, float* acc,
                         int dim, float softening, int flags, cudaStream_t stream, const DirectFused* fused, bool plan_only) {
    if (!ctx) return NB_ERR_INVALID;
    if (dim != 2 && dim != 3) return nb_fail(ctx, NB_ERR_INVALID, "dim must be 2 or 3");
    if (!nb_softening_ok(softening)) return nb_fail(ctx, NB_ERR_INVALID, "softening must be >= 0");
    softening = nb_effective_softening(softening);
    if (n < 0 || target_begin < 0 || target_count < 0 || (long long)target_begin + target_count > n)
        return nb_fail(ctx, NB_ERR_INVALID, "target slice outside [0, n)");
    if (target_count == 0) return NB_OK;
    if (!pos || (!acc && !fused)) return nb_fail(ctx, NB_ERR_INVALID, "null device pointer");
    NB_DEVICE(ctx);

    const int ntiles = (n + TILE - 1) / TILE;
    int rc = NB_OK;

    // choose the register blocking: 2 * BIG targets/thread when there are enough targets to fill the GPU with
    // them, else 4 (finer granularity); NB_DIRECT_PAIRS (2 = small, 4 = big) overrides for tuning.
    int pairs = env_int("NB_DIRECT_PAIRS", 0);
    if (pairs != 2 && pairs != 4) {
        // the big variant as soon as its target blocks times the admissible source-range splits fill two blocks per SM
        // (a 1/8 slice of a million bodies on 8 GPUs qualifies; 16,384 bodies do not)
        const long long big_blocks = (target_count + NT * 2 * BIG - 1) / (NT * 2 * BIG);
        const int max_js = (flags & NB_DIRECT_ORDERED) ? 1 : (ntiles / 2 > 64 ? 64 : (ntiles / 2 < 1 ? 1 : ntiles / 2));
        pairs = (big_blocks * max_js >= 2LL * ctx->sm_count) ? 4 : 2;
    }
    // Hoisting the common mass out of the sum changes the rounding (not the order) of the sum: exact mode keeps
    // the reference arithmetic; NB_DIRECT_NO_HOIST switches it off for tuning.
    const bool hoist = !(flags & (NB_DIRECT_ORDERED | NB_DIRECT_NO_HOIST)) && env_int("NB_DIRECT_NO_HOIST", 0) == 0;
    const Variant& v = g_variants[hoist ? 1 : 0][dim - 2][pairs == 4 ? 1 : 0];
    int& occupancy = ctx->direct_occ[hoist ? 1 : 0][dim - 2][pairs == 4 ? 1 : 0];
    if (occupancy == 0) {   // first use of this variant on the handle's device (the current device, set above)
        NB_CUDA(ctx, cudaFuncSetAttribute(v.fn, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes(v.pairs)));
        int occ = 0;
        NB_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, v.fn, NT, smem_bytes(v.pairs)));
        occupancy = occ > 0 ? occ : 1;
    }
    const int per_block = NT * 2 * v.pairs;
    const int tblocks = (target_count + per_block - 1) / per_block;
    const int slots = ctx->sm_count * occupancy;

    // split the source range so that the grid is close to a whole number of waves
    int jsplit = 1;
    if (!(flags & NB_DIRECT_ORDERED)) {
        int forced = env_int("NB_DIRECT_JSPLIT", 0);
        if (forced > 0) {
            jsplit = forced < ntiles ? forced : ntiles;
        } else {
            int max_js = ntiles / 2;  // keep >= 2 tiles (2048 sources) per CTA
            if (max_js > 64) max_js = 64;
            if (max_js < 1) max_js = 1;
            double best = -1.0;
            for (int js = 1; js <= max_js; js++) {
                if (js > 1 && (size_t)js * target_count * sizeof(float4) > ((size_t)1 << 30)) break;   // partial sums: <= 1 GiB
                long long blocks = (long long)tblocks * js;
                long long waves = (blocks + slots - 1) / slots;
                double eff = (double)blocks / (double)(waves * slots);
                if (eff > best + 1e-9) {
                    best = eff;
                    jsplit = js;
                }
                if (eff >= 0.97) break;
            }
        }
    }
    if (jsplit < 1) jsplit = 1;

    float4* out = reinterpret_cast<float4*>(acc);
    if (jsplit > 1 || fused) {   // the fused step always goes through the scratch: its last kernel is the one that moves bodies
        rc = nb_reserve(ctx, &ctx->partial, &ctx->partial_bytes, (size_t)jsplit * target_count * sizeof(float4));
        if (rc) return rc;
        out = reinterpret_cast<float4*>(ctx->partial);
    }

    int launches = 1;
    unsigned char* flags_d = nullptr;
    float* mref_d = nullptr;
    if (hoist) {
        // layout: 64-byte header (m_ref, 1/m_ref, hoisting on, ticket) | one byte per tile | one float per tile
        const size_t flag_bytes = ((size_t)ntiles + 63) & ~(size_t)63;
        const size_t had = ctx->tile_flags_bytes;
        rc = nb_reserve(ctx, &ctx->tile_flags, &ctx->tile_flags_bytes, 64 + flag_bytes + (size_t)ntiles * sizeof(float));
        if (rc) return rc;
        if (ctx->tile_flags_bytes != had) NB_CUDA(ctx, cudaMemset(ctx->tile_flags, 0, 64));   // the ticket starts at 0
        mref_d = reinterpret_cast<float*>(ctx->tile_flags);
        flags_d = reinterpret_cast<unsigned char*>(ctx->tile_flags) + 64;
        if (!plan_only) {
            DirectFused none{};
            classify_tiles_kernel<<<ntiles, 256, 0, stream>>>(reinterpret_cast<const float4*>(pos), n, flags_d, mref_d,
                                                              reinterpret_cast<unsigned int*>(ctx->tile_flags) + 4,
                                                              reinterpret_cast<float*>(flags_d + flag_bytes), fused ? *fused : none);
        }
        launches++;
    } else if (fused && fused->pre.drift != NB_DRIFT_NONE) {
        if (!plan_only) pre_drift_kernel<<<(fused->count + 255) / 256, 256, 0, stream>>>(*fused);
        launches++;
    }
    if (plan_only) return NB_OK;
    dim3 grid(tblocks, jsplit);
    const int two_level = (flags & NB_DIRECT_ORDERED) ? 0 : 1;
    v.fn<<<grid, NT, smem_bytes(v.pairs), stream>>>(reinterpret_cast<const float4*>(pos), n, target_begin, target_count,
                                                    softening, out, jsplit, flags_d, mref_d, two_level);
    if (fused) {
        reduce_step_kernel<<<(target_count + 255) / 256, 256, 0, stream>>>(reinterpret_cast<const float4*>(ctx->partial), jsplit,
                                                                          *fused);
        launches++;
    } else if (jsplit > 1) {
        reduce_partials_kernel<<<(target_count + 255) / 256, 256, 0, stream>>>(
            reinterpret_cast<const float4*>(ctx->partial), target_count, jsplit, reinterpret_cast<float4*>(acc));
        launches++;
    }
    NB_CUDA(ctx, cudaGetLastError());
    ctx->last_jsplit = jsplit;
    ctx->last_blocks = tblocks * jsplit;
    ctx->last_launches = launches;
    return NB_OK;
}

// ------------------------------------------------------------------------------------------------------
// nb_direct_step_slice -- one [pre-force drift; force; kick; drift] group of a stepper on a target slice, 3 launches
// instead of 5 (include/nbody_b200.h).  Small systems are launch-bound: the launches are captured once into a CUDA graph
// and replayed while the call's arguments stay the same.
// ------------------------------------------------------------------------------------------------------
extern "C" int nb_direct_step_slice(nb_context* ctx, float* pos, float* vel, int n, const nb_direct_step* step, void* stream_) {
    if (!ctx) return NB_ERR_INVALID;
    if (!step) return nb_fail(ctx, NB_ERR_INVALID, "step description is NULL");
    auto drift_ok = [](int k) { return k >= NB_DRIFT_NONE && k <= NB_DRIFT_VDT_F64; };   // NB_DRIFT_VV_F64 needs the old acceleration
    auto kick_ok = [](int k) { return k >= NB_KICK_NONE && k <= NB_KICK_DTA_F64; };      // NB_KICK_MEAN_F64 needs two
    if (!drift_ok(step->pre_drift_kind) || !drift_ok(step->drift_kind) || !kick_ok(step->kick_kind))
        return nb_fail(ctx, NB_ERR_INVALID, "kick/drift kind not available in the fused step");
    if (step->target_count > 0 && !vel) return nb_fail(ctx, NB_ERR_INVALID, "null device pointer");
    cudaStream_t stream = (cudaStream_t)stream_;
    DirectFused f{};
    f.pos_rw = reinterpret_cast<float4*>(pos);
    f.vel = reinterpret_cast<float4*>(vel);
    f.begin = step->target_begin;
    f.count = step->target_count;
    f.pre = SubstepArgs{NB_KICK_NONE, step->pre_drift_kind, step->dt, 0.0, step->pre_drift_coef};
    f.post = SubstepArgs{step->kick_kind, step->drift_kind, step->dt, step->kick_coef, step->drift_coef};
    f.acc_out = reinterpret_cast<float4*>(step->acc_out);
    f.stats = step->stats_out;
    f.dim = step->dim;
    static const int graph_max_n = env_int("NB_DIRECT_GRAPH_MAX_N", 65536);
    if (n > graph_max_n || step->target_count == 0)
        return direct_launch(ctx, pos, n, step->target_begin, step->target_count, nullptr, step->dim, step->softening, step->flags,
                             stream, &f, false);
    // launch-bound sizes: replay a captured graph of the same call
    // attributes + scratch first, outside any capture; it may move a scratch buffer, which the key below then reflects
    int rc = direct_launch(ctx, pos, n, step->target_begin, step->target_count, nullptr, step->dim, step->softening, step->flags,
                           stream, &f, true);
    if (rc) return rc;
    struct Key { const float* pos; const float* vel; int n; unsigned int epoch; nb_direct_step st; } key;
    std::memset(&key, 0, sizeof(key));
    key.pos = pos; key.vel = vel; key.n = n; key.epoch = ctx->scratch_epoch; key.st = *step;
    static_assert(sizeof(Key) <= sizeof(ctx->direct_graph_key), "graph key storage");
    if (!ctx->direct_graph || std::memcmp(&key, ctx->direct_graph_key, sizeof(key)) != 0) {
        if (ctx->direct_graph) { cudaGraphExecDestroy((cudaGraphExec_t)ctx->direct_graph); ctx->direct_graph = nullptr; }
        if (!ctx->direct_capture_stream)
            NB_CUDA(ctx, cudaStreamCreateWithFlags((cudaStream_t*)&ctx->direct_capture_stream, cudaStreamNonBlocking));
        cudaStream_t cs = (cudaStream_t)ctx->direct_capture_stream;
        NB_CUDA(ctx, cudaStreamBeginCapture(cs, cudaStreamCaptureModeThreadLocal));
        rc = direct_launch(ctx, pos, n, step->target_begin, step->target_count, nullptr, step->dim, step->softening, step->flags,
                           cs, &f, false);
        cudaGraph_t graph = nullptr;
        cudaError_t e = cudaStreamEndCapture(cs, &graph);
        if (rc) { if (graph) cudaGraphDestroy(graph); return rc; }
        if (e != cudaSuccess) return nb_fail(ctx, NB_ERR_CUDA, "cudaStreamEndCapture", e);
        cudaGraphExec_t exec = nullptr;
        e = cudaGraphInstantiate(&exec, graph, 0);
        cudaGraphDestroy(graph);
        if (e != cudaSuccess) return nb_fail(ctx, NB_ERR_CUDA, "cudaGraphInstantiate", e);
        ctx->direct_graph = exec;
        std::memcpy(ctx->direct_graph_key, &key, sizeof(key));
    }
    NB_CUDA(ctx, cudaGraphLaunch((cudaGraphExec_t)ctx->direct_graph, stream));
    return NB_OK;
}

// ------------------------------------------------------------------------------------------------------
// nb_direct_accel_noalloc -- the split-source direct sum for callers that are not allowed to allocate.
//
// Some hosts of this library cannot call cudaMalloc once their time loop is running: a simulation that has
// already handed every free byte of device memory to its own pools, a real-time or in-situ driver whose
// step must never hit the implicit device-wide synchronisation of cudaMalloc/cudaFree, or a plug-in that is
// only given a stream and two device pointers and no right to own memory.  For them the partial sums of the
// split-source kernel live in the fixed array below, which is part of the loaded module: every GPU that
// loads libnbody_b200.so carries these 16 MiB until the process exits.
//
//   * limit: n <= NB_NOALLOC_MAX_BODIES (262,144); larger systems are refused with NB_ERR_TOO_LARGE;
//   * nothing in the array has to start at zero -- every partial that the reduction reads is written by the
//     force kernel of the same call -- so no clearing pass is launched;
//   * all calls share the one array and it keeps the previous call's partials: two calls whose kernels can
//     overlap on the GPU (different streams, no dependency) race; serialise them;
//   * the first call on each device looks up the array's device address and raises the kernel's shared-memory
//     limit; make that first call outside any stream capture.  The device that is current at the call is used.
// ------------------------------------------------------------------------------------------------------
#include <mutex>

namespace {

constexpr size_t NOALLOC_SCRATCH_BYTES = (size_t)16 << 20;
constexpr int NOALLOC_MAX_PARTIALS = (int)(NOALLOC_SCRATCH_BYTES / sizeof(float4));   // 1,048,576 float4
constexpr int NOALLOC_MAX_DEVICES = 64;

static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_SCRATCH_BYTES];

struct NoAllocDevice {
    bool ready = false;
    float4* scratch = nullptr;   // device address of g_noalloc_scratch on this device
    int sm_count = 0;
    int occupancy[2] = {0, 0};   // [dim - 2]
};
NoAllocDevice g_noalloc_dev[NOALLOC_MAX_DEVICES];
std::mutex g_noalloc_mutex;

int noalloc_fail(int code, const char* what, cudaError_t e = cudaSuccess) { return nb_fail(nullptr, code, what, e); }

}  // namespace

extern "C" int nb_direct_accel_noalloc(const float* pos, float* acc, int n, int dim, float softening, void* stream) {
    if (dim != 2 && dim != 3) return noalloc_fail(NB_ERR_INVALID, "nb_direct_accel_noalloc: dim must be 2 or 3");
    if (n < 0) return noalloc_fail(NB_ERR_INVALID, "nb_direct_accel_noalloc: n must be >= 0");
    if (!nb_softening_ok(softening)) return noalloc_fail(NB_ERR_INVALID, "nb_direct_accel_noalloc: softening must be >= 0");
    softening = nb_effective_softening(softening);
    if (n > NB_NOALLOC_MAX_BODIES)
        return noalloc_fail(NB_ERR_TOO_LARGE,
                            "nb_direct_accel_noalloc: n exceeds NB_NOALLOC_MAX_BODIES (262144), the static scratch "
                            "cannot hold the partial sums; use nb_direct_accel");
    if (n == 0) return NB_OK;
    if (!pos || !acc) return noalloc_fail(NB_ERR_INVALID, "nb_direct_accel_noalloc: null device pointer");

    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return noalloc_fail(NB_ERR_NO_DEVICE, "nb_direct_accel_noalloc: cudaGetDevice", e);
    if (dev < 0 || dev >= NOALLOC_MAX_DEVICES) return noalloc_fail(NB_ERR_INVALID, "nb_direct_accel_noalloc: device index");
    NoAllocDevice& d = g_noalloc_dev[dev];
    kernel_fn fn = dim == 2 ? (kernel_fn)direct_kernel<2, 2, 3, false> : (kernel_fn)direct_kernel<3, 2, 3, false>;
    {
        std::lock_guard<std::mutex> lock(g_noalloc_mutex);
        if (!d.ready) {
            // once per device, with that device current: the symbol's name is only a host-side stand-in, the
            // kernel needs the address the array has on THIS device
            void* p = nullptr;
            e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
            if (e != cudaSuccess) return noalloc_fail(NB_ERR_CUDA, "nb_direct_accel_noalloc: cudaGetSymbolAddress", e);
            e = cudaDeviceGetAttribute(&d.sm_count, cudaDevAttrMultiProcessorCount, dev);
            if (e != cudaSuccess) return noalloc_fail(NB_ERR_CUDA, "nb_direct_accel_noalloc: cudaDeviceGetAttribute", e);
            kernel_fn fns[2] = {(kernel_fn)direct_kernel<2, 2, 3, false>, (kernel_fn)direct_kernel<3, 2, 3, false>};
            for (int k = 0; k < 2; k++) {
                e = cudaFuncSetAttribute(fns[k], cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes(2));
                if (e != cudaSuccess) return noalloc_fail(NB_ERR_CUDA, "nb_direct_accel_noalloc: cudaFuncSetAttribute", e);
                int occ = 0;
                e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, fns[k], NT, smem_bytes(2));
                if (e != cudaSuccess) return noalloc_fail(NB_ERR_CUDA, "nb_direct_accel_noalloc: occupancy query", e);
                d.occupancy[k] = occ > 0 ? occ : 1;
            }
            d.scratch = reinterpret_cast<float4*>(p);
            d.ready = true;
        }
    }

    // same wave-filling rule as nb_direct_accel, bounded by what the static array can hold
    const int ntiles = (n + TILE - 1) / TILE;
    const int per_block = NT * 2 * 2;
    const int tblocks = (n + per_block - 1) / per_block;
    const int slots = d.sm_count * d.occupancy[dim - 2];
    int max_js = ntiles / 2;
    if (max_js > 64) max_js = 64;
    if (max_js > NOALLOC_MAX_PARTIALS / n) max_js = NOALLOC_MAX_PARTIALS / n;
    if (max_js < 1) max_js = 1;
    int jsplit = 1;
    double best = -1.0;
    for (int js = 1; js <= max_js; js++) {
        const long long blocks = (long long)tblocks * js;
        const long long waves = (blocks + slots - 1) / slots;
        const double eff = (double)blocks / (double)(waves * slots);
        if (eff > best + 1e-9) {
            best = eff;
            jsplit = js;
        }
        if (eff >= 0.97) break;
    }

    cudaStream_t s = (cudaStream_t)stream;
    float4* out = jsplit > 1 ? d.scratch : reinterpret_cast<float4*>(acc);
    fn<<<dim3(tblocks, jsplit), NT, smem_bytes(2), s>>>(reinterpret_cast<const float4*>(pos), n, 0, n, softening, out, jsplit,
                                                        nullptr, nullptr, 1);
    if (jsplit > 1)
        reduce_partials_kernel<<<(n + 255) / 256, 256, 0, s>>>(d.scratch, n, jsplit, reinterpret_cast<float4*>(acc));
    e = cudaGetLastError();
    if (e != cudaSuccess) return noalloc_fail(NB_ERR_CUDA, "nb_direct_accel_noalloc: kernel launch", e);
    return NB_OK;
}
