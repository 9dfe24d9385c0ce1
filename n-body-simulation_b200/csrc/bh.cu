// bh.cu -- Barnes-Hut force pipeline for sm_100a (2D quadtree, like the reference).
//
// Replaces barnesHutCompute and its kernels (reference src/BarnesHutKernels.cu:11-70,108-451) with one
// stream-ordered pipeline (no host synchronisation, no per-call allocation):
//
//   bounds  -> keys -> sort(SFC) , sort(tree) -> gather + FP64 prefix sums -> boundaries -> node ids -> emit -> walk
//
// Parity contract (BASELINE.json north_star):
//   * bounds, SFC keys and the sort permutation are bit-exact against the reference (float bisection with
//     the same comparisons, fmaf padding, stable LSD radix sort);
//   * the tree has the reference's topology -- the bucket-1 region quadtree over the padded box obtained
//     by '>=' bisection (src/BarnesHutKernels.cu:266-293) -- but it is built WITHOUT locks, atomics or
//     insertion order: bodies are sorted by their 64-bit tree-path key and every internal node is found
//     from the sorted keys alone (common-prefix searches), so node ids and monopoles are deterministic.
//     Chains of single-child cells are stored once (compressed) with the level at which they finally split;
//     nb_bh_export_tree expands them again for the canonical comparison;
//   * centres of mass come from FP64 prefix sums over the sorted bodies (no float atomicAdd);
//   * the walk keeps the reference's acceptance rule: groups of 32 consecutive entries of the SFC
//     permutation vote unanimously, a cell is tested against its parent's width, level-1 cells are always
//     opened, bodies are always summed (src/BarnesHutKernels.cu:398-445), same LIFO order, same FP32
//     arithmetic per interaction.
#include "common.cuh"
#include "integrate.cuh"

#include <algorithm>
#include <cfloat>
#include <vector>

namespace {

constexpr int MAX_LEVEL = 32;             // depth cap: 32 levels of bisection = 64-bit keys; deeper cells become buckets
constexpr uint32_t LINK_EMPTY = 0xFFFFFFFFu;
constexpr uint32_t LINK_BODY = 0xFFFFFFFEu;
constexpr uint32_t LINK_BUCKET = 0x80000000u;  // link = node id, bit 31 set when the node is a depth-cap bucket
constexpr float FAR_AWAY = 3.0e18f;            // position of an empty slot: zero mass, contributes exactly 0
constexpr int WALK_STACK = 128;           // >= 3 * 33 + 1

struct D3 {
    double m, mx, my;
};
__host__ __device__ inline D3 operator+(const D3& a, const D3& b) { return D3{a.m + b.m, a.mx + b.mx, a.my + b.my}; }

}  // namespace

struct nb_bh_state {
    int cap = 0;           // bodies the buffers are sized for
    bool bounds_valid = false;
    float4* bounds = nullptr;          // (minX, minY, maxX, maxY)
    float4* bounds_part = nullptr;     // per-block partials
    unsigned int* ticket = nullptr;
    uint32_t* key32[2] = {nullptr, nullptr};
    int32_t* idx32[2] = {nullptr, nullptr};
    uint64_t* key64[2] = {nullptr, nullptr};
    int32_t* idx64[2] = {nullptr, nullptr};
    uint32_t* ghist = nullptr;         // [12][256] global digit histograms, then 12 tile tickets, then the status words
    uint32_t* os_status = nullptr;     // [12][os_stride] look-back status, os_stride = tiles * 256
    unsigned int* os_ticket = nullptr; // [12]
    size_t os_stride = 0, os_bytes = 0;
    int32_t* fix_work = nullptr;       // positions of bodies in runs the five-pass tree-key sort leaves unsorted
    // The five-pass sort reports (one store into mapped page-locked memory, no copy, no sync) when it had to fall
    // back to the conditional full sort; the host then asks for eight plain passes for the next evaluations and
    // probes again later.  A stale reading only costs time, never correctness.
    uint32_t* h_full_flag = nullptr;   // host view
    uint32_t* d_full_flag = nullptr;   // device alias of the same word
    int full_sort_left = 0;
    bool sort_full = false;
    float4* spos = nullptr;            // bodies in tree order: (x, y, m, -)
    D3* psum = nullptr;                // chunk-local inclusive prefix sums of (m, m x, m y)
    D3* psum_blk = nullptr;            // exclusive prefix of the chunk totals
    int32_t* ext_a = nullptr;          // per boundary: first body of its node
    int32_t* ext_c = nullptr;          //               first body of the child that ends at the boundary
    int32_t* ext_r = nullptr;          //               last body of the child that starts after it
    int32_t* flag_scan = nullptr;      // chunk-local inclusive scan of leader flags
    int32_t* flag_blk = nullptr;
    float4* cblk = nullptr;            // child blocks, 4 x (cx, cy, m, acceptance threshold) per node
    uint4* clink = nullptr;            // the four child links of a node (LINK_EMPTY / LINK_BODY / bucket flag | node id)
    int32_t* node_first = nullptr;
    int32_t* node_last = nullptr;
    int8_t* node_level = nullptr;
    int8_t* node_top = nullptr;        // first level of the chain this node compresses
    int32_t* root = nullptr;           // [0] root node id
    float4* root_rec = nullptr;        // the root cell as a child record (cx, cy, m, link)
    unsigned long long* counters = nullptr;  // opened, examined, interactions, max stack, overflow flag
    int last_n = 0;
    int group_work_n = 0;      // > 0: ext_a holds the per-group record counts of an instrumented walk over that many bodies
    bool perm_valid = false;   // the SFC permutation of the last evaluation is intact (nb_bh_step_groups: rows_in)
    // result locations of the last sorts
    int sfc_sel = 0, tree_sel = 0;
    // small systems are launch-bound (20 launches + 1 memset per evaluation): the whole evaluation is captured once
    // into a CUDA graph and replayed while the call's arguments stay the same
    struct GraphKey {
        const float* pos; float* acc; int n, flags, group_begin, group_count, accumulate; float theta, eps; bool sort_full;
        int qcap, sched;
        bool operator==(const GraphKey& o) const {
            return pos == o.pos && acc == o.acc && n == o.n && flags == o.flags && group_begin == o.group_begin &&
                   group_count == o.group_count && accumulate == o.accumulate && theta == o.theta && eps == o.eps &&
                   sort_full == o.sort_full && qcap == o.qcap && sched == o.sched;
        }
    };
    GraphKey graph_key{};
    cudaGraphExec_t graph_exec = nullptr;
    GraphKey seen_key{};       // above 262,144 bodies: the last uncaptured call, captured if it comes again
    bool seen_valid = false;
    cudaStream_t capture_stream = nullptr;
    // the fused step (nb_bh_step_groups) at launch-bound sizes: a few captured graphs, matched on every argument of the
    // call (a sharded stepper alternates between two rows arrays, a Forest-Ruth step between three coefficient sets)
    // Up to STEP_GRAPH_EAGER_N bodies a new argument set is captured at once; above, only when it comes a second time
    // (seen && !exec marks the first sighting), so a caller whose arguments change every call never pays for a capture.
    struct StepGraph {
        unsigned char key[448];
        cudaGraphExec_t exec = nullptr;
        bool seen = false;
    };
    StepGraph step_graphs[8];
    int step_graph_next = 0;
    // the SFC sort (target order) and the tree-key sort + tree build do not depend on each other: the former runs on
    // a side stream between a fork and a join event (each radix pass fills only a fraction of the machine)
    cudaStream_t side_stream = nullptr;
    cudaEvent_t fork_ev = nullptr, join_ev = nullptr;
};

namespace {

// What the first kernel of an evaluation does to the bodies before it measures the box (nb_bh_step_groups):
//   mode 0  nothing (plain force evaluation);
//   mode 1  the stepper's pre-force drift, in place, on all n bodies (one GPU);
//   mode 2  sharded run: `rows` holds (x', y', vx', vy') of every body after the previous evaluation's kick + drift, in
//           THAT evaluation's sorted order (`perm`: body of each sorted slot), as all-gathered from the ranks that walked
//           them; they are scattered back to body order, z takes the same drift (its velocity never changes in the 2D
//           path: a_z = 0), then the pre-force drift.  Every rank does this for all n bodies and ends up with the same
//           replicated state.
struct PrepArgs {
    int mode;
    float4* pos_rw;
    float4* vel_rw;
    const float4* rows;
    const int32_t* perm;
    SubstepArgs post;   // mode 2: drift of the previous evaluation's epilogue (applied to z here)
    SubstepArgs pre;    // modes 1, 2: drift before this evaluation (NB_DRIFT_NONE: none)
    int planar;         // mode 2: the caller vouches for v_z = 0 everywhere, so z never moves: x, y, vx, vy are stored as two
                        // 8-byte halves and nothing is read back (half the scattered traffic)
    // mode 2, rows stored by peers (nb_bh_step::sync_local): this rank's arrival words; the rows are complete once words
    // [0, sync_n) have all reached word NB_SYNC_COUNT (this rank's own number of finished evaluations)
    const unsigned int* sync_wait;
    int sync_n;
};

constexpr int NB_SYNC_COUNT = 8;            // word of the sync array that counts this rank's finished evaluations
constexpr unsigned long long SYNC_TIMEOUT_NS = 10ull * 1000ull * 1000ull * 1000ull;

__device__ __forceinline__ unsigned int ld_acquire_sys(const unsigned int* p) {
    unsigned int v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(unsigned int* p, unsigned int v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned long long global_timer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// After a rank's walk (same stream, so its row stores to the peers have completed): count the evaluation and tell every
// rank, this one included, by storing the count into word `rank` of its sync array.  Plain stores to peer memory.
struct SyncPeers {
    unsigned int* local;
    unsigned int* peer[NB_MAX_PEERS];
    int n, rank;
};
__global__ void peer_signal_kernel(const SyncPeers sp) {
    __shared__ unsigned int count_s;
    if (threadIdx.x == 0) {
        count_s = sp.local[NB_SYNC_COUNT] + 1u;
        sp.local[NB_SYNC_COUNT] = count_s;
    }
    __syncthreads();
    if ((int)threadIdx.x < sp.n) {
        __threadfence_system();
        st_release_sys(sp.peer[threadIdx.x] + sp.rank, count_s);
    }
}

// What the walk does with a target's acceleration (nb_bh_step_groups): the kick and drift that follow the force
// evaluation in the reference's steppers (src/IntegrationMethods.cu:81-93 and the other K+D kernels) run in the walk's
// epilogue, where the acceleration still sits in registers.
//   mode 0  acc[body] = a;
//   mode 1  vel[body], pos[body] updated in place (one GPU);
//   mode 2  rows[sorted slot - row_base] = (x', y', vx', vy') -- the rank's contribution to the all-gather.
// `stats` (optional): per group, sum m|v| and sum m v^2 / 2 of the kicked velocities (src/SimulationStats.cu:29-44).
struct WalkEpilogue {
    int mode;
    float4* pos_rw;
    float4* vel_rw;
    float4* rows;
    int row_base;
    SubstepArgs s;
    double* stats;
    // mode 2 with n_peers > 0: the row goes straight into EVERY rank's rows array (float4[32 ceil(n/32)], indexed by the
    // global sorted slot) -- plain coalesced stores to peer memory over NVLink, so the all-gather happens inside the
    // walk instead of after it.  (Stores only; no bulk / TMA copy touches peer memory.)
    int n_peers;
    float4* rows_peer[NB_MAX_PEERS];
};

// ------------------------------------------------------------------------------------------------------
// bounds: global min/max of x and y, padded by 0.1 % of |coord|   (src/BarnesHutKernels.cu:152-195)
// ------------------------------------------------------------------------------------------------------
__global__ void bounds_kernel(const float4* __restrict__ pos, int n, float4* part,
                              unsigned int* __restrict__ ticket, float4* __restrict__ bounds, int accumulate,
                              unsigned long long* __restrict__ walk_status, const PrepArgs prep) {
    float mnx = FLT_MAX, mny = FLT_MAX, mxx = -FLT_MAX, mxy = -FLT_MAX;
    if (prep.mode == 2 && prep.sync_wait) {
        // the rows come from the peers' walks: wait until every rank has reported the evaluation this rank finished last
        // (each block polls for itself -- no block depends on another one being resident).  Bounded: a peer that died
        // leaves a count in walk_status[2] (nb_bh_last_counters status bit 1) instead of a hung GPU.
        if (threadIdx.x == 0) {
            const unsigned int want = prep.sync_wait[NB_SYNC_COUNT];
            const unsigned long long t0 = global_timer_ns();
            for (int r = 0; r < prep.sync_n; r++) {
                while ((int)(ld_acquire_sys(prep.sync_wait + r) - want) < 0) {
                    if (global_timer_ns() - t0 > SYNC_TIMEOUT_NS) {
                        atomicAdd(walk_status + 2, 1ull);
                        break;
                    }
                }
            }
        }
        __syncthreads();
    }
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        float4 p;
        if (prep.mode == 0) {
            p = pos[i];
        } else if (prep.mode == 1) {
            p = prep.pos_rw[i];
            const float4 v = prep.vel_rw[i];
            p.x = drift1g(prep.pre, p.x, v.x, 0.f);
            p.y = drift1g(prep.pre, p.y, v.y, 0.f);
            p.z = drift1g(prep.pre, p.z, v.z, 0.f);
            prep.pos_rw[i] = p;
        } else {
            const int body = prep.perm[i];
            const float4 r = __ldcg(prep.rows + i);   // possibly written by a peer GPU: read at L2, this GPU's point of coherence
            if (prep.planar) {
                p.x = drift1g(prep.pre, r.x, r.z, 0.f);
                p.y = drift1g(prep.pre, r.y, r.w, 0.f);
                *reinterpret_cast<float2*>(&prep.pos_rw[body]) = make_float2(p.x, p.y);
                *reinterpret_cast<float2*>(&prep.vel_rw[body]) = make_float2(r.z, r.w);
                mnx = fminf(mnx, p.x);
                mny = fminf(mny, p.y);
                mxx = fmaxf(mxx, p.x);
                mxy = fmaxf(mxy, p.y);
                continue;
            }
            p = prep.pos_rw[body];
            float4 v = prep.vel_rw[body];
            p.x = r.x; p.y = r.y; v.x = r.z; v.y = r.w;
            p.z = drift1g(prep.post, p.z, v.z, 0.f);
            p.x = drift1g(prep.pre, p.x, v.x, 0.f);
            p.y = drift1g(prep.pre, p.y, v.y, 0.f);
            p.z = drift1g(prep.pre, p.z, v.z, 0.f);
            prep.pos_rw[body] = p;
            prep.vel_rw[body] = v;
        }
        mnx = fminf(mnx, p.x);
        mny = fminf(mny, p.y);
        mxx = fmaxf(mxx, p.x);
        mxy = fmaxf(mxy, p.y);
    }
    __shared__ float4 sh[32];
    __shared__ bool last;
    auto warp_red = [&]() {
        for (int o = 16; o > 0; o >>= 1) {
            mnx = fminf(mnx, __shfl_xor_sync(0xffffffffu, mnx, o));
            mny = fminf(mny, __shfl_xor_sync(0xffffffffu, mny, o));
            mxx = fmaxf(mxx, __shfl_xor_sync(0xffffffffu, mxx, o));
            mxy = fmaxf(mxy, __shfl_xor_sync(0xffffffffu, mxy, o));
        }
    };
    auto block_red = [&]() {
        warp_red();
        if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = make_float4(mnx, mny, mxx, mxy);
        __syncthreads();
        if (threadIdx.x < 32) {
            float4 v = threadIdx.x < (blockDim.x >> 5) ? sh[threadIdx.x] : make_float4(FLT_MAX, FLT_MAX, -FLT_MAX, -FLT_MAX);
            mnx = v.x; mny = v.y; mxx = v.z; mxy = v.w;
            warp_red();
        }
        __syncthreads();
    };
    block_red();
    if (threadIdx.x == 0) {
        part[blockIdx.x] = make_float4(mnx, mny, mxx, mxy);
        __threadfence();
        last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
    }
    __syncthreads();
    if (!last) return;
    __threadfence();
    mnx = FLT_MAX; mny = FLT_MAX; mxx = -FLT_MAX; mxy = -FLT_MAX;
    for (int b = threadIdx.x; b < gridDim.x; b += blockDim.x) {
        const float4 v = __ldcg(&part[b]);   // written by other blocks of this launch: read through L2
        mnx = fminf(mnx, v.x); mny = fminf(mny, v.y); mxx = fmaxf(mxx, v.z); mxy = fmaxf(mxy, v.w);
    }
    block_red();
    if (threadIdx.x == 0) {
        // pad: the reference's `v - fabsf(v) * 0.001f` is one FFMA on sm_100
        float b0 = __fmaf_rn(fabsf(mnx), -0.001f, mnx);
        float b1 = __fmaf_rn(fabsf(mny), -0.001f, mny);
        float b2 = __fmaf_rn(fabsf(mxx), 0.001f, mxx);
        float b3 = __fmaf_rn(fabsf(mxy), 0.001f, mxy);
        if (accumulate) {  // reference behaviour: running union with the previous box
            const float4 o = *bounds;
            b0 = fminf(b0, o.x); b1 = fminf(b1, o.y); b2 = fmaxf(b2, o.z); b3 = fmaxf(b3, o.w);
        }
        *bounds = make_float4(b0, b1, b2, b3);
        *ticket = 0u;
        *walk_status = 0ull;   // the status flags of nb_bh_last_counters describe THIS evaluation
    }
}

// ------------------------------------------------------------------------------------------------------
// keys: SFC key (src/BarnesHutKernels.cu:197-240, strict '>') and 64-bit tree-path key (the '>=' descent of
// :266-285, digit = path ^ 2 so that both orders agree away from exact midpoint hits)
// ------------------------------------------------------------------------------------------------------
#ifndef NB_KEYS_THREADS
#define NB_KEYS_THREADS 256
#endif
#ifndef NB_KEYS_PER_BLOCK
#define NB_KEYS_PER_BLOCK 1024
#endif
constexpr int KEYS_THREADS = NB_KEYS_THREADS;
constexpr int KEYS_PER_BLOCK = NB_KEYS_PER_BLOCK;
constexpr int N_DIGITS = 12;   // 4 digits of the 32-bit SFC key, then 8 digits of the 64-bit tree key

__global__ void __launch_bounds__(KEYS_THREADS) keys_kernel(const float4* __restrict__ pos, int n,
                                                           const float4* __restrict__ bounds,
                                                           uint32_t* __restrict__ key32, uint64_t* __restrict__ key64,
                                                           uint32_t* __restrict__ ghist /* [12][256] or null */) {
    __shared__ uint32_t h[N_DIGITS][256];
    if (ghist) {
        for (int k = threadIdx.x; k < N_DIGITS * 256; k += KEYS_THREADS) (&h[0][0])[k] = 0;
        __syncthreads();
    }
    const float4 bb = *bounds;
    for (int i = blockIdx.x * KEYS_PER_BLOCK + threadIdx.x; i < min(n, (blockIdx.x + 1) * KEYS_PER_BLOCK); i += KEYS_THREADS) {
        const float4 p = pos[i];
        uint32_t k32 = 0;
        {
            float l = bb.x, r = bb.z, b = bb.y, t = bb.w;
#pragma unroll
            for (int s = 30; s >= 0; s -= 2) {
                const float midX = (r + l) * 0.5f;
                const float midY = (t + b) * 0.5f;
                if (p.x > midX) {
                    l = midX;
                    if (p.y > midY) { b = midY; k32 |= 1u << s; }
                    else            { t = midY; k32 |= 3u << s; }
                } else {
                    r = midX;
                    if (p.y > midY) { b = midY; }
                    else            { t = midY; k32 |= 2u << s; }
                }
            }
        }
        uint64_t k64 = 0;
        {
            float minX = bb.x, maxX = bb.z, minY = bb.y, maxY = bb.w;
#pragma unroll 8
            for (int lev = 0; lev < MAX_LEVEL; lev++) {
                unsigned path = 0;
                const float midX = (maxX + minX) * 0.5f;
                if (p.x >= midX) { minX = midX; path += 1; } else { maxX = midX; }
                const float midY = (maxY + minY) * 0.5f;
                if (p.y >= midY) { minY = midY; path += 2; } else { maxY = midY; }
                k64 |= (uint64_t)(path ^ 2u) << (62 - 2 * lev);
            }
        }
        key32[i] = k32;
        key64[i] = k64;
        if (ghist) {
#pragma unroll
            for (int d = 0; d < 4; d++) atomicAdd(&h[d][(k32 >> (8 * d)) & 255u], 1u);
#pragma unroll
            for (int d = 0; d < 8; d++) atomicAdd(&h[4 + d][(uint32_t)(k64 >> (8 * d)) & 255u], 1u);
        }
    }
    if (ghist) {
        __syncthreads();
        for (int k = threadIdx.x; k < N_DIGITS * 256; k += KEYS_THREADS) {
            const uint32_t v = (&h[0][0])[k];
            if (v) atomicAdd(&ghist[k], v);
        }
    }
}

#include "bh_shared.cuh"   // radix sort + chunked scans (shared with bh3d.cu), included inside this anonymous namespace

// ------------------------------------------------------------------------------------------------------
// Tree-key sort in five passes instead of eight.  The LSD passes over the upper 40 bits (20 tree levels) leave
// the bodies ordered except inside runs that share those 40 bits -- bodies closer than 2^-20 of the box, a
// handful in a million for anything but a collapsed system.  mark_runs_kernel collects the members of such runs,
// fix_runs_kernel (one block) sorts them by the full key and puts them back into the same slots, stable by
// current position, which is exactly what three more LSD passes would have produced.  More than FIX_CAP members:
// a device flag switches on eight conditional passes (a full LSD sort from the current order -- ties keep their
// order, so the result is again the stable sort) that otherwise return at once.
// ------------------------------------------------------------------------------------------------------
// bits left to the fix-up: 24 (five passes over 20 levels).  Measured at N = 1M: four passes (32 bits left) save one 23 us
// pass but put 1300 bodies instead of a handful into the one-block repair sort, which then takes 37 us instead of 5
inline int hybrid_low_bits(int) { return 24; }
constexpr int FIX_CAP = 2048;            // members of short-of-sorted runs one block can repair
constexpr int N_SLOTS = N_DIGITS + 8;    // ticket / status areas: 12 regular passes + 8 conditional ones
constexpr int FIX_COUNT_WORD = 32, FIX_FLAG_WORD = 33;   // in the 64-word ticket area (cleared with it)

__global__ void __launch_bounds__(256) mark_runs_kernel(const uint64_t* __restrict__ K, int n, int32_t* __restrict__ work,
                                                       uint32_t* __restrict__ count, uint32_t* __restrict__ full_sort,
                                                       const int HYBRID_LOW_BITS) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    bool f = false;
    if (i < n) {
        const uint64_t hi = K[i] >> HYBRID_LOW_BITS;
        f = (i > 0 && (K[i - 1] >> HYBRID_LOW_BITS) == hi) || (i + 1 < n && (K[i + 1] >> HYBRID_LOW_BITS) == hi);
    }
    const unsigned m = __ballot_sync(0xffffffffu, f);
    if (m == 0u) return;
    if (*reinterpret_cast<volatile uint32_t*>(full_sort)) return;   // already beyond repair: no more bookkeeping
    const int lane = threadIdx.x & 31;
    uint32_t base = 0;
    if (lane == 0) base = atomicAdd(count, (uint32_t)__popc(m));
    base = __shfl_sync(0xffffffffu, base, 0);
    if (base + __popc(m) > FIX_CAP) {
        if (lane == 0) *full_sort = 1u;
        return;
    }
    if (f) work[base + __popc(m & ((1u << lane) - 1u))] = i;
}

template <typename T, typename Less>
__device__ __forceinline__ void bitonic_sort_shared(T* a, int p2, Less less) {   // p2: power of two, whole block
    for (int k = 2; k <= p2; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int t = threadIdx.x; t < p2; t += blockDim.x) {
                const int l = t ^ j;
                if (l > t) {
                    const bool up = (t & k) == 0;
                    const T x = a[t], y = a[l];
                    if (less(y, x) == up) { a[t] = y; a[l] = x; }
                }
            }
            __syncthreads();
        }
    }
}

struct FixRec { uint64_t key; int32_t pos; int32_t val; };

__global__ void __launch_bounds__(1024) fix_runs_kernel(uint64_t* __restrict__ K, int32_t* __restrict__ V,
                                                       const int32_t* __restrict__ work, const uint32_t* __restrict__ count,
                                                       const uint32_t* __restrict__ full_sort, uint32_t* host_flag) {
    __shared__ FixRec rec[FIX_CAP];
    __shared__ int32_t slot[FIX_CAP];
    const int m = (int)*count;
    if (*full_sort) {   // too many for one block: the conditional passes that follow sort everything; tell the host
        if (threadIdx.x == 0) *host_flag = 1u;
        return;
    }
    if (m == 0) return;
    int p2 = 2;
    while (p2 < m) p2 <<= 1;
    for (int j = threadIdx.x; j < p2; j += blockDim.x) {
        if (j < m) {
            const int p = work[j];
            rec[j] = FixRec{K[p], p, V[p]};
            slot[j] = p;
        } else {
            rec[j] = FixRec{~0ull, 0x7fffffff, 0};
            slot[j] = 0x7fffffff;
        }
    }
    __syncthreads();
    bitonic_sort_shared(rec, p2, [](const FixRec& a, const FixRec& b) { return a.key < b.key || (a.key == b.key && a.pos < b.pos); });
    bitonic_sort_shared(slot, p2, [](int32_t a, int32_t b) { return a < b; });
    for (int j = threadIdx.x; j < m; j += blockDim.x) {
        K[slot[j]] = rec[j].key;
        V[slot[j]] = rec[j].val;
    }
}

template <>
__device__ __forceinline__ D3 shfl_up_t<D3>(D3 v, int o) {
    return D3{__shfl_up_sync(0xffffffffu, v.m, o), __shfl_up_sync(0xffffffffu, v.mx, o), __shfl_up_sync(0xffffffffu, v.my, o)};
}
__device__ __forceinline__ D3 zero_of(D3) { return D3{0.0, 0.0, 0.0}; }

// gather bodies into tree order and scan (m, m*x, m*y) in FP64
__global__ void __launch_bounds__(SCAN_THREADS) gather_scan_kernel(const float4* __restrict__ pos,
                                                                  const int32_t* __restrict__ order, int n,
                                                                  float4* __restrict__ spos, D3* __restrict__ psum,
                                                                  D3* __restrict__ tot) {
    D3 v[SCAN_ITEMS];
    const int i0 = blockIdx.x * SCAN_CHUNK + threadIdx.x * SCAN_ITEMS;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) {
        const int i = i0 + k;
        if (i < n) {
            const float4 p = pos[order[i]];
            spos[i] = make_float4(p.x, p.y, p.w, 0.f);
            v[k] = D3{(double)p.w, (double)p.w * (double)p.x, (double)p.w * (double)p.y};
        } else {
            v[k] = D3{0.0, 0.0, 0.0};
        }
    }
    block_scan_chunk<D3>(v, &tot[blockIdx.x]);
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++)
        if (i0 + k < n) psum[i0 + k] = v[k];
}

// ------------------------------------------------------------------------------------------------------
// tree from sorted keys
// ------------------------------------------------------------------------------------------------------
__device__ __forceinline__ int cpl(uint64_t a, uint64_t b) {  // common prefix, in levels
    const uint64_t x = a ^ b;
    return x ? (__clzll((long long)x) >> 1) : MAX_LEVEL;
}

struct Level2 {   // common prefix of two tree keys in quadtree levels, for the shared boundary kernel
    static __device__ __forceinline__ int of(uint64_t a, uint64_t b) { return cpl(a, b); }
};

struct TreeView {
    const uint64_t* K;
    const float4* spos;
    const D3* psum;
    const D3* psum_blk;
    const int32_t* flag_scan;
    const int32_t* flag_blk;
    const int32_t* ext_a;
    const int32_t* ext_c;
    const int32_t* ext_r;
    int n;
};

__device__ __forceinline__ D3 prefix_incl(const TreeView& t, int i) {  // sum over sorted bodies [0, i]
    const D3 a = t.psum[i], b = t.psum_blk[i >> SCAN_SHIFT];
    return D3{a.m + b.m, a.mx + b.mx, a.my + b.my};
}
__device__ __forceinline__ int flag_incl(const TreeView& t, int i) { return t.flag_scan[i] + t.flag_blk[i >> SCAN_SHIFT]; }
__device__ __forceinline__ int level_at(const TreeView& t, int i) {   // d(i), -1 outside [0, n-1)
    return (i < 0 || i >= t.n - 1) ? -1 : cpl(t.K[i], t.K[i + 1]);
}
// id of the node that separator j (level L) belongs to: walk back over the node's earlier separators to its leader
__device__ __forceinline__ int node_of(const TreeView& t, int j, int L) {
    if (L >= MAX_LEVEL) j = t.ext_a[j];                  // bucket: every boundary separates, the first one leads
    else while (t.ext_c[j] != t.ext_a[j]) j = t.ext_c[j] - 1;   // at most three hops
    return flag_incl(t, j) - 1;
}

// Acceptance threshold (squared size / theta^2) a cell that splits at level L is tested with:
// s0 * 0.25^(L-1), s0 = (dim_size/theta)^2 (src/BarnesHutKernels.cu:403-404,411; the scaling is exact).
// Level-0/1 cells are never tested (:398-407) -> +inf.  Bodies are always summed (:430) -> 0.
__device__ __forceinline__ float threshold_of(float s0, int L) {
    if (L < 2) return __int_as_float(0x7f800000);
    return s0 * __int_as_float((127 - 2 * (L - 1)) << 23);
}

// One thread per separator, no searches: it writes the body children next to it and the empty slots between its
// two children; the LAST separator of a node (the only one that knows the node's whole range) writes the node's
// own record -- monopole from the FP64 prefix sums, acceptance threshold, link -- into its parent's child block.
__global__ void __launch_bounds__(256) emit_kernel(TreeView t, const float4* __restrict__ bounds, float theta,
                                                  float4* __restrict__ cblk, uint32_t* __restrict__ clink,
                                                  int32_t* __restrict__ node_first, int32_t* __restrict__ node_last,
                                                  int8_t* __restrict__ node_level, int8_t* __restrict__ node_top,
                                                  int32_t* __restrict__ root, float4* __restrict__ root_rec) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= t.n - 1) return;
    const uint64_t* K = t.K;
    const uint64_t ki = K[i], kn = K[i + 1];
    const int L = cpl(ki, kn);
    const int a = t.ext_a[i], c = t.ext_c[i], r = t.ext_r[i];
    const int id = node_of(t, i, L);
    const bool first = (c == a);
    const bool last = (r == t.n - 1) || (cpl(K[r], K[r + 1]) < L);
    if (first) {
        node_first[id] = a;
        node_level[id] = (int8_t)L;
    }
    if (L < MAX_LEVEL) {
        float4* blk = cblk + 4 * (size_t)id;
        uint32_t* lnk = clink + 4 * (size_t)id;
        const int ql = (int)((ki >> (62 - 2 * L)) & 3ull), qn = (int)((kn >> (62 - 2 * L)) & 3ull);   // key digits, ql < qn
        const float4 empty = make_float4(FAR_AWAY, FAR_AWAY, 0.f, 0.f);   // massless, always "accepted", adds exactly 0
        // empty slots: before the first child, between my two children, after the last child (slot = digit ^ 2)
#pragma unroll
        for (int q = 0; q < 4; q++) {
            if ((q > ql && q < qn) || (first && q < ql) || (last && q > qn)) {
                blk[q ^ 2] = empty;
                lnk[q ^ 2] = LINK_EMPTY;
            }
        }
        if (c == i) {              // the child that ends at i is the single body i
            const float4 p = t.spos[i];
            blk[ql ^ 2] = make_float4(p.x, p.y, p.z, 0.f);
            lnk[ql ^ 2] = LINK_BODY;
        }
        if (last && r == i + 1) {  // the last child is the single body i+1
            const float4 p = t.spos[i + 1];
            blk[qn ^ 2] = make_float4(p.x, p.y, p.z, 0.f);
            lnk[qn ^ 2] = LINK_BODY;
        }
    }
    if (!last) return;
    // ---- the node [a, r] as a whole ------------------------------------------------------------------------
    const int b = r;
    node_last[id] = b;
    if (L >= MAX_LEVEL)      // bucket: all keys equal; the walk sums bodies [a, b] directly
        cblk[4 * (size_t)id] = make_float4(__int_as_float(a), __int_as_float(b - a + 1), 0.f, 0.f);
    const float4 bb = *bounds;
    const float dim_size = fmaxf(bb.z - bb.x, bb.w - bb.y);       // src/BarnesHutKernels.cu:384
    float s0 = dim_size / theta;                                   // :403
    s0 *= s0;                                                      // :404
    const D3 hi = prefix_incl(t, b);
    D3 lo = D3{0.0, 0.0, 0.0};
    if (a > 0) lo = prefix_incl(t, a - 1);
    const double m = hi.m - lo.m;
    float cx = 0.f, cy = 0.f;   // centre_of_mass (src/BarnesHutKernels.cu:367-379): only where mass > 0
    if (m > 0.0) { cx = (float)((hi.mx - lo.mx) / m); cy = (float)((hi.my - lo.my) / m); }
    const float4 rec = make_float4(cx, cy, (float)m, threshold_of(s0, L));
    const uint32_t link = (uint32_t)id | (L >= MAX_LEVEL ? LINK_BUCKET : 0u);
    const int dl = level_at(t, a - 1), dr = level_at(t, b);
    if (dl < 0 && dr < 0) {       // the root cell
        root[0] = id;
        node_top[id] = 0;
        root_rec[0] = rec;
        root_rec[1] = make_float4(__uint_as_float(link), 0.f, 0.f, 0.f);
        return;
    }
    const int Lp = max(dl, dr);
    const int p = (dl >= dr) ? a - 1 : b;           // a separator of the parent, next to this node
    const int pid = node_of(t, p, Lp);
    const int slot = (int)((K[a] >> (62 - 2 * Lp)) & 3ull) ^ 2;
    node_top[id] = (int8_t)(Lp + 1);
    cblk[4 * (size_t)pid + slot] = rec;
    clink[4 * (size_t)pid + slot] = link;
}

// ------------------------------------------------------------------------------------------------------
// walk  (src/BarnesHutKernels.cu:381-451)
// ------------------------------------------------------------------------------------------------------
constexpr int WALK_WARPS = 8;

__device__ __forceinline__ float rcp_approx(float x) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

template <bool COUNT>
__global__ void __launch_bounds__(WALK_WARPS * 32) walk_kernel(const float4* __restrict__ pos,
                                                              const int32_t* __restrict__ sorted_idx, int n,
                                                              const float4* __restrict__ cblk, const uint4* __restrict__ clink,
                                                              const float4* __restrict__ spos,
                                                              const float4* __restrict__ root_rec, float eps,
                                                              int group_begin, int group_count, float4* __restrict__ acc,
                                                              unsigned long long* __restrict__ counters) {
    __shared__ uint32_t stack_s[WALK_WARPS][WALK_STACK];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int g = blockIdx.x * WALK_WARPS + warp;
    if (g >= group_count) return;
    const int bi = (group_begin + g) * 32 + lane;
    const bool valid = bi < n;
    const unsigned act = __ballot_sync(0xffffffffu, valid);
    if (!valid) return;
    const int leader = __ffs(act) - 1;
    const int body = sorted_idx[bi];
    const float4 me = pos[body];
    const float x = me.x, y = me.y;
    float ax = 0.f, ay = 0.f;
    uint32_t* stack = stack_s[warp];
    unsigned long long c_open = 0, c_exam = 0, c_inter = 0;
    int c_depth = 0;

    // One candidate (cx, cy, m, thr): summed when every target of the group has thr <= r (src/BarnesHutKernels.cu:430;
    // thr = 0 for bodies and empty slots, +inf for level-1 cells), pushed otherwise.  k = m / r  (:431).
    auto consider = [&](const float4 c, const uint32_t link, int& wr) {
        const float dx = c.x - x, dy = c.y - y;
        const float r = __fmaf_rn(dy, dy, __fmaf_rn(dx, dx, eps));
        if (__all_sync(act, c.w <= r)) {
            const float k = rcp_approx(r) * c.z;
            ax = __fmaf_rn(k, dx, ax);
            ay = __fmaf_rn(k, dy, ay);
            if (COUNT) { if (link != LINK_EMPTY) { c_exam++; c_inter++; } }
        } else {
            if (lane == leader) stack[wr] = link;
            wr++;
            if (COUNT) c_exam++;
        }
    };

    int top;
    {
        int wr = 0;
        consider(root_rec[0], __float_as_uint(root_rec[1].x), wr);
        top = wr - 1;
    }
    __syncwarp(act);
    while (top >= 0) {
        if (COUNT) { c_depth = max(c_depth, top + 1); c_open++; }
        const uint32_t ent = stack[top];
        __syncwarp(act);
        const uint32_t id = ent & 0x7FFFFFFFu;
        const float4* blk = cblk + 4 * (size_t)id;
        if (ent & LINK_BUCKET) {
            const float4 h = blk[0];
            const int first = __float_as_int(h.x), cnt = __float_as_int(h.y);
            for (int k = 0; k < cnt; k++) {
                const float4 p = spos[first + k];
                const float dx = p.x - x, dy = p.y - y;
                const float r = __fmaf_rn(dy, dy, __fmaf_rn(dx, dx, eps));
                const float kk = rcp_approx(r) * p.z;
                ax = __fmaf_rn(kk, dx, ax);
                ay = __fmaf_rn(kk, dy, ay);
            }
            if (COUNT) { c_exam += cnt; c_inter += cnt; }
            top--;
            continue;
        }
        const float4 r0 = blk[0], r1 = blk[1], r2 = blk[2], r3 = blk[3];
        const uint4 lk = clink[id];
        int wr = top;  // the first rejected child overwrites the popped slot (:436-440), LIFO order
        consider(r0, lk.x, wr);
        consider(r1, lk.y, wr);
        consider(r2, lk.z, wr);
        consider(r3, lk.w, wr);
        __syncwarp(act);
        top = wr - 1;
    }
    acc[body] = make_float4(ax, ay, 0.f, 0.f);
    if (COUNT) {
        if (lane == leader) {
            atomicAdd(&counters[0], c_open);
            atomicAdd(&counters[1], c_exam);
            atomicMax(&counters[3], (unsigned long long)c_depth);
        }
        atomicAdd(&counters[2], c_inter);
    }
}

// ------------------------------------------------------------------------------------------------------
// walk, list-building form.  Same accepted set as walk_kernel (and as the reference), different schedule:
// the 32 lanes of a group classify 32 DIFFERENT child records at a time instead of all testing the same one.
//   * a cell fails the unanimous vote as soon as ONE target fails it: every lane tests its record against the
//     group's first target with the reference's own arithmetic -> "rejected" is exact;
//   * a cell certainly passes if thr <= (eps + squared distance to the group's bounding box) with a 1e-6 safety
//     margin for FP32 rounding -> "accepted" is exact;
//   * whatever falls between the two is decided by the real 32-lane vote (src/BarnesHutKernels.cu:430).
// Accepted records go to a shared-memory list that all lanes then evaluate for their own target (k = m / r,
// :431-433); rejected nodes go to a LIFO of node ids that is popped eight nodes (32 child records) at a time.
// Summation order differs from the depth-first walk (FP32 rounding only).
// ------------------------------------------------------------------------------------------------------
#ifndef NB_WALK_UNROLL
#define NB_WALK_UNROLL 8
#endif
constexpr int WL_UNROLL = NB_WALK_UNROLL;   // packed pairs per iteration of the list evaluation
#ifndef NB_WALK_MINB
#define NB_WALK_MINB 7
#endif
constexpr int WL_MINB = NB_WALK_MINB;   // resident blocks per SM the walk is compiled for: 7 = 72 registers per thread (measured
                                        // against 8 / 6 / 5: 1.160 / 1.149 / 1.197 ms per evaluation, 7: 1.119)
constexpr int WL_WARPS = 4;
constexpr int WL_QCAP = 1024;   // node ids per group (measured peak at N = 1M, theta = 0.5: see bench counters)
constexpr int WL_LCAP = 64;     // pending accepted records per group
// A group whose queue would pass `qcap` entries (default WL_QCAP - WL_DFS_ROOM) finishes depth-first instead: one
// node per pop, all lanes voting on the same child, as walk_kernel does.  From any queue state that order adds at most
// 3 entries per remaining tree level (3 * 32 = 96) on top of the batch being pushed (<= 32), so the array cannot
// overflow and the accepted set is unchanged -- only the schedule and the order of the sum differ.
constexpr int WL_DFS_ROOM = 128;

// SPLIT > 1 (small systems, where one warp per group leaves most of the machine idle): SPLIT warps of a block share
// one group.  They walk the top of the tree in lockstep until the queue holds 8 * SPLIT nodes, then each keeps every
// SPLIT-th node (the others discard what they summed so far) and walks its share; the partial sums are added in
// warp order, so the result stays deterministic.
template <bool COUNT, int SPLIT>
__global__ void __launch_bounds__(WL_WARPS * 32, WL_MINB) walk_lists_kernel(const float4* __restrict__ pos,
                                                                  const int32_t* __restrict__ sorted_idx, int n,
                                                                  const float4* __restrict__ cblk,
                                                                  const uint32_t* __restrict__ clink,   // 4 links per node
                                                                  const float4* __restrict__ spos,
                                                                  const float4* __restrict__ root_rec, float eps,
                                                                  int group_begin, int group_count, float4* __restrict__ acc,
                                                                  unsigned long long* __restrict__ counters, int qcap,
                                                                  const WalkEpilogue epi, int32_t* __restrict__ group_work) {
    __shared__ uint32_t q_s[WL_WARPS][WL_QCAP];
    // pending accepted records, two per slot so that the evaluation runs on FP32x2: (cx0, cx1, cy0, cy1) and (m0, m1)
    __shared__ float4 lxy_s[WL_WARPS][WL_LCAP / 2];
    __shared__ float2 lm_s[WL_WARPS][WL_LCAP / 2];
    __shared__ float4 amb_s[WL_WARPS][32];   // undecided records of the current batch
    __shared__ float2 wit_s[WL_WARPS][4];
    __shared__ float2 part_s[SPLIT > 1 ? WL_WARPS : 1][32];
    static_assert(SPLIT == 1 || SPLIT == 2 || SPLIT == 4, "SPLIT warps of one block share a group");
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const unsigned FULLM = 0xffffffffu;
    const int g = (blockIdx.x * WL_WARPS + warp) / SPLIT;
    const int sub = warp % SPLIT;
    if (g >= group_count) {
        if (SPLIT > 1) __syncthreads();   // pairs with the barrier before the partial sums are added
        return;
    }
    bool lockstep = SPLIT > 1;
    const int bi = (group_begin + g) * 32 + lane;
    const bool valid = bi < n;
    // lanes past the end of the sorted order shadow the group's first target: they vote like it and store nothing
    const int body = sorted_idx[valid ? bi : (group_begin + g) * 32];
    const float4 me = pos[body];
    const float x = me.x, y = me.y;
    const float xr = __shfl_sync(FULLM, x, 0), yr = __shfl_sync(FULLM, y, 0);
    float bx0 = x, bx1 = x, by0 = y, by1 = y;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        bx0 = fminf(bx0, __shfl_xor_sync(FULLM, bx0, o));
        bx1 = fmaxf(bx1, __shfl_xor_sync(FULLM, bx1, o));
        by0 = fminf(by0, __shfl_xor_sync(FULLM, by0, o));
        by1 = fmaxf(by1, __shfl_xor_sync(FULLM, by1, o));
    }
    // the targets that touch the four sides of the box: witnesses for early rejection
    float2* const wit = wit_s[warp];
    {
        const unsigned m0 = __ballot_sync(FULLM, x == bx0), m1 = __ballot_sync(FULLM, x == bx1);
        const unsigned m2 = __ballot_sync(FULLM, y == by0), m3 = __ballot_sync(FULLM, y == by1);
        if (lane == __ffs(m0) - 1) wit[0] = make_float2(x, y);
        if (lane == __ffs(m1) - 1) wit[1] = make_float2(x, y);
        if (lane == __ffs(m2) - 1) wit[2] = make_float2(x, y);
        if (lane == __ffs(m3) - 1) wit[3] = make_float2(x, y);
        __syncwarp();
    }
    float ax = 0.f, ay = 0.f;
    uint64_t AX2 = 0ull, AY2 = 0ull;          // packed partial sums of the list evaluation (even / odd records)
    const uint64_t X2 = pk2(x, x), Y2 = pk2(y, y), E2 = pk2(eps, eps);
    uint32_t* const q = q_s[warp];
    float* const lxy = reinterpret_cast<float*>(lxy_s[warp]);
    float* const lm = reinterpret_cast<float*>(lm_s[warp]);
    float4* const ambuf = amb_s[warp];
    int qtop = 0, lcount = 0;
    unsigned long long c_open = 0, c_exam = 0, c_inter = 0;
    int c_depth = 0;
    bool depth_first = false;

    auto add = [&](float cx, float cy, float m) {
        const float dx = cx - x, dy = cy - y;
        const float r = __fmaf_rn(dy, dy, __fmaf_rn(dx, dx, eps));
        const float k = rcp_approx(r) * m;
        ax = __fmaf_rn(k, dx, ax);
        ay = __fmaf_rn(k, dy, ay);
    };
    auto append = [&](int e, float cx, float cy, float m) {   // record e of the list
        lxy[(e >> 1) * 4 + (e & 1)] = cx;
        lxy[(e >> 1) * 4 + 2 + (e & 1)] = cy;
        lm[e] = m;
    };
    auto flush = [&]() {
        if ((lcount & 1) && lane == 0) append(lcount, FAR_AWAY, FAR_AWAY, 0.f);   // odd count: a massless partner
        __syncwarp();
        const int pairs = (lcount + 1) >> 1;
#pragma unroll WL_UNROLL
        for (int k = 0; k < pairs; k++) {
            const ulonglong2 c = reinterpret_cast<const ulonglong2*>(lxy)[k];     // (cx0, cx1) (cy0, cy1)
            const uint64_t m2 = reinterpret_cast<const uint64_t*>(lm)[k];          // (m0, m1)
            const uint64_t dx = sub2(c.x, X2), dy = sub2(c.y, Y2);
            const uint64_t r2 = fma2(dy, dy, fma2(dx, dx, E2));
            float r0, r1;
            up2(r2, r0, r1);
            const uint64_t k2 = mul2(pk2(rcp_approx(r0), rcp_approx(r1)), m2);
            AX2 = fma2(k2, dx, AX2);
            AY2 = fma2(k2, dy, AY2);
        }
        __syncwarp();
        lcount = 0;
    };

    // the lane's current record; lane 0 starts with the root cell itself
    float4 rec = make_float4(0.f, 0.f, 0.f, 0.f);
    uint32_t link = LINK_EMPTY;
    if (lane == 0) {
        rec = root_rec[0];
        link = __float_as_uint(root_rec[1].x);
    }
    for (;;) {
        // ---- classify this lane's record ---------------------------------------------------------------
        const bool present = link != LINK_EMPTY;
        bool rej = false, amb = false, accp = false;
        if (present) {
            const float dxr = rec.x - xr, dyr = rec.y - yr;
            const float rr = __fmaf_rn(dyr, dyr, __fmaf_rn(dxr, dxr, eps));     // the first target's own vote
            if (!(rec.w <= rr)) {
                rej = true;
            } else {
                const float ddx = fmaxf(fmaxf(bx0 - rec.x, rec.x - bx1), 0.f);
                const float ddy = fmaxf(fmaxf(by0 - rec.y, rec.y - by1), 0.f);
                const float dmin = __fmaf_rn(ddy, ddy, __fmaf_rn(ddx, ddx, eps));
                if (rec.w <= dmin * 0.999999f) {
                    accp = true;
                } else {   // a second witness: the group's extreme target on the side the record lies on
                    const float2 wa = wit[rec.x < bx0 ? 0 : 1], wb = wit[rec.y < by0 ? 2 : 3];
                    const float dxa = rec.x - wa.x, dya = rec.y - wa.y, dxb = rec.x - wb.x, dyb = rec.y - wb.y;
                    const float ra = __fmaf_rn(dya, dya, __fmaf_rn(dxa, dxa, eps));
                    const float rb = __fmaf_rn(dyb, dyb, __fmaf_rn(dxb, dxb, eps));
                    if (!(rec.w <= fminf(ra, rb))) rej = true; else amb = true;
                }
            }
        }
        // ---- undecided records: the real unanimous vote --------------------------------------------------
        // They are compacted into shared memory and read back as broadcasts (one LDS.128 per record instead of
        // four shuffles and a find-first-set); rejections come back as a bit per compacted slot.
        const unsigned ambm = __ballot_sync(FULLM, amb);
        const unsigned below = (1u << lane) - 1u;
        if (ambm) {
            const int slot = __popc(ambm & below);
            if (amb) ambuf[slot] = rec;
            __syncwarp();
            const int na = __popc(ambm);
            unsigned rejk = 0u, bit = 1u;
#pragma unroll 2
            for (int k = 0; k < na; k++, bit += bit) {
                const float4 c = ambuf[k];
                const float dx = c.x - x, dy = c.y - y;
                const float r = __fmaf_rn(dy, dy, __fmaf_rn(dx, dx, eps));
                if (__all_sync(FULLM, c.w <= r)) {
                    const float kk = rcp_approx(r) * c.z;
                    ax = __fmaf_rn(kk, dx, ax);
                    ay = __fmaf_rn(kk, dy, ay);
                    if (COUNT) c_inter++;
                } else {
                    rejk |= bit;
                }
            }
            if (amb && ((rejk >> slot) & 1u)) rej = true;
        }
        // ---- accepted -> interaction list, rejected -> node LIFO -----------------------------------------
        const unsigned accm = __ballot_sync(FULLM, accp);
        const unsigned rejm = __ballot_sync(FULLM, rej);
        if (COUNT) { c_exam += __popc(__ballot_sync(FULLM, present)); c_inter += __popc(accm); }
        if (accp) append(lcount + __popc(accm & below), rec.x, rec.y, rec.z);
        lcount += __popc(accm);
        const int nrej = __popc(rejm);
        if (qtop + nrej > qcap) depth_first = true;   // still room for this batch (WL_DFS_ROOM), none for the schedule
        if (rej) q[qtop + __popc(rejm & below)] = link;
        qtop += nrej;
        __syncwarp();
        if (COUNT) c_depth = max(c_depth, qtop);
        if (depth_first) {
            // the rest of this group's walk in depth-first order (src/BarnesHutKernels.cu:409-445): bounded queue growth
            while (qtop > 0) {
                const uint32_t ent = q[qtop - 1];
                __syncwarp();
                qtop--;
                if (COUNT) c_open++;
                const uint32_t id = ent & 0x7FFFFFFFu;
                if (ent & LINK_BUCKET) {
                    const float4 h = cblk[4 * (size_t)id];
                    const int first = __float_as_int(h.x), cnt = __float_as_int(h.y);
                    for (int t = 0; t < cnt; t++) {
                        const float4 p = spos[first + t];
                        add(p.x, p.y, p.z);
                    }
                    if (COUNT) { c_exam += cnt; c_inter += cnt; }
                    continue;
                }
#pragma unroll
                for (int c = 0; c < 4; c++) {
                    const float4 r4 = cblk[4 * (size_t)id + c];
                    const uint32_t lk = clink[4 * (size_t)id + c];
                    const float dx = r4.x - x, dy = r4.y - y;
                    const float r = __fmaf_rn(dy, dy, __fmaf_rn(dx, dx, eps));
                    if (__all_sync(FULLM, r4.w <= r)) {
                        const float kk = rcp_approx(r) * r4.z;
                        ax = __fmaf_rn(kk, dx, ax);
                        ay = __fmaf_rn(kk, dy, ay);
                        if (COUNT && lk != LINK_EMPTY) { c_exam++; c_inter++; }
                    } else {
                        if (lane == 0) q[qtop] = lk;
                        qtop++;
                        if (COUNT) c_exam++;
                    }
                }
                __syncwarp();
                if (COUNT) c_depth = max(c_depth, qtop);
            }
            break;
        }
        if (SPLIT > 1 && lockstep && (qtop >= 8 * SPLIT || qtop == 0)) {
            // end of the common prefix: warp `sub` keeps queue entries sub, sub + SPLIT, ...; only warp 0 keeps the sums
            lockstep = false;
            if (sub != 0) {
                ax = ay = 0.f;
                AX2 = AY2 = 0ull;
                lcount = 0;
                if (COUNT) c_open = c_exam = c_inter = 0;
            }
            for (int base = 0; base < qtop; base += 32) {
                const int j = base + lane;
                const uint32_t e = j < qtop ? q[j] : 0u;
                __syncwarp();
                if (j < qtop && j % SPLIT == sub) q[j / SPLIT] = e;
                __syncwarp();
            }
            qtop = (qtop - sub + SPLIT - 1) / SPLIT;
        }
        if (qtop == 0) break;
        // ---- pop up to eight nodes: lane l takes child (l & 3) of the (l >> 2)-th node from the top ----------
        const int k = min(8, qtop);
        const int slot = lane >> 2;
        uint32_t ent = LINK_EMPTY;
        if (slot < k) ent = q[qtop - 1 - slot];
        __syncwarp();
        qtop -= k;
        if (COUNT) c_open += k;
        link = LINK_EMPTY;
        const bool is_bucket = (ent != LINK_EMPTY) && (ent & LINK_BUCKET);
        if (ent != LINK_EMPTY && !is_bucket) {
            const uint32_t id = ent & 0x7FFFFFFFu;
            rec = cblk[4 * (size_t)id + (lane & 3)];
            link = clink[4 * (size_t)id + (lane & 3)];
        }
        if (lcount >= 32) flush();   // the loads just issued are in flight while the list is evaluated
        unsigned bm = __ballot_sync(FULLM, is_bucket && (lane & 3) == 0);
        while (bm) {   // depth-cap buckets: every body of the cell, summed by all lanes
            const int src = __ffs(bm) - 1;
            bm &= bm - 1;
            const uint32_t id = __shfl_sync(FULLM, ent, src) & 0x7FFFFFFFu;
            const float4 h = cblk[4 * (size_t)id];
            const int first = __float_as_int(h.x), cnt = __float_as_int(h.y);
            for (int t = 0; t < cnt; t++) {
                const float4 p = spos[first + t];
                add(p.x, p.y, p.z);
            }
            if (COUNT) { c_exam += cnt; c_inter += cnt; }
        }
    }
    flush();
    {
        float a0, a1, b0, b1;
        up2(AX2, a0, a1);
        up2(AY2, b0, b1);
        ax += a0 + a1;
        ay += b0 + b1;
    }
    if (SPLIT > 1) {
        if (lockstep && sub != 0) ax = ay = 0.f;   // left the loop inside the common prefix (went depth-first there)
        if (sub != 0) part_s[warp][lane] = make_float2(ax, ay);
        __syncthreads();
        if (sub == 0) {
#pragma unroll
            for (int w = 1; w < SPLIT; w++) {
                const float2 p = part_s[warp + w][lane];
                ax += p.x;
                ay += p.y;
            }
        }
    }
    if (epi.mode == 0) {
        if (valid && sub == 0) acc[body] = make_float4(ax, ay, 0.f, 0.f);
    } else if (sub == 0) {
        // the stepper's kick + drift on this target (a_z = 0 in the 2D path), src/IntegrationMethods.cu:81-93 and kin
        float4 v = epi.vel_rw[body];
        float4 p = epi.pos_rw[body];   // z and mass were not kept in registers through the walk
        v.x = kick1(epi.s, v.x, ax, 0.f);
        v.y = kick1(epi.s, v.y, ay, 0.f);
        p.x = drift1g(epi.s, p.x, v.x, ax);
        p.y = drift1g(epi.s, p.y, v.y, ay);
        if (epi.mode == 1) {
            v.z = kick1(epi.s, v.z, 0.f, 0.f);
            p.z = drift1g(epi.s, p.z, v.z, 0.f);
            if (valid) {
                epi.vel_rw[body] = v;
                epi.pos_rw[body] = p;
            }
            if (acc && valid) acc[body] = make_float4(ax, ay, 0.f, 0.f);
        } else if (valid) {
            const float4 row = make_float4(p.x, p.y, v.x, v.y);
            if (epi.n_peers > 0) {
#pragma unroll
                for (int r = 0; r < NB_MAX_PEERS; r++)     // constant indices: the pointers come straight from the parameter bank
                    if (r < epi.n_peers) epi.rows_peer[r][bi] = row;
            } else {
                epi.rows[bi - epi.row_base] = row;
            }
        }
        if (epi.stats) {   // SimulationStats "moment" and kinetic energy of the kicked velocities (R2 formula, :34-36)
            double vv = (double)sqrtf(__fmaf_rn(v.x, v.x, __fmul_rn(v.y, v.y)));
            double lm = valid ? vv * (double)p.w : 0.0, lk = valid ? 0.5 * vv * vv * (double)p.w : 0.0;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                lm += __shfl_xor_sync(FULLM, lm, o);
                lk += __shfl_xor_sync(FULLM, lk, o);
            }
            if (lane == 0) {
                epi.stats[2 * (size_t)(group_begin + g)] = lm;
                epi.stats[2 * (size_t)(group_begin + g) + 1] = lk;
            }
        }
    }
    if (depth_first && lane == 0) atomicOr(&counters[4], 1ull);   // informative: some group left the batched schedule
    if (COUNT && SPLIT > 1 && lockstep && sub != 0) c_open = c_exam = c_inter = 0;   // a redundant copy of warp 0's walk
    if (COUNT && lane == 0) {
        atomicAdd(&counters[0], c_open);
        atomicAdd(&counters[1], c_exam);
        atomicMax(&counters[3], (unsigned long long)c_depth);
    }
    if (COUNT && valid) atomicAdd(&counters[2], c_inter);   // lane interactions: per valid target
    if (COUNT && lane == 0 && group_work) atomicAdd(&group_work[group_begin + g], (int)c_inter);   // records this group summed
}

__global__ void zero_acc_kernel(float4* acc, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) acc[i] = make_float4(0.f, 0.f, 0.f, 0.f);
}

// a system of one body: zero acceleration, the epilogue still applies (the drift moves it)
__global__ void lone_body_kernel(const float4* __restrict__ pos, float4* __restrict__ acc, const WalkEpilogue epi) {
    if (threadIdx.x != 0) return;
    float4 v = epi.vel_rw[0];
    float4 p = pos[0];
    v.x = kick1(epi.s, v.x, 0.f, 0.f);
    v.y = kick1(epi.s, v.y, 0.f, 0.f);
    p.x = drift1g(epi.s, p.x, v.x, 0.f);
    p.y = drift1g(epi.s, p.y, v.y, 0.f);
    if (epi.mode == 1) {
        v.z = kick1(epi.s, v.z, 0.f, 0.f);
        p.z = drift1g(epi.s, p.z, v.z, 0.f);
        epi.vel_rw[0] = v;
        epi.pos_rw[0] = p;
        if (acc) acc[0] = make_float4(0.f, 0.f, 0.f, 0.f);
    } else if (epi.n_peers > 0) {
#pragma unroll
        for (int r = 0; r < NB_MAX_PEERS; r++)
            if (r < epi.n_peers) epi.rows_peer[r][0] = make_float4(p.x, p.y, v.x, v.y);
    } else {
        epi.rows[0 - epi.row_base] = make_float4(p.x, p.y, v.x, v.y);
    }
    if (epi.stats) {
        const double vv = (double)sqrtf(__fmaf_rn(v.x, v.x, __fmul_rn(v.y, v.y)));
        epi.stats[0] = vv * (double)p.w;
        epi.stats[1] = 0.5 * vv * vv * (double)p.w;
    }
}

template <typename T>
int alloc(nb_context* ctx, T** p, size_t count) {
    NB_CUDA(ctx, cudaMalloc((void**)p, count * sizeof(T)));
    ctx->bh_bytes += count * sizeof(T);
    return NB_OK;
}

void free_state(nb_bh_state* st) {
    if (!st) return;
    if (st->graph_exec) cudaGraphExecDestroy(st->graph_exec);
    for (auto& g : st->step_graphs)
        if (g.exec) { cudaGraphExecDestroy(g.exec); g.exec = nullptr; }
    if (st->capture_stream) cudaStreamDestroy(st->capture_stream);
    if (st->h_full_flag) cudaFreeHost(st->h_full_flag);
    st->h_full_flag = st->d_full_flag = nullptr;
    if (st->side_stream) cudaStreamDestroy(st->side_stream);
    if (st->fork_ev) cudaEventDestroy(st->fork_ev);
    if (st->join_ev) cudaEventDestroy(st->join_ev);
    st->graph_exec = nullptr;
    st->capture_stream = nullptr;
    st->side_stream = nullptr;
    st->fork_ev = st->join_ev = nullptr;
    void* ptrs[] = {st->bounds, st->bounds_part, st->ticket, st->key32[0], st->key32[1], st->idx32[0], st->idx32[1],
                    st->key64[0], st->key64[1], st->idx64[0], st->idx64[1], st->ghist, st->fix_work, st->clink, st->spos, st->psum, st->psum_blk,
                    st->ext_a, st->ext_c, st->ext_r, st->flag_scan, st->flag_blk, st->cblk, st->node_first, st->node_last, st->node_level,
                    st->node_top, st->root, st->root_rec, st->counters};
    for (void* p : ptrs)
        if (p) cudaFree(p);
}

int ensure_state(nb_context* ctx, int n) {
    if (ctx->bh && ctx->bh->cap >= n) return NB_OK;
    NB_CUDA(ctx, cudaDeviceSynchronize());
    bool keep_bounds = false;
    float4 old_bounds;
    if (ctx->bh) {
        if (ctx->bh->bounds_valid) {
            cudaMemcpy(&old_bounds, ctx->bh->bounds, sizeof(float4), cudaMemcpyDeviceToHost);
            keep_bounds = true;
        }
        free_state(ctx->bh);
        delete ctx->bh;
        ctx->bh = nullptr;
        ctx->bh_bytes = 0;
    }
    nb_bh_state* st = new nb_bh_state();
    ctx->bh = st;
    const size_t cap = (size_t)std::max(n, std::max(ctx->max_bodies, 1024));
    const size_t chunks = (cap + SCAN_CHUNK - 1) / SCAN_CHUNK + 1;
    const size_t sblocks = (cap + SORT_TILE - 1) / SORT_TILE;
    int rc = 0;
    rc |= alloc(ctx, &st->bounds, 1);
    rc |= alloc(ctx, &st->bounds_part, 1024);
    rc |= alloc(ctx, &st->ticket, 1);
    for (int k = 0; k < 2; k++) {
        rc |= alloc(ctx, &st->key32[k], cap);
        rc |= alloc(ctx, &st->idx32[k], cap);
        rc |= alloc(ctx, &st->key64[k], cap);
        rc |= alloc(ctx, &st->idx64[k], cap);
    }
    const size_t os_words = (size_t)N_DIGITS * 256 + 64 + (size_t)N_SLOTS * 256 * sblocks;
    rc |= alloc(ctx, &st->fix_work, (size_t)FIX_CAP);
    st->os_bytes = os_words * sizeof(uint32_t);
    rc |= alloc(ctx, &st->ghist, os_words);
    st->os_ticket = reinterpret_cast<unsigned int*>(st->ghist + N_DIGITS * 256);
    st->os_status = st->ghist + N_DIGITS * 256 + 64;
    rc |= alloc(ctx, &st->spos, cap);
    rc |= alloc(ctx, &st->psum, cap);
    rc |= alloc(ctx, &st->psum_blk, chunks);
    rc |= alloc(ctx, &st->ext_a, cap);
    rc |= alloc(ctx, &st->ext_c, cap);
    rc |= alloc(ctx, &st->ext_r, cap);
    rc |= alloc(ctx, &st->flag_scan, cap);
    rc |= alloc(ctx, &st->flag_blk, chunks);
    rc |= alloc(ctx, &st->cblk, 4 * cap);
    rc |= alloc(ctx, &st->clink, cap);
    rc |= alloc(ctx, &st->node_first, cap);
    rc |= alloc(ctx, &st->node_last, cap);
    rc |= alloc(ctx, &st->node_level, cap);
    rc |= alloc(ctx, &st->node_top, cap);
    rc |= alloc(ctx, &st->root, 2);
    rc |= alloc(ctx, &st->root_rec, 2);
    rc |= alloc(ctx, &st->counters, 8);
    if (rc) return NB_ERR_CUDA;
    NB_CUDA(ctx, cudaHostAlloc((void**)&st->h_full_flag, sizeof(uint32_t), cudaHostAllocMapped));
    *st->h_full_flag = 0u;
    NB_CUDA(ctx, cudaHostGetDevicePointer((void**)&st->d_full_flag, st->h_full_flag, 0));
    NB_CUDA(ctx, cudaStreamCreateWithFlags(&st->side_stream, cudaStreamNonBlocking));
    NB_CUDA(ctx, cudaEventCreateWithFlags(&st->fork_ev, cudaEventDisableTiming));
    NB_CUDA(ctx, cudaEventCreateWithFlags(&st->join_ev, cudaEventDisableTiming));
    NB_CUDA(ctx, cudaMemset(st->ticket, 0, sizeof(unsigned int)));
    NB_CUDA(ctx, cudaMemset(st->counters, 0, 8 * sizeof(unsigned long long)));
    if (keep_bounds) {
        NB_CUDA(ctx, cudaMemcpy(st->bounds, &old_bounds, sizeof(float4), cudaMemcpyHostToDevice));
        st->bounds_valid = true;
    }
    st->cap = (int)cap;
    return NB_OK;
}

// once per evaluation, before run_build (and before the graph key is formed: the choice is part of it)
void pick_sort_mode(nb_bh_state* st) {
    static const bool env_full = getenv("NB_BH_FULL_SORT") != nullptr;   // measurement switch: eight plain passes
    if (*reinterpret_cast<volatile uint32_t*>(st->h_full_flag)) {
        *st->h_full_flag = 0u;
        st->full_sort_left = 64;
    }
    st->sort_full = env_full || st->full_sort_left > 0;
    if (st->full_sort_left > 0) st->full_sort_left--;
}

// stages: 1 = bounds, 2 = keys, 3 = sorts, 4 = tree
int run_build(nb_context* ctx, const float* pos_, int n, int flags, int upto, float theta, cudaStream_t stream,
              const PrepArgs& prep = PrepArgs{}) {
    int rc = ensure_state(ctx, n);
    if (rc) return rc;
    nb_bh_state* st = ctx->bh;
    const float4* pos = reinterpret_cast<const float4*>(pos_);
    st->last_n = n;
    if (upto >= 3) st->perm_valid = false;   // the sorts below overwrite the permutation a pending rows_in refers to
    if (upto >= 4) st->group_work_n = 0;     // ... and the tree stage the per-group counts of an earlier instrumented walk
    const int accumulate = ((flags & NB_BH_ACCUMULATE_BOUNDS) && st->bounds_valid) ? 1 : 0;
    int bblocks = std::min(1024, std::max(1, (n + 1023) / 1024));
    bounds_kernel<<<bblocks, 256, 0, stream>>>(pos, n, st->bounds_part, st->ticket, st->bounds, accumulate, st->counters + 4, prep);
    st->bounds_valid = true;
    if (upto >= 2) {
        // one memset clears the digit histograms, the tile tickets and the look-back status words of all 12 passes
        const int tiles = (n + SORT_TILE - 1) / SORT_TILE;
        st->os_stride = (size_t)tiles * 256;   // compact for this n, so that the words to clear are contiguous
        NB_CUDA(ctx, cudaMemsetAsync(st->ghist, 0, ((size_t)N_DIGITS * 256 + 64 + N_SLOTS * st->os_stride) * sizeof(uint32_t),
                                     stream));
        keys_kernel<<<(n + KEYS_PER_BLOCK - 1) / KEYS_PER_BLOCK, KEYS_THREADS, 0, stream>>>(pos, n, st->bounds, st->key32[0],
                                                                                          st->key64[0], st->ghist);
    }
    static const bool no_fork = getenv("NB_BH_NO_FORK") != nullptr;   // measurement switch: everything on one stream
    const bool fork = upto >= 3 && !no_fork;
    if (upto >= 3) {
        cudaStream_t sfc_stream = stream;
        if (fork) {
            NB_CUDA(ctx, cudaEventRecord(st->fork_ev, stream));
            NB_CUDA(ctx, cudaStreamWaitEvent(st->side_stream, st->fork_ev, 0));
            sfc_stream = st->side_stream;
        }
        rc = radix_sort<uint32_t>(ctx, st->ghist, st->os_status, st->os_stride, st->os_ticket, st->key32, st->idx32, n, 0, 32, 0, 0,
                                  true, 0, nullptr, &st->sfc_sel, sfc_stream);
        if (fork) {   // recorded even after a failed launch, so that a capture in progress can still be joined
            cudaEventRecord(st->join_ev, st->side_stream);
        }
        if (rc) { if (fork) cudaStreamWaitEvent(stream, st->join_ev, 0); return rc; }
        if (st->sort_full) {
            rc = radix_sort<uint64_t>(ctx, st->ghist, st->os_status, st->os_stride, st->os_ticket, st->key64, st->idx64, n, 0, 64,
                                      4, 4, true, 0, nullptr, &st->tree_sel, stream);
        } else {
            uint32_t* const fix_count = st->os_ticket + FIX_COUNT_WORD;
            uint32_t* const fix_flag = st->os_ticket + FIX_FLAG_WORD;
            int sel = 0;
            const int low_bits = hybrid_low_bits(n);
            rc = radix_sort<uint64_t>(ctx, st->ghist, st->os_status, st->os_stride, st->os_ticket, st->key64, st->idx64, n,
                                      low_bits, 64, 4, 4 + low_bits / 8, true, 0, nullptr, &sel, stream);
            if (!rc) {
                mark_runs_kernel<<<(n + 255) / 256, 256, 0, stream>>>(st->key64[sel], n, st->fix_work, fix_count, fix_flag,
                                                                      low_bits);
                fix_runs_kernel<<<1, 1024, 0, stream>>>(st->key64[sel], st->idx64[sel], st->fix_work, fix_count, fix_flag,
                                                        st->d_full_flag);
                rc = radix_sort<uint64_t>(ctx, st->ghist, st->os_status, st->os_stride, st->os_ticket, st->key64, st->idx64, n, 0,
                                          64, 4, N_DIGITS, false, sel, fix_flag, &sel, stream);
            }
            st->tree_sel = sel;
        }
        if (rc) { if (fork) cudaStreamWaitEvent(stream, st->join_ev, 0); return rc; }
    }
    if (upto >= 4 && n >= 2) {
        const int chunks = (n + SCAN_CHUNK - 1) / SCAN_CHUNK;
        const uint64_t* K = st->key64[st->tree_sel];
        // the mass prefix sums (side stream, behind the SFC sort) and the node boundaries both follow from the sorted
        // tree keys alone; the emit kernel needs both
        cudaStream_t sums_stream = stream;
        if (fork) {
            NB_CUDA(ctx, cudaEventRecord(st->fork_ev, stream));
            NB_CUDA(ctx, cudaStreamWaitEvent(st->side_stream, st->fork_ev, 0));
            sums_stream = st->side_stream;
        }
        gather_scan_kernel<<<chunks, SCAN_THREADS, 0, sums_stream>>>(pos, st->idx64[st->tree_sel], n, st->spos, st->psum,
                                                                    st->psum_blk);
        scan_totals_kernel<D3><<<1, 1024, 0, sums_stream>>>(st->psum_blk, chunks);
        if (fork) cudaEventRecord(st->join_ev, st->side_stream);
        const int bchunks = (n - 1 + SCAN_CHUNK - 1) / SCAN_CHUNK;
        boundary_kernel<Level2><<<bchunks, SCAN_THREADS, 0, stream>>>(K, n, st->ext_a, st->ext_c, st->ext_r, st->flag_scan, st->flag_blk);
        scan_totals_kernel<int><<<1, 1024, 0, stream>>>(st->flag_blk, bchunks);
        if (fork) NB_CUDA(ctx, cudaStreamWaitEvent(stream, st->join_ev, 0));
        TreeView tv{K, st->spos, st->psum, st->psum_blk, st->flag_scan, st->flag_blk, st->ext_a, st->ext_c, st->ext_r, n};
        emit_kernel<<<(n - 1 + 255) / 256, 256, 0, stream>>>(tv, st->bounds, theta, st->cblk,
                                                             reinterpret_cast<uint32_t*>(st->clink), st->node_first,
                                                             st->node_last, st->node_level, st->node_top, st->root,
                                                             st->root_rec);
    } else if (fork) {
        NB_CUDA(ctx, cudaStreamWaitEvent(stream, st->join_ev, 0));   // no tree stage: join behind the SFC sort
    }
    NB_CUDA(ctx, cudaGetLastError());
    return NB_OK;
}

// schedule switch (tests, measurements), read at every call: NB_BH_WALK_SPLIT = warps per group (1, 2, 4; default by
// the number of groups)
int walk_forced_split() { const char* v = getenv("NB_BH_WALK_SPLIT"); return v ? atoi(v) : 0; }

// NB_BH_WALK_QCAP (tests): a smaller queue threshold sends groups down the depth-first path early
int walk_qcap() {
    int qcap = WL_QCAP - WL_DFS_ROOM;
    if (const char* v = getenv("NB_BH_WALK_QCAP")) qcap = std::max(0, std::min(qcap, atoi(v)));
    return qcap;
}

int run_walk(nb_context* ctx, const float* pos_, float* acc_, int n, float theta, float eps, int group_begin,
             int group_count, bool count, cudaStream_t stream, const WalkEpilogue& epi = WalkEpilogue{}) {
    nb_bh_state* st = ctx->bh;
    const float4* pos = reinterpret_cast<const float4*>(pos_);
    float4* acc = reinterpret_cast<float4*>(acc_);
    if (group_count <= 0) return NB_OK;
    if (n == 1) {
        if (epi.mode == 0) zero_acc_kernel<<<1, 32, 0, stream>>>(acc, 1);
        else lone_body_kernel<<<1, 32, 0, stream>>>(pos, acc, epi);
        return NB_OK;
    }
    static const int use_dfs = [] { const char* v = getenv("NB_BH_WALK"); return (v && v[0] == 'd') ? 1 : 0; }();
    if (use_dfs && epi.mode != 0) return nb_fail(ctx, NB_ERR_INVALID, "NB_BH_WALK=dfs has no fused kick/drift epilogue");
    if (count) {
        cudaMemsetAsync(st->counters, 0, 8 * sizeof(unsigned long long), stream);
        // per-group work of the instrumented walk lands in ext_a, which the tree stage no longer needs once emit has run
        cudaMemsetAsync(st->ext_a, 0, (size_t)((n + 31) / 32) * sizeof(int32_t), stream);
        st->group_work_n = n;
    }
    if (use_dfs) {   // depth-first walk in the reference's own order (kept for comparison: NB_BH_WALK=dfs)
        const int blocks = (group_count + WALK_WARPS - 1) / WALK_WARPS;
        if (count)
            walk_kernel<true><<<blocks, WALK_WARPS * 32, 0, stream>>>(pos, st->idx32[st->sfc_sel], n, st->cblk, st->clink, st->spos,
                                                                     st->root_rec, eps, group_begin, group_count, acc, st->counters);
        else
            walk_kernel<false><<<blocks, WALK_WARPS * 32, 0, stream>>>(pos, st->idx32[st->sfc_sel], n, st->cblk, st->clink, st->spos,
                                                                      st->root_rec, eps, group_begin, group_count, acc, st->counters);
    } else {
        // fewer groups than it takes to keep the SMs busy (16 warps each): 2 or 4 warps share a group
        const int forced_split = walk_forced_split();
        const int busy = ctx->sm_count * 16;
        int split = group_count >= busy ? 1 : (2 * group_count >= busy ? 2 : 4);
        if (forced_split == 1 || forced_split == 2 || forced_split == 4) split = forced_split;
        const int blocks = (group_count * split + WL_WARPS - 1) / WL_WARPS;
        const int qcap = walk_qcap();
        const uint32_t* links = reinterpret_cast<const uint32_t*>(st->clink);
#define NB_LAUNCH_WALK(COUNTING, SPLIT)                                                                               \
    walk_lists_kernel<COUNTING, SPLIT><<<blocks, WL_WARPS * 32, 0, stream>>>(pos, st->idx32[st->sfc_sel], n, st->cblk, links, \
                                                                            st->spos, st->root_rec, eps, group_begin,         \
                                                                            group_count, acc, st->counters, qcap, epi,   \
                                                                            count ? st->ext_a : nullptr)
        if (count) {
            if (split == 1) NB_LAUNCH_WALK(true, 1); else if (split == 2) NB_LAUNCH_WALK(true, 2); else NB_LAUNCH_WALK(true, 4);
        } else {
            if (split == 1) NB_LAUNCH_WALK(false, 1); else if (split == 2) NB_LAUNCH_WALK(false, 2); else NB_LAUNCH_WALK(false, 4);
        }
#undef NB_LAUNCH_WALK
    }
    NB_CUDA(ctx, cudaGetLastError());
    return NB_OK;
}

int check_args(nb_context* ctx, const void* pos, const void* acc, int n, float theta) {
    if (!ctx) return NB_ERR_INVALID;
    if (n < 0) return nb_fail(ctx, NB_ERR_INVALID, "n must be >= 0");
    if (n > (1 << 26) - 2) return nb_fail(ctx, NB_ERR_INVALID, "n exceeds the 26-bit node id range (67,108,862)");
    if (n > 0 && (!pos || !acc)) return nb_fail(ctx, NB_ERR_INVALID, "null device pointer");
    if (!(theta > 0.f)) return nb_fail(ctx, NB_ERR_INVALID, "theta must be > 0");
    return NB_OK;
}

}  // namespace

void nb_bh_destroy(nb_context* ctx) {
    if (ctx->bh) {
        free_state(ctx->bh);
        delete ctx->bh;
        ctx->bh = nullptr;
    }
}

#define NB_BH_COUNT_FLAG NB_BH_COUNT_WALK

extern "C" {

int nb_bh_accel_groups(nb_context* ctx, const float* pos, float* acc, int n, float theta, float softening, int flags,
                       int group_begin, int group_count, void* stream) {
    int rc = check_args(ctx, pos, acc, n, theta);
    if (rc) return rc;
    if (!nb_softening_ok(softening)) return nb_fail(ctx, NB_ERR_INVALID, "softening must be >= 0");
    softening = nb_effective_softening(softening);
    if (n == 0) return NB_OK;
    const int groups = (n + 31) / 32;
    if (group_begin < 0 || group_count < 0 || group_begin + group_count > groups)
        return nb_fail(ctx, NB_ERR_INVALID, "group range outside [0, ceil(n/32))");
    NB_DEVICE(ctx);
    static const int graph_max_n = [] { const char* v = getenv("NB_BH_GRAPH_MAX_N"); return v ? atoi(v) : 2097152; }();
    const bool counting = (flags & NB_BH_COUNT_FLAG) != 0;
    rc = ensure_state(ctx, n);   // allocation must not happen inside a capture
    if (rc) return rc;
    pick_sort_mode(ctx->bh);
    nb_bh_state* st = ctx->bh;
    const int accumulate = ((flags & NB_BH_ACCUMULATE_BOUNDS) && st->bounds_valid) ? 1 : 0;
    const nb_bh_state::GraphKey key{pos, acc, n, flags, group_begin, group_count, accumulate, theta, softening, st->sort_full,
                                      walk_qcap(), walk_forced_split()};
    bool direct = n > graph_max_n || n < 2 || counting;
    // a slice of the groups means a sharded run, where a collective sits between two evaluations: measured at N = 1M on
    // two GPUs, graph -> NCCL kernel -> graph transitions cost more (+35 us per step) than the launch gaps they remove
    if (n > 262144 && group_count != groups) direct = true;
    if (!direct && n > 262144 && !(st->graph_exec && st->graph_key == key)) {
        // past the launch-bound sizes a capture costs more than it saves unless the call repeats: wait for the second one
        direct = !(st->seen_valid && st->seen_key == key);
        st->seen_key = key;
        st->seen_valid = true;
    }
    if (direct) {
        rc = run_build(ctx, pos, n, flags, 4, theta, (cudaStream_t)stream);
        if (rc) return rc;
        return run_walk(ctx, pos, acc, n, theta, softening, group_begin, group_count, counting, (cudaStream_t)stream);
    }
    // replay a captured graph of the same evaluation
    if (!st->graph_exec || !(st->graph_key == key)) {
        if (st->graph_exec) { cudaGraphExecDestroy(st->graph_exec); st->graph_exec = nullptr; }
        if (!st->capture_stream) NB_CUDA(ctx, cudaStreamCreateWithFlags(&st->capture_stream, cudaStreamNonBlocking));
        NB_CUDA(ctx, cudaStreamBeginCapture(st->capture_stream, cudaStreamCaptureModeThreadLocal));
        rc = run_build(ctx, pos, n, flags, 4, theta, st->capture_stream);
        if (!rc) rc = run_walk(ctx, pos, acc, n, theta, softening, group_begin, group_count, false, st->capture_stream);
        cudaGraph_t graph = nullptr;
        cudaError_t e = cudaStreamEndCapture(st->capture_stream, &graph);
        if (rc) { if (graph) cudaGraphDestroy(graph); return rc; }
        if (e != cudaSuccess) return nb_fail(ctx, NB_ERR_CUDA, "cudaStreamEndCapture", e);
        e = cudaGraphInstantiate(&st->graph_exec, graph, 0);
        cudaGraphDestroy(graph);
        if (e != cudaSuccess) { st->graph_exec = nullptr; return nb_fail(ctx, NB_ERR_CUDA, "cudaGraphInstantiate", e); }
        st->graph_key = key;
    }
    st->last_n = n;
    st->bounds_valid = true;
    st->perm_valid = false;
    NB_CUDA(ctx, cudaGraphLaunch(st->graph_exec, (cudaStream_t)stream));
    return NB_OK;
}

int nb_bh_step_groups(nb_context* ctx, float* pos, float* vel, int n, const nb_bh_step* step, void* stream_) {
    if (!ctx) return NB_ERR_INVALID;
    if (!step) return nb_fail(ctx, NB_ERR_INVALID, "step description is NULL");
    int rc = check_args(ctx, pos, vel, n, step->theta);
    if (rc) return rc;
    if (!nb_softening_ok(step->softening)) return nb_fail(ctx, NB_ERR_INVALID, "softening must be >= 0");
    const float eps = nb_effective_softening(step->softening);
    auto drift_ok = [](int k) { return k >= NB_DRIFT_NONE && k <= NB_DRIFT_VDT_F64; };   // NB_DRIFT_VV_F64 needs the old acceleration
    auto kick_ok = [](int k) { return k >= NB_KICK_NONE && k <= NB_KICK_DTA_F64; };      // NB_KICK_MEAN_F64 needs two
    if (!drift_ok(step->pre_drift_kind) || !drift_ok(step->drift_kind) || !kick_ok(step->kick_kind) ||
        !drift_ok(step->prev_drift_kind))
        return nb_fail(ctx, NB_ERR_INVALID, "kick/drift kind not available in the fused step (Velocity Verlet needs nb_integrate_substep)");
    if (n == 0) return NB_OK;
    const int groups = (n + 31) / 32;
    if (step->group_begin < 0 || step->group_count < 0 || step->group_begin + step->group_count > groups)
        return nb_fail(ctx, NB_ERR_INVALID, "group range outside [0, ceil(n/32))");
    if (step->flags & NB_BH_COUNT_WALK) return nb_fail(ctx, NB_ERR_INVALID, "the fused step has no instrumented walk");
    NB_DEVICE(ctx);
    cudaStream_t stream = (cudaStream_t)stream_;
    rc = ensure_state(ctx, n);
    if (rc) return rc;
    nb_bh_state* st = ctx->bh;
    PrepArgs prep{};
    prep.pos_rw = reinterpret_cast<float4*>(pos);
    prep.vel_rw = reinterpret_cast<float4*>(vel);
    prep.pre = SubstepArgs{NB_KICK_NONE, step->pre_drift_kind, step->dt, 0.0, step->pre_drift_coef};
    if (step->rows_in) {
        // the rows are in the order of the LAST evaluation on this handle: its permutation is still in the sort buffers
        if (!st->perm_valid || st->last_n != n)
            return nb_fail(ctx, NB_ERR_INVALID, "rows_in needs a previous nb_bh_step_groups call with the same n on this handle");
        prep.mode = 2;
        prep.rows = reinterpret_cast<const float4*>(step->rows_in);
        prep.perm = st->idx32[st->sfc_sel];
        prep.post = SubstepArgs{NB_KICK_NONE, step->prev_drift_kind, step->dt, 0.0, step->prev_drift_coef};
        prep.planar = (step->flags & NB_BH_PLANAR) ? 1 : 0;
    } else {
        prep.mode = step->pre_drift_kind != NB_DRIFT_NONE ? 1 : 0;
    }
    if (step->n_rows_peers < 0 || step->n_rows_peers > NB_MAX_PEERS)
        return nb_fail(ctx, NB_ERR_INVALID, "n_rows_peers outside [0, NB_MAX_PEERS]");
    WalkEpilogue epi{};
    epi.mode = (step->rows_out || step->n_rows_peers > 0) ? 2 : 1;
    epi.n_peers = step->n_rows_peers;
    for (int r = 0; r < step->n_rows_peers; r++) {
        if (!step->rows_peers[r]) return nb_fail(ctx, NB_ERR_INVALID, "rows_peers holds a NULL pointer");
        epi.rows_peer[r] = reinterpret_cast<float4*>(step->rows_peers[r]);
    }
    epi.pos_rw = reinterpret_cast<float4*>(pos);
    epi.vel_rw = reinterpret_cast<float4*>(vel);
    epi.rows = reinterpret_cast<float4*>(step->rows_out);
    epi.row_base = step->group_begin * 32;
    epi.s = SubstepArgs{step->kick_kind, step->drift_kind, step->dt, step->kick_coef, step->drift_coef};
    epi.stats = step->stats_out;
    // cross-rank ordering by arrival words instead of a barrier collective between two calls
    SyncPeers sync{};
    if (step->sync_local) {
        if (step->n_rows_peers <= 0) return nb_fail(ctx, NB_ERR_INVALID, "sync_local needs rows_peers");
        if (step->sync_rank < 0 || step->sync_rank >= step->n_rows_peers)
            return nb_fail(ctx, NB_ERR_INVALID, "sync_rank outside [0, n_rows_peers)");
        sync.local = step->sync_local;
        sync.n = step->n_rows_peers;
        sync.rank = step->sync_rank;
        for (int r = 0; r < step->n_rows_peers; r++) {
            if (!step->sync_peers[r]) return nb_fail(ctx, NB_ERR_INVALID, "sync_peers holds a NULL pointer");
            sync.peer[r] = step->sync_peers[r];
        }
        if (prep.mode == 2) {
            prep.sync_wait = step->sync_local;
            prep.sync_n = step->n_rows_peers;
        }
    }
    pick_sort_mode(st);
    st->perm_valid = false;
    // replay pays up to a few million bodies (N = 1M: 1.126 -> 1.104 ms per step; the ~25 launches' gaps)
    static const int graph_max_n = [] { const char* v = getenv("NB_BH_GRAPH_MAX_N"); return v ? atoi(v) : 2097152; }();
    constexpr int STEP_GRAPH_EAGER_N = 262144;
    const bool nothing_to_do = step->group_count == 0 && step->kick_kind == NB_KICK_NONE && step->drift_kind == NB_DRIFT_NONE;
    // the sharded form (rows in or out) has a collective or a barrier between two calls: above the launch-bound sizes the
    // graph -> NCCL kernel -> graph transitions cost more than the gaps a graph removes (N = 1M, 2 GPUs: 0.802 -> 0.838 ms)
    // (the same holds with the arrival words in place of the barrier: 0.797 direct against 0.824 ms replayed)
    const bool sharded_form = step->rows_in || step->rows_out || step->n_rows_peers > 0;
    if (n <= graph_max_n && n >= 2 && !nothing_to_do && !(sharded_form && n > STEP_GRAPH_EAGER_N)) {
        // replay a captured graph of the same call
        struct Key {
            const float* pos; const float* vel; int n, accumulate, sort_full, qcap, split; const void* perm; nb_bh_step st;
        } key;
        static_assert(sizeof(Key) <= sizeof(st->step_graphs[0].key), "step graph key storage");
        memset(&key, 0, sizeof(key));
        key.pos = pos; key.vel = vel; key.n = n;
        key.accumulate = ((step->flags & NB_BH_ACCUMULATE_BOUNDS) && st->bounds_valid) ? 1 : 0;
        key.sort_full = st->sort_full ? 1 : 0; key.qcap = walk_qcap(); key.split = walk_forced_split();
        key.perm = prep.perm; key.st = *step;
        nb_bh_state::StepGraph* hit = nullptr;
        for (auto& g : st->step_graphs)
            if (g.seen && memcmp(g.key, &key, sizeof(key)) == 0) hit = &g;
        bool capture = hit && !hit->exec;
        if (!hit) {
            hit = &st->step_graphs[st->step_graph_next];
            st->step_graph_next = (st->step_graph_next + 1) % 8;
            if (hit->exec) { cudaGraphExecDestroy(hit->exec); hit->exec = nullptr; }
            memcpy(hit->key, &key, sizeof(key));
            hit->seen = true;
            capture = n <= STEP_GRAPH_EAGER_N;
        }
        if (capture) {
            if (!st->capture_stream) NB_CUDA(ctx, cudaStreamCreateWithFlags(&st->capture_stream, cudaStreamNonBlocking));
            NB_CUDA(ctx, cudaStreamBeginCapture(st->capture_stream, cudaStreamCaptureModeThreadLocal));
            rc = run_build(ctx, pos, n, step->flags, 4, step->theta, st->capture_stream, prep);
            if (!rc && step->group_count > 0)
                rc = run_walk(ctx, pos, step->acc_out, n, step->theta, eps, step->group_begin, step->group_count, false,
                              st->capture_stream, epi);
            if (!rc && sync.local) peer_signal_kernel<<<1, 32, 0, st->capture_stream>>>(sync);
            cudaGraph_t graph = nullptr;
            cudaError_t e = cudaStreamEndCapture(st->capture_stream, &graph);
            if (rc) { if (graph) cudaGraphDestroy(graph); return rc; }
            if (e != cudaSuccess) return nb_fail(ctx, NB_ERR_CUDA, "cudaStreamEndCapture", e);
            e = cudaGraphInstantiate(&hit->exec, graph, 0);
            cudaGraphDestroy(graph);
            if (e != cudaSuccess) { hit->exec = nullptr; return nb_fail(ctx, NB_ERR_CUDA, "cudaGraphInstantiate", e); }
        }
        if (hit->exec) {
            st->last_n = n;
            st->bounds_valid = true;
            NB_CUDA(ctx, cudaGraphLaunch(hit->exec, stream));
            st->perm_valid = true;
            return NB_OK;
        }
    }
    // nothing to walk, kick or drift: only the first kernel runs (it brings pos / vel up to date from rows_in).  A rank
    // that merely owns no groups still sorts: the next call needs this evaluation's permutation.
    const bool only_prep = step->group_count == 0 && step->kick_kind == NB_KICK_NONE && step->drift_kind == NB_DRIFT_NONE;
    rc = run_build(ctx, pos, n, step->flags, only_prep ? 1 : 4, step->theta, stream, prep);
    if (rc) return rc;
    if (only_prep) return NB_OK;
    rc = run_walk(ctx, pos, step->acc_out, n, step->theta, eps, step->group_begin, step->group_count, false, stream, epi);
    if (rc) return rc;
    if (sync.local) peer_signal_kernel<<<1, 32, 0, stream>>>(sync);   // also for a rank that owns no groups
    st->perm_valid = true;
    return NB_OK;
}

int nb_bh_accel(nb_context* ctx, const float* pos, float* acc, int n, float theta, float softening, int flags,
                void* stream) {
    return nb_bh_accel_groups(ctx, pos, acc, n, theta, softening, flags, 0, (n + 31) / 32, stream);
}

int nb_bh_accel_host(nb_context* ctx, const float* host_pos, float* host_acc, int n, float theta, float softening,
                     int flags) {
    if (!ctx) return NB_ERR_INVALID;
    if (n < 0 || (n > 0 && (!host_pos || !host_acc))) return nb_fail(ctx, NB_ERR_INVALID, "bad host buffers");
    if (n == 0) return NB_OK;
    NB_DEVICE(ctx);
    const size_t bytes = (size_t)n * 16;
    int rc = nb_reserve(ctx, &ctx->dev_pos, &ctx->dev_pos_bytes, bytes);
    if (rc) return rc;
    rc = nb_reserve(ctx, &ctx->dev_acc, &ctx->dev_acc_bytes, bytes);
    if (rc) return rc;
    cudaStream_t s = ctx->host_stream;
    NB_CUDA(ctx, cudaMemcpyAsync(ctx->dev_pos, host_pos, bytes, cudaMemcpyHostToDevice, s));
    rc = nb_bh_accel(ctx, (const float*)ctx->dev_pos, (float*)ctx->dev_acc, n, theta, softening, flags, s);
    if (rc) return rc;
    NB_CUDA(ctx, cudaMemcpyAsync(host_acc, ctx->dev_acc, bytes, cudaMemcpyDeviceToHost, s));
    NB_CUDA(ctx, cudaStreamSynchronize(s));
    return NB_OK;
}

int nb_bh_stages_host(nb_context* ctx, const float* pos, int n, int flags, float* bounds, uint32_t* keys,
                      int32_t* sorted_idx, uint64_t* tree_keys) {
    if (!ctx) return NB_ERR_INVALID;
    if (n <= 0 || !pos) return nb_fail(ctx, NB_ERR_INVALID, "need n > 0 and a device pointer");
    NB_DEVICE(ctx);
    int rc = run_build(ctx, pos, n, flags, 3, 1.0f, 0);
    if (rc) return rc;
    nb_bh_state* st = ctx->bh;
    NB_CUDA(ctx, cudaDeviceSynchronize());
    if (bounds) NB_CUDA(ctx, cudaMemcpy(bounds, st->bounds, 16, cudaMemcpyDeviceToHost));
    if (sorted_idx) NB_CUDA(ctx, cudaMemcpy(sorted_idx, st->idx32[st->sfc_sel], (size_t)n * 4, cudaMemcpyDeviceToHost));
    if (keys || tree_keys) {
        // keys in body order: rerun the key kernel into the scratch halves the sorts no longer need
        keys_kernel<<<(n + KEYS_PER_BLOCK - 1) / KEYS_PER_BLOCK, KEYS_THREADS>>>(reinterpret_cast<const float4*>(pos), n, st->bounds,
                                                                               st->key32[st->sfc_sel ^ 1],
                                                                               st->key64[st->tree_sel ^ 1], nullptr);
        NB_CUDA(ctx, cudaDeviceSynchronize());
        if (keys) NB_CUDA(ctx, cudaMemcpy(keys, st->key32[st->sfc_sel ^ 1], (size_t)n * 4, cudaMemcpyDeviceToHost));
        if (tree_keys) NB_CUDA(ctx, cudaMemcpy(tree_keys, st->key64[st->tree_sel ^ 1], (size_t)n * 8, cudaMemcpyDeviceToHost));
    }
    return NB_OK;
}

int nb_bh_export_tree(nb_context* ctx, int* n_cells, uint64_t* prefix, int32_t* level, int32_t* count, double* mass,
                      double* cx, double* cy) {
    if (!ctx || !n_cells) return NB_ERR_INVALID;
    nb_bh_state* st = ctx->bh;
    if (!st || st->last_n < 1) return nb_fail(ctx, NB_ERR_INVALID, "no tree has been built on this handle");
    NB_DEVICE(ctx);
    NB_CUDA(ctx, cudaDeviceSynchronize());
    const int n = st->last_n;
    if (n < 2) { *n_cells = 0; return NB_OK; }
    const int chunks = (n - 1 + SCAN_CHUNK - 1) / SCAN_CHUNK;
    // node count = total number of leader flags
    std::vector<int32_t> fblk(chunks), fs(n - 1);
    NB_CUDA(ctx, cudaMemcpy(fblk.data(), st->flag_blk, chunks * 4, cudaMemcpyDeviceToHost));
    NB_CUDA(ctx, cudaMemcpy(fs.data(), st->flag_scan, (size_t)(n - 1) * 4, cudaMemcpyDeviceToHost));
    const int K = fs[n - 2] + fblk[(n - 2) >> SCAN_SHIFT];
    std::vector<int32_t> first(K), last(K);
    std::vector<int8_t> lev(K), top(K);
    std::vector<uint64_t> keys(n);
    std::vector<D3> ps(n), pb((n + SCAN_CHUNK - 1) / SCAN_CHUNK);
    NB_CUDA(ctx, cudaMemcpy(first.data(), st->node_first, (size_t)K * 4, cudaMemcpyDeviceToHost));
    NB_CUDA(ctx, cudaMemcpy(last.data(), st->node_last, (size_t)K * 4, cudaMemcpyDeviceToHost));
    NB_CUDA(ctx, cudaMemcpy(lev.data(), st->node_level, (size_t)K, cudaMemcpyDeviceToHost));
    NB_CUDA(ctx, cudaMemcpy(top.data(), st->node_top, (size_t)K, cudaMemcpyDeviceToHost));
    NB_CUDA(ctx, cudaMemcpy(keys.data(), st->key64[st->tree_sel], (size_t)n * 8, cudaMemcpyDeviceToHost));
    NB_CUDA(ctx, cudaMemcpy(ps.data(), st->psum, (size_t)n * sizeof(D3), cudaMemcpyDeviceToHost));
    NB_CUDA(ctx, cudaMemcpy(pb.data(), st->psum_blk, pb.size() * sizeof(D3), cudaMemcpyDeviceToHost));
    auto pre = [&](int i) { return ps[i] + pb[i >> SCAN_SHIFT]; };
    struct Cell { uint64_t prefix; int level, count; double m, cx, cy; };
    std::vector<Cell> cells;
    for (int id = 0; id < K; id++) {
        const int a = first[id], b = last[id];
        D3 hi = pre(b), lo = a > 0 ? pre(a - 1) : D3{0, 0, 0};
        const double m = hi.m - lo.m;
        const double ccx = m > 0 ? (hi.mx - lo.mx) / m : 0.0, ccy = m > 0 ? (hi.my - lo.my) / m : 0.0;
        const uint64_t path = keys[a] ^ 0xAAAAAAAAAAAAAAAAull;   // digit ^ 2 = path index
        for (int l = top[id]; l <= lev[id]; l++) {               // expand the compressed chain
            const uint64_t mask = l == 0 ? 0ull : (~0ull << (64 - 2 * l));
            cells.push_back(Cell{path & mask, l, b - a + 1, m, ccx, ccy});
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
            mass[k] = cells[k].m; cx[k] = cells[k].cx; cy[k] = cells[k].cy;
        }
    }
    return NB_OK;
}

int nb_bh_build_info(nb_context* ctx, int64_t info[4]) {
    if (!ctx || !info) return NB_ERR_INVALID;
    nb_bh_state* st = ctx->bh;
    if (!st || st->last_n < 1) return nb_fail(ctx, NB_ERR_INVALID, "no tree has been built on this handle");
    NB_DEVICE(ctx);
    NB_CUDA(ctx, cudaDeviceSynchronize());
    const int n = st->last_n;
    int nodes = 0;
    if (n >= 2) {
        int32_t a = 0, b = 0;
        NB_CUDA(ctx, cudaMemcpy(&a, st->flag_scan + (n - 2), 4, cudaMemcpyDeviceToHost));
        NB_CUDA(ctx, cudaMemcpy(&b, st->flag_blk + ((n - 2) >> SCAN_SHIFT), 4, cudaMemcpyDeviceToHost));
        nodes = a + b;
    }
    uint32_t words[2] = {0, 0};
    NB_CUDA(ctx, cudaMemcpy(words, st->os_ticket + FIX_COUNT_WORD, sizeof(words), cudaMemcpyDeviceToHost));
    std::vector<int8_t> lev(nodes > 0 ? nodes : 1, 0);
    if (nodes > 0) NB_CUDA(ctx, cudaMemcpy(lev.data(), st->node_level, (size_t)nodes, cudaMemcpyDeviceToHost));
    int deepest = 0;
    for (int k = 0; k < nodes; k++) deepest = std::max(deepest, (int)lev[k]);
    info[0] = nodes;
    info[1] = words[0];
    info[2] = (words[1] ? 1 : 0) | (st->sort_full ? 2 : 0);
    info[3] = deepest;
    return NB_OK;
}

int nb_bh_group_work(nb_context* ctx, int32_t* work, int groups) {
    if (!ctx || !work) return NB_ERR_INVALID;
    nb_bh_state* st = ctx->bh;
    if (!st || st->group_work_n <= 0 || st->group_work_n != st->last_n)
        return nb_fail(ctx, NB_ERR_INVALID, "the last evaluation on this handle was not an NB_BH_COUNT_WALK one");
    if (groups != (st->group_work_n + 31) / 32) return nb_fail(ctx, NB_ERR_INVALID, "groups must be ceil(n / 32)");
    NB_DEVICE(ctx);
    NB_CUDA(ctx, cudaDeviceSynchronize());
    NB_CUDA(ctx, cudaMemcpy(work, st->ext_a, (size_t)groups * sizeof(int32_t), cudaMemcpyDeviceToHost));
    return NB_OK;
}

int nb_bh_last_counters(nb_context* ctx, int64_t counters[6]) {
    if (!ctx || !counters) return NB_ERR_INVALID;
    nb_bh_state* st = ctx->bh;
    if (!st) return nb_fail(ctx, NB_ERR_INVALID, "no Barnes-Hut evaluation has run on this handle");
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
    counters[4] = nodes; counters[5] = (int64_t)(c[4] | (c[6] ? 2ull : 0ull));
    return NB_OK;
}

}  // extern "C"
