// bh_shared.cuh -- pieces shared by the 2D (bh.cu) and 3D (bh3d.cu) Barnes-Hut pipelines: the one-kernel-per-digit
// stable radix sort and the chunked scans.  No include guard games: each translation unit includes this file once,
// INSIDE its anonymous namespace, after common.cuh; a unit that scans its own sum type T provides
//   template <> T shfl_up_t<T>(T, int)   and   T zero_of(T)   before the first use.
#pragma once

// ------------------------------------------------------------------------------------------------------
// stable LSD radix sort, 8-bit digits, one kernel per digit (replaces thrust::sort_by_key,
// src/BarnesHutKernels.cu:45).  The global digit histograms come from keys_kernel; each pass ranks its tile
// with warp-level match.any (stable), and tiles chain their per-digit prefixes through a status word per
// (tile, digit) -- count << 2 | {1: tile total, 2: inclusive prefix} -- read backwards until an inclusive
// prefix is met ("decoupled look-back").  Tile numbers are handed out by an atomic ticket, so a tile only
// ever waits on tiles that already started.
// ------------------------------------------------------------------------------------------------------
#ifndef NB_SORT_THREADS
#define NB_SORT_THREADS 256
#endif
#ifndef NB_SORT_ITEMS
#define NB_SORT_ITEMS 8
#endif
constexpr int SORT_THREADS = NB_SORT_THREADS;
constexpr int SORT_ITEMS = NB_SORT_ITEMS;
constexpr int SORT_TILE = SORT_THREADS * SORT_ITEMS;  // keys per block
#ifndef NB_LOOKBACK_WINDOW
#define NB_LOOKBACK_WINDOW 8
#endif
constexpr int LOOKBACK_WINDOW = NB_LOOKBACK_WINDOW;

template <typename K>
__global__ void __launch_bounds__(SORT_THREADS) onesweep_kernel(const K* __restrict__ keys_in,
                                                               const int32_t* __restrict__ vals_in,   // null: identity
                                                               int n, int shift, const uint32_t* __restrict__ ghist,
                                                               uint32_t* status, unsigned int* ticket,
                                                               K* __restrict__ keys_out, int32_t* __restrict__ vals_out,
                                                               const uint32_t* __restrict__ enable /* null: always */) {
    static_assert(SORT_THREADS == 256, "one thread per digit");
    if (enable && *enable == 0u) return;   // a pass of the conditional full re-sort that is not needed
    constexpr int WARPS = SORT_THREADS / 32;
    constexpr unsigned FULLM = 0xffffffffu;
    __shared__ uint32_t cnt[WARPS][256];   // per warp and digit: count, then exclusive offset inside the tile's digit bin
    __shared__ uint32_t dbase[256];        // global position of the tile's digit bin minus its tile-local offset
    __shared__ uint32_t lexcl[256];        // tile-local exclusive offset of the digit bin
    __shared__ uint32_t dw[WARPS], lw[WARPS];
    __shared__ unsigned int tile_s;
    __shared__ K skey[SORT_TILE];
    __shared__ int32_t sval[SORT_TILE];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) tile_s = atomicAdd(ticket, 1u);
    for (int d = lane; d < 256; d += 32) cnt[warp][d] = 0;
    __syncthreads();
    const int tile = (int)tile_s;
    const int tile_count = min(SORT_TILE, n - tile * SORT_TILE);
    // tile order = warp-major, round-major, lane: warp w owns a contiguous 32*ITEMS run of the tile
    const int wbase = tile * SORT_TILE + warp * (32 * SORT_ITEMS);
    K key[SORT_ITEMS];
    int32_t val[SORT_ITEMS];
    uint32_t rank[SORT_ITEMS];
#pragma unroll
    for (int k = 0; k < SORT_ITEMS; k++) {
        const int i = wbase + k * 32 + lane;
        key[k] = i < n ? keys_in[i] : (K)0;
        val[k] = (i < n && vals_in) ? vals_in[i] : i;
    }
    // where each digit's bin starts in the output: exclusive prefix of the 256 global digit counts
    uint32_t gstart;
    {
        const uint32_t v = ghist[threadIdx.x];
        uint32_t inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(FULLM, inc, o);
            if (lane >= o) inc += t;
        }
        if (lane == 31) dw[warp] = inc;
        gstart = inc - v;
    }
    // stable ranks: all warp primitives run converged with the full mask (variable masks cost a WARPSYNC each)
#pragma unroll
    for (int k = 0; k < SORT_ITEMS; k++) {
        const bool valid = wbase + k * 32 + lane < n;
        const uint32_t d = (uint32_t)(key[k] >> shift) & 255u;
        unsigned peers = __ballot_sync(FULLM, valid);
#pragma unroll
        for (int b = 0; b < 8; b++) {
            const unsigned m = __ballot_sync(FULLM, (d >> b) & 1u);
            peers &= ((d >> b) & 1u) ? m : ~m;
        }
        const int leader = __ffs(peers) - 1;           // lowest valid lane holding this digit (-1 only if !valid)
        uint32_t basev = 0;
        if (valid && lane == leader) {
            basev = cnt[warp][d];
            cnt[warp][d] = basev + __popc(peers);
        }
        basev = __shfl_sync(FULLM, basev, leader & 31);
        rank[k] = basev + __popc(peers & ((1u << lane) - 1u));
        __syncwarp();
    }
    __syncthreads();
    {
        const int d = threadIdx.x;
        // per digit: exclusive prefix over the warps of this tile ...
        uint32_t run = 0;
#pragma unroll
        for (int w = 0; w < WARPS; w++) {
            const uint32_t c = cnt[w][d];
            cnt[w][d] = run;
            run += c;
        }
        // ... publish the tile total, then chain the prefix over the tiles before it
        volatile uint32_t* st = status;
        st[(size_t)tile * 256 + d] = (run << 2) | (tile == 0 ? 2u : 1u);
        // tile-local exclusive offsets of the digit bins (where the bin sits in the staged, tile-sorted order)
        uint32_t linc = run;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(FULLM, linc, o);
            if (lane >= o) linc += t;
        }
        if (lane == 31) lw[warp] = linc;
        uint32_t excl = 0;
        int p = tile - 1;
        bool done = false;
        while (p >= 0 && !done) {
            uint32_t v[LOOKBACK_WINDOW];
#pragma unroll
            for (int j = 0; j < LOOKBACK_WINDOW; j++) v[j] = (p - j >= 0) ? st[(size_t)(p - j) * 256 + d] : 2u;
#pragma unroll
            for (int j = 0; j < LOOKBACK_WINDOW; j++) {
                if (done || p - j < 0) continue;
                uint32_t x = v[j];
                while ((x & 3u) == 0u) x = st[(size_t)(p - j) * 256 + d];
                excl += x >> 2;
                if ((x & 3u) == 2u) done = true;
            }
            p -= LOOKBACK_WINDOW;
        }
        if (tile > 0) st[(size_t)tile * 256 + d] = ((excl + run) << 2) | 2u;
        __syncthreads();   // dw, lw complete
        uint32_t g = gstart + excl, l = linc - run;
        for (int w = 0; w < warp; w++) { g += dw[w]; l += lw[w]; }
        lexcl[d] = l;
        dbase[d] = g - l;
    }
    __syncthreads();
    // stage the tile in sorted order in shared memory, then write it out in runs (coalesced per digit bin)
#pragma unroll
    for (int k = 0; k < SORT_ITEMS; k++) {
        if (wbase + k * 32 + lane < n) {
            const uint32_t d = (uint32_t)(key[k] >> shift) & 255u;
            const uint32_t lp = lexcl[d] + cnt[warp][d] + rank[k];
            skey[lp] = key[k];
            sval[lp] = val[k];
        }
    }
    __syncthreads();
    for (int j = threadIdx.x; j < tile_count; j += SORT_THREADS) {
        const K kk = skey[j];
        const uint32_t d = (uint32_t)(kk >> shift) & 255u;
        const uint32_t dst = dbase[d] + j;
        keys_out[dst] = kk;
        vals_out[dst] = sval[j];
    }
}

// Passes over the key bits [shift_begin, shift_end).  `hist0` is the index, in the [..][256] global histogram layout,
// of the histogram of the key's LOWEST digit; `slot0` the index of the first pass's ticket and look-back status
// area (every pass of an evaluation needs its own).  `sel_in` says which buffer holds the input; the first pass
// takes the identity permutation as values when `identity` is set.  `enable` (device flag, may be null) turns
// every pass into a no-op while it reads 0.
template <typename K>
int radix_sort(nb_context* ctx, uint32_t* ghist, uint32_t* os_status, size_t os_stride, unsigned int* os_ticket, K* keys[2],
               int32_t* vals[2], int n, int shift_begin, int shift_end, int hist0, int slot0, bool identity, int sel_in,
               const uint32_t* enable, int* sel_out, cudaStream_t stream) {
    const int nblocks = (n + SORT_TILE - 1) / SORT_TILE;
    int sel = sel_in, slot = slot0;
    for (int shift = shift_begin; shift < shift_end; shift += 8, slot++) {
        const bool first = identity && shift == shift_begin;
        onesweep_kernel<K><<<nblocks, SORT_THREADS, 0, stream>>>(keys[sel], first ? nullptr : vals[sel], n, shift,
                                                                ghist + 256 * (hist0 + shift / 8),
                                                                os_status + (size_t)slot * os_stride, os_ticket + slot,
                                                                keys[sel ^ 1], vals[sel ^ 1], enable);
        sel ^= 1;
    }
    *sel_out = sel;
    NB_CUDA(ctx, cudaGetLastError());
    return NB_OK;
}

// ------------------------------------------------------------------------------------------------------
// chunked scans (chunk = 2048 elements per block): local inclusive prefix + exclusive prefix of chunk totals
// ------------------------------------------------------------------------------------------------------
constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = 8;
constexpr int SCAN_CHUNK = SCAN_THREADS * SCAN_ITEMS;  // 2048
constexpr int SCAN_SHIFT = 11;

template <typename T>
__device__ __forceinline__ T shfl_up_t(T v, int o);
template <>
__device__ __forceinline__ int shfl_up_t<int>(int v, int o) { return __shfl_up_sync(0xffffffffu, v, o); }
__device__ __forceinline__ int zero_of(int) { return 0; }

// v[] holds the thread's SCAN_ITEMS consecutive inputs; on return they hold the chunk-local inclusive scan.
template <typename T>
__device__ __forceinline__ void block_scan_chunk(T (&v)[SCAN_ITEMS], T* total_out) {
    __shared__ T wtot[SCAN_THREADS / 32];
    T s = v[0];
#pragma unroll
    for (int k = 1; k < SCAN_ITEMS; k++) { s = s + v[k]; v[k] = s; }
    T inc = s;
    for (int o = 1; o < 32; o <<= 1) {
        T t = shfl_up_t<T>(inc, o);
        if ((threadIdx.x & 31) >= o) inc = t + inc;
    }
    if ((threadIdx.x & 31) == 31) wtot[threadIdx.x >> 5] = inc;
    __syncthreads();
    T off = zero_of(T());
    const int w = threadIdx.x >> 5;
    for (int k = 0; k < w; k++) off = off + wtot[k];
    T excl = shfl_up_t<T>(inc, 1);
    if ((threadIdx.x & 31) == 0) excl = zero_of(T());
    off = off + excl;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) v[k] = off + v[k];
    if (threadIdx.x == SCAN_THREADS - 1 && total_out) *total_out = v[SCAN_ITEMS - 1];
}

// one block: exclusive scan of the chunk totals in place
template <typename T>
__global__ void __launch_bounds__(1024) scan_totals_kernel(T* __restrict__ tot, int count) {
    __shared__ T wtot[32];
    __shared__ T carry_s;
    if (threadIdx.x == 0) carry_s = zero_of(T());
    __syncthreads();
    for (int base = 0; base < count; base += 1024) {
        const int i = base + threadIdx.x;
        const T v = i < count ? tot[i] : zero_of(T());
        T inc = v;
        for (int o = 1; o < 32; o <<= 1) {
            T t = shfl_up_t<T>(inc, o);
            if ((threadIdx.x & 31) >= o) inc = t + inc;
        }
        if ((threadIdx.x & 31) == 31) wtot[threadIdx.x >> 5] = inc;
        __syncthreads();
        T off = carry_s;
        for (int k = 0; k < (int)(threadIdx.x >> 5); k++) off = off + wtot[k];
        T excl = shfl_up_t<T>(inc, 1);
        if ((threadIdx.x & 31) == 0) excl = zero_of(T());
        const T mine = off + excl;  // exclusive
        if (i < count) tot[i] = mine;
        __syncthreads();
        if (threadIdx.x == 1023) carry_s = mine + v;
        __syncthreads();
    }
}


// ------------------------------------------------------------------------------------------------------
// node boundaries from sorted keys (shared by the quadtree and the octree build)
// ------------------------------------------------------------------------------------------------------
// Every boundary i (between sorted bodies i and i+1) is a separator between two children of exactly one node:
// the node at level L = LV::of(K[i], K[i+1]) (common prefix of two keys, in levels).  Three extents describe its neighbourhood:
//   A = first body of that node,  C = first body of the child that ends at i,  R = last body of the child that
//   starts at i+1.
// i is the node's first separator ("leader", gives the node its id) iff C == A.
// With d(j) = LV::of(K[j], K[j+1]) the three are runs of d: C = start of the run of d >= L+1 that ends at i, A = start of the
// run of d >= L it continues into, R = end of the run of d >= L+1 that starts at i+1.  A block keeps d of its 2048
// separators in shared memory together with the minima over every aligned-by-start power-of-two window (12 x 2 KB), so a
// run is walked in 12 uniform steps -- take the window if its minimum still clears the level -- with no divergence
// between lanes; only a run that reaches the edge of the chunk goes on with the galloping key comparisons of
// left_extent / right_extent (a few separators per chunk: those of nodes that straddle it).
constexpr int BND_LOG = SCAN_SHIFT;   // 2^11 separators per block

// continue a leftward / rightward run in global memory from a body known to be inside it
template <typename LV>
__device__ __forceinline__ int left_extent_from(const uint64_t* __restrict__ K, int good, uint64_t ki, int L) {
    int step = 1;
    while (good - step >= 0 && LV::of(K[good - step], ki) >= L) { good -= step; step <<= 1; }
    int bad = good - step;
    if (bad < -1) bad = -1;
    while (good - bad > 1) {
        const int mid = (good + bad) >> 1;
        if (LV::of(K[mid], ki) >= L) good = mid; else bad = mid;
    }
    return good;
}
template <typename LV>
__device__ __forceinline__ int right_extent_from(const uint64_t* __restrict__ K, int good, int limit, uint64_t ki, int L) {
    int step = 1;
    while (good + step <= limit && LV::of(ki, K[good + step]) >= L) { good += step; step <<= 1; }
    int bad = good + step;
    if (bad > limit + 1) bad = limit + 1;
    while (bad - good > 1) {
        const int mid = (good + bad) >> 1;
        if (LV::of(ki, K[mid]) >= L) good = mid; else bad = mid;
    }
    return good;
}

template <typename LV>
__global__ void __launch_bounds__(SCAN_THREADS) boundary_kernel(const uint64_t* __restrict__ K, int n,
                                                               int32_t* __restrict__ ext_a, int32_t* __restrict__ ext_c,
                                                               int32_t* __restrict__ ext_r, int32_t* __restrict__ flag_scan,
                                                               int32_t* __restrict__ tot) {
    // T[k][j] = 1 + min d over separators [j, j + 2^k) of this chunk (0 where no separator exists: stops every run)
    __shared__ __align__(16) uint8_t T[BND_LOG + 1][SCAN_CHUNK];
    const int base = blockIdx.x * SCAN_CHUNK;
    const int j0 = threadIdx.x * SCAN_ITEMS;
    {
        uint64_t kprev = base + j0 < n ? K[base + j0] : 0ull;
#pragma unroll
        for (int k = 0; k < SCAN_ITEMS; k++) {
            const int i = base + j0 + k;
            uint8_t lev = 0;
            if (i < n - 1) {
                const uint64_t knext = K[i + 1];
                lev = (uint8_t)(LV::of(kprev, knext) + 1);
                kprev = knext;
            }
            T[0][j0 + k] = lev;
        }
    }
    __syncthreads();
#pragma unroll
    for (int k = 1; k <= BND_LOG; k++) {
        const int half = 1 << (k - 1);
        if (half >= 4) {   // eight bytes at a time: two packed unsigned byte minima
            const uint2 lo = *reinterpret_cast<const uint2*>(&T[k - 1][j0]);
            uint2 hi = make_uint2(0u, 0u);
            if (j0 + half < SCAN_CHUNK) hi = *reinterpret_cast<const uint2*>(&T[k - 1][j0 + half]);
            *reinterpret_cast<uint2*>(&T[k][j0]) = make_uint2(__vminu4(lo.x, hi.x), __vminu4(lo.y, hi.y));
        } else {
#pragma unroll
            for (int t = 0; t < SCAN_ITEMS; t++) {
                const int j = j0 + t;
                const uint8_t lo = T[k - 1][j];
                const uint8_t hi = j + half < SCAN_CHUNK ? T[k - 1][j + half] : (uint8_t)0;
                T[k][j] = lo < hi ? lo : hi;
            }
        }
        __syncthreads();
    }
    // the thread's eight separators advance together, one window size at a time: sixteen independent chains of
    // shared-memory loads instead of one long dependent one
    int v[SCAN_ITEMS], pc[SCAN_ITEMS], pa[SCAN_ITEMS], pr[SCAN_ITEMS];
    uint8_t v1[SCAN_ITEMS];   // threshold for d >= L+1 in the +1 coding; d >= L is v1 - 1
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) {
        v1[k] = (uint8_t)(T[0][j0 + k] + 1);   // = L + 2 (1 for a slot without separator: its runs are never used)
        pc[k] = j0 + k;
        pr[k] = j0 + k + 1;
    }
#pragma unroll
    for (int b = BND_LOG; b >= 0; b--) {
        const int w = 1 << b;
#pragma unroll
        for (int k = 0; k < SCAN_ITEMS; k++) {
            if (pc[k] - w >= 0 && T[b][pc[k] - w] >= v1[k]) pc[k] -= w;          // run of d >= L+1 ending at the separator
            if (pr[k] + w <= SCAN_CHUNK && T[b][pr[k]] >= v1[k]) pr[k] += w;     // run of d >= L+1 starting behind it
        }
    }
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) pa[k] = pc[k];
#pragma unroll
    for (int b = BND_LOG; b >= 0; b--) {
        const int w = 1 << b;
#pragma unroll
        for (int k = 0; k < SCAN_ITEMS; k++)
            if (pa[k] - w >= 0 && T[b][pa[k] - w] >= v1[k] - 1) pa[k] -= w;      // ... continued as a run of d >= L
    }
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) {
        const int j = j0 + k, i = base + j;
        int f = 0;
        if (i < n - 1) {
            const int L = (int)v1[k] - 2;
            int c = base + pc[k], a = base + pa[k], r = base + pr[k];
            // a run that reached the edge of the chunk may go on beyond it
            if ((pc[k] == 0 || pa[k] == 0) && base > 0) {
                const uint64_t ki = K[i];
                if (pc[k] == 0) c = left_extent_from<LV>(K, c, ki, L + 1);
                a = left_extent_from<LV>(K, pc[k] == 0 ? c : a, ki, L);
            }
            if (pr[k] == SCAN_CHUNK && r < n - 1) r = right_extent_from<LV>(K, r, n - 1, K[i + 1], L + 1);
            ext_a[i] = a;
            ext_c[i] = c;
            ext_r[i] = r;
            f = (c == a);
        }
        v[k] = f;
    }
    block_scan_chunk<int>(v, &tot[blockIdx.x]);
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++)
        if (base + j0 + k < n - 1) flag_scan[base + j0 + k] = v[k];
}

