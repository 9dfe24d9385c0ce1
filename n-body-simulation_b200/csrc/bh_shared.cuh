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

