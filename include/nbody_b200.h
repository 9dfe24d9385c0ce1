/*
 * nbody_b200.h -- C ABI of the B200-native gravitational force engine.
 *
 * Drop-in boundary for the force path of MoonArchitect/N-body-simulation (reference paths below are
 * relative to that repository).  The reference reaches its force kernels through three free functions
 * taking raw device pointers:
 *
 *     void direct2DCompute (float4* pos, float4* acc, int n, const float SOFTENING);   src/DirectComputeKernels.cuh:7
 *     void direct3DCompute (float4* pos, float4* acc, int n, const float SOFTENING);   src/DirectComputeKernels.cuh:12
 *     void barnesHutCompute(float4* pos, float4* acc, float4* d_bounds, int* index, int* nodes,
 *                           int* sortedIdx, int* SFCkeys, int n, int m, float theta,
 *                           const float SOFTENING);                                     src/BarnesHutKernels.cuh:5
 *
 * called only from ComputeMethods::Direct::computeAcc / BarnesHut::computeAcc (src/ComputeMethods.cpp:21,23,48).
 * The entry points here replace exactly those calls.  Conventions:
 *   - plain C: opaque handle, int status (0 = NB_OK), pointers and sizes only, no C++ or torch types;
 *   - body layout = the reference's float4 (x, y, z, mass), G = 1, softening added un-squared to r^2
 *     (softening must be >= 0; 0 is raised to 1e-20 so that the self term, which the reference includes, stays
 *     0 instead of the reference's 0 * inf = NaN -- no other term changes at FP32 resolution)
 *     (src/DirectComputeKernels.cu:26,64); `dim` 2 selects the reference's 2D law m*d/(eps+r^2), 3 the
 *     Newtonian m*d/(eps+r^2)^(3/2);
 *   - `pos`/`acc` device pointers are owned by the caller (NbodySystem::device, src/NbodySystem.h:17-26);
 *     scratch memory is owned by the handle (the reference allocates its scratch in
 *     BarnesHut::setSystem, src/ComputeMethods.cpp:40-44);
 *   - `stream` is a cudaStream_t passed as void* (NULL = the legacy default stream the reference
 *     uses); calls are asynchronous on that stream unless stated otherwise;
 *   - acc[i] is written as (ax, ay, az, 0); in 2D az = 0 as well (the reference leaves .z/.w
 *     untouched, src/DirectComputeKernels.cu:34-35 -- SURVEY.md section 9);
 *   - no CPU fallback exists: without a CUDA device every compute entry point returns an error.
 */
#ifndef NBODY_B200_H
#define NBODY_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NB_OK 0
#define NB_ERR_INVALID 1   /* bad argument */
#define NB_ERR_CUDA 2      /* a CUDA runtime call failed; see nb_last_error */
#define NB_ERR_OVERFLOW 3  /* reserved.  No entry point returns it today: the capacities the reference can exceed silently
                              (node pool src/BarnesHutKernels.cu:308-311, 64-entry stack :388) cannot be exceeded here --
                              the tree has at most n-1 nodes by construction, cells at the 32-level depth cap become
                              buckets, and a walk whose node queue fills up finishes its group depth-first in bounded
                              space (status bit 0 of nb_bh_last_counters reports that this happened; the result is
                              the same accepted set) */
#define NB_ERR_NO_DEVICE 4
#define NB_ERR_TOO_LARGE 5  /* problem size beyond a fixed capacity (nb_direct_accel_noalloc) */

/* flags for nb_direct_accel* */
#define NB_DIRECT_AUTO 0      /* split the source range across CTAs when that fills the GPU better */
#define NB_DIRECT_ORDERED 1   /* one accumulator per target, sources summed in index order 0..n-1: the
                                 summation order of src/DirectComputeKernels.cu:16-30,52-68, bit-identical
                                 to the reference kernel's output */
#define NB_DIRECT_NO_HOIST 2  /* keep the 12-operation arithmetic for every source tile: never take a common mass out
                                 of the sum (the default does so for tiles of 1024 equal masses -- same summation order,
                                 one extra rounding, 11 operations) */

/* flags for nb_bh_accel* */
#define NB_BH_FRESH_BOUNDS 0       /* bounding box recomputed from scratch every call (fix) */
#define NB_BH_ACCUMULATE_BOUNDS 1  /* reference behaviour: d_bounds is the running union over all calls,
                                      src/BarnesHutKernels.cu:190-193 + src/ComputeMethods.cpp:40 */

#define NB_BH_PLANAR 0x200          /* nb_bh_step_groups, sharded form: the caller vouches that every body has v_z = 0 (the 2D path
                                      never changes v_z), so z never moves and rows_in can be scattered without reading the
                                      old positions and velocities back */

#define NB_BH_COUNT_WALK 0x100      /* instrumented walk: also fills the counters of nb_bh_last_counters
                                      (nodes opened, children examined, ...); slower, for roofline accounting */

typedef struct nb_context nb_context;

/* ---- lifetime ------------------------------------------------------------------------------------ */

/* Creates an engine bound to CUDA device `device`.  Replaces the cudaMalloc block of BarnesHut::setSystem
 * (src/ComputeMethods.cpp:40-44).  Scratch memory is allocated by the FIRST call of each path (direct: tile flags and,
 * when the source range is split, partial sums; Barnes-Hut: keys, tree, prefix sums, sized for max(max_bodies, n)) and
 * grows -- behind a device synchronisation -- only when a later call passes more bodies than any before it; steady-state
 * calls allocate nothing.  Make the first call of a path outside any stream capture. */
int nb_create(nb_context** out, int device, int max_bodies);
int nb_destroy(nb_context* ctx);
/* Human-readable description of the last failure on this handle (or of the last failed nb_create when
 * ctx is NULL).  Replaces the printf(cudaGetErrorString()) pattern of src/BarnesHutKernels.cu:15-17. */
const char* nb_last_error(const nb_context* ctx);
/* Device the handle is bound to, number of SMs, and bytes of device scratch currently held (all paths). */
int nb_device_info(const nb_context* ctx, int* device, int* sm_count, uint64_t* scratch_bytes);

/* ---- direct all-pairs ---------------------------------------------------------------------------- */

/* Replaces direct2DCompute / direct3DCompute (src/DirectComputeKernels.cu:5-8,41-44).
 * pos: device float4[n]; acc: device float4[n]; any n >= 0 (the reference needs n % 128 == 0). */
int nb_direct_accel(nb_context* ctx, const float* pos, float* acc, int n, int dim, float softening,
                    int flags, void* stream);

/* Target-slice form used for multi-GPU sharding: accelerations of bodies [target_begin,
 * target_begin + target_count) due to ALL n sources; acc is device float4[target_count] (slice-local).
 * No counterpart in the single-GPU reference; with target_begin = 0, target_count = n it is
 * nb_direct_accel. */
int nb_direct_accel_slice(nb_context* ctx, const float* pos, int n, int target_begin, int target_count,
                          float* acc, int dim, float softening, int flags, void* stream);

/* Host-buffer form: copies host float4[n] positions to the device, evaluates, copies float4[n]
 * accelerations back and synchronises.  What NbodySystem::updateDeviceData + computeAcceleration(true)
 * + updateHostData do around the reference kernels (src/NbodySystem.cpp:70-93). */
int nb_direct_accel_host(nb_context* ctx, const float* host_pos, float* host_acc, int n, int dim,
                         float softening, int flags);

/* No counterpart in the reference.  The same direct sum for callers that must not allocate device memory --
 * a host whose time loop has already claimed the GPU's memory, or whose step may never run into the
 * device-wide synchronisation of cudaMalloc/cudaFree.  Takes no handle and calls no allocator: the partial
 * sums of the split-source kernel live in a fixed 16 MiB array inside the library (every GPU that loads it
 * holds that array until exit).  n <= NB_NOALLOC_MAX_BODIES, else NB_ERR_TOO_LARGE (text via
 * nb_last_error(NULL)).  Runs on the CURRENT device, on `stream`.  All calls share the array and it keeps the
 * last call's partials (nothing in it needs clearing: each call overwrites what it reads), so calls that
 * can overlap on the GPU race -- serialise them.  The first call on a device looks up the array's device
 * address; make that call before any stream capture. */
#define NB_NOALLOC_MAX_BODIES 262144
int nb_direct_accel_noalloc(const float* pos, float* acc, int n, int dim, float softening, void* stream);

/* Number of source-range splits (1 = ordered) and thread blocks the last direct call used. */
int nb_direct_last_launch(const nb_context* ctx, int* jsplit, int* blocks, int* launches);

/* ---- Barnes-Hut (2D quadtree, as the reference: src/ComputeMethods.cpp:35-38 forces R2) ------------ */

/* Replaces barnesHutCompute (src/BarnesHutKernels.cu:11-70): bounds -> keys -> sort -> tree -> centre of
 * mass -> walk, one stream-ordered pipeline without host synchronisation.  pos: device float4[n],
 * acc: device float4[n].  The tree lives in the handle's scratch; pos[n..m) is never touched. */
int nb_bh_accel(nb_context* ctx, const float* pos, float* acc, int n, float theta, float softening,
                int flags, void* stream);

/* Sharded form: builds the tree over all n bodies, walks only the 32-target groups
 * [group_begin, group_begin + group_count) of the sorted order (group g = sorted positions
 * [32 g, 32 g + 32)); acc is still indexed by body id (device float4[n], only the walked bodies are
 * written).  No counterpart in the single-GPU reference. */
int nb_bh_accel_groups(nb_context* ctx, const float* pos, float* acc, int n, float theta, float softening,
                       int flags, int group_begin, int group_count, void* stream);

int nb_bh_accel_host(nb_context* ctx, const float* host_pos, float* host_acc, int n, float theta,
                     float softening, int flags);

/* Stage entry points (tests and profiling).  Each runs the pipeline up to and including the named
 * stage on `pos` and copies the result to HOST memory; they synchronise.
 *   bounds     : float[4] (minX, minY, maxX, maxY)      src/BarnesHutKernels.cu:152-195
 *   keys       : uint32[n] SFC keys in body order       src/BarnesHutKernels.cu:197-240
 *   sorted_idx : int32[n] permutation after the sort    src/BarnesHutKernels.cu:45
 *   tree_keys  : uint64[n] tree-path keys in body order (the descent of src/BarnesHutKernels.cu:266-293)
 * Any output pointer may be NULL. */
int nb_bh_stages_host(nb_context* ctx, const float* pos, int n, int flags, float* bounds, uint32_t* keys,
                      int32_t* sorted_idx, uint64_t* tree_keys);

/* Canonical topology of the tree built by the last nb_bh_* call: one record per cell holding >= 2
 * bodies, (prefix, level, count, mass, cx, cy) sorted by (prefix, level); prefix packs the path digits
 * (x>=midX) + 2*(y>=midY) of src/BarnesHutKernels.cu:267-285 from bit 63 down.  Returns the number of
 * records through *n_cells; call with NULL arrays to size the buffers (host memory). */
int nb_bh_export_tree(nb_context* ctx, int* n_cells, uint64_t* prefix, int32_t* level, int32_t* count,
                      double* mass, double* cx, double* cy);

/* Counters of the last walk: nodes opened, children examined, lane interactions, max queue depth, nodes
 * in the tree, status flags (bit 0 = at least one group's node queue filled up and that group finished its walk in
 * depth-first order -- slower, same result; bit 1 = a wait for the peers' rows (nb_bh_step::sync_local) timed out; the
 * counters [0..3] need NB_BH_COUNT_WALK, the status bits do not). */
int nb_bh_last_counters(nb_context* ctx, int64_t counters[6]);

/* Per-group work of the last NB_BH_COUNT_WALK evaluation: work[g] = records summed for each target of group g (bodies and
 * accepted cells; host int32[ceil(n/32)]).  The spread max / mean over groups -- and over blocks of four consecutive
 * groups, the unit the kernel schedules -- is the walk's load imbalance (profiles/bh_counters.py).  Synchronises. */
int nb_bh_group_work(nb_context* ctx, int32_t* work, int groups);

/* Facts about the tree built by the last nb_bh_* call (synchronises; tests and profiling): info[0] = nodes,
 * info[1] = bodies the five-pass tree-key sort left to its repair block (runs sharing the upper 40 key bits),
 * info[2] = bit 0: the repair block overflowed and the conditional full sort ran on the device, bit 1: the host had
 * already asked for eight plain passes, info[3] = deepest level at which a node splits (32 = depth-cap buckets exist).
 * The reference prints nothing comparable; its node pool overflow message is src/BarnesHutKernels.cu:308-311. */
int nb_bh_build_info(nb_context* ctx, int64_t info[4]);

/* ---- Barnes-Hut in 3D (octree) -------------------------------------------------------------------- */

/* NO COUNTERPART IN THE REFERENCE, which runs Barnes-Hut in R2 only (src/ComputeMethods.cpp:35-38; the adapter keeps
 * that downgrade).  The 2D pipeline one dimension up, same rules: fmaf-padded box, 21 levels of '>=' bisection packed
 * in a 63-bit key, one stable radix sort, lock-free octree, unanimous-vote walk over 32 key-consecutive targets with
 * the parent-width criterion, Newtonian force law of direct3DCompute.  pos/acc: device float4[n]; acc = (ax, ay, az, 0).
 * Only NB_BH_COUNT_WALK is honoured in `flags`.  Checked against the CPU octree of oracle/ and the FP64 direct sum. */
int nb_bh3d_accel(nb_context* ctx, const float* pos, float* acc, int n, float theta, float softening, int flags,
                  void* stream);
int nb_bh3d_accel_groups(nb_context* ctx, const float* pos, float* acc, int n, float theta, float softening, int flags,
                         int group_begin, int group_count, void* stream);
/* bounds6 = padded (minX, minY, minZ, maxX, maxY, maxZ); tree_keys in body order; sorted_idx = the sort permutation. */
int nb_bh3d_stages_host(nb_context* ctx, const float* pos, int n, float* bounds6, uint64_t* tree_keys,
                        int32_t* sorted_idx);
/* As nb_bh_export_tree, with a third centre-of-mass coordinate; prefix digits are (x>=midX)+2(y>=midY)+4(z>=midZ),
 * three bits per level from bit 62 down. */
int nb_bh3d_export_tree(nb_context* ctx, int* n_cells, uint64_t* prefix, int32_t* level, int32_t* count, double* mass,
                        double* cx, double* cy, double* cz);
int nb_bh3d_last_counters(nb_context* ctx, int64_t counters[6]);

/* ---- Leapfrog kick/drift on a target slice (callers of the path) ------------ */

/* The two kernels Verlet_DKD::integrate launches around computeAcceleration (src/IntegrationMethods.cu:
 * 72-79 DKD_Verlet_1 and :81-93 DKD_Verlet_2), restricted to bodies [begin, begin+count).  `pos` is the
 * full device float4[n] array (the slice is updated in place); `vel` and `acc` are slice-local
 * float4[count].  Same mixed FP32/FP64 arithmetic as the reference (the 0.5 literal is a double). */
int nb_verlet_dkd_drift(nb_context* ctx, float* pos, const float* vel, int begin, int count, float dt,
                        void* stream);
int nb_verlet_dkd_kick_drift(nb_context* ctx, float* pos, float* vel, const float* acc, int begin, int count,
                             float dt, void* stream);

/* Any kick/drift kernel of the reference's steppers (src/IntegrationMethods.cu:13-349) on a target slice: an
 * optional velocity update followed by an optional position update that uses the new velocity.  The kinds
 * name the reference's expression types (a double literal promotes the product to FP64, rounded once):
 *   NB_KICK_F32       vel += dt*acc                      Euler KD/DK :16-18,52-54, Verlet DKD :84-86
 *   NB_KICK_HALF_F64  vel += dt*0.5*acc                  Verlet KDK :113-115,126-128
 *   NB_KICK_COEF_F64  vel += kick_coef*acc               PEFRL :293-295.. (kick_coef = the double c*dt)
 *   NB_KICK_DTA_F64   vel += dt*acc*kick_coef            Forest-Ruth :214-216..
 *   NB_KICK_MEAN_F64  vel += 0.5*dt*(acc+acc_prev)       Velocity Verlet :162-164
 *   NB_DRIFT_F32      pos += vel*dt                      Euler :20-22,43-45, Verlet KDK :117-119
 *   NB_DRIFT_HALF_F64 pos += vel*(dt*0.5)                Verlet DKD :76-78,88-90
 *   NB_DRIFT_COEF_F64 pos += drift_coef*vel              PEFRL :281-283..
 *   NB_DRIFT_VDT_F64  pos += vel*dt*drift_coef*0.5       Forest-Ruth :203-205..
 *   NB_DRIFT_VV_F64   pos += vel*dt + acc*0.5*dt*dt      Velocity Verlet :152-154
 * pos is the full float4[n] array, vel/acc/acc_prev are slice-local float4[count]. */
#define NB_KICK_NONE 0
#define NB_KICK_F32 1
#define NB_KICK_HALF_F64 2
#define NB_KICK_COEF_F64 3
#define NB_KICK_DTA_F64 4
#define NB_KICK_MEAN_F64 5
#define NB_DRIFT_NONE 0
#define NB_DRIFT_F32 1
#define NB_DRIFT_HALF_F64 2
#define NB_DRIFT_COEF_F64 3
#define NB_DRIFT_VDT_F64 4
#define NB_DRIFT_VV_F64 5
int nb_integrate_substep(nb_context* ctx, float* pos, float* vel, const float* acc, const float* acc_prev, int begin,
                         int count, float dt, int kick_kind, double kick_coef, int drift_kind, double drift_coef,
                         void* stream);

/* ---- one stepper sub-step around the Barnes-Hut force, fused ----------------------
 *
 * Every stepper of the reference except Velocity Verlet is a pre-force drift followed by one or more
 * [force evaluation; kick; drift] groups (src/IntegrationMethods.cu:26-35 KD, :58-67 DK, :96-105 Verlet DKD,
 * :245-270 Forest-Ruth, :343-374 PEFRL).  nb_bh_step_groups runs ONE such group -- optional drift of all bodies, tree
 * build, walk of the groups [group_begin, group_begin + group_count) of the sorted order, and the kick + drift of the
 * walked bodies in the walk's epilogue, where the acceleration is still in registers -- with the reference's mixed
 * FP32/FP64 arithmetic (the kinds of nb_integrate_substep; NB_KICK_MEAN_F64 and NB_DRIFT_VV_F64 are not available
 * here).  No counterpart in the reference, which launches separate kernels on one GPU.
 *
 *   pos, vel : device float4[n] in body order, read AND written.
 *   rows_out == NULL (one GPU): pos[b], vel[b] of every walked body are updated in place; acc_out (device float4[n], may
 *       be NULL) also receives the accelerations.
 *   rows_out != NULL (sharded): pos/vel of the walked bodies are NOT updated; instead row (s - 32 group_begin) of
 *       rows_out (device float4[32 group_count]) receives (x', y', vx', vy') of the body in sorted slot s.  The caller
 *       all-gathers the rows of all ranks into one float4[n] array in sorted order and hands it to the NEXT call as
 *       rows_in (with the drift kind of THIS call as prev_drift_*): that call first scatters the rows back to body order
 *       (it still holds this call's permutation), applies the same drift to z (a_z = 0 in the 2D path, so v_z never
 *       changes), then its own pre-force drift.  Every rank does that for all n bodies, so positions and velocities stay
 *       replicated and bit-identical on all ranks with ONE collective per force evaluation (16 bytes per body) instead
 *       of an all-gather of positions plus an exchange of accelerations.  A last call with group_count = 0 and
 *       kick = drift = none brings pos/vel up to date.
 *   stats_out (device double[2 ceil(n/32)], may be NULL): per walked group g, [2g] = sum m|v| and [2g+1] = sum m v^2/2
 *       of the kicked velocities, lanes added in a fixed order -- the two sums of SimulationStats::computeStats
 *       (src/SimulationStats.cu:29-44) without a second pass over the bodies. */
#define NB_MAX_PEERS 8
typedef struct nb_bh_step {
    float theta, softening, dt;
    int flags;                 /* NB_BH_FRESH_BOUNDS / NB_BH_ACCUMULATE_BOUNDS */
    int group_begin, group_count;
    int pre_drift_kind;        /* NB_DRIFT_*: drift of ALL bodies before the force (NB_DRIFT_NONE: none) */
    double pre_drift_coef;
    int kick_kind;             /* NB_KICK_*: after the force, walked bodies only */
    double kick_coef;
    int drift_kind;            /* NB_DRIFT_*: after the kick, walked bodies only */
    double drift_coef;
    const float* rows_in;      /* sharded: all-gathered rows of the previous call, or NULL */
    int prev_drift_kind;       /* the drift_kind / drift_coef of the call that produced rows_in */
    double prev_drift_coef;
    float* rows_out;           /* sharded: this rank's rows; NULL = update pos/vel in place */
    float* acc_out;            /* in-place mode only, may be NULL */
    double* stats_out;         /* may be NULL */
    /* Sharded without a collective call: with n_rows_peers > 0 the epilogue stores each row into rows_peers[r][s] for
     * every r -- s the GLOBAL sorted slot, each rows_peers[r] a float4[32 ceil(n/32)] array on rank r's GPU mapped into this
     * process (nb_peer_open; this rank's own array is among them) -- so the all-gather happens inside the walk, as coalesced
     * stores over NVLink.  rows_out is then ignored.  The caller double-buffers the arrays and separates consecutive
     * force evaluations by one cross-rank barrier (integrators.FusedBarnesHut(p2p=True)). */
    int n_rows_peers;
    float* rows_peers[NB_MAX_PEERS];
    /* ... and without a barrier either: sync_local != NULL names this rank's sync array (unsigned int[NB_SYNC_WORDS] from
     * nb_peer_alloc, zeroed once by the caller before any rank's first call), sync_peers[r] rank r's array mapped here
     * (this rank's own among them) and sync_rank this rank's index.  After its walk the call counts the evaluation in
     * word 8 of its own array and stores the count into word sync_rank of every rank's array; a call with rows_in waits,
     * on the device and before it reads the rows, until words [0, n_rows_peers) of its own array have reached its own
     * count.  No host synchronisation and no collective between consecutive force evaluations; every rank must make
     * the same sequence of calls.  A wait that lasts 10 s gives up (nb_bh_last_counters status bit 1). */
    int sync_rank;
    unsigned int* sync_local;
    unsigned int* sync_peers[NB_MAX_PEERS];
} nb_bh_step;
#define NB_SYNC_WORDS 16
int nb_bh_step_groups(nb_context* ctx, float* pos, float* vel, int n, const nb_bh_step* step, void* stream);

/* The same for the direct sum, on a target slice: optional drift of the slice's bodies (done by the kernel that classifies
 * the source tiles), the force, and the kick + drift of the slice's bodies (done by the kernel that adds up the source-range
 * partial sums: positions may only move once every block has read them as sources).  3 launches per sub-step instead of 5;
 * for n <= 65,536 (launch-bound) the launches are captured into a CUDA graph once and replayed while the arguments stay
 * the same.  pos: device float4[n], read and written; vel: device float4[target_count], slice-local; acc_out (optional,
 * slice-local) receives the accelerations; stats_out (optional, device double[2 ceil(target_count/256)]): per block of 256
 * targets sum m|v| and sum m v^2/2 of the kicked velocities (src/SimulationStats.cu:29-60).  A sharded run drifts with
 * nb_integrate_substep, all-gathers the positions, and passes pre_drift_kind = NB_DRIFT_NONE here.  NB_KICK_MEAN_F64 is not
 * available (it needs two accelerations).  No counterpart in the reference (separate launches, one GPU). */
typedef struct nb_direct_step {
    float softening, dt;
    int dim, flags;            /* NB_DIRECT_* */
    int target_begin, target_count;
    int pre_drift_kind;
    double pre_drift_coef;
    int kick_kind;
    double kick_coef;
    int drift_kind;
    double drift_coef;
    float* acc_out;
    double* stats_out;
} nb_direct_step;
int nb_direct_step_slice(nb_context* ctx, float* pos, float* vel, int n, const nb_direct_step* step, void* stream);

/* Device memory other processes of the same box can map (one process per GPU): nb_peer_alloc allocates `bytes` on the
 * handle's device and fills `handle` (64 bytes, a CUDA IPC handle: send it to the other ranks by any means); nb_peer_open
 * maps a buffer allocated by ANOTHER process into this one and returns a device pointer this GPU's kernels can store
 * through (NVLink peer access).  nb_peer_close unmaps, nb_peer_free releases.  No counterpart in the single-GPU reference. */
int nb_peer_alloc(nb_context* ctx, uint64_t bytes, void** ptr, unsigned char handle[64]);
int nb_peer_open(nb_context* ctx, const unsigned char handle[64], void** ptr);
int nb_peer_close(nb_context* ctx, void* ptr);
int nb_peer_free(nb_context* ctx, void* ptr);

/* ---- diagnostics (callers of the path) ------------------------------------- */

/* SimulationStats::computeStats (src/SimulationStats.cu:7-27) for the bodies [target_begin, target_begin +
 * target_count) against all n bodies: out[0] = sum m|v| ("moment"), out[1] = kinetic energy, out[2] = FP64
 * potential energy with the reference's formula (every ordered pair including i = j, epsilon 1e-6, 1/r law).
 * pos: device float4[n]; vel: device float4[target_count] (slice-local).  Sums of slices add up to the
 * whole-system values (sharding: all-reduce the three doubles).  Synchronises `stream`; out is host memory. */
int nb_stats(nb_context* ctx, const float* pos, const float* vel, int n, int target_begin, int target_count, int dim,
             double out[3], void* stream);

/* ---- device-side initial conditions (callers of the path) -------------------- */

/* UniformEllipsoid (src/InitialConditions.cu:111-144) with a counter-based generator (splitmix64 of seed, stream,
 * index: the same bodies on any host, any GPU, any shard): bodies [offset, offset+count) of `pos` get
 * centre + radius*lambda*(cos phi, sin phi), phi ~ U[0,2pi), lambda ~ U[0.02,1), z = 0 or U[-h,h], mass `mass`;
 * body `offset` becomes the central mass (DiskModel, :157) unless central_mass < 0; `vel` (slice-local, may be NULL)
 * is set to the bulk velocity. */
int nb_ic_uniform_ellipsoid(nb_context* ctx, float* pos, float* vel, int offset, int count, uint64_t seed, double cx, double cy,
                            double rx, double ry, double z_half_height, float mass, float central_mass, float vx, float vy,
                            void* stream);
/* UniformBox (src/InitialConditions.cu:78-107) with the same counter-based generator: bodies [offset, offset+count)
 * uniform in the box of size (sx, sy) centred on (cx, cy) -- in R3 (dim = 3) z is drawn from the y range, as the
 * reference does (:85) -- mass ~ U[mass_lo, mass_hi); `vel` (slice-local, may be NULL) gets a speed ~ U[speed_lo,
 * speed_hi) in a random in-plane direction plus the bulk velocity (vx, vy). */
int nb_ic_uniform_box(nb_context* ctx, float* pos, float* vel, int offset, int count, uint64_t seed, double cx, double cy, double sx,
                      double sy, int dim, double mass_lo, double mass_hi, double speed_lo, double speed_hi, float vx, float vy,
                      void* stream);
/* The velocity loop of DiskModel::initialize (src/InitialConditions.cu:169-192) after one force evaluation:
 * v = sqrt(|a| r), tangential, in the reference's FP32 arithmetic; acc and vel are slice-local float4[count]. */
int nb_ic_disk_velocities(nb_context* ctx, const float* pos, const float* acc, float* vel, int begin, int count, float cx, float cy,
                          int dim, float bulk_vx, float bulk_vy, void* stream);

/* Standard (src/InitialConditions.cu:13-76, the reference's debug generator) with the counter-based generator: bodies
 * [offset, offset+count) at centre + r (cos theta, sin theta), theta ~ U[0, 2 pi), r = (Exp(12) + 0.02) * 50000,
 * z = sin(r/2000) * 2000, mass 100; body `offset` is the central mass 10^6.  nb_ic_standard_velocities is its second half
 * (:49-73) after one force evaluation: v = sqrt(|F| dist) along rotation * (sin theta, -cos theta), body `offset` at rest;
 * acc and vel are slice-local float4[count]. */
int nb_ic_standard(nb_context* ctx, float* pos, int offset, int count, uint64_t seed, double cx, double cy, void* stream);
int nb_ic_standard_velocities(nb_context* ctx, const float* pos, const float* acc, float* vel, int offset, int count,
                              uint64_t seed, float rotation, void* stream);

/* ---- calibration --------------------------------------------------------------------------------- */

/* Measures this GPU's FP32 roofline denominators with register-resident microbenchmarks:
 * out[0] = scalar FFMA TFLOP/s, out[1] = packed FFMA2 TFLOP/s, out[2] = MUFU.RSQ Gop/s,
 * out[3] = SM clock in MHz derived from clock64 during the FFMA run. */
int nb_calibrate(nb_context* ctx, double out[4]);

#ifdef __cplusplus
}
#endif
#endif /* NBODY_B200_H */
