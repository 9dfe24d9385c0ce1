/*
 * nbody_oracle.c -- CPU restatement of the reference force path.  TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this library; the product path (n-body-simulation_b200/) never does.
 *
 * Parity status: the reference (MoonArchitect/N-body-simulation) has no tests and no CPU path
 * (SURVEY.md section 4), so this oracle is pinned against outputs of the reference's own CUDA kernels
 * run on a B200 (oracle/ref_shim.cu -> oracle/_ref/libnbody_ref.so, fixtures under tests/golden/
 * written by tests/golden/make_ref_golden.py).
 *
 * Every function cites the reference lines it restates (paths relative to /root/reference).
 * Body layout everywhere: float4 = (x, y, z, mass), G = 1.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <float.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define ORACLE_API __attribute__((visibility("default")))

ORACLE_API int oracle_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* ------------------------------------------------------------------------------------------ */
/* Direct all-pairs                                                                            */
/* ------------------------------------------------------------------------------------------ */

/*
 * FP32 direct sum, arithmetic and j-order of the reference kernels:
 *   3D src/DirectComputeKernels.cu:60-68   distSqr = S + dx*dx + dy*dy + dz*dz (FMA-contracted left to
 *      right, as the sm_100 SASS shows: 3 FFMA), invDist3 = inv*inv*inv*m, F += d*invDist3 (FFMA)
 *   2D src/DirectComputeKernels.cu:23-30   invDist3 = inv*inv*m  (a 1/r force law), x/y only
 * Self term included (contributes 0).  rsqrtf is replaced by 1/sqrtf (MUFU.RSQ is not available
 * on the host; the two agree to ~2 ulp).  Targets [t0, t0+nt), sources [0, n).  acc is float4 per
 * target, .w (and .z in 2D) written as 0 (SURVEY.md section 9 "fix").
 */
ORACLE_API void oracle_direct_f32(const float *pos, int n, int t0, int nt, int dim, float eps,
                                  float *acc, int threads)
{
#ifdef _OPENMP
    if (threads <= 0) threads = omp_get_max_threads();
#endif
#pragma omp parallel for schedule(static) num_threads(threads)
    for (int ii = 0; ii < nt; ii++) {
        const float bx = pos[4 * (size_t)(t0 + ii) + 0];
        const float by = pos[4 * (size_t)(t0 + ii) + 1];
        const float bz = pos[4 * (size_t)(t0 + ii) + 2];
        float fx = 0.f, fy = 0.f, fz = 0.f;
        if (dim == 3) {
            for (int j = 0; j < n; j++) {
                float dx = pos[4 * (size_t)j + 0] - bx;
                float dy = pos[4 * (size_t)j + 1] - by;
                float dz = pos[4 * (size_t)j + 2] - bz;
                float d2 = fmaf(dz, dz, fmaf(dy, dy, fmaf(dx, dx, eps)));
                float inv = 1.0f / sqrtf(d2);
                float s = inv * inv * inv * pos[4 * (size_t)j + 3];
                fx = fmaf(dx, s, fx);
                fy = fmaf(dy, s, fy);
                fz = fmaf(dz, s, fz);
            }
        } else {
            for (int j = 0; j < n; j++) {
                float dx = pos[4 * (size_t)j + 0] - bx;
                float dy = pos[4 * (size_t)j + 1] - by;
                float d2 = fmaf(dy, dy, fmaf(dx, dx, eps));
                float inv = 1.0f / sqrtf(d2);
                float s = inv * inv * pos[4 * (size_t)j + 3];
                fx = fmaf(dx, s, fx);
                fy = fmaf(dy, s, fy);
            }
        }
        acc[4 * (size_t)ii + 0] = fx;
        acc[4 * (size_t)ii + 1] = fy;
        acc[4 * (size_t)ii + 2] = fz;
        acc[4 * (size_t)ii + 3] = 0.f;
    }
}

/*
 * Same loops written for the host vector units (no explicit fmaf ordering, restrict pointers, SoA
 * copies) -- this is the "straightforward OpenMP C++ transcription" that bench.py times as the host
 * baseline.  Results agree with oracle_direct_f32 to rounding.
 */
__attribute__((optimize("fp-contract=fast")))
ORACLE_API void oracle_direct_f32_fast(const float *pos, int n, int t0, int nt, int dim, float eps,
                                       float *acc, int threads)
{
#ifdef _OPENMP
    if (threads <= 0) threads = omp_get_max_threads();
#endif
    float *sx = (float *)aligned_alloc(64, ((size_t)n * 4 + 15) / 16 * 16 * sizeof(float));
    float *sy = sx + n, *sz = sy + n, *sm = sz + n;
    for (int j = 0; j < n; j++) {
        sx[j] = pos[4 * (size_t)j];
        sy[j] = pos[4 * (size_t)j + 1];
        sz[j] = pos[4 * (size_t)j + 2];
        sm[j] = pos[4 * (size_t)j + 3];
    }
#pragma omp parallel for schedule(static) num_threads(threads)
    for (int ii = 0; ii < nt; ii++) {
        const float bx = sx[t0 + ii], by = sy[t0 + ii], bz = sz[t0 + ii];
        float fx = 0.f, fy = 0.f, fz = 0.f;
        if (dim == 3) {
#pragma omp simd reduction(+ : fx, fy, fz)
            for (int j = 0; j < n; j++) {
                float dx = sx[j] - bx, dy = sy[j] - by, dz = sz[j] - bz;
                float d2 = eps + dx * dx + dy * dy + dz * dz;
                float inv = 1.0f / sqrtf(d2);
                float s = inv * inv * inv * sm[j];
                fx += dx * s;
                fy += dy * s;
                fz += dz * s;
            }
        } else {
#pragma omp simd reduction(+ : fx, fy)
            for (int j = 0; j < n; j++) {
                float dx = sx[j] - bx, dy = sy[j] - by;
                float d2 = eps + dx * dx + dy * dy;
                float inv = 1.0f / sqrtf(d2);
                float s = inv * inv * sm[j];
                fx += dx * s;
                fy += dy * s;
            }
        }
        acc[4 * (size_t)ii + 0] = fx;
        acc[4 * (size_t)ii + 1] = fy;
        acc[4 * (size_t)ii + 2] = fz;
        acc[4 * (size_t)ii + 3] = 0.f;
    }
    free(sx);
}

/*
 * FP64 direct sum with the same force law (accuracy reference).  Output is
 * double[3] per target.
 */
ORACLE_API void oracle_direct_f64(const float *pos, int n, int t0, int nt, int dim, double eps,
                                  double *acc, int threads)
{
#ifdef _OPENMP
    if (threads <= 0) threads = omp_get_max_threads();
#endif
#pragma omp parallel for schedule(static) num_threads(threads)
    for (int ii = 0; ii < nt; ii++) {
        const double bx = pos[4 * (size_t)(t0 + ii) + 0];
        const double by = pos[4 * (size_t)(t0 + ii) + 1];
        const double bz = pos[4 * (size_t)(t0 + ii) + 2];
        double fx = 0, fy = 0, fz = 0;
        if (dim == 3) {
#pragma omp simd reduction(+ : fx, fy, fz)
            for (int j = 0; j < n; j++) {
                double dx = pos[4 * (size_t)j + 0] - bx;
                double dy = pos[4 * (size_t)j + 1] - by;
                double dz = pos[4 * (size_t)j + 2] - bz;
                double d2 = eps + dx * dx + dy * dy + dz * dz;
                double inv = 1.0 / sqrt(d2);
                double s = inv * inv * inv * pos[4 * (size_t)j + 3];
                fx += dx * s;
                fy += dy * s;
                fz += dz * s;
            }
        } else {
#pragma omp simd reduction(+ : fx, fy)
            for (int j = 0; j < n; j++) {
                double dx = pos[4 * (size_t)j + 0] - bx;
                double dy = pos[4 * (size_t)j + 1] - by;
                double d2 = eps + dx * dx + dy * dy;
                double s = pos[4 * (size_t)j + 3] / d2;
                fx += dx * s;
                fy += dy * s;
            }
        }
        acc[3 * (size_t)ii + 0] = fx;
        acc[3 * (size_t)ii + 1] = fy;
        acc[3 * (size_t)ii + 2] = fz;
    }
}

/* ------------------------------------------------------------------------------------------ */
/* Barnes-Hut (2D quadtree), stage by stage                                                    */
/* ------------------------------------------------------------------------------------------ */

/*
 * Bounds: src/BarnesHutKernels.cu:152-195.  Global min/max of x and y, padded by 0.1 % of |coord|;
 * the sm_100 build of the reference evaluates `v - fabsf(v)*0.001f` as one FFMA(|v|, -0.001f, v)
 * (SURVEY.md section 3.3 step 2), hence fmaf here.  The pad is monotone, so pad(global min) equals the
 * reference's atomicMin over per-block padded minima.  bounds = (minX, minY, maxX, maxY).
 * `accumulate` != 0 reproduces the reference's never-reset d_bounds (running union with the
 * incoming bounds); 0 is the fixed behaviour (fresh box).
 */
ORACLE_API void oracle_bh_bounds(const float *pos, int n, float *bounds, int accumulate)
{
    float mnx = FLT_MAX, mny = FLT_MAX, mxx = -FLT_MAX, mxy = -FLT_MAX;
    for (int i = 0; i < n; i++) {
        float x = pos[4 * (size_t)i], y = pos[4 * (size_t)i + 1];
        mnx = fminf(mnx, x);
        mny = fminf(mny, y);
        mxx = fmaxf(mxx, x);
        mxy = fmaxf(mxy, y);
    }
    float b0 = fmaf(fabsf(mnx), -0.001f, mnx);
    float b1 = fmaf(fabsf(mny), -0.001f, mny);
    float b2 = fmaf(fabsf(mxx), 0.001f, mxx);
    float b3 = fmaf(fabsf(mxy), 0.001f, mxy);
    if (accumulate) {
        b0 = fminf(b0, bounds[0]);
        b1 = fminf(b1, bounds[1]);
        b2 = fmaxf(b2, bounds[2]);
        b3 = fmaxf(b3, bounds[3]);
    }
    bounds[0] = b0;
    bounds[1] = b1;
    bounds[2] = b2;
    bounds[3] = b3;
}

/*
 * Space-filling-curve keys: src/BarnesHutKernels.cu:197-240.  16 rounds of float midpoint bisection
 * with STRICT '>' comparisons; 2-bit code per level 01/11/00/10, first code in bits 31:30.
 */
ORACLE_API void oracle_bh_keys(const float *pos, int n, const float *bounds, uint32_t *keys)
{
    for (int b = 0; b < n; b++) {
        float l = bounds[0], r = bounds[2], bo = bounds[1], t = bounds[3];
        float x = pos[4 * (size_t)b], y = pos[4 * (size_t)b + 1];
        uint32_t key = 0;
        for (int i = 30; i >= 0; i -= 2) {
            float midX = (r + l) * 0.5f;
            float midY = (t + bo) * 0.5f;
            if (x > midX) {
                l = midX;
                if (y > midY) { bo = midY; key |= 1u << i; }
                else          { t = midY;  key |= 3u << i; }
            } else {
                r = midX;
                if (y > midY) { bo = midY; }
                else          { t = midY;  key |= 2u << i; }
            }
        }
        keys[b] = key;
    }
}

/*
 * Tree-path keys (no direct counterpart: the reference never materialises them).  They encode
 * the descent of src/BarnesHutKernels.cu:266-293: per level midX=(maxX+minX)*0.5f, x >= midX -> bit 0,
 * y >= midY -> bit 1 ('>=' , unlike the SFC keys).  32 levels, first level in bits 63:62.  The digit
 * stored is path ^ 2 (y flipped) so that the numeric order matches the SFC order wherever a body
 * does not sit exactly on a midpoint; path = digit ^ 2.
 */
ORACLE_API void oracle_bh_tree_keys(const float *pos, int n, const float *bounds, uint64_t *keys)
{
    for (int b = 0; b < n; b++) {
        float minX = bounds[0], maxX = bounds[2], minY = bounds[1], maxY = bounds[3];
        float x = pos[4 * (size_t)b], y = pos[4 * (size_t)b + 1];
        uint64_t key = 0;
        for (int lev = 0; lev < 32; lev++) {
            unsigned path = 0;
            float midX = (maxX + minX) * 0.5f;
            if (x >= midX) { minX = midX; path += 1; } else { maxX = midX; }
            float midY = (maxY + minY) * 0.5f;
            if (y >= midY) { minY = midY; path += 2; } else { maxY = midY; }
            key |= (uint64_t)(path ^ 2u) << (62 - 2 * lev);
        }
        keys[b] = key;
    }
}

/*
 * Stable ascending sort of (key, index): src/BarnesHutKernels.cu:45 (thrust::sort_by_key on unsigned
 * keys with values 0..n-1 written at :237).  LSD radix sort, 8-bit digits == std::stable_sort.
 */
ORACLE_API void oracle_sort_pairs_u32(const uint32_t *keys_in, int n, uint32_t *keys_out, int32_t *idx_out)
{
    uint32_t *k0 = (uint32_t *)malloc(sizeof(uint32_t) * (size_t)n), *k1 = (uint32_t *)malloc(sizeof(uint32_t) * (size_t)n);
    int32_t *v0 = (int32_t *)malloc(sizeof(int32_t) * (size_t)n), *v1 = (int32_t *)malloc(sizeof(int32_t) * (size_t)n);
    memcpy(k0, keys_in, sizeof(uint32_t) * (size_t)n);
    for (int i = 0; i < n; i++) v0[i] = i;
    for (int pass = 0; pass < 4; pass++) {
        size_t cnt[257] = {0};
        int sh = 8 * pass;
        for (int i = 0; i < n; i++) cnt[((k0[i] >> sh) & 255u) + 1]++;
        for (int d = 0; d < 256; d++) cnt[d + 1] += cnt[d];
        for (int i = 0; i < n; i++) {
            size_t p = cnt[(k0[i] >> sh) & 255u]++;
            k1[p] = k0[i];
            v1[p] = v0[i];
        }
        uint32_t *tk = k0; k0 = k1; k1 = tk;
        int32_t *tv = v0; v0 = v1; v1 = tv;
    }
    memcpy(keys_out, k0, sizeof(uint32_t) * (size_t)n);
    memcpy(idx_out, v0, sizeof(int32_t) * (size_t)n);
    free(k0); free(k1); free(v0); free(v1);
}

ORACLE_API void oracle_sort_pairs_u64(const uint64_t *keys_in, int n, uint64_t *keys_out, int32_t *idx_out)
{
    uint64_t *k0 = (uint64_t *)malloc(sizeof(uint64_t) * (size_t)n), *k1 = (uint64_t *)malloc(sizeof(uint64_t) * (size_t)n);
    int32_t *v0 = (int32_t *)malloc(sizeof(int32_t) * (size_t)n), *v1 = (int32_t *)malloc(sizeof(int32_t) * (size_t)n);
    memcpy(k0, keys_in, sizeof(uint64_t) * (size_t)n);
    for (int i = 0; i < n; i++) v0[i] = i;
    for (int pass = 0; pass < 8; pass++) {
        size_t cnt[257] = {0};
        int sh = 8 * pass;
        for (int i = 0; i < n; i++) cnt[((k0[i] >> sh) & 255u) + 1]++;
        for (int d = 0; d < 256; d++) cnt[d + 1] += cnt[d];
        for (int i = 0; i < n; i++) {
            size_t p = cnt[(k0[i] >> sh) & 255u]++;
            k1[p] = k0[i];
            v1[p] = v0[i];
        }
        uint64_t *tk = k0; k0 = k1; k1 = tk;
        int32_t *tv = v0; v0 = v1; v1 = tv;
    }
    memcpy(keys_out, k0, sizeof(uint64_t) * (size_t)n);
    memcpy(idx_out, v0, sizeof(int32_t) * (size_t)n);
    free(k0); free(k1); free(v0); free(v1);
}

/*
 * Region quadtree by sequential insertion: src/BarnesHutKernels.cu:242-365.
 *   - descent with '>=' against float midpoints (:266-285), box narrowed per level;
 *   - bucket size 1: a slot holding a body is split until the two separate (:304-352);
 *   - monopole sums (m, m*x, m*y) added to every node on the way down (:287-289, :327-329); kept in
 *     double here (the reference uses float atomics in scheduler order, which is not reproducible,
 *     SURVEY.md section 0 item 7);
 *   - `prebuilt` != 0 emulates prebuild_tree (:125-150, depth 4): every cell at levels 0..5 is a node
 *     whether or not it holds two bodies;
 *   - fixes: every body is inserted (no 8192-chunk tail rule, :243-245); depth is
 *     capped at ORACLE_MAX_LEVEL, cells at the cap become buckets (the reference has no cap and its
 *     duplicate rule :298-299 drops the resident body).
 * Node ids are oracle-private; compare trees through oracle_bh_tree_canonical().
 */
#define ORACLE_MAX_LEVEL 32

typedef struct {
    int child[4];      /* -1 empty, >=0 body id, <= -2 : node id = -(v) - 2 */
    int level;         /* depth of this cell; root = 0 */
    int count;         /* bodies below */
    uint64_t prefix;   /* path digits (tree convention path = (x>=mx)+2*(y>=my)), first level in bits 63:62 */
    double m, mx, my;  /* monopole sums */
    int bucket_head;   /* level == MAX: singly linked list of bodies through bucket_next */
} onode;

typedef struct {
    onode *nodes;
    int n_nodes, cap;
    int *bucket_next;
    int n;
    float bounds[4];
    const float *pos;
    int prebuilt;
} otree;

static int otree_new_node(otree *t, int level, uint64_t prefix)
{
    if (t->n_nodes == t->cap) {
        t->cap = t->cap * 2 + 16;
        t->nodes = (onode *)realloc(t->nodes, sizeof(onode) * (size_t)t->cap);
    }
    onode *nd = &t->nodes[t->n_nodes];
    nd->child[0] = nd->child[1] = nd->child[2] = nd->child[3] = -1;
    nd->level = level;
    nd->count = 0;
    nd->prefix = prefix;
    nd->m = nd->mx = nd->my = 0.0;
    nd->bucket_head = -1;
    return t->n_nodes++;
}

#define NODE_REF(id) (-(id) - 2)
#define IS_NODE(v) ((v) <= -2)
#define NODE_ID(v) (-(v) - 2)

static inline int path_of(float x, float y, float *minX, float *minY, float *maxX, float *maxY)
{
    int path = 0;
    float midX = (*maxX + *minX) * 0.5f;
    if (x >= midX) { *minX = midX; path += 1; } else { *maxX = midX; }
    float midY = (*maxY + *minY) * 0.5f;
    if (y >= midY) { *minY = midY; path += 2; } else { *maxY = midY; }
    return path;
}

static void otree_insert(otree *t, int body)
{
    const float *pos = t->pos;
    float x = pos[4 * (size_t)body], y = pos[4 * (size_t)body + 1];
    double w = pos[4 * (size_t)body + 3];
    float minX = t->bounds[0], minY = t->bounds[1], maxX = t->bounds[2], maxY = t->bounds[3];
    int node = 0;
    for (;;) {
        onode *nd = &t->nodes[node];
        nd->m += w; nd->mx += w * x; nd->my += w * y; nd->count++;
        if (nd->level == ORACLE_MAX_LEVEL) { /* bucket at the depth cap */
            t->bucket_next[body] = nd->bucket_head;
            nd->bucket_head = body;
            return;
        }
        float cminX = minX, cminY = minY, cmaxX = maxX, cmaxY = maxY;
        int path = path_of(x, y, &cminX, &cminY, &cmaxX, &cmaxY);
        int slot = nd->child[path];
        int clevel = nd->level + 1;
        uint64_t cprefix = nd->prefix | ((uint64_t)path << (62 - 2 * nd->level));
        if (IS_NODE(slot)) {
            node = NODE_ID(slot);
            minX = cminX; minY = cminY; maxX = cmaxX; maxY = cmaxY;
            continue;
        }
        if (slot == -1) {
            if (t->prebuilt && clevel <= 5) {           /* prebuilt top tree: the cell is a node already */
                int nn = otree_new_node(t, clevel, cprefix);
                t->nodes[node].child[path] = NODE_REF(nn);
                node = nn;
                minX = cminX; minY = cminY; maxX = cmaxX; maxY = cmaxY;
                continue;
            }
            nd->child[path] = body;
            return;
        }
        /* slot holds a resident body: push it one level down into a new node, then keep descending */
        {
            int resident = slot;
            int nn = otree_new_node(t, clevel, cprefix);
            nd = &t->nodes[node]; /* realloc may have moved the pool */
            nd->child[path] = NODE_REF(nn);
            onode *c = &t->nodes[nn];
            float rx = pos[4 * (size_t)resident], ry = pos[4 * (size_t)resident + 1];
            double rw = pos[4 * (size_t)resident + 3];
            c->m += rw; c->mx += rw * rx; c->my += rw * ry; c->count++;
            if (clevel == ORACLE_MAX_LEVEL) {
                t->bucket_next[resident] = c->bucket_head;
                c->bucket_head = resident;
            } else {
                float a = cminX, b = cminY, cc = cmaxX, d = cmaxY;
                int rpath = path_of(rx, ry, &a, &b, &cc, &d);
                c->child[rpath] = resident;
            }
            node = nn;
            minX = cminX; minY = cminY; maxX = cmaxX; maxY = cmaxY;
        }
    }
}

ORACLE_API void *oracle_bh_tree_build(const float *pos, int n, const float *bounds, const int32_t *order,
                                      int prebuilt)
{
    otree *t = (otree *)calloc(1, sizeof(otree));
    t->pos = pos;
    t->n = n;
    t->prebuilt = prebuilt;
    memcpy(t->bounds, bounds, sizeof(float) * 4);
    t->bucket_next = (int *)malloc(sizeof(int) * (size_t)(n > 0 ? n : 1));
    for (int i = 0; i < n; i++) t->bucket_next[i] = -1;
    otree_new_node(t, 0, 0);
    for (int i = 0; i < n; i++) otree_insert(t, order ? order[i] : i);
    return t;
}

ORACLE_API void oracle_bh_tree_free(void *h)
{
    otree *t = (otree *)h;
    if (!t) return;
    free(t->nodes);
    free(t->bucket_next);
    free(t);
}

ORACLE_API int oracle_bh_tree_num_nodes(void *h) { return ((otree *)h)->n_nodes; }

/*
 * Canonical topology: one record per cell that holds >= 2 bodies -- (prefix, level, count) plus the
 * monopole (total mass, centre of mass = sum(m*x)/sum(m), src/BarnesHutKernels.cu:367-379), sorted by
 * (prefix, level).  Cells kept alive only by the prebuilt top tree (count < 2) are skipped.  Returns
 * the number of records; pass NULL arrays to query the count.
 */
typedef struct { uint64_t prefix; int level; int count; double m, cx, cy; } ocell;
static int ocell_cmp(const void *a, const void *b)
{
    const ocell *p = (const ocell *)a, *q = (const ocell *)b;
    if (p->prefix != q->prefix) return p->prefix < q->prefix ? -1 : 1;
    return p->level - q->level;
}

ORACLE_API int oracle_bh_tree_canonical(void *h, uint64_t *prefix, int32_t *level, int32_t *count,
                                        double *mass, double *cx, double *cy)
{
    otree *t = (otree *)h;
    int k = 0;
    for (int i = 0; i < t->n_nodes; i++) if (t->nodes[i].count >= 2) k++;
    if (!prefix) return k;
    ocell *cells = (ocell *)malloc(sizeof(ocell) * (size_t)(k > 0 ? k : 1));
    k = 0;
    for (int i = 0; i < t->n_nodes; i++) {
        onode *nd = &t->nodes[i];
        if (nd->count < 2) continue;
        cells[k].prefix = nd->prefix;
        cells[k].level = nd->level;
        cells[k].count = nd->count;
        cells[k].m = nd->m;
        cells[k].cx = nd->m > 0 ? nd->mx / nd->m : 0.0;
        cells[k].cy = nd->m > 0 ? nd->my / nd->m : 0.0;
        k++;
    }
    qsort(cells, (size_t)k, sizeof(ocell), ocell_cmp);
    for (int i = 0; i < k; i++) {
        prefix[i] = cells[i].prefix;
        level[i] = cells[i].level;
        count[i] = cells[i].count;
        mass[i] = cells[i].m;
        cx[i] = cells[i].cx;
        cy[i] = cells[i].cy;
    }
    free(cells);
    return k;
}

/*
 * Tree walk: src/BarnesHutKernels.cu:381-451.  Groups of 32 consecutive entries of `sorted_idx` share
 * one traversal (the reference's warp); a cell is accepted only if `size <= r` holds for EVERY
 * active target of the group (:430, __all_sync), bodies are always summed directly (child < n).
 * size starts as (dim_size/theta)^2 for the root's children (:403-404) and is multiplied by 0.25 at
 * every pop (:411) -- so a cell is tested against its PARENT's width (effective opening angle
 * theta/2) and level-1 cells are never tested.  r = S + dx*dx + dy*dy (:428, FMA-contracted),
 * k = rsqrt(r*r)*m (:431), a += k*d (:432-433).  LIFO order as in :436-444: children rejected
 * later are visited first.  Child order = path index (x>=mid) + 2*(y>=mid).
 * A bucket (depth-cap cell) that is not accepted is summed body by body.
 * counters[0] += nodes opened, [1] += children examined, [2] += lane interactions (accepted cell or
 * body, per active target), [3] = max stack depth seen.
 */
ORACLE_API void oracle_bh_walk(void *h, const int32_t *sorted_idx, int g0, int g1, float theta, float eps,
                               float *acc, int64_t *counters, int threads)
{
    otree *t = (otree *)h;
    const float *pos = t->pos;
    const int n = t->n;
    const float dim_size = fmaxf(t->bounds[2] - t->bounds[0], t->bounds[3] - t->bounds[1]);
    float s0 = dim_size / theta;
    s0 *= s0;
    int64_t c_open = 0, c_exam = 0, c_inter = 0, c_depth = 0;
#ifdef _OPENMP
    if (threads <= 0) threads = omp_get_max_threads();
#endif
#pragma omp parallel for schedule(dynamic, 16) num_threads(threads) reduction(+ : c_open, c_exam, c_inter) reduction(max : c_depth)
    for (int g = g0; g < g1; g++) {
        int base = g * 32;
        int lanes = n - base < 32 ? n - base : 32;
        if (lanes <= 0) continue;
        float x[32], y[32], ax[32], ay[32];
        for (int l = 0; l < lanes; l++) {
            int b = sorted_idx[base + l];
            x[l] = pos[4 * (size_t)b];
            y[l] = pos[4 * (size_t)b + 1];
            ax[l] = ay[l] = 0.f;
        }
        int st_node[4 * ORACLE_MAX_LEVEL + 8];
        float st_size[4 * ORACLE_MAX_LEVEL + 8];
        int top = -1;
        /* :398-407 push the root's non-empty children untested */
        for (int i = 0; i < 4; i++) {
            int c = t->nodes[0].child[i];
            if (c == -1) continue;
            top++;
            st_node[top] = c;
            st_size[top] = s0;
        }
        while (top >= 0) {
            if (top + 1 > c_depth) c_depth = top + 1;
            int ref = st_node[top];
            float size = st_size[top] * 0.25f;
            if (!IS_NODE(ref)) {
                /* a body sitting directly in a root slot: the reference would treat `node` < n as a
                   node index and read garbage; cannot happen for N >= 2 with the prebuilt tree.  Sum it. */
                float m = pos[4 * (size_t)ref + 3];
                for (int l = 0; l < lanes; l++) {
                    float dx = pos[4 * (size_t)ref] - x[l], dy = pos[4 * (size_t)ref + 1] - y[l];
                    float r = fmaf(dy, dy, fmaf(dx, dx, eps));
                    float k = (1.0f / sqrtf(r * r)) * m;
                    ax[l] = fmaf(k, dx, ax[l]);
                    ay[l] = fmaf(k, dy, ay[l]);
                }
                c_inter += lanes;
                top--;
                continue;
            }
            onode *nd = &t->nodes[NODE_ID(ref)];
            c_open++;
            if (nd->level == ORACLE_MAX_LEVEL) { /* opened bucket */
                for (int b = nd->bucket_head; b >= 0; b = t->bucket_next[b]) {
                    float m = pos[4 * (size_t)b + 3];
                    for (int l = 0; l < lanes; l++) {
                        float dx = pos[4 * (size_t)b] - x[l], dy = pos[4 * (size_t)b + 1] - y[l];
                        float r = fmaf(dy, dy, fmaf(dx, dx, eps));
                        float k = (1.0f / sqrtf(r * r)) * m;
                        ax[l] = fmaf(k, dx, ax[l]);
                        ay[l] = fmaf(k, dy, ay[l]);
                    }
                    c_inter += lanes;
                    c_exam++;
                }
                top--;
                continue;
            }
            for (int i = 0; i < 4; i++) {
                int c = nd->child[i];
                if (c == -1) continue;
                c_exam++;
                float cx, cy, cm;
                int is_body = !IS_NODE(c);
                if (is_body) {
                    cx = pos[4 * (size_t)c]; cy = pos[4 * (size_t)c + 1]; cm = pos[4 * (size_t)c + 3];
                } else {
                    onode *cn = &t->nodes[NODE_ID(c)];
                    if (cn->m > 0) { cx = (float)(cn->mx / cn->m); cy = (float)(cn->my / cn->m); }
                    else { cx = 0.f; cy = 0.f; }
                    cm = (float)cn->m;
                }
                int accept = 1;
                float dxs[32], dys[32], rs[32];
                for (int l = 0; l < lanes; l++) {
                    dxs[l] = cx - x[l];
                    dys[l] = cy - y[l];
                    rs[l] = fmaf(dys[l], dys[l], fmaf(dxs[l], dxs[l], eps));
                    if (!(size <= rs[l])) accept = 0;
                }
                if (is_body || accept) {
                    for (int l = 0; l < lanes; l++) {
                        float k = (1.0f / sqrtf(rs[l] * rs[l])) * cm;
                        ax[l] = fmaf(k, dxs[l], ax[l]);
                        ay[l] = fmaf(k, dys[l], ay[l]);
                    }
                    c_inter += lanes;
                } else {
                    st_node[top] = c;
                    st_size[top] = size;
                    top++;
                }
            }
            top--;
        }
        for (int l = 0; l < lanes; l++) {
            int b = sorted_idx[base + l];
            acc[4 * (size_t)b + 0] = ax[l];
            acc[4 * (size_t)b + 1] = ay[l];
            acc[4 * (size_t)b + 2] = 0.f;
            acc[4 * (size_t)b + 3] = 0.f;
        }
    }
    if (counters) {
        counters[0] += c_open;
        counters[1] += c_exam;
        counters[2] += c_inter;
        if (c_depth > counters[3]) counters[3] = c_depth;
    }
}

/*
 * Whole pipeline in the order of barnesHutCompute (src/BarnesHutKernels.cu:11-70): bounds -> keys ->
 * sort -> build -> (centre of mass folded into the walk) -> walk.  Optional outputs may be NULL.
 * This is also the timed host baseline for the Barnes-Hut metric (serial-equivalent insert, OpenMP
 * walk), BASELINE.md section 2.
 */
ORACLE_API int oracle_bh_accel(const float *pos, int n, float theta, float eps, int prebuilt, float *acc,
                               float *bounds_out, uint32_t *keys_out, int32_t *sorted_idx_out,
                               int64_t *counters, int threads)
{
    float bounds[4];
    oracle_bh_bounds(pos, n, bounds, 0);
    uint32_t *keys = (uint32_t *)malloc(sizeof(uint32_t) * (size_t)(n > 0 ? n : 1));
    uint32_t *skeys = (uint32_t *)malloc(sizeof(uint32_t) * (size_t)(n > 0 ? n : 1));
    int32_t *sidx = (int32_t *)malloc(sizeof(int32_t) * (size_t)(n > 0 ? n : 1));
    oracle_bh_keys(pos, n, bounds, keys);
    oracle_sort_pairs_u32(keys, n, skeys, sidx);
    void *tree = oracle_bh_tree_build(pos, n, bounds, sidx, prebuilt);
    int groups = (n + 31) / 32;
    oracle_bh_walk(tree, sidx, 0, groups, theta, eps, acc, counters, threads);
    int k = oracle_bh_tree_num_nodes(tree);
    oracle_bh_tree_free(tree);
    if (bounds_out) memcpy(bounds_out, bounds, sizeof(bounds));
    if (keys_out) memcpy(keys_out, keys, sizeof(uint32_t) * (size_t)n);
    if (sorted_idx_out) memcpy(sorted_idx_out, sidx, sizeof(int32_t) * (size_t)n);
    free(keys); free(skeys); free(sidx);
    return k;
}

/* ------------------------------------------------------------------------------------------ */
/* Barnes-Hut in 3D (octree).  NO REFERENCE COUNTERPART: the reference forces Barnes-Hut to R2 */
/* (src/ComputeMethods.cpp:35-38).  These functions extend the restated 2D rules one dimension */
/* up -- same padding, same '>=' bisection, same bucket-1 region tree, same unanimous-vote     */
/* parent-width criterion, the 3D force law of src/DirectComputeKernels.cu:60-68 -- and serve  */
/* as the checker of the 3D CUDA path (parity unpinned by the reference; accuracy is checked   */
/* against the FP64 direct sum).  Keys: 21 levels x 3 bits = 63 bits, first level in 62:60.    */
/* ------------------------------------------------------------------------------------------ */
#define ORACLE3_MAX_LEVEL 21

ORACLE_API void oracle_bh3_bounds(const float *pos, int n, float *b6)
{
    float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (int i = 0; i < n; i++)
        for (int k = 0; k < 3; k++) {
            mn[k] = fminf(mn[k], pos[4 * (size_t)i + k]);
            mx[k] = fmaxf(mx[k], pos[4 * (size_t)i + k]);
        }
    for (int k = 0; k < 3; k++) {
        b6[k] = fmaf(fabsf(mn[k]), -0.001f, mn[k]);
        b6[3 + k] = fmaf(fabsf(mx[k]), 0.001f, mx[k]);
    }
}

static inline int path3_of(const float *p, float *lo, float *hi)
{
    int path = 0;
    for (int k = 0; k < 3; k++) {
        float mid = (hi[k] + lo[k]) * 0.5f;
        if (p[k] >= mid) { lo[k] = mid; path += 1 << k; } else { hi[k] = mid; }
    }
    return path;
}

ORACLE_API void oracle_bh3_tree_keys(const float *pos, int n, const float *b6, uint64_t *keys)
{
    for (int b = 0; b < n; b++) {
        float lo[3] = {b6[0], b6[1], b6[2]}, hi[3] = {b6[3], b6[4], b6[5]};
        uint64_t key = 0;
        for (int lev = 0; lev < ORACLE3_MAX_LEVEL; lev++)
            key |= (uint64_t)path3_of(pos + 4 * (size_t)b, lo, hi) << (60 - 3 * lev);
        keys[b] = key;
    }
}

typedef struct {
    int child[8];
    int level, count;
    uint64_t prefix;
    double m, mx, my, mz;
    int bucket_head;
} o3node;

typedef struct {
    o3node *nodes;
    int n_nodes, cap;
    int *bucket_next;
    int n;
    float b6[6];
    const float *pos;
} o3tree;

static int o3_new(o3tree *t, int level, uint64_t prefix)
{
    if (t->n_nodes == t->cap) {
        t->cap = t->cap * 2 + 16;
        t->nodes = (o3node *)realloc(t->nodes, sizeof(o3node) * (size_t)t->cap);
    }
    o3node *nd = &t->nodes[t->n_nodes];
    for (int k = 0; k < 8; k++) nd->child[k] = -1;
    nd->level = level; nd->count = 0; nd->prefix = prefix;
    nd->m = nd->mx = nd->my = nd->mz = 0.0;
    nd->bucket_head = -1;
    return t->n_nodes++;
}

static void o3_insert(o3tree *t, int body)
{
    const float *p = t->pos + 4 * (size_t)body;
    double w = p[3];
    float lo[3] = {t->b6[0], t->b6[1], t->b6[2]}, hi[3] = {t->b6[3], t->b6[4], t->b6[5]};
    int node = 0;
    for (;;) {
        o3node *nd = &t->nodes[node];
        nd->m += w; nd->mx += w * p[0]; nd->my += w * p[1]; nd->mz += w * p[2]; nd->count++;
        if (nd->level == ORACLE3_MAX_LEVEL) {
            t->bucket_next[body] = nd->bucket_head;
            nd->bucket_head = body;
            return;
        }
        float clo[3] = {lo[0], lo[1], lo[2]}, chi[3] = {hi[0], hi[1], hi[2]};
        int path = path3_of(p, clo, chi);
        int slot = nd->child[path];
        int clevel = nd->level + 1;
        uint64_t cprefix = nd->prefix | ((uint64_t)path << (60 - 3 * nd->level));
        if (IS_NODE(slot)) {
            node = NODE_ID(slot);
            memcpy(lo, clo, sizeof(lo)); memcpy(hi, chi, sizeof(hi));
            continue;
        }
        if (slot == -1) { nd->child[path] = body; return; }
        int resident = slot;
        int nn = o3_new(t, clevel, cprefix);
        nd = &t->nodes[node];
        nd->child[path] = NODE_REF(nn);
        o3node *c = &t->nodes[nn];
        const float *rp = t->pos + 4 * (size_t)resident;
        double rw = rp[3];
        c->m += rw; c->mx += rw * rp[0]; c->my += rw * rp[1]; c->mz += rw * rp[2]; c->count++;
        if (clevel == ORACLE3_MAX_LEVEL) {
            t->bucket_next[resident] = c->bucket_head;
            c->bucket_head = resident;
        } else {
            float a[3] = {clo[0], clo[1], clo[2]}, b[3] = {chi[0], chi[1], chi[2]};
            c->child[path3_of(rp, a, b)] = resident;
        }
        node = nn;
        memcpy(lo, clo, sizeof(lo)); memcpy(hi, chi, sizeof(hi));
    }
}

ORACLE_API void *oracle_bh3_tree_build(const float *pos, int n, const float *b6, const int32_t *order)
{
    o3tree *t = (o3tree *)calloc(1, sizeof(o3tree));
    t->pos = pos; t->n = n;
    memcpy(t->b6, b6, sizeof(float) * 6);
    t->bucket_next = (int *)malloc(sizeof(int) * (size_t)(n > 0 ? n : 1));
    for (int i = 0; i < n; i++) t->bucket_next[i] = -1;
    o3_new(t, 0, 0);
    for (int i = 0; i < n; i++) o3_insert(t, order ? order[i] : i);
    return t;
}

ORACLE_API void oracle_bh3_tree_free(void *h)
{
    o3tree *t = (o3tree *)h;
    if (!t) return;
    free(t->nodes); free(t->bucket_next); free(t);
}

typedef struct { uint64_t prefix; int level, count; double m, cx, cy, cz; } o3cell;
static int o3cell_cmp(const void *a, const void *b)
{
    const o3cell *p = (const o3cell *)a, *q = (const o3cell *)b;
    if (p->prefix != q->prefix) return p->prefix < q->prefix ? -1 : 1;
    return p->level - q->level;
}

ORACLE_API int oracle_bh3_tree_canonical(void *h, uint64_t *prefix, int32_t *level, int32_t *count, double *mass,
                                         double *cx, double *cy, double *cz)
{
    o3tree *t = (o3tree *)h;
    int k = 0;
    for (int i = 0; i < t->n_nodes; i++) if (t->nodes[i].count >= 2) k++;
    if (!prefix) return k;
    o3cell *cells = (o3cell *)malloc(sizeof(o3cell) * (size_t)(k > 0 ? k : 1));
    k = 0;
    for (int i = 0; i < t->n_nodes; i++) {
        o3node *nd = &t->nodes[i];
        if (nd->count < 2) continue;
        cells[k].prefix = nd->prefix; cells[k].level = nd->level; cells[k].count = nd->count; cells[k].m = nd->m;
        cells[k].cx = nd->m > 0 ? nd->mx / nd->m : 0.0;
        cells[k].cy = nd->m > 0 ? nd->my / nd->m : 0.0;
        cells[k].cz = nd->m > 0 ? nd->mz / nd->m : 0.0;
        k++;
    }
    qsort(cells, (size_t)k, sizeof(o3cell), o3cell_cmp);
    for (int i = 0; i < k; i++) {
        prefix[i] = cells[i].prefix; level[i] = cells[i].level; count[i] = cells[i].count; mass[i] = cells[i].m;
        cx[i] = cells[i].cx; cy[i] = cells[i].cy; cz[i] = cells[i].cz;
    }
    free(cells);
    return k;
}

/* Walk: the rules of oracle_bh_walk with a third coordinate and the Newtonian law
 * k = inv*inv*inv*m, inv = 1/sqrt(S + |d|^2) (src/DirectComputeKernels.cu:64-67). */
ORACLE_API void oracle_bh3_walk(void *h, const int32_t *sorted_idx, int g0, int g1, float theta, float eps, float *acc,
                                int64_t *counters, int threads)
{
    o3tree *t = (o3tree *)h;
    const float *pos = t->pos;
    const int n = t->n;
    const float dim_size = fmaxf(fmaxf(t->b6[3] - t->b6[0], t->b6[4] - t->b6[1]), t->b6[5] - t->b6[2]);
    float s0 = dim_size / theta;
    s0 *= s0;
    int64_t c_open = 0, c_exam = 0, c_inter = 0, c_depth = 0;
#ifdef _OPENMP
    if (threads <= 0) threads = omp_get_max_threads();
#endif
#pragma omp parallel for schedule(dynamic, 16) num_threads(threads) reduction(+ : c_open, c_exam, c_inter) reduction(max : c_depth)
    for (int g = g0; g < g1; g++) {
        int base = g * 32;
        int lanes = n - base < 32 ? n - base : 32;
        if (lanes <= 0) continue;
        float x[32], y[32], z[32], ax[32], ay[32], az[32];
        for (int l = 0; l < lanes; l++) {
            int b = sorted_idx[base + l];
            x[l] = pos[4 * (size_t)b]; y[l] = pos[4 * (size_t)b + 1]; z[l] = pos[4 * (size_t)b + 2];
            ax[l] = ay[l] = az[l] = 0.f;
        }
        int st_node[8 * ORACLE3_MAX_LEVEL + 16];
        float st_size[8 * ORACLE3_MAX_LEVEL + 16];
        int top = -1;
        for (int i = 0; i < 8; i++) {
            int c = t->nodes[0].child[i];
            if (c == -1) continue;
            top++; st_node[top] = c; st_size[top] = s0;
        }
#define O3_ADD(CX, CY, CZ, CM)                                                           \
        for (int l = 0; l < lanes; l++) {                                                \
            float dx = (CX) - x[l], dy = (CY) - y[l], dz = (CZ) - z[l];                  \
            float r = fmaf(dz, dz, fmaf(dy, dy, fmaf(dx, dx, eps)));                     \
            float inv = 1.0f / sqrtf(r);                                                 \
            float k = inv * inv * inv * (CM);                                            \
            ax[l] = fmaf(k, dx, ax[l]); ay[l] = fmaf(k, dy, ay[l]); az[l] = fmaf(k, dz, az[l]); \
        }
        while (top >= 0) {
            if (top + 1 > c_depth) c_depth = top + 1;
            int ref = st_node[top];
            float size = st_size[top] * 0.25f;
            if (!IS_NODE(ref)) {   /* a body directly under the root */
                O3_ADD(pos[4 * (size_t)ref], pos[4 * (size_t)ref + 1], pos[4 * (size_t)ref + 2], pos[4 * (size_t)ref + 3]);
                c_inter += lanes;
                top--;
                continue;
            }
            o3node *nd = &t->nodes[NODE_ID(ref)];
            c_open++;
            if (nd->level == ORACLE3_MAX_LEVEL) {
                for (int b = nd->bucket_head; b >= 0; b = t->bucket_next[b]) {
                    O3_ADD(pos[4 * (size_t)b], pos[4 * (size_t)b + 1], pos[4 * (size_t)b + 2], pos[4 * (size_t)b + 3]);
                    c_inter += lanes; c_exam++;
                }
                top--;
                continue;
            }
            for (int i = 0; i < 8; i++) {
                int c = nd->child[i];
                if (c == -1) continue;
                c_exam++;
                float cx, cy, cz, cm;
                int is_body = !IS_NODE(c);
                if (is_body) {
                    cx = pos[4 * (size_t)c]; cy = pos[4 * (size_t)c + 1]; cz = pos[4 * (size_t)c + 2]; cm = pos[4 * (size_t)c + 3];
                } else {
                    o3node *cn = &t->nodes[NODE_ID(c)];
                    if (cn->m > 0) { cx = (float)(cn->mx / cn->m); cy = (float)(cn->my / cn->m); cz = (float)(cn->mz / cn->m); }
                    else { cx = cy = cz = 0.f; }
                    cm = (float)cn->m;
                }
                int accept = 1;
                for (int l = 0; l < lanes; l++) {
                    float dx = cx - x[l], dy = cy - y[l], dz = cz - z[l];
                    float r = fmaf(dz, dz, fmaf(dy, dy, fmaf(dx, dx, eps)));
                    if (!(size <= r)) accept = 0;
                }
                if (is_body || accept) {
                    O3_ADD(cx, cy, cz, cm);
                    c_inter += lanes;
                } else {
                    st_node[top] = c; st_size[top] = size; top++;
                }
            }
            top--;
        }
#undef O3_ADD
        for (int l = 0; l < lanes; l++) {
            int b = sorted_idx[base + l];
            acc[4 * (size_t)b + 0] = ax[l]; acc[4 * (size_t)b + 1] = ay[l]; acc[4 * (size_t)b + 2] = az[l]; acc[4 * (size_t)b + 3] = 0.f;
        }
    }
    if (counters) {
        counters[0] += c_open; counters[1] += c_exam; counters[2] += c_inter;
        if (c_depth > counters[3]) counters[3] = c_depth;
    }
}
