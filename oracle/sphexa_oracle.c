/* TEST INFRASTRUCTURE ONLY — CPU restatement (plain C11, all-double "f64 mode") of the reference
 * algorithms on the hot path.  It is the checker, never the product: only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline leg may load liboracle.so.
 *
 * Parity: PINNED.  tests/test_oracle.py checks every function below against (i) the reference's own
 * known answers (sph/test/ve.cpp:112-233 + example_data.txt, domain/test/unit/sfc/hilbert.cpp curve
 * table/inversion) stored in tests/golden/, and (ii) outputs of the reference itself
 * (oracle/_ref/libsphexa_ref.so = the reference headers compiled where they lie) on seeded inputs,
 * bit-exactly for integer results and to 1e-13 for floating point.
 *
 * Each function cites the reference lines it restates (paths relative to the reference root,
 * D = domain/include/cstone, S = sph/include/sph).  Compile with -ffp-contract=off.
 */
#define _GNU_SOURCE
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef uint64_t KeyT;
#define MAXLEVEL 21

typedef struct
{
    double lim[6], len[3], inv[3];
    int bc[3]; /* 0 open, 1 periodic, 2 fixed */
} OBox;

void orc_make_box(const double* lim, const int* bc, OBox* b)
{
    /* D/sfc/box.hpp:128-141 */
    for (int k = 0; k < 6; ++k) b->lim[k] = lim[k];
    for (int d = 0; d < 3; ++d)
    {
        b->len[d] = lim[2 * d + 1] - lim[2 * d];
        b->inv[d] = 1.0 / (lim[2 * d + 1] - lim[2 * d]);
        b->bc[d]  = bc[d];
    }
}

/* ------------------------------------------------------------------ Hilbert keys */
/* D/sfc/hilbert.hpp:58-110 */
KeyT orc_hilbert(unsigned px, unsigned py, unsigned pz)
{
    static const unsigned m2h[8] = {0, 1, 3, 2, 7, 6, 4, 5};
    KeyT key = 0;
    for (int level = MAXLEVEL - 1; level >= 0; --level)
    {
        unsigned xi = (px >> level) & 1u, yi = (py >> level) & 1u, zi = (pz >> level) & 1u;
        key = (key << 3) + m2h[(xi << 2) | (yi << 1) | zi];
        px ^= -(xi & ((!yi) | zi));
        py ^= -((xi & (yi | zi)) | (yi & (!zi)));
        pz ^= -((xi & (!yi) & (!zi)) | (yi & (!zi)));
        if (zi) { unsigned t = px; px = py; py = pz; pz = t; }
        else if (!yi) { unsigned t = px; px = pz; pz = t; }
    }
    return key;
}

/* D/sfc/hilbert.hpp:146-188 */
void orc_hilbert_decode(KeyT key, unsigned* ox, unsigned* oy, unsigned* oz)
{
    unsigned px = 0, py = 0, pz = 0;
    for (unsigned level = 0; level < MAXLEVEL; ++level)
    {
        unsigned octant = (key >> (3 * level)) & 7u;
        unsigned xi = octant >> 2u, yi = (octant >> 1u) & 1u, zi = octant & 1u;
        if (yi ^ zi) { unsigned t = px; px = pz; pz = py; py = t; }
        else if ((!xi & !yi & !zi) || (xi & yi & zi)) { unsigned t = px; px = pz; pz = t; }
        unsigned mask = (1u << level) - 1;
        px ^= mask & (-(xi & (yi | zi)));
        py ^= mask & (-((xi & ((!yi) | (!zi))) | ((!xi) & yi & zi)));
        pz ^= mask & (-((xi & (!yi) & (!zi)) | (yi & zi)));
        px |= (xi << level);
        py |= ((xi ^ yi) << level);
        pz |= ((yi ^ zi) << level);
    }
    *ox = px; *oy = py; *oz = pz;
}

/* D/sfc/sfc.hpp:157-193,283-291 */
void orc_sfc_keys(const double* x, const double* y, const double* z, size_t n, const OBox* b, KeyT* keys)
{
    const int mcoord  = (1 << MAXLEVEL) - 1;
    const double cube = (double)(1u << MAXLEVEL);
    double mx = cube * b->inv[0], my = cube * b->inv[1], mz = cube * b->inv[2];
#pragma omp parallel for schedule(static)
    for (size_t i = 0; i < n; ++i)
    {
        int ix = floor(x[i] * mx) - b->lim[0] * mx;
        int iy = floor(y[i] * my) - b->lim[2] * my;
        int iz = floor(z[i] * mz) - b->lim[4] * mz;
        if (ix > mcoord) ix = mcoord;
        if (iy > mcoord) iy = mcoord;
        if (iz > mcoord) iz = mcoord;
        keys[i] = orc_hilbert((unsigned)ix, (unsigned)iy, (unsigned)iz);
    }
}

/* ------------------------------------------------------------------ sort (D/primitives/gather.hpp:58-90) */
typedef struct { KeyT k; uint32_t v; } KV;
static int kv_cmp(const void* a, const void* b)
{
    const KV *p = a, *q = b;
    if (p->k != q->k) return p->k < q->k ? -1 : 1;
    return p->v < q->v ? -1 : (p->v > q->v); /* stable tie-break (the GPU reference sort is stable) */
}
void orc_sort_by_key(KeyT* keys, size_t n, uint32_t* perm)
{
    KV* kv = malloc(n * sizeof(KV));
    for (size_t i = 0; i < n; ++i) { kv[i].k = keys[i]; kv[i].v = (uint32_t)i; }
    qsort(kv, n, sizeof(KV), kv_cmp);
    for (size_t i = 0; i < n; ++i) { keys[i] = kv[i].k; perm[i] = kv[i].v; }
    free(kv);
}

/* ------------------------------------------------------------------ cornerstone leaf array */
static size_t lower_bound_u64(const KeyT* a, size_t n, KeyT v)
{
    size_t lo = 0, hi = n;
    while (lo < hi) { size_t mid = (lo + hi) / 2; if (a[mid] < v) lo = mid + 1; else hi = mid; }
    return lo;
}
static KeyT node_range(unsigned level) { return (KeyT)1 << (3u * (MAXLEVEL - level)); } /* D/sfc/common.hpp:125-132 */
static unsigned tree_level(KeyT range) { return (__builtin_clzll(range - 1) - 1) / 3; }    /* common.hpp:173-178 */
static unsigned octal_digit(KeyT code, unsigned pos) { return (code >> (3u * (MAXLEVEL - pos))) & 7u; }

/* D/tree/csarray.hpp:94-103,203-257 */
void orc_node_counts(const KeyT* leaves, int nl, const KeyT* keys, size_t n, unsigned* counts)
{
#pragma omp parallel for schedule(static)
    for (int i = 0; i < nl; ++i)
        counts[i] = (unsigned)(lower_bound_u64(keys, n, leaves[i + 1]) - lower_bound_u64(keys, n, leaves[i]));
}

/* D/tree/csarray.hpp:273-312 */
int orc_node_op(const KeyT* tree, int i, const unsigned* counts, unsigned bucket)
{
    KeyT node = tree[i];
    unsigned level = tree_level(tree[i + 1] - node);
    int sib = -1;
    if (level > 0)
    {
        sib = (int)octal_digit(node, level);
        if (!(tree[i - sib + 8] == tree[i - sib] + node_range(level - 1))) sib = -1;
    }
    if (sib > 0)
    {
        const unsigned* g = counts + i - sib;
        size_t pc = (size_t)g[0] + g[1] + g[2] + g[3] + g[4] + g[5] + g[6] + g[7];
        if (pc <= (size_t)bucket) return 0;
    }
    if (counts[i] > bucket * 512 && level + 3 < MAXLEVEL) return 4096;
    if (counts[i] > bucket * 64 && level + 2 < MAXLEVEL) return 512;
    if (counts[i] > bucket * 8 && level + 1 < MAXLEVEL) return 64;
    if (counts[i] > bucket && level < MAXLEVEL) return 8;
    return 1;
}

/* one updateOctree step (csarray.hpp:331-451); newLeaves must hold the result; returns new leaf count, *converged */
int orc_update_octree(const KeyT* keys, size_t n, unsigned bucket, const KeyT* leaves, const unsigned* counts, int nl,
                      KeyT* newLeaves, unsigned* newCounts, int capacity, int* converged)
{
    int* ops = malloc((size_t)(nl + 1) * sizeof(int));
    int conv = 1;
    for (int i = 0; i < nl; ++i) { ops[i] = orc_node_op(leaves, i, counts, bucket); if (ops[i] != 1) conv = 0; }
    int total = 0;
    for (int i = 0; i < nl; ++i) { int o = ops[i]; ops[i] = total; total += o; } /* exclusive scan */
    ops[nl] = total;
    if (total > capacity) { free(ops); return -1; }
    for (int i = 0; i < nl; ++i)
    {
        int op = ops[i + 1] - ops[i];
        KeyT node = leaves[i], range = leaves[i + 1] - node;
        for (int s = 0; s < op; ++s) newLeaves[ops[i] + s] = node + (KeyT)s * (range / (KeyT)op); /* processNode :362-388 */
    }
    newLeaves[total] = leaves[nl];
    free(ops);
    orc_node_counts(newLeaves, total, keys, n, newCounts);
    *converged = conv;
    return total;
}

int orc_compute_octree(const KeyT* keys, size_t n, unsigned bucket, KeyT* leaves, unsigned* counts, int capacity,
                       int* iterations)
{
    KeyT* a = malloc((size_t)(capacity + 2) * sizeof(KeyT));
    unsigned* ca = malloc((size_t)(capacity + 2) * sizeof(unsigned));
    leaves[0] = 0; leaves[1] = node_range(0); counts[0] = (unsigned)n;
    int nl = 1, conv = 0, it = 0;
    while (!conv)
    {
        int nn = orc_update_octree(keys, n, bucket, leaves, counts, nl, a, ca, capacity, &conv);
        if (nn < 0) { free(a); free(ca); return -1; }
        memcpy(leaves, a, (size_t)(nn + 1) * sizeof(KeyT));
        memcpy(counts, ca, (size_t)nn * sizeof(unsigned));
        nl = nn; ++it;
    }
    free(a); free(ca);
    if (iterations) *iterations = it;
    return nl;
}

/* ------------------------------------------------------------------ internal octree (D/tree/octree.hpp:94-214) */
static KeyT enc_placeholder(KeyT code, int plen) { return ((KeyT)1 << plen) | (code >> (3 * MAXLEVEL - plen)); }
static unsigned dec_prefix_len(KeyT code) { return 63 - __builtin_clzll(code); }
static KeyT dec_placeholder(KeyT code)
{
    int pl = (int)dec_prefix_len(code);
    return (code ^ ((KeyT)1 << pl)) << (3 * MAXLEVEL - pl);
}
static int u64_cmp(const void* a, const void* b) { KeyT p = *(const KeyT*)a, q = *(const KeyT*)b; return p < q ? -1 : p > q; }

/* Builds the level-sorted node array directly: the set of internal nodes is the set of proper ancestors of the
 * leaves; sorting Warren-Salmon keys orders by (level, key) — the same final arrays buildOctreeCpu produces. */
void orc_build_octree(const KeyT* leaves, int nl, KeyT* prefixes, int* childOffsets, int* parents, int* levelRange,
                      int* internalToLeaf, int* leafToInternal)
{
    int ni = (nl - 1) / 7, nn = nl + ni;
    KV* nodes = malloc((size_t)nn * sizeof(KV));
    int cnt = 0;
    /* leaves: unsorted index ni + i */
    for (int i = 0; i < nl; ++i)
    {
        unsigned lvl = tree_level(leaves[i + 1] - leaves[i]);
        nodes[cnt].k = enc_placeholder(leaves[i], 3 * (int)lvl);
        nodes[cnt].v = (uint32_t)(ni + i);
        cnt++;
    }
    /* internal nodes: every ancestor prefix that starts at a leaf key and is not itself a leaf; enumerate by
       walking up from each leaf while the leaf is the first child */
    KeyT* ik = malloc((size_t)(ni + 1) * sizeof(KeyT));
    int nik = 0;
    for (int i = 0; i < nl; ++i)
    {
        KeyT key = leaves[i];
        unsigned lvl = tree_level(leaves[i + 1] - key);
        while (lvl > 0 && octal_digit(key, lvl) == 0)
        {
            lvl--;
            ik[nik++] = enc_placeholder(key, 3 * (int)lvl);
        }
    }
    qsort(ik, (size_t)nik, sizeof(KeyT), u64_cmp);
    for (int i = 0; i < nik; ++i) { nodes[cnt].k = ik[i]; nodes[cnt].v = (uint32_t)i; cnt++; } /* any index in [0,ni) */
    free(ik);
    qsort(nodes, (size_t)nn, sizeof(KV), kv_cmp);
    /* unsorted ids: leaves ni + j; internal nodes by the binary/octree index correspondence of the reference
       (octree.hpp:71-81,108-115), they surface in leafToInternal[0..ni) and as negative internalToLeaf entries */
    for (int i = 0; i < nn; ++i) prefixes[i] = nodes[i].k;
    for (int i = 0; i < nn; ++i)
    {
        KeyT p = prefixes[i];
        KeyT key = dec_placeholder(p);
        unsigned plen = dec_prefix_len(p);
        int j = (int)lower_bound_u64(leaves, (size_t)nl, key);
        unsigned leafLevel = tree_level(leaves[j + 1] - leaves[j]);
        int isLeaf = (leaves[j] == key) && (3 * leafLevel == plen);
        int uid;
        if (isLeaf) uid = ni + j;
        else
        {
            /* the only leaf boundary whose common prefix is exactly this node's prefix lies between children 3|4
               (octal 011|100 share no bit): tid = last leaf of child 3 */
            KeyT childEnd = key + 4 * node_range(plen / 3 + 1);
            int tid = (int)lower_bound_u64(leaves, (size_t)nl, childEnd) - 1;
            int w = 0;
            for (unsigned l = 1; l <= plen / 3 + 1; ++l)
            {
                int d = (int)octal_digit(leaves[tid], l);
                w += (d >= 4) ? (7 - d) : -d; /* digitWeight, common.hpp:258-262 */
            }
            uid = (tid + w) / 7;
        }
        internalToLeaf[i]   = uid - ni;
        leafToInternal[uid] = i;
    }
    free(nodes);
    for (unsigned level = 0; level <= MAXLEVEL; ++level)
        levelRange[level] = (int)lower_bound_u64(prefixes, (size_t)nn, enc_placeholder(0, 3 * (int)level));
    levelRange[MAXLEVEL + 1] = nn;
    for (int i = 0; i <= nn; ++i) childOffsets[i] = 0;
    for (int i = 0; i < nn; ++i)
    {
        KeyT p = prefixes[i];
        unsigned plen = dec_prefix_len(p), level = plen / 3;
        if (level >= MAXLEVEL) continue;
        KeyT childPrefix = enc_placeholder(dec_placeholder(p), (int)plen + 3);
        int s = levelRange[level + 1], e = levelRange[level + 2];
        int c = s + (int)lower_bound_u64(prefixes + s, (size_t)(e - s), childPrefix);
        if (c != e && prefixes[c] == childPrefix) { childOffsets[i] = c; parents[(c - 1) / 8] = i; }
    }
}

/* D/focus/source_center.hpp:145-158, D/sfc/hilbert.hpp:264-281, D/sfc/box.hpp:334-359 */
void orc_node_centers(const KeyT* prefixes, int nn, const OBox* b, double* centers, double* sizes)
{
    const double uL = 1.0 / (double)(1u << MAXLEVEL);
    for (int i = 0; i < nn; ++i)
    {
        KeyT start = dec_placeholder(prefixes[i]);
        unsigned level = dec_prefix_len(prefixes[i]) / 3;
        unsigned cube = (1u << MAXLEVEL) >> level, mask = ~(cube - 1), c[3];
        orc_hilbert_decode(start, &c[0], &c[1], &c[2]);
        for (int d = 0; d < 3; ++d)
        {
            int lo = (int)(c[d] & mask), hi = lo + (int)cube;
            double half = 0.5 * uL * b->len[d];
            centers[3 * i + d] = b->lim[2 * d] + (hi + lo) * half;
            sizes[3 * i + d]   = (hi - lo) * half;
        }
    }
}

/* ------------------------------------------------------------------ neighbor search */
typedef struct
{
    const int* childOffsets;
    const int* internalToLeaf;
    const uint32_t* layout;
    const double* centers;
    const double* sizes;
} OTree;

static double dist_sq(int pbc, double xj, double yj, double zj, double xi, double yi, double zi, const OBox* b)
{
    /* D/findneighbors.hpp:51-79 */
    double dx = xj - xi, dy = yj - yi, dz = zj - zi;
    if (pbc)
    {
        dx -= (b->bc[0] == 1) * b->len[0] * rint(dx * b->inv[0]);
        dy -= (b->bc[1] == 1) * b->len[1] * rint(dy * b->inv[1]);
        dz -= (b->bc[2] == 1) * b->len[2] * rint(dz * b->inv[2]);
    }
    return dx * dx + dy * dy + dz * dz;
}

/* D/findneighbors.hpp:95-165 with D/traversal/traversal.hpp:68-108 and D/traversal/boxoverlap.hpp:184-218 */
unsigned orc_find_neighbors_one(uint32_t i, const double* x, const double* y, const double* z, const double* h,
                                const OTree* t, const OBox* b, unsigned ngmax, uint32_t* nb)
{
    double xi = x[i], yi = y[i], zi = z[i], hi = h[i];
    double radiusSq = 4.0 * hi * hi;
    int anyPbc = b->bc[0] == 1 || b->bc[1] == 1 || b->bc[2] == 1;
    double r2h = 2.0 * hi;
    int inside = xi - r2h >= b->lim[0] && yi - r2h >= b->lim[2] && zi - r2h >= b->lim[4] && xi + r2h <= b->lim[1] &&
                 yi + r2h <= b->lim[3] && zi + r2h <= b->lim[5];
    int usePbc = anyPbc && !inside;
    unsigned count = 0;
    int stack[256], sp = 0;
    stack[sp++] = 0;
    const double P[3] = {xi, yi, zi};
    while (sp > 0)
    {
        int node = stack[--sp];
        /* overlap: norm2(minDistance(point, center, size[, box])) < radiusSq */
        double d2 = 0;
        for (int d = 0; d < 3; ++d)
        {
            double dx = t->centers[3 * node + d] - P[d];
            if (usePbc && b->bc[d] == 1) dx -= b->len[d] * rint(dx * b->inv[d]);
            dx = fabs(dx) - t->sizes[3 * node + d];
            dx += fabs(dx);
            dx *= 0.5;
            d2 += dx * dx;
        }
        if (!(d2 < radiusSq)) continue;
        int c0 = t->childOffsets[node];
        if (c0 == 0)
        {
            int leaf = t->internalToLeaf[node];
            for (uint32_t j = t->layout[leaf]; j < t->layout[leaf + 1]; ++j)
            {
                if (j == i) continue;
                if (dist_sq(usePbc, x[j], y[j], z[j], xi, yi, zi, b) < radiusSq)
                {
                    if (count < ngmax) nb[count] = j;
                    count++;
                }
            }
        }
        else
            for (int o = 7; o >= 0; --o) stack[sp++] = c0 + o; /* ascending leaf order */
    }
    return count;
}

static double update_h(unsigned ng0, unsigned nc, double h)
{
    /* S/kernels.hpp:26-32 */
    return h * 0.5 * pow(1.0 + 1023.0 * ng0 / (double)nc, 1.0 / 10.0);
}

/* S/find_neighbors.hpp:10-45; nc = count + 1 */
void orc_find_neighbors(const double* x, const double* y, const double* z, double* h, uint32_t first, uint32_t last,
                        const OBox* b, const OTree* t, unsigned ng0, unsigned ngmax, int iterateH, uint32_t* nb,
                        unsigned* nc)
{
    unsigned ngmin = ng0 / 4;
#pragma omp parallel for schedule(dynamic, 64)
    for (uint32_t i = first; i < last; ++i)
    {
        uint32_t* row = nb + (size_t)(i - first) * ngmax;
        unsigned ncSph = 1 + orc_find_neighbors_one(i, x, y, z, h, t, b, ngmax, row);
        int it = 0;
        while (iterateH && (ngmin > ncSph || (ncSph - 1) > ngmax) && it++ < 10)
        {
            h[i]  = update_h(ng0, ncSph, h[i]);
            ncSph = 1 + orc_find_neighbors_one(i, x, y, z, h, t, b, ngmax, row);
        }
        nc[i - first] = ncSph;
    }
}

/* ------------------------------------------------------------------ SPH-VE */
#define TABLE 20000
static double lookup(const double* tab, double v)
{
    /* S/table_lookup.hpp:13-25 */
    const int numIntervals = TABLE - 1;
    const double dx = 2.0 / numIntervals, invDx = 1.0 / dx;
    int idx = v * invDx;
    if (idx >= numIntervals) return 0.0;
    double der = (tab[idx + 1] - tab[idx]) * invDx;
    return tab[idx] + der * (v - (double)idx * dx);
}

static void apply_pbc(const OBox* b, double r, double* xx, double* yy, double* zz)
{
    /* D/sfc/box.hpp:234-255 */
    if (b->bc[0] == 1 && *xx > r) *xx -= b->len[0]; else if (b->bc[0] == 1 && *xx < -r) *xx += b->len[0];
    if (b->bc[1] == 1 && *yy > r) *yy -= b->len[1]; else if (b->bc[1] == 1 && *yy < -r) *yy += b->len[1];
    if (b->bc[2] == 1 && *zz > r) *zz -= b->len[2]; else if (b->bc[2] == 1 && *zz < -r) *zz += b->len[2];
}

typedef struct
{
    double *x, *y, *z, *h, *m, *vx, *vy, *vz, *x_m1, *y_m1, *z_m1, *temp, *du_m1, *alpha;
    double *xm, *kx, *gradh, *prho, *c, *c11, *c12, *c13, *c22, *c23, *c33, *divv, *curlv, *ax, *ay, *az, *du;
    double *dV11, *dV12, *dV13, *dV22, *dV23, *dV33;
} OF;

#define NBLOOP                                                                                                        \
    const uint32_t* row = nb + (size_t)(i - first) * ngmax;                                                          \
    unsigned cnt        = nc[i - first] - 1 < ngmax ? nc[i - first] - 1 : ngmax;                                      \
    for (unsigned pj = 0; pj < cnt; ++pj)

/* S/hydro_ve/xmass_kern.hpp:50-78 */
void orc_xmass(uint32_t first, uint32_t last, unsigned ngmax, double K, const OBox* b, const uint32_t* nb,
               const unsigned* nc, const OF* f, const double* wh)
{
#pragma omp parallel for schedule(static)
    for (uint32_t i = first; i < last; ++i)
    {
        double hi = f->h[i], hInv = 1.0 / hi, h3 = hInv * hInv * hInv, rho0 = f->m[i];
        NBLOOP
        {
            uint32_t j = row[pj];
            double xx = f->x[i] - f->x[j], yy = f->y[i] - f->y[j], zz = f->z[i] - f->z[j];
            apply_pbc(b, 2.0 * hi, &xx, &yy, &zz);
            double dist = sqrt(xx * xx + yy * yy + zz * zz);
            rho0 += lookup(wh, dist * hInv) * f->m[j];
        }
        f->xm[i] = f->m[i] / (rho0 * K * h3);
    }
}

/* S/hydro_ve/ve_def_gradh_kern.hpp:43-90 */
void orc_ve_def_gradh(uint32_t first, uint32_t last, unsigned ngmax, double K, const OBox* b, const uint32_t* nb,
                      const unsigned* nc, const OF* f, const double* wh, const double* whd)
{
#pragma omp parallel for schedule(static)
    for (uint32_t i = first; i < last; ++i)
    {
        double hi = f->h[i], mi = f->m[i], xmi = f->xm[i], hInv = 1.0 / hi, h3 = hInv * hInv * hInv;
        double kxi = xmi, whom = -3.0 * xmi, wrho = -3.0 * mi;
        NBLOOP
        {
            uint32_t j = row[pj];
            double xx = f->x[i] - f->x[j], yy = f->y[i] - f->y[j], zz = f->z[i] - f->z[j];
            apply_pbc(b, 2.0 * hi, &xx, &yy, &zz);
            double v = sqrt(xx * xx + yy * yy + zz * zz) * hInv;
            double w = lookup(wh, v), dw = lookup(whd, v), dterh = -(3.0 * w + v * dw);
            kxi += w * f->xm[j];
            whom += dterh * f->xm[j];
            wrho += dterh * f->m[j];
        }
        kxi *= K * h3;
        whom *= K * h3 * hInv;
        wrho *= K * h3 * hInv;
        whom = whom * mi / xmi + (kxi - K * xmi * h3) * wrho;
        double rhoi = kxi * mi / xmi, dhdrho = -hi / (rhoi * 3.0);
        f->kx[i] = kxi;
        f->gradh[i] = 1.0 - dhdrho * whom;
    }
}

/* S/hydro_ve/eos.hpp:52-76, S/eos.hpp:13-40 */
void orc_eos(uint32_t first, uint32_t last, const OF* f, double mui, double gamma)
{
    double cv = 8.317e7 / mui / (gamma - 1.0);
    for (uint32_t i = first; i < last; ++i)
    {
        double rho = f->kx[i] * f->m[i] / f->xm[i];
        double tmp = cv * f->temp[i] * (gamma - 1.0);
        f->prho[i] = rho * tmp / (f->kx[i] * f->m[i] * f->m[i] * f->gradh[i]);
        f->c[i] = sqrt(tmp);
    }
}

/* S/hydro_ve/iad_kern.hpp:43-109 + divv_curlv_kern.hpp:44-123 (driver iad_divv_curlv.hpp:80-92) */
void orc_iad_divv_curlv(uint32_t first, uint32_t last, unsigned ngmax, double K, const OBox* b, const uint32_t* nb,
                        const unsigned* nc, const OF* f, const double* wh)
{
#pragma omp parallel for schedule(static)
    for (uint32_t i = first; i < last; ++i)
    {
        double xi = f->x[i], yi = f->y[i], zi = f->z[i], hi = f->h[i], hInv = 1.0 / hi;
        double t11 = 0, t12 = 0, t13 = 0, t22 = 0, t23 = 0, t33 = 0;
        {
            NBLOOP
            {
                uint32_t j = row[pj];
                double rx = xi - f->x[j], ry = yi - f->y[j], rz = zi - f->z[j];
                apply_pbc(b, 2.0 * hi, &rx, &ry, &rz);
                double w = lookup(wh, sqrt(rx * rx + ry * ry + rz * rz) * hInv);
                double vw = f->xm[j] / f->kx[j] * w;
                t11 += rx * rx * vw; t12 += rx * ry * vw; t13 += rx * rz * vw;
                t22 += ry * ry * vw; t23 += ry * rz * vw; t33 += rz * rz * vw;
            }
        }
#define GETEXP(v) ((v) == 0.0 ? 0 : ilogb(v))
        int es = GETEXP(t11) + GETEXP(t12) + GETEXP(t13) + GETEXP(t22) + GETEXP(t23) + GETEXP(t33);
        double norm = ldexp(1.0, -es / 6);
        t11 *= norm; t12 *= norm; t13 *= norm; t22 *= norm; t23 *= norm; t33 *= norm;
        double det = t11 * t22 * t33 + 2.0 * t12 * t23 * t13 - t11 * t23 * t23 - t22 * t13 * t13 - t33 * t12 * t12;
        double fac = norm * (hi * hi * hi) / (det * K);
        double c11 = (t22 * t33 - t23 * t23) * fac, c12 = (t13 * t23 - t33 * t12) * fac, c13 = (t12 * t23 - t22 * t13) * fac;
        double c22 = (t11 * t33 - t13 * t13) * fac, c23 = (t13 * t12 - t11 * t23) * fac, c33 = (t11 * t22 - t12 * t12) * fac;
        f->c11[i] = c11; f->c12[i] = c12; f->c13[i] = c13; f->c22[i] = c22; f->c23[i] = c23; f->c33[i] = c33;

        double dVx[3] = {0, 0, 0}, dVy[3] = {0, 0, 0}, dVz[3] = {0, 0, 0}, h3 = hInv * hInv * hInv;
        {
            NBLOOP
            {
                uint32_t j = row[pj];
                double rx = xi - f->x[j], ry = yi - f->y[j], rz = zi - f->z[j];
                apply_pbc(b, 2.0 * hi, &rx, &ry, &rz);
                double Wi = lookup(wh, sqrt(rx * rx + ry * ry + rz * rz) * hInv);
                double A0 = -(c11 * rx + c12 * ry + c13 * rz) * Wi, A1 = -(c12 * rx + c22 * ry + c23 * rz) * Wi,
                       A2 = -(c13 * rx + c23 * ry + c33 * rz) * Wi;
                double sx = (f->vx[j] - f->vx[i]) * f->xm[j], sy = (f->vy[j] - f->vy[i]) * f->xm[j],
                       sz = (f->vz[j] - f->vz[i]) * f->xm[j];
                dVx[0] += sx * A0; dVx[1] += sx * A1; dVx[2] += sx * A2;
                dVy[0] += sy * A0; dVy[1] += sy * A1; dVy[2] += sy * A2;
                dVz[0] += sz * A0; dVz[1] += sz * A1; dVz[2] += sz * A2;
            }
        }
        double nk = K * h3 / f->kx[i];
        f->divv[i] = nk * (dVx[0] + dVy[1] + dVz[2]);
        double cx = dVz[1] - dVy[2], cy = dVx[2] - dVz[0], cz = dVy[0] - dVx[1];
        if (f->curlv) f->curlv[i] = nk * sqrt(cx * cx + cy * cy + cz * cz);
        if (f->dV11)
        {
            f->dV11[i] = nk * dVx[0]; f->dV12[i] = nk * (dVx[1] + dVy[0]); f->dV13[i] = nk * (dVx[2] + dVz[0]);
            f->dV22[i] = nk * dVy[1]; f->dV23[i] = nk * (dVy[2] + dVz[1]); f->dV33[i] = nk * dVz[2];
        }
    }
}

/* S/hydro_ve/av_switches_kern.hpp:44-137 */
void orc_av_switches(uint32_t first, uint32_t last, unsigned ngmax, double K, const OBox* b, const uint32_t* nb,
                     const unsigned* nc, const OF* f, const double* wh, double dt, double alphamin, double alphamax,
                     double decay)
{
#pragma omp parallel for schedule(static)
    for (uint32_t i = first; i < last; ++i)
    {
        double xi = f->x[i], yi = f->y[i], zi = f->z[i], hi = f->h[i], ci = f->c[i], hInv = 1.0 / hi, h3 = hInv * hInv * hInv;
        double c11 = f->c11[i], c12 = f->c12[i], c13 = f->c13[i], c22 = f->c22[i], c23 = f->c23[i], c33 = f->c33[i];
        double vsig = 1.e-40 * ci, divvi = f->divv[i], gx = 0, gy = 0, gz = 0, alpha = f->alpha[i];
        NBLOOP
        {
            uint32_t j = row[pj];
            double rx = xi - f->x[j], ry = yi - f->y[j], rz = zi - f->z[j];
            apply_pbc(b, 2.0 * hi, &rx, &ry, &rz);
            double dist = sqrt(rx * rx + ry * ry + rz * rz);
            double rv = rx * (f->vx[i] - f->vx[j]) + ry * (f->vy[i] - f->vy[j]) + rz * (f->vz[i] - f->vz[j]);
            double vij = 0.0;
            if (rv < 0.0) vij = ci + f->c[j] - 3.0 * rv / dist;
            vsig = vsig > vij ? vsig : vij;
            double Wi = K * h3 * lookup(wh, dist * hInv);
            double A1 = -(c11 * rx + c12 * ry + c13 * rz) * Wi, A2 = -(c12 * rx + c22 * ry + c23 * rz) * Wi,
                   A3 = -(c13 * rx + c23 * ry + c33 * rz) * Wi;
            double fac = f->xm[j] / f->kx[j] * (divvi - f->divv[j]);
            gx += fac * A1; gy += fac * A2; gz += fac * A3;
        }
        double gd = sqrt(gx * gx + gy * gy + gz * gz), aloc = 0.0;
        if (divvi < 0.0)
        {
            double a = hi * hi * gd;
            aloc = alphamax * a / (a + hi * fabs(divvi) + 0.05 * ci);
        }
        if (aloc >= alpha) alpha = aloc;
        else
        {
            double dec = hi / (decay * vsig);
            double adot = (aloc >= alphamin) ? (aloc - alpha) / dec : (alphamin - alpha) / dec;
            alpha += adot * dt;
        }
        f->alpha[i] = alpha;
    }
}

/* S/hydro_ve/momentum_energy_kern.hpp:43-222, S/kernels.hpp:10-16,70-94; returns min Courant dt */
double orc_momentum_energy(uint32_t first, uint32_t last, unsigned ngmax, double K, const OBox* b, const uint32_t* nb,
                           const unsigned* nc, const OF* f, const double* wh, double Atmin, double Atmax, double ramp,
                           float Kcour, int avClean)
{
    double minDt = INFINITY;
#pragma omp parallel for schedule(static) reduction(min : minDt)
    for (uint32_t i = first; i < last; ++i)
    {
        double xi = f->x[i], yi = f->y[i], zi = f->z[i], vxi = f->vx[i], vyi = f->vy[i], vzi = f->vz[i];
        double hi = f->h[i], mi = f->m[i], ci = f->c[i], kxi = f->kx[i], ai = f->alpha[i], xmi = f->xm[i];
        double rhoi = kxi * mi / xmi, prhoi = f->prho[i], hInv = 1.0 / hi, h3 = hInv * hInv * hInv;
        double maxv = 0, mx = 0, my = 0, mz = 0, en = 0, ave = 0;
        double c11 = f->c11[i], c12 = f->c12[i], c13 = f->c13[i], c22 = f->c22[i], c23 = f->c23[i], c33 = f->c33[i];
        const uint32_t* row0 = nb + (size_t)(i - first) * ngmax;
        unsigned cnt0 = nc[i - first] - 1 < ngmax ? nc[i - first] - 1 : ngmax;
        double eta_crit = cbrt(32.0 * M_PI / 3.0 / (double)(cnt0 + 1));
        (void)row0;
        NBLOOP
        {
            uint32_t j = row[pj];
            double rx = xi - f->x[j], ry = yi - f->y[j], rz = zi - f->z[j];
            apply_pbc(b, 2.0 * hi, &rx, &ry, &rz);
            double dist = sqrt(rx * rx + ry * ry + rz * rz);
            double vx = vxi - f->vx[j], vy = vyi - f->vy[j], vz = vzi - f->vz[j];
            double hj = f->h[j], hjInv = 1.0 / hj, v1 = dist * hInv, v2 = dist * hjInv, hj3 = hjInv * hjInv * hjInv;
            double Wi = h3 * lookup(wh, v1), Wj = hj3 * lookup(wh, v2);
            double A1i = -(c11 * rx + c12 * ry + c13 * rz) * Wi, A2i = -(c12 * rx + c22 * ry + c23 * rz) * Wi,
                   A3i = -(c13 * rx + c23 * ry + c33 * rz) * Wi;
            double A1j = -(f->c11[j] * rx + f->c12[j] * ry + f->c13[j] * rz) * Wj,
                   A2j = -(f->c12[j] * rx + f->c22[j] * ry + f->c23[j] * rz) * Wj,
                   A3j = -(f->c13[j] * rx + f->c23[j] * ry + f->c33[j] * rz) * Wj;
            double mj = f->m[j], cj = f->c[j], xmj = f->xm[j], rhoj = f->kx[j] * mj / xmj;
            double rv = rx * vx + ry * vy + rz * vz;
            if (avClean)
            {
                /* avRvCorrection :43-63 with symv (kernels.hpp:86-94) */
                double gi[6] = {f->dV11[i], f->dV12[i], f->dV13[i], f->dV22[i], f->dV23[i], f->dV33[i]};
                double gj[6] = {f->dV11[j], f->dV12[j], f->dV13[j], f->dV22[j], f->dV23[j], f->dV33[j]};
                double d1 = rx * (gi[0] * rx + gi[1] * ry + gi[2] * rz) + ry * (gi[3] * ry + gi[4] * rz) + rz * (gi[5] * rz);
                double d2 = rx * (gj[0] * rx + gj[1] * ry + gj[2] * rz) + ry * (gj[3] * ry + gj[4] * rz) + rz * (gj[5] * rz);
                double eta = v1 < v2 ? v1 : v2, d3 = 1.0;
                if (eta < eta_crit) { double e = 5.0 * (eta - eta_crit); d3 = exp(-e * e); }
                double A = (d2 != 0.0) ? d1 / d2 : 0.0, Ap = 1.0 + A;
                double q = 4.0 * A / (Ap * Ap);
                q = q < 1.0 ? q : 1.0;
                q = q > 0.0 ? q : 0.0;
                rv += -(0.5 * d3 * q) * (d1 + d2);
            }
            double wij = rv / dist, visc = 0.0;
            if (wij < 0.0) { double vs = (ai + f->alpha[j]) / 4.0 * (ci + cj) - 2.0 * wij; visc = -vs * wij; }
            double vsig = 0.5 * (ci + cj) - 2.0 * wij;
            maxv = vsig > maxv ? vsig : maxv;
            double am, bm, At = fabs(rhoi - rhoj) / (rhoi + rhoj);
            if (At < Atmin) { am = xmi * xmi; bm = xmj * xmj; }
            else if (At > Atmax) { am = xmi * xmj; bm = am; }
            else
            {
                double s = ramp * (At - Atmin);
                am = pow(xmi, 2.0 - s) * pow(xmj, s);
                bm = pow(xmj, 2.0 - s) * pow(xmi, s);
            }
            double av = mj / rhoi * visc, bv = mj / rhoj * visc;
            double avx = 0.5 * (av * A1i + bv * A1j), avy = 0.5 * (av * A2i + bv * A2j), avz = 0.5 * (av * A3i + bv * A3j);
            ave += avx * vx + avy * vy + avz * vz;
            en += mj * am * (vx * A1i + vy * A2i + vz * A3i);
            double mi_ = mj * prhoi * am, mj_ = mj * f->prho[j] * bm;
            mx += mi_ * A1i + mj_ * A1j + avx;
            my += mi_ * A2i + mj_ * A2j + avy;
            mz += mi_ * A3i + mj_ * A3j + avz;
        }
        ave = ave > 0.0 ? ave : 0.0;
        f->du[i] = K * (prhoi * en + 0.5 * ave);
        f->ax[i] = -K * mx; f->ay[i] = -K * my; f->az[i] = -K * mz;
        double v = maxv > 0.0 ? maxv : ci;
        double dti = Kcour * hi / v;
        minDt = dti < minDt ? dti : minDt;
    }
    return minDt;
}

/* S/ts_global.hpp:71-94 */
double orc_rho_timestep(uint32_t first, uint32_t last, const double* divv, double Krho)
{
    double m = -INFINITY;
    for (uint32_t i = first; i < last; ++i) m = divv[i] > m ? divv[i] : m;
    return Krho / fabs(m);
}

/* S/positions.hpp:54-138 (value semantics for du_m1) + S/update_h.hpp:12-24 */
void orc_positions_update_h(uint32_t first, uint32_t last, double dt, double dt_m1, const OBox* b, const OF* f,
                            double mui, double gamma, unsigned ng0, const unsigned* nc)
{
    double cv = 8.317e7 / mui / (gamma - 1.0);
    for (uint32_t i = first; i < last; ++i)
    {
        double* X[3]  = {&f->x[i], &f->y[i], &f->z[i]};
        double* V[3]  = {&f->vx[i], &f->vy[i], &f->vz[i]};
        double* D[3]  = {&f->x_m1[i], &f->y_m1[i], &f->z_m1[i]};
        double A[3]   = {f->ax[i], f->ay[i], f->az[i]};
        for (int d = 0; d < 3; ++d)
        {
            double Vh = *D[d] * (1.0 / dt_m1), Vn = Vh + 0.5 * dt_m1 * A[d];
            double dX = (Vn + 0.5 * A[d] * fabs(dt)) * dt, Xn = *X[d] + dX;
            if (b->bc[d] == 1 && Xn > b->lim[2 * d + 1]) Xn -= b->len[d];
            else if (b->bc[d] == 1 && Xn < b->lim[2 * d]) Xn += b->len[d];
            *V[d] = Vn + A[d] * dt; *D[d] = dX; *X[d] = Xn;
        }
        double u_old = cv * f->temp[i], du = f->du[i];
        double u_new = u_old + du * dt + 0.5 * (du - f->du_m1[i]) / dt_m1 * fabs(dt) * dt;
        if (u_new < 0.) u_new = u_old * exp(u_new * dt / u_old);
        f->temp[i] = u_new / cv;
        f->du_m1[i] = du;
        f->h[i] = update_h(ng0, nc[i - first], f->h[i]);
    }
}

/* S/sph_kernel_tables.hpp:28-101 (sinc^n), S/kernels.hpp:35-61 */
static double wharm(double v) { if (v == 0.0) return 1.0; double pv = M_PI_2 * v; return sin(pv) / pv; }
static double wharm_d(double v)
{
    if (v == 0.0) return 0.0;
    double pv = M_PI_2 * v, s = sin(pv) / pv;
    return s * M_PI_2 * ((cos(pv) / sin(pv)) - 1.0 / pv);
}
static int dbl_cmp(const void* a, const void* b) { double p = *(const double*)a, q = *(const double*)b; return p < q ? -1 : p > q; }
void orc_tables(double n, double* wh, double* whd, double* K)
{
    double dx = 2.0 / (TABLE - 1);
    for (int i = 0; i < TABLE; ++i)
    {
        double x = 0.0 + i * dx;
        wh[i]  = pow(wharm(x), n);
        whd[i] = n * pow(wharm(x), n - 1) * wharm_d(x);
    }
    const int m = 2000, nodd = m / 2, neven = nodd - 1;
    double h = 2.0 / m, so = 0, se = 0, odd[1000], even[999];
    for (int i = 0; i < nodd; ++i) { double x = 0.0 + (double)(2 * (i + 1) - 1) * h; odd[i] = 4.0 * M_PI * x * x * pow(wharm(x), n); }
    for (int i = 0; i < neven; ++i) { double x = 0.0 + (double)(2 * (i + 1)) * h; even[i] = 4.0 * M_PI * x * x * pow(wharm(x), n); }
    qsort(odd, nodd, sizeof(double), dbl_cmp);
    qsort(even, neven, sizeof(double), dbl_cmp);
    for (int i = 0; i < nodd; ++i) so += odd[i];
    for (int i = 0; i < neven; ++i) se += even[i];
    double fa = 0.0, fb = 4.0 * M_PI * 4.0 * pow(wharm(2.0), n);
    *K = 1.0 / (h / 3.0 * (fa + fb + 4.0 * so + 2.0 * se));
}

/* ---- observables: localConservedQuantities (main/src/observables/conserved_quantities.hpp:48-107), f64 types.
 * out = { eKin, eInt, 0 (egrav), linmom[3], angmom[3], sum nc }; sequential sums in index order */
void orc_conserved_quantities(uint32_t first, uint32_t last, const double* x, const double* y, const double* z,
                              const double* vx, const double* vy, const double* vz, const double* m, const double* temp,
                              const uint32_t* nc, double mui, double gamma, double* out)
{
    const double R = 8.317e7;                 /* sph/eos.hpp:13-21 idealGasCv */
    double cv = R / mui / (gamma - 1.0);
    double eKin = 0, eInt = 0, lin[3] = {0, 0, 0}, ang[3] = {0, 0, 0}, ncsum = 0;
    for (uint32_t i = first; i < last; ++i)
    {
        double mi = m[i];
        eKin += mi * (vx[i] * vx[i] + vy[i] * vy[i] + vz[i] * vz[i]);           /* :75 mi * norm2(V) */
        lin[0] += mi * vx[i]; lin[1] += mi * vy[i]; lin[2] += mi * vz[i];         /* :76 */
        ang[0] += mi * (y[i] * vz[i] - z[i] * vy[i]);                             /* :77 mi * cross(X, V) */
        ang[1] += mi * (z[i] * vx[i] - x[i] * vz[i]);
        ang[2] += mi * (x[i] * vy[i] - y[i] * vx[i]);
        eInt += cv * temp[i] * mi;                                                /* :101 */
        if (nc) ncsum += (double)nc[i];                                           /* :137-141 */
    }
    out[0] = 0.5 * eKin; out[1] = eInt; out[2] = 0.0;
    for (int k = 0; k < 3; ++k) { out[3 + k] = lin[k]; out[6 + k] = ang[k]; }
    out[9] = ncsum;
}

/* ---- turbulence stirring, f64 types ----
 * computePhases (sph/include/sph/hydro_turb/phases.hpp:46-72): projection of the OU phases */
void orc_turb_phases(int numModes, int numDim, const double* ou, double solWeight, const double* modes,
                     double* phasesReal, double* phasesImag)
{
    for (int i = 0; i < numModes; ++i)
    {
        double ka = 0.0, kb = 0.0, kk = 0.0;
        for (int j = 0; j < numDim; ++j)
        {
            kk = kk + modes[3 * i + j] * modes[3 * i + j];
            ka = ka + modes[3 * i + j] * ou[6 * i + 2 * j + 1];
            kb = kb + modes[3 * i + j] * ou[6 * i + 2 * j];
        }
        for (int j = 0; j < numDim; ++j)
        {
            double diva  = modes[3 * i + j] * ka / kk;
            double divb  = modes[3 * i + j] * kb / kk;
            double curla = ou[6 * i + 2 * j] - divb;
            double curlb = ou[6 * i + 2 * j + 1] - diva;
            phasesReal[3 * i + j] = solWeight * curla + (1.0 - solWeight) * divb;
            phasesImag[3 * i + j] = solWeight * curlb + (1.0 - solWeight) * diva;
        }
    }
}

/* stirParticle + computeStirring (stirring.hpp:43-122) */
void orc_stirring(uint32_t first, uint32_t last, const double* x, const double* y, const double* z, double* ax,
                  double* ay, double* az, int numModes, const double* modes, const double* phaseReal,
                  const double* phaseImag, const double* amplitudes, double solWeightNorm)
{
    for (uint32_t i = first; i < last; ++i)
    {
        double tx = 0.0, ty = 0.0, tz = 0.0;
        for (int m = 0; m < numModes; ++m)
        {
            double ck = cos(modes[3 * m + 2] * z[i]), sk = sin(modes[3 * m + 2] * z[i]);
            double cj = cos(modes[3 * m + 1] * y[i]), sj = sin(modes[3 * m + 1] * y[i]);
            double ci = cos(modes[3 * m] * x[i]), si = sin(modes[3 * m] * x[i]);
            double re = (ci * cj - si * sj) * ck - (si * cj + ci * sj) * sk;
            double im = ci * (cj * sk + sj * ck) + si * (cj * ck - sj * sk);
            tx += amplitudes[m] * (phaseReal[3 * m] * re - phaseImag[3 * m] * im);
            ty += amplitudes[m] * (phaseReal[3 * m + 1] * re - phaseImag[3 * m + 1] * im);
            tz += amplitudes[m] * (phaseReal[3 * m + 2] * re - phaseImag[3 * m + 2] * im);
        }
        ax[i] += solWeightNorm * tx;
        ay[i] += solWeightNorm * ty;
        az[i] += solWeightNorm * tz;
    }
}
