// bvhlab.c -- CPU design tool (NOT product code, NOT the oracle): how many wide-node visits and triangle tests does a ray
// cost on a given mesh for different builders and node widths?  Mirrors the device pipeline's decisions -- Morton LBVH
// (rtb_build.cu: 30-bit keys, radix-tree splits), bottom-up SAH leaf decision (k_fit), greedy largest-area collapse to
// W-wide nodes (k_collapse), front-to-back traversal with tmax culling and a cull-on-pop stack (rtb_trace_state.cuh) --
// next to a binned-SAH binary builder, so that a builder change can be judged before it is written for the GPU.
// Driven by tools/bvh_lab.py (ctypes).
#include <float.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct { float lo[3], hi[3]; } Box;
typedef struct { Box box; int left, right, first, count; float area, cost; int leaf; } BNode;   // binary node; leaf=1: emit as one leaf
typedef struct { Box cb[8]; int ref[8]; int n; } WNode;                                        // ref >= 0 wide node, < 0: ~(first<<4 | count-1)

static const float* g_tri;      // n x 9 floats
static int* g_order;            // leaf order -> triangle
static BNode* g_b; static int g_nb;
static WNode* g_w; static int g_nw;

static float half_area(const Box* b) { float x = b->hi[0] - b->lo[0], y = b->hi[1] - b->lo[1], z = b->hi[2] - b->lo[2]; return x * (y + z) + y * z; }
static void box_init(Box* b) { for (int a = 0; a < 3; a++) { b->lo[a] = FLT_MAX; b->hi[a] = -FLT_MAX; } }
static void box_grow_tri(Box* b, int t)
{
    for (int v = 0; v < 3; v++) for (int a = 0; a < 3; a++) {
        float x = g_tri[9 * t + 3 * v + a];
        if (x < b->lo[a]) b->lo[a] = x;
        if (x > b->hi[a]) b->hi[a] = x;
    }
}
static void box_merge(Box* d, const Box* s) { for (int a = 0; a < 3; a++) { if (s->lo[a] < d->lo[a]) d->lo[a] = s->lo[a]; if (s->hi[a] > d->hi[a]) d->hi[a] = s->hi[a]; } }
static float centroid(int t, int a) { return (g_tri[9 * t + a] + g_tri[9 * t + 3 + a] + g_tri[9 * t + 6 + a]) * (1.0f / 3.0f); }

// ------------------------------------------------------------------ LBVH
static uint32_t expand10(uint32_t v)
{
    v = (v * 0x00010001u) & 0xFF0000FFu; v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u; v = (v * 0x00000005u) & 0x49249249u;
    return v;
}
typedef struct { uint64_t key; int tri; } KV;
static int kv_cmp(const void* a, const void* b) { uint64_t x = ((const KV*)a)->key, y = ((const KV*)b)->key; return x < y ? -1 : (x > y ? 1 : 0); }
static uint64_t* g_keys;

static int new_bnode(void) { return g_nb++; }

// radix-tree split of sorted range [lo, hi]: first position where the highest differing bit flips (ties: the middle)
static int lbvh_build(int lo, int hi)
{
    int id = new_bnode();
    BNode* nd = &g_b[id];
    nd->first = lo; nd->count = hi - lo + 1; nd->left = nd->right = -1;
    if (lo == hi) { box_init(&nd->box); box_grow_tri(&nd->box, g_order[lo]); return id; }
    uint64_t a = g_keys[lo], b = g_keys[hi];
    int split;
    if (a == b) split = (lo + hi) >> 1;
    else {
        int bit = 63 - __builtin_clzll(a ^ b);
        int l = lo, r = hi;                              // last index whose bit `bit` equals a's
        while (l < r) { int m = (l + r + 1) >> 1; if (((g_keys[m] >> bit) & 1) == ((a >> bit) & 1)) l = m; else r = m - 1; }
        split = l;
    }
    int L = lbvh_build(lo, split), R = lbvh_build(split + 1, hi);
    nd = &g_b[id];
    nd->left = L; nd->right = R;
    nd->box = g_b[L].box; box_merge(&nd->box, &g_b[R].box);
    return id;
}

// ------------------------------------------------------------------ binned SAH (16 bins on the centroid bounds, all three axes)
static int sah_build(int lo, int hi)
{
    int id = new_bnode();
    BNode* nd = &g_b[id];
    nd->first = lo; nd->count = hi - lo + 1; nd->left = nd->right = -1;
    box_init(&nd->box);
    Box cb; box_init(&cb);
    for (int i = lo; i <= hi; i++) {
        box_grow_tri(&nd->box, g_order[i]);
        for (int a = 0; a < 3; a++) { float c = centroid(g_order[i], a); if (c < cb.lo[a]) cb.lo[a] = c; if (c > cb.hi[a]) cb.hi[a] = c; }
    }
    if (lo == hi) return id;
    enum { NB = 16 };
    float best = FLT_MAX; int baxis = -1, bbin = -1;
    for (int a = 0; a < 3; a++) {
        float ext = cb.hi[a] - cb.lo[a];
        if (!(ext > 0)) continue;
        Box bb[NB]; int cnt[NB];
        for (int k = 0; k < NB; k++) { box_init(&bb[k]); cnt[k] = 0; }
        for (int i = lo; i <= hi; i++) {
            int k = (int)((centroid(g_order[i], a) - cb.lo[a]) / ext * NB); if (k >= NB) k = NB - 1; if (k < 0) k = 0;
            box_grow_tri(&bb[k], g_order[i]); cnt[k]++;
        }
        float ra[NB]; int rc[NB]; Box acc; box_init(&acc); int c = 0;
        for (int k = NB - 1; k > 0; k--) { if (cnt[k]) box_merge(&acc, &bb[k]); c += cnt[k]; ra[k] = c ? half_area(&acc) : 0; rc[k] = c; }
        box_init(&acc); c = 0;
        for (int k = 0; k < NB - 1; k++) {
            if (cnt[k]) box_merge(&acc, &bb[k]); c += cnt[k];
            if (c == 0 || rc[k + 1] == 0) continue;
            float cost = half_area(&acc) * c + ra[k + 1] * rc[k + 1];
            if (cost < best) { best = cost; baxis = a; bbin = k; }
        }
    }
    int mid;
    if (baxis < 0) mid = (lo + hi) >> 1;
    else {
        float ext = cb.hi[baxis] - cb.lo[baxis];
        int i = lo, j = hi;
        while (i <= j) {
            int k = (int)((centroid(g_order[i], baxis) - cb.lo[baxis]) / ext * NB); if (k >= NB) k = NB - 1; if (k < 0) k = 0;
            if (k <= bbin) i++; else { int t = g_order[i]; g_order[i] = g_order[j]; g_order[j] = t; j--; }
        }
        mid = i - 1;
        if (mid < lo || mid >= hi) mid = (lo + hi) >> 1;
    }
    int L = sah_build(lo, mid), R = sah_build(mid + 1, hi);
    nd = &g_b[id];
    nd->left = L; nd->right = R;
    return id;
}

// ------------------------------------------------------------------ k_fit's leaf decision, k_collapse's greedy opening
static void fit(int id, int max_leaf)
{
    BNode* nd = &g_b[id];
    nd->area = half_area(&nd->box);
    if (nd->left < 0) { nd->cost = nd->area; nd->leaf = 1; return; }
    fit(nd->left, max_leaf); fit(nd->right, max_leaf);
    float c_split = nd->area + g_b[nd->left].cost + g_b[nd->right].cost;
    float c_leaf = nd->count <= max_leaf ? nd->area * (float)nd->count : FLT_MAX;
    nd->leaf = nd->count <= max_leaf && !(c_leaf > c_split);
    nd->cost = nd->leaf ? c_leaf : c_split;
}

static int g_quant_bits = 0;     // > 0: child boxes snapped outwards to a 2^bits grid on the node's own box (power-of-two cell size per axis,
                                 // as in compressed wide BVHs), to count what quantised nodes cost in extra visits
void lab_set_quant(int bits) { g_quant_bits = bits; }

static void quantise(WNode* w)
{
    Box nb; box_init(&nb);
    for (int k = 0; k < w->n; k++) box_merge(&nb, &w->cb[k]);
    const float cells = (float)((1 << g_quant_bits) - 1);
    for (int a = 0; a < 3; a++) {
        float ext = nb.hi[a] - nb.lo[a];
        if (!(ext > 0)) continue;
        float cell = exp2f(ceilf(log2f(ext / cells)));
        for (int k = 0; k < w->n; k++) {
            w->cb[k].lo[a] = nb.lo[a] + floorf((w->cb[k].lo[a] - nb.lo[a]) / cell) * cell;
            w->cb[k].hi[a] = nb.lo[a] + ceilf((w->cb[k].hi[a] - nb.lo[a]) / cell) * cell;
        }
    }
}

// Compressed-wide-BVH child placement (Ylitie et al. 2017): child c goes to the slot s whose octant direction (sign bits of s)
// best matches (centre_c - node centre); greedy on the 8 x 8 cost table.  Traversal order for a ray of octant o = slots by
// descending (s ^ o'), no sort.  Empty slots get ref = EMPTY.
#define EMPTY_SLOT ((int)0x80000000)
static int g_slot_assign = 0;
void lab_set_slots(int on) { g_slot_assign = on; }
static void assign_slots(WNode* w)
{
    Box nb; box_init(&nb);
    for (int k = 0; k < w->n; k++) box_merge(&nb, &w->cb[k]);
    float cost[8][8]; int used_c[8] = {0}, used_s[8] = {0};
    for (int c = 0; c < w->n; c++) for (int s = 0; s < 8; s++) {
        float v = 0;
        for (int a = 0; a < 3; a++) {
            float rel = (w->cb[c].lo[a] + w->cb[c].hi[a]) - (nb.lo[a] + nb.hi[a]);
            v += ((s >> a) & 1) ? rel : -rel;          /* slot bit a set = child on the + side of axis a */
        }
        cost[c][s] = v;
    }
    Box cb[8]; int ref[8];
    for (int s = 0; s < 8; s++) { ref[s] = EMPTY_SLOT; box_init(&cb[s]); }
    for (int it = 0; it < w->n; it++) {
        int bc = -1, bs = -1; float bv = -FLT_MAX;
        for (int c = 0; c < w->n; c++) if (!used_c[c]) for (int s = 0; s < 8; s++) if (!used_s[s] && cost[c][s] > bv) { bv = cost[c][s]; bc = c; bs = s; }
        used_c[bc] = used_s[bs] = 1; cb[bs] = w->cb[bc]; ref[bs] = w->ref[bc];
    }
    for (int s = 0; s < 8; s++) { w->cb[s] = cb[s]; w->ref[s] = ref[s]; }
    w->n = 8;
}

static int collapse(int bid, int width)
{
    int wid = g_nw++;
    int slot[8], ns;
    if (g_b[bid].leaf) { slot[0] = bid; ns = 1; }
    else { slot[0] = g_b[bid].left; slot[1] = g_b[bid].right; ns = 2; }
    while (ns < width) {
        int best = -1; float ba = -1.f;
        for (int k = 0; k < ns; k++) if (!g_b[slot[k]].leaf && g_b[slot[k]].area > ba) { ba = g_b[slot[k]].area; best = k; }
        if (best < 0) break;
        int id = slot[best];
        slot[best] = g_b[id].left; slot[ns++] = g_b[id].right;
    }
    g_w[wid].n = ns;
    for (int k = 0; k < ns; k++) {
        g_w[wid].cb[k] = g_b[slot[k]].box;
        if (g_b[slot[k]].leaf) g_w[wid].ref[k] = ~((g_b[slot[k]].first << 4) | (g_b[slot[k]].count - 1));
        else { int c = collapse(slot[k], width); g_w[wid].ref[k] = c; }
    }
    if (g_quant_bits > 0) quantise(&g_w[wid]);
    if (g_slot_assign) assign_slots(&g_w[wid]);
    return wid;
}

// ------------------------------------------------------------------ traversal (closest hit), counting
static int g_last_tri = -1;
static int tri_hit(int t, const float* o, const float* d, float* tmax)
{
    const float* v = g_tri + 9 * t;
    float e1[3] = { v[3] - v[0], v[4] - v[1], v[5] - v[2] }, e2[3] = { v[6] - v[0], v[7] - v[1], v[8] - v[2] };
    float p[3] = { d[1] * e2[2] - d[2] * e2[1], d[2] * e2[0] - d[0] * e2[2], d[0] * e2[1] - d[1] * e2[0] };
    float det = e1[0] * p[0] + e1[1] * p[1] + e1[2] * p[2];
    if (fabsf(det) < 1e-12f) return 0;
    float inv = 1.0f / det, s[3] = { o[0] - v[0], o[1] - v[1], o[2] - v[2] };
    float u = (s[0] * p[0] + s[1] * p[1] + s[2] * p[2]) * inv;
    if (u < 0 || u > 1) return 0;
    float q[3] = { s[1] * e1[2] - s[2] * e1[1], s[2] * e1[0] - s[0] * e1[2], s[0] * e1[1] - s[1] * e1[0] };
    float w = (d[0] * q[0] + d[1] * q[1] + d[2] * q[2]) * inv;
    if (w < 0 || u + w > 1) return 0;
    float t_ = (e2[0] * q[0] + e2[1] * q[1] + e2[2] * q[2]) * inv;
    if (t_ <= 0 || t_ >= *tmax) return 0;
    *tmax = t_;
    g_last_tri = t;
    return 1;
}

typedef struct { float t; int ref; } SE;
static int g_order_mode = 0;     // 0: hit children fully sorted by entry distance (the device's 5-comparator network); 1: only the nearest is
                                 // picked, the others are pushed in slot order; 2: no ordering at all (slot order)
void lab_set_order(int mode) { g_order_mode = mode; }

// rays: n x 6 (origin, direction).  out: per ray [nodes, tris, hit].  tio (optional): per ray [tmax in / closest t out, hit triangle
// or -1 (as a float)] so that a caller can chain several instances of the mesh along one ray
void lab_trace_t(const float* rays, int n, int32_t* out, float* tio)
{
    SE stack[256];
    for (int r = 0; r < n; r++) {
        const float* o = rays + 6 * r; const float* d = o + 3;
        float id[3]; for (int a = 0; a < 3; a++) id[a] = 1.0f / (fabsf(d[a]) < 1e-20f ? 1e-20f : d[a]);
        float tmax = tio ? tio[2 * r] : FLT_MAX; int sp = 0, nodes = 0, tris = 0, hit = 0;
        g_last_tri = -1;
        int cur = 0;
        for (;;) {
            if (cur >= 0) {
                const WNode* w = &g_w[cur];
                nodes++;
                SE h[8]; int nh = 0;
                for (int k = 0; k < w->n; k++) {
                    float tn = 0, tf = tmax;
                    for (int a = 0; a < 3; a++) {
                        float t0 = (w->cb[k].lo[a] - o[a]) * id[a], t1 = (w->cb[k].hi[a] - o[a]) * id[a];
                        if (t0 > t1) { float t = t0; t0 = t1; t1 = t; }
                        if (t0 > tn) tn = t0;
                        if (t1 < tf) tf = t1;
                    }
                    if (tn <= tf) { h[nh].t = tn; h[nh].ref = w->ref[k]; nh++; }
                }
                if (g_order_mode == 0) { for (int i = 1; i < nh; i++) { SE x = h[i]; int j = i - 1; while (j >= 0 && h[j].t > x.t) { h[j + 1] = h[j]; j--; } h[j + 1] = x; } }
                else if (g_order_mode == 3) {          // by the child centre projected on the ray's octant diagonal: the best an octant-indexed, sort-free order can do
                    float key[8];
                    for (int i = 0; i < nh; i++) {
                        int k = -1; for (int q = 0; q < w->n; q++) if (w->ref[q] == h[i].ref) k = q;
                        key[i] = 0; for (int a = 0; a < 3; a++) key[i] += (w->cb[k].lo[a] + w->cb[k].hi[a]) * (d[a] < 0 ? -1.f : 1.f);
                    }
                    for (int i = 1; i < nh; i++) { SE x = h[i]; float kx = key[i]; int j = i - 1; while (j >= 0 && key[j] > kx) { h[j + 1] = h[j]; key[j + 1] = key[j]; j--; } h[j + 1] = x; key[j + 1] = kx; }
                }
                else if (g_order_mode == 1) { int m = 0; for (int i = 1; i < nh; i++) if (h[i].t < h[m].t) m = i; if (nh) { SE x = h[m]; h[m] = h[0]; h[0] = x; } }
                for (int i = nh - 1; i >= 1; i--) stack[sp++] = h[i];
                if (nh) { cur = h[0].ref; continue; }
            } else {
                unsigned b = ~(unsigned)cur; int first = (int)(b >> 4), count = (int)(b & 15u) + 1;
                for (int i = 0; i < count; i++) { tris++; hit |= tri_hit(g_order[first + i], o, d, &tmax); }
            }
            int got = 0;
            while (sp > 0) { SE e = stack[--sp]; if (e.t < tmax) { cur = e.ref; got = 1; break; } }
            if (!got) break;
        }
        out[3 * r] = nodes; out[3 * r + 1] = tris; out[3 * r + 2] = hit;
        if (tio) { tio[2 * r] = tmax; tio[2 * r + 1] = (float)g_last_tri; }
    }
}
void lab_trace(const float* rays, int n, int32_t* out) { lab_trace_t(rays, n, out, 0); }

// Compressed-wide-BVH traversal: children visited in (slot ^ octant) order, the stack holds GROUPS (node, mask of children still
// to visit) and cannot cull on pop (cull = 1: per-child entry distances kept, to measure what that costs); triangle leaves of a
// node are tested right after the node (postpone = 0) .  out as lab_trace_t.
void lab_trace_cw(const float* rays, int n, int32_t* out, float* tio, int cull)
{
    typedef struct { int node; unsigned mask; float tn[8]; } GE;
    static GE stack[256];
    for (int r = 0; r < n; r++) {
        const float* o = rays + 6 * r; const float* d = o + 3;
        float id[3]; for (int a = 0; a < 3; a++) id[a] = 1.0f / (fabsf(d[a]) < 1e-20f ? 1e-20f : d[a]);
        int oct = (d[0] < 0 ? 1 : 0) | (d[1] < 0 ? 2 : 0) | (d[2] < 0 ? 4 : 0);     /* slots with bit set = + side: visited first when d < 0 ... */
        float tmax = tio ? tio[2 * r] : FLT_MAX; int sp = 0, nodes = 0, tris = 0, hit = 0;
        g_last_tri = -1;
        int cur = 0;
        for (;;) {
            const WNode* w = &g_w[cur];
            nodes++;
            GE g; g.node = cur; g.mask = 0;
            unsigned leafmask = 0;
            for (int k = 0; k < 8; k++) {
                if (w->ref[k] == EMPTY_SLOT) continue;
                float tn = 0, tf = tmax;
                for (int a = 0; a < 3; a++) {
                    float t0 = (w->cb[k].lo[a] - o[a]) * id[a], t1 = (w->cb[k].hi[a] - o[a]) * id[a];
                    if (t0 > t1) { float t = t0; t0 = t1; t1 = t; }
                    if (t0 > tn) tn = t0;
                    if (t1 < tf) tf = t1;
                }
                if (tn <= tf) { if (w->ref[k] >= 0) { g.mask |= 1u << k; g.tn[k] = tn; } else leafmask |= 1u << k; }
            }
            /* leaves of this node first, in traversal order (nearest slots first) */
            for (int p = 0; p < 8; p++) {
                int k = p ^ oct;                           /* priority p = 0 first: slot whose sign bits equal the ray's negative bits -> - side for + dirs */
                if (!((leafmask >> k) & 1)) continue;
                unsigned b = ~(unsigned)w->ref[k]; int first = (int)(b >> 4), count = (int)(b & 15u) + 1;
                for (int i = 0; i < count; i++) { tris++; hit |= tri_hit(g_order[first + i], o, d, &tmax); }
            }
            int got = 0;
            for (;;) {
                if (g.mask == 0) { if (sp == 0) break; g = stack[--sp]; }
                int k = -1;
                for (int p = 0; p < 8; p++) { int q = p ^ oct; if ((g.mask >> q) & 1) { k = q; break; } }
                g.mask &= ~(1u << k);
                if (cull && !(g.tn[k] < tmax)) continue;
                cur = g_w[g.node].ref[k];
                if (g.mask) stack[sp++] = g;
                got = 1; break;
            }
            if (!got) break;
        }
        out[3 * r] = nodes; out[3 * r + 1] = tris; out[3 * r + 2] = hit;
        if (tio) { tio[2 * r] = tmax; tio[2 * r + 1] = (float)g_last_tri; }
    }
}

// builder: 0 = Morton LBVH (bits per axis), 1 = binned SAH.  Returns the wide-node count; stats[0] = leaves, stats[1] = SAH cost of the root
int lab_build(const float* tris, int n, int builder, int bits, int max_leaf, int width, float* stats)
{
    g_tri = tris;
    free(g_order); free(g_b); free(g_w); free(g_keys);
    g_order = (int*)malloc(sizeof(int) * n); g_b = (BNode*)calloc(2 * (size_t)n, sizeof(BNode)); g_w = (WNode*)calloc((size_t)n + 1, sizeof(WNode));
    g_keys = (uint64_t*)malloc(sizeof(uint64_t) * n);
    g_nb = g_nw = 0;
    if (builder == 0) {
        float lo[3] = { FLT_MAX, FLT_MAX, FLT_MAX }, hi[3] = { -FLT_MAX, -FLT_MAX, -FLT_MAX };
        float* c = (float*)malloc(sizeof(float) * 3 * n);
        for (int t = 0; t < n; t++) {
            Box b; box_init(&b); box_grow_tri(&b, t);
            for (int a = 0; a < 3; a++) { c[3 * t + a] = 0.5f * (b.lo[a] + b.hi[a]); if (c[3 * t + a] < lo[a]) lo[a] = c[3 * t + a]; if (c[3 * t + a] > hi[a]) hi[a] = c[3 * t + a]; }
        }
        KV* kv = (KV*)malloc(sizeof(KV) * n);
        const float scale = (float)(1u << bits);
        for (int t = 0; t < n; t++) {
            uint64_t k = 0; uint32_t q[3];
            for (int a = 0; a < 3; a++) {
                float e = hi[a] - lo[a], x = e > 0 ? (c[3 * t + a] - lo[a]) / e : 0;
                uint32_t v = (uint32_t)(x * scale); if (v > (1u << bits) - 1) v = (1u << bits) - 1;
                q[a] = v;
            }
            for (int b = bits - 1; b >= 0; b--) k = (k << 3) | (((q[0] >> b) & 1) << 2) | (((q[1] >> b) & 1) << 1) | ((q[2] >> b) & 1);
            kv[t].key = k; kv[t].tri = t;
        }
        (void)expand10;
        qsort(kv, n, sizeof(KV), kv_cmp);
        for (int i = 0; i < n; i++) { g_order[i] = kv[i].tri; g_keys[i] = kv[i].key; }
        free(kv); free(c);
        lbvh_build(0, n - 1);
    } else {
        for (int i = 0; i < n; i++) g_order[i] = i;
        sah_build(0, n - 1);
    }
    fit(0, max_leaf);
    collapse(0, width);
    int leaves = 0;
    for (int i = 0; i < g_nw; i++) for (int k = 0; k < g_w[i].n; k++) leaves += g_w[i].ref[k] < 0;
    stats[0] = (float)leaves; stats[1] = g_b[0].cost / g_b[0].area;
    return g_nw;
}
