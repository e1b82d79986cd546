/*
 * rt_oracle.c -- CPU restatement of the AFIDclan/cuda-raytracing render path (TEST INFRASTRUCTURE).
 * See rt_oracle.h for the rules that apply to this directory and for the parity status.
 * Build:  gcc -O2 -ffp-contract=off -pthread -fPIC -shared rt_oracle.c -lm   (oracle/Makefile)
 *
 * All float expressions keep the reference's operand order and associativity; there is no FMA
 * contraction.  References are to /root/reference/CudaRaytracer/<file>:<line>.
 */
#include "rt_oracle.h"
#include <float.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <ctype.h>
#include <pthread.h>

/* ------------------------------------------------------------------ utils.hpp */

/* utils.hpp:12-27  Q_rsqrt; `long` taken as 32 bits (the author's MSVC target; SURVEY Q1 / patch P2). */
float o_Q_rsqrt(float number)
{
    int32_t i;
    float x2, y;
    const float threehalfs = 1.5F;
    x2 = number * 0.5F;
    y = number;
    memcpy(&i, &y, 4);
    i = 0x5f3759df - (i >> 1);
    memcpy(&y, &i, 4);
    y = y * (threehalfs - (x2 * y * y));
    return y;
}

static inline o_f3 f3(float x, float y, float z) { o_f3 r = { x, y, z }; return r; }
/* utils.hpp:49-55,61-63 */
static inline float dot3(o_f3 a, o_f3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static inline o_f3 sub3(o_f3 a, o_f3 b) { return f3(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline o_f3 add3(o_f3 a, o_f3 b) { return f3(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline o_f3 scale3(o_f3 a, float b) { return f3(a.x * b, a.y * b, a.z * b); }
static inline o_f3 cross3(o_f3 a, o_f3 b)
{
    return f3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
/* utils.hpp:29-31 */
static inline float magnitude3(o_f3 v) { return sqrtf(v.x * v.x + v.y * v.y + v.z * v.z); }
/* utils.hpp:37-47 */
o_f3 o_normalize(o_f3 v)
{
    float inv_mag = o_Q_rsqrt(v.x * v.x + v.y * v.y + v.z * v.z);
    return f3(v.x * inv_mag, v.y * inv_mag, v.z * inv_mag);
}
/* utils.hpp:142-160 */
void o_invert_intrinsic(const float K[9], float Ki[9])
{
    float fx_inv = 1.0f / K[0];
    float fy_inv = 1.0f / K[4];
    float cx = K[2], cy = K[5];
    Ki[0] = fx_inv; Ki[1] = 0.0f; Ki[2] = -cx * fx_inv;
    Ki[3] = 0.0f; Ki[4] = fy_inv; Ki[5] = -cy * fy_inv;
    Ki[6] = 0.0f; Ki[7] = 0.0f; Ki[8] = 1.0f;
}

/* the libm calls of the path, exposed so device-side restatements can be checked against them */
void o_libm_atanf(const float* in, float* out, int n) { for (int i = 0; i < n; i++) out[i] = atanf(in[i]); }

/* ------------------------------------------------------------------ transforms.hpp */

typedef struct { float m[3][3]; } m33;
typedef struct { float m[4][4]; } m44;

/* transforms.hpp:129-144 */
static m33 euler2rotmat(o_f3 e)
{
    float sy = sinf(e.x), cy = cosf(e.x), sp = sinf(e.y), cp = cosf(e.y), sr = sinf(e.z), cr = cosf(e.z);
    m33 r = { { { cr * cy + sr * sp * sy, -cr * sy + sr * sp * cy, -sr * cp },
                { cp * sy, cp * cy, sp },
                { sr * cy - cr * sp * sy, -sr * sy - cr * sp * cy, cr * cp } } };
    return r;
}
/* transforms.hpp:119-126 */
static o_f3 rotmat2euler(m33 r)
{
    float a = r.m[1][2];
    if (a > 1) a = 1; else if (a < -1) a = -1;
    return f3(atan2f(r.m[1][0], r.m[1][1]), asinf(a), atan2f(-r.m[0][2], r.m[2][2]));
}
/* transforms.hpp:58-73 */
static m33 invert_rotmat(m33 r)
{
    m33 o;
    for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) o.m[i][j] = r.m[j][i];
    return o;
}
static o_f3 apply_rotmat(m33 r, o_f3 v)
{
    return f3(r.m[0][0] * v.x + r.m[0][1] * v.y + r.m[0][2] * v.z,
              r.m[1][0] * v.x + r.m[1][1] * v.y + r.m[1][2] * v.z,
              r.m[2][0] * v.x + r.m[2][1] * v.y + r.m[2][2] * v.z);
}
/* transforms.hpp:178-193 */
static m44 lre2homo(o_lre v)
{
    o_f3 shift = f3(-v.x, -v.y, -v.z);
    m33 R = euler2rotmat(f3(v.yaw, v.pitch, v.roll));
    o_f3 rs = apply_rotmat(R, shift);
    m44 h = { { { R.m[0][0], R.m[0][1], R.m[0][2], rs.x },
                { R.m[1][0], R.m[1][1], R.m[1][2], rs.y },
                { R.m[2][0], R.m[2][1], R.m[2][2], rs.z },
                { 0.0f, 0.0f, 0.0f, 1.0f } } };
    return h;
}
/* transforms.hpp:76-100 */
static m44 invert_homo(m44 H)
{
    m33 R;
    for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) R.m[i][j] = H.m[i][j];
    m33 Ri = invert_rotmat(R);
    o_f3 t = f3(-H.m[0][3], -H.m[1][3], -H.m[2][3]);
    o_f3 ti = apply_rotmat(Ri, t);
    m44 o;
    for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) o.m[i][j] = Ri.m[i][j];
    o.m[0][3] = ti.x; o.m[1][3] = ti.y; o.m[2][3] = ti.z;
    o.m[3][0] = 0.0f; o.m[3][1] = 0.0f; o.m[3][2] = 0.0f; o.m[3][3] = 1.0f;
    return o;
}
/* transforms.hpp:195-216 */
static o_lre homo2lre(m44 H)
{
    m33 R;
    for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) R.m[i][j] = H.m[i][j];
    o_f3 e = rotmat2euler(R);
    o_f3 shift = f3(H.m[0][3], H.m[1][3], H.m[2][3]);
    shift = apply_rotmat(invert_rotmat(R), shift);
    o_lre o;
    o.x = -shift.x; o.y = -shift.y; o.z = -shift.z;
    o.yaw = e.x; o.pitch = e.y; o.roll = e.z;
    return o;
}
/* transforms.hpp:232-235 */
o_lre o_invert_lre(o_lre l) { return homo2lre(invert_homo(lre2homo(l))); }

/* transforms.hpp:148-163.  `euler.x * 0.5` is a double product narrowed to float for sinf: exact. */
o_f4 o_euler2quat(o_f3 e)
{
    float sy = sinf((float)(e.x * 0.5)), cy = cosf((float)(e.x * 0.5));
    float sp = sinf((float)(e.y * 0.5)), cp = cosf((float)(e.y * 0.5));
    float sr = sinf((float)(e.z * 0.5)), cr = cosf((float)(e.z * 0.5));
    o_f4 q = { sy * sp * sr + cy * cp * cr,
               cy * sp * cr + sy * cp * sr,
               -sy * sp * cr + cy * cp * sr,
               cy * sp * sr - sy * cp * cr };
    return q;
}
/* transforms.hpp:165-176 */
o_f3 o_apply_quat(o_f4 q, o_f3 v)
{
    float a = -v.x * q.y - v.y * q.z - v.z * q.w;
    float b = v.x * q.x + v.y * q.w - v.z * q.z;
    float c = v.y * q.x + v.z * q.y - v.x * q.w;
    float d = v.z * q.x + v.x * q.z - v.y * q.y;
    return f3(q.x * b - q.y * a - q.z * d + q.w * c,
              q.x * c - q.z * a - q.w * b + q.y * d,
              q.x * d - q.w * a - q.y * c + q.z * b);
}
/* transforms.hpp:219-221 */
o_f3 o_apply_euler(o_f3 e, o_f3 v) { return o_apply_quat(o_euler2quat(e), v); }
/* transforms.hpp:223-226 */
o_f3 o_apply_lre(o_lre l, o_f3 v)
{
    o_f3 s = f3(v.x - l.x, v.y - l.y, v.z - l.z);
    return o_apply_euler(f3(l.yaw, l.pitch, l.roll), s);
}
/* MeshInstance.hpp:39-46 */
void o_instance_build_inv(o_instance* in)
{
    in->inv_pose = o_invert_lre(in->pose);
    in->inv_scale = f3(1 / in->scale.x, 1 / in->scale.y, 1 / in->scale.z);
    in->rotation = f3(in->pose.yaw, in->pose.pitch, in->pose.roll);
    in->inv_rotation = f3(in->inv_pose.yaw, in->inv_pose.pitch, in->inv_pose.roll);
}

/* ------------------------------------------------------------------ TrianglePrimitive.hpp */

/* OBJLoader.hpp:143-147: normal = normalize(cross(v[i]-v[0], v[i+1]-v[0])), then the 4/7-arg ctor
 * (TrianglePrimitive.hpp:26-48) stores it verbatim.  UVs zeroed (the reference leaves them
 * uninitialised when the face has no vt -- SURVEY Q21). */
o_tri o_make_triangle(o_f3 a, o_f3 b, o_f3 c)
{
    o_tri t;
    memset(&t, 0, sizeof t);
    t.vertices[0] = a; t.vertices[1] = b; t.vertices[2] = c;
    t.normal = o_normalize(cross3(sub3(b, a), sub3(c, a)));
    return t;
}

/* TrianglePrimitive.hpp:81-83 */
static inline o_f3 tri_center(const o_tri* t)
{
    o_f3 s = add3(add3(t->vertices[0], t->vertices[1]), t->vertices[2]);
    return f3(s.x / 3.0f, s.y / 3.0f, s.z / 3.0f);
}

/* ------------------------------------------------------------------ BVHTree.hpp (host builder) */

typedef struct {
    o_node* nodes; int n_nodes, cap_nodes;
    int* indices; int n_indices;
    const o_f3 *tmin, *tmax, *tcen;     /* per-triangle bounds/centroid: pure functions of the triangle,  */
} build_ctx;                            /* hoisted out of the 15 re-scans per node (BVHTree.hpp:294-361)  */

static int push_node(build_ctx* c)
{
    if (c->n_nodes == c->cap_nodes) {
        c->cap_nodes = c->cap_nodes ? c->cap_nodes * 2 : 1024;
        c->nodes = (o_node*)realloc(c->nodes, sizeof(o_node) * (size_t)c->cap_nodes);
    }
    o_node* n = &c->nodes[c->n_nodes];
    n->min = f3(FLT_MAX, FLT_MAX, FLT_MAX);              /* BVHTree.hpp:76-84 */
    n->max = f3(-FLT_MAX, -FLT_MAX, -FLT_MAX);
    n->child_a = n->child_b = -1;
    n->tri_offset = n->tri_count = 0;
    return c->n_nodes++;
}

typedef struct { o_f3 mn, mx; size_t cnt; } sbox;
static inline void sbox_init(sbox* b)
{
    b->mn = f3(FLT_MAX, FLT_MAX, FLT_MAX); b->mx = f3(-FLT_MAX, -FLT_MAX, -FLT_MAX); b->cnt = 0;
}
/* BVHTree.hpp:175-190 */
static inline void sbox_grow(sbox* b, o_f3 mn, o_f3 mx)
{
    b->mn.x = fminf(b->mn.x, mn.x); b->mn.y = fminf(b->mn.y, mn.y); b->mn.z = fminf(b->mn.z, mn.z);
    b->mx.x = fmaxf(b->mx.x, mx.x); b->mx.y = fmaxf(b->mx.y, mx.y); b->mx.z = fmaxf(b->mx.z, mx.z);
    b->cnt++;
}
/* BVHTree.hpp:192-201 */
static inline float box_cost(o_f3 mn, o_f3 mx, size_t cnt)
{
    if (cnt == 0) return FLT_MAX;
    o_f3 size = sub3(mx, mn);
    float half_area = size.x * (size.y + size.z) + size.y * size.z;
    return half_area * cnt;
}

static inline float axis_of(o_f3 v, int a) { return a == 0 ? v.x : (a == 1 ? v.y : v.z); }

/* BVHTree.hpp:203-292 (fill) with evaluate_split (294-361) folded into one scan per node:
 * the 15 candidate (axis, plane) boxes only take min/max/count, so scan order does not matter. */
static void fill(build_ctx* c, int node, int* idx, int n, int depth, int max_depth)
{
    o_f3 mn = f3(FLT_MAX, FLT_MAX, FLT_MAX), mx = f3(-FLT_MAX, -FLT_MAX, -FLT_MAX);
    for (int i = 0; i < n; i++) {
        o_f3 a = c->tmin[idx[i]], b = c->tmax[idx[i]];
        mn.x = fminf(mn.x, a.x); mn.y = fminf(mn.y, a.y); mn.z = fminf(mn.z, a.z);
        mx.x = fmaxf(mx.x, b.x); mx.y = fmaxf(mx.y, b.y); mx.z = fmaxf(mx.z, b.z);
    }
    c->nodes[node].min = mn; c->nodes[node].max = mx;

    int make_leaf = 1;
    int *left = NULL, *right = NULL, nl = 0, nr = 0;
    if (depth < max_depth && n > 1) {
        const float tests_per_axis = 5;
        float pos[3][5];
        sbox L[3][5], R[3][5];
        for (int a = 0; a < 3; a++)
            for (int s = 0; s < 5; s++) {
                float split_t = ((float)s + 1) / (tests_per_axis + 1);
                float lo = axis_of(mn, a), hi = axis_of(mx, a);
                pos[a][s] = lo + (hi - lo) * split_t;
                sbox_init(&L[a][s]); sbox_init(&R[a][s]);
            }
        for (int i = 0; i < n; i++) {
            int t = idx[i];
            o_f3 cen = c->tcen[t];
            for (int a = 0; a < 3; a++) {
                float chk = axis_of(cen, a);
                for (int s = 0; s < 5; s++) {
                    if (chk <= pos[a][s]) sbox_grow(&L[a][s], c->tmin[t], c->tmax[t]);
                    else sbox_grow(&R[a][s], c->tmin[t], c->tmax[t]);
                }
            }
        }
        float ev_cost[3], ev_pos[3];
        for (int a = 0; a < 3; a++) {
            float best_cost = FLT_MAX, best_split = 0.0f;
            for (int s = 0; s < 5; s++) {
                float cost = box_cost(L[a][s].mn, L[a][s].mx, L[a][s].cnt) + box_cost(R[a][s].mn, R[a][s].mx, R[a][s].cnt);
                if (cost < best_cost) { best_cost = cost; best_split = pos[a][s]; }
            }
            ev_cost[a] = best_cost; ev_pos[a] = best_split;
        }
        int axis; float split_pos, best_cost;                  /* BVHTree.hpp:225-243: ties fall to z */
        if (ev_cost[0] < ev_cost[1] && ev_cost[0] < ev_cost[2]) { axis = 0; }
        else if (ev_cost[1] < ev_cost[0] && ev_cost[1] < ev_cost[2]) { axis = 1; }
        else { axis = 2; }
        split_pos = ev_pos[axis]; best_cost = ev_cost[axis];

        if (!(best_cost >= box_cost(mn, mx, (size_t)n))) {      /* BVHTree.hpp:246-247 */
            left = (int*)malloc(sizeof(int) * (size_t)n);
            right = (int*)malloc(sizeof(int) * (size_t)n);
            for (int i = 0; i < n; i++) {                       /* BVHTree.hpp:253-277 */
                int t = idx[i];
                float chk = axis_of(c->tcen[t], axis);
                if (chk <= split_pos) left[nl++] = t; else right[nr++] = t;
            }
            if (nl != 0 && nr != 0) make_leaf = 0;              /* BVHTree.hpp:279-280 */
        }
    }
    if (make_leaf) {
        c->nodes[node].tri_offset = c->n_indices;
        c->nodes[node].tri_count = n;
        memcpy(c->indices + c->n_indices, idx, sizeof(int) * (size_t)n);
        c->n_indices += n;
    } else {                                                    /* BVHTree.hpp:283-289: DFS pre-order */
        int a = push_node(c);
        c->nodes[node].child_a = a;
        fill(c, a, left, nl, depth + 1, max_depth);
        int b = push_node(c);
        c->nodes[node].child_b = b;
        fill(c, b, right, nr, depth + 1, max_depth);
    }
    free(left); free(right);
}

/* MeshPrimitive.cpp:5-15 (copy triangles) and 38-56 (build_bvh: fill(1, 32)) */
o_mesh* o_mesh_create(const o_tri* tris, int n)
{
    o_mesh* m = (o_mesh*)calloc(1, sizeof(o_mesh));
    m->num_triangles = n;
    m->triangles = (o_tri*)malloc(sizeof(o_tri) * (size_t)(n > 0 ? n : 1));
    memcpy(m->triangles, tris, sizeof(o_tri) * (size_t)n);
    o_f3* tmin = (o_f3*)malloc(sizeof(o_f3) * (size_t)(n + 1));
    o_f3* tmax = (o_f3*)malloc(sizeof(o_f3) * (size_t)(n + 1));
    o_f3* tcen = (o_f3*)malloc(sizeof(o_f3) * (size_t)(n + 1));
    int* idx = (int*)malloc(sizeof(int) * (size_t)(n + 1));
    for (int i = 0; i < n; i++) {
        const o_tri* t = &tris[i];
        o_f3 mn = f3(FLT_MAX, FLT_MAX, FLT_MAX), mx = f3(-FLT_MAX, -FLT_MAX, -FLT_MAX);
        for (int k = 0; k < 3; k++) {
            o_f3 v = t->vertices[k];
            mn.x = fminf(mn.x, v.x); mn.y = fminf(mn.y, v.y); mn.z = fminf(mn.z, v.z);
            mx.x = fmaxf(mx.x, v.x); mx.y = fmaxf(mx.y, v.y); mx.z = fmaxf(mx.z, v.z);
        }
        tmin[i] = mn; tmax[i] = mx; tcen[i] = tri_center(t); idx[i] = i;
    }
    build_ctx c;
    memset(&c, 0, sizeof c);
    c.tmin = tmin; c.tmax = tmax; c.tcen = tcen;
    c.indices = (int*)malloc(sizeof(int) * (size_t)(n + 1));
    int root = push_node(&c);
    fill(&c, root, idx, n, 1, 32);
    m->nodes = c.nodes; m->num_nodes = c.n_nodes;
    m->indices = c.indices; m->num_indices = c.n_indices;
    free(tmin); free(tmax); free(tcen); free(idx);
    return m;
}

void o_mesh_destroy(o_mesh* m)
{
    if (!m) return;
    free(m->triangles); free(m->nodes); free(m->indices); free(m);
}

void o_free(void* p) { free(p); }

/* BVHTree.hpp:117-172 print_stats (values only).  "depth" is the reference's figure: the largest
 * size of its explicit DFS stack just after a pop. */
void o_mesh_stats(const o_mesh* m, int out[6])
{
    int count_nodes = 0, max_t = 0, min_t = 1000000, max_depth = 0, leaves = 0;
    int* stack = (int*)malloc(sizeof(int) * (size_t)(m->num_nodes + 2));
    int sp = 0;
    stack[sp++] = 0;
    while (sp > 0) {
        int n = stack[--sp];
        if (sp > max_depth) max_depth = sp;
        count_nodes++;
        const o_node* nd = &m->nodes[n];
        if (nd->child_a == -1) {
            if (nd->tri_count > max_t) max_t = nd->tri_count;
            if (nd->tri_count < min_t) min_t = nd->tri_count;
            leaves++;
        } else {
            stack[sp++] = nd->child_a;
            stack[sp++] = nd->child_b;
        }
    }
    free(stack);
    out[0] = count_nodes; out[1] = max_t; out[2] = min_t; out[3] = max_depth; out[4] = leaves; out[5] = m->num_triangles;
}

/* ------------------------------------------------------------------ OBJLoader.hpp */

typedef struct { char** tok; int n; } toks;
static int split_ws(char* line, char** tok, int max)
{
    int n = 0;
    char* p = line;
    while (*p) {
        while (*p && isspace((unsigned char)*p)) p++;
        if (!*p) break;
        if (n < max) tok[n++] = p; else break;
        while (*p && !isspace((unsigned char)*p)) p++;
        if (*p) *p++ = 0;
    }
    return n;
}

/* OBJLoader.hpp:15-179.  Grammar: `v x y z`, `vt u v`, (`vn` parsed and ignored), `f` with tokens
 * `v`, `v/vt`, `v/vt/vn`; fan triangulation 0,i,i+1 (140-168); face normal normalize(cross()).
 * stof/stoi -> strtof/strtol (same rounding).  Returns triangle count, -1 if the file cannot be opened
 * (the reference prints and exit(1)s, OBJLoader.hpp:23-27). */
int o_obj_load(const char* path, o_tri** out)
{
    FILE* f = fopen(path, "r");
    if (!f) return -1;
    size_t cap_v = 1024, nv = 0, cap_t = 1024, nt = 0, cap_tri = 1024, ntri = 0;
    o_f3* V = (o_f3*)malloc(sizeof(o_f3) * cap_v);
    o_f2* T = (o_f2*)malloc(sizeof(o_f2) * cap_t);
    o_tri* tris = (o_tri*)malloc(sizeof(o_tri) * cap_tri);
    char* line = NULL; size_t lcap = 0;
    char* tok[256];
    while (getline(&line, &lcap, f) > 0) {                       /* pass 1: OBJLoader.hpp:36-71 */
        int n = split_ws(line, tok, 256);
        if (n == 0) continue;
        if (!strcmp(tok[0], "v") && n >= 4) {
            if (nv == cap_v) { cap_v *= 2; V = (o_f3*)realloc(V, sizeof(o_f3) * cap_v); }
            V[nv++] = f3(strtof(tok[1], NULL), strtof(tok[2], NULL), strtof(tok[3], NULL));
        } else if (!strcmp(tok[0], "vt") && n >= 3) {
            if (nt == cap_t) { cap_t *= 2; T = (o_f2*)realloc(T, sizeof(o_f2) * cap_t); }
            T[nt].x = strtof(tok[1], NULL); T[nt].y = strtof(tok[2], NULL); nt++;
        }
    }
    rewind(f);
    while (getline(&line, &lcap, f) > 0) {                       /* pass 2: OBJLoader.hpp:80-171 */
        int n = split_ws(line, tok, 256);
        if (n == 0 || strcmp(tok[0], "f")) continue;
        int vi[256], ti[256], nvi = 0, nti = 0;
        for (int i = 1; i < n; i++) {
            char* s1 = strchr(tok[i], '/');
            vi[nvi++] = (int)strtol(tok[i], NULL, 10) - 1;
            if (s1) ti[nti++] = (int)strtol(s1 + 1, NULL, 10) - 1;
        }
        for (int i = 1; i + 1 < nvi; i++) {
            o_f3 a = V[vi[0]], b = V[vi[i]], c = V[vi[i + 1]];
            o_tri t = o_make_triangle(a, b, c);
            if (nti > 0) { t.uv_coords[0] = T[ti[0]]; t.uv_coords[1] = T[ti[i]]; t.uv_coords[2] = T[ti[i + 1]]; }
            if (ntri == cap_tri) { cap_tri *= 2; tris = (o_tri*)realloc(tris, sizeof(o_tri) * cap_tri); }
            tris[ntri++] = t;
        }
    }
    free(line); free(V); free(T); fclose(f);
    *out = tris;
    return (int)ntri;
}

/* ------------------------------------------------------------------ raycast.cu */

typedef struct {                 /* Ray.hpp:5-23 (fields the path reads) */
    o_f3 origin, direction, direction_inv;
} ray_t;

static inline ray_t make_ray(o_f3 o, o_f3 d)
{
    ray_t r; r.origin = o; r.direction = d;
    r.direction_inv = f3(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);   /* Ray.hpp:20 */
    return r;
}

typedef struct {                 /* raycast.cu:10-18 HitInfo + the additive AOVs of SURVEY 8c patch P3 */
    float min;
    o_f3 location, normal;
    o_f2 uv;
    int prim, inst, material;
} hit_t;

/* BVHTree.hpp:40-54 */
static inline float node_ray_intersects(const o_node* n, const ray_t* ray)
{
    o_f3 tmin = f3((n->min.x - ray->origin.x) * ray->direction_inv.x, (n->min.y - ray->origin.y) * ray->direction_inv.y,
                   (n->min.z - ray->origin.z) * ray->direction_inv.z);
    o_f3 tmax = f3((n->max.x - ray->origin.x) * ray->direction_inv.x, (n->max.y - ray->origin.y) * ray->direction_inv.y,
                   (n->max.z - ray->origin.z) * ray->direction_inv.z);
    o_f3 t1 = f3(fminf(tmin.x, tmax.x), fminf(tmin.y, tmax.y), fminf(tmin.z, tmax.z));
    o_f3 t2 = f3(fmaxf(tmin.x, tmax.x), fmaxf(tmin.y, tmax.y), fmaxf(tmin.z, tmax.z));
    float dst_far = fminf(fminf(t2.x, t2.y), t2.z);
    float dst_near = fmaxf(fmaxf(t1.x, t1.y), t1.z);
    int hit = dst_far >= dst_near && dst_far > 0.0f;
    return hit ? dst_near : FLT_MAX;
}

/* TrianglePrimitive.hpp:62-79.  Returns 0 for the (FLT_MAX,..) sentinel. */
static inline int tri_ray_intersect(const o_tri* t, const ray_t* ray, o_f3* point)
{
    float denom = dot3(ray->direction, t->normal);
    if ((double)fabsf(denom) < 1e-6) return 0;
    float tt = dot3(sub3(t->vertices[0], ray->origin), t->normal) / denom;
    if (tt < 0.0f) return 0;
    *point = add3(ray->origin, scale3(ray->direction, tt));
    return 1;
}

/* TrianglePrimitive.hpp:151-185.  Returns 0 for the (FLT_MAX, FLT_MAX) sentinel. */
static inline int tri_point_inside(const o_tri* t, o_f3 point, o_f2* uv)
{
    o_f3 v0 = sub3(t->vertices[2], t->vertices[0]);
    o_f3 v1 = sub3(t->vertices[1], t->vertices[0]);
    o_f3 v2 = sub3(point, t->vertices[0]);
    float dot00 = dot3(v0, v0), dot01 = dot3(v0, v1), dot02 = dot3(v0, v2), dot11 = dot3(v1, v1), dot12 = dot3(v1, v2);
    float invDenom = 1.0f / (dot00 * dot11 - dot01 * dot01);
    float u = (dot11 * dot02 - dot01 * dot12) * invDenom;
    float v = (dot00 * dot12 - dot01 * dot02) * invDenom;
    if ((u >= 0.0f) && (v >= 0.0f) && (u + v <= 1.0f)) {
        float w = 1.0f - u - v;
        o_f2 a = t->uv_coords[0], b = t->uv_coords[1], c = t->uv_coords[2];
        uv->x = (w * a.x) + (v * b.x) + (u * c.x);
        uv->y = (w * a.y) + (v * b.y) + (u * c.y);
        return 1;
    }
    return 0;
}

typedef struct { uint64_t nodes, boxes, tris; } ctr_t;

/* raycast.cu:21-142 cast_ray.  `any_hit` is the commented-out lighting_pass early return
 * (raycast.cu:128-132): the shadow query only asks whether min stayed FLT_MAX. */
static void cast_ray_exhaustive_hit(const o_scene* sc, o_f3 origin, o_f3 direction, hit_t* hit, int any_hit);
static int g_exhaustive = 0;     /* test support, see o_set_exhaustive below */

static void cast_ray(const o_scene* sc, o_f3 origin, o_f3 direction, hit_t* hit, int any_hit, ctr_t* ctr)
{
    hit->min = FLT_MAX; hit->prim = -1; hit->inst = -1; hit->material = -1;
    hit->location = f3(0, 0, 0); hit->normal = f3(0, 0, 0); hit->uv.x = hit->uv.y = 0;
    if (g_exhaustive) { cast_ray_exhaustive_hit(sc, origin, direction, hit, any_hit); return; }
    for (int mesh_idx = 0; mesh_idx < sc->num_instances; mesh_idx++) {
        const o_instance* mi = &sc->instances[mesh_idx];
        const o_mesh* mesh = sc->meshes[mi->mesh_index];
        o_f3 r_direction = o_apply_euler(mi->rotation, direction);             /* raycast.cu:33-37 */
        r_direction.x *= mi->inv_scale.x; r_direction.y *= mi->inv_scale.y; r_direction.z *= mi->inv_scale.z;
        o_f3 r_origin = o_apply_lre(mi->pose, origin);                          /* raycast.cu:40-45 */
        r_origin.x *= mi->inv_scale.x; r_origin.y *= mi->inv_scale.y; r_origin.z *= mi->inv_scale.z;
        ray_t r_ray = make_ray(r_origin, r_direction);

        int stack[64];
        int sp = 0;
        stack[sp++] = 0;
        while (sp > 0) {                                                         /* raycast.cu:60-138 */
            int node_index = stack[--sp];
            const o_node* cur = &mesh->nodes[node_index];
            if (ctr) ctr->nodes++;
            if (cur->child_a > 0) {
                float dist_a = node_ray_intersects(&mesh->nodes[cur->child_a], &r_ray);
                float dist_b = node_ray_intersects(&mesh->nodes[cur->child_b], &r_ray);
                if (ctr) ctr->boxes += 2;
                if (dist_a < dist_b) {
                    if (dist_b < hit->min) stack[sp++] = cur->child_b;
                    if (dist_a < hit->min) stack[sp++] = cur->child_a;
                } else {
                    if (dist_a < hit->min) stack[sp++] = cur->child_a;
                    if (dist_b < hit->min) stack[sp++] = cur->child_b;
                }
            } else {
                for (int i = 0; i < cur->tri_count; i++) {
                    int index = mesh->indices[cur->tri_offset + i];
                    const o_tri* tri = &mesh->triangles[index];
                    if (ctr) ctr->tris++;
                    o_f3 isect;
                    if (!tri_ray_intersect(tri, &r_ray, &isect)) continue;
                    o_f2 uv;
                    if (!tri_point_inside(tri, isect, &uv)) continue;
                    o_f3 loc = f3(isect.x * mi->scale.x, isect.y * mi->scale.y, isect.z * mi->scale.z);
                    loc = o_apply_lre(mi->inv_pose, loc);                        /* raycast.cu:98-102 */
                    float distance = magnitude3(sub3(loc, origin));              /* raycast.cu:104 */
                    float same_dir = dot3(r_ray.direction, tri->normal);         /* raycast.cu:107 */
                    if (same_dir < 0 && (hit->min == FLT_MAX || distance < hit->min)) {
                        hit->min = distance;
                        o_f3 nrm = o_apply_euler(mi->inv_rotation, tri->normal); /* raycast.cu:115-122 */
                        nrm.x *= mi->scale.x; nrm.y *= mi->scale.y; nrm.z *= mi->scale.z;
                        hit->normal = o_normalize(nrm);
                        hit->uv = uv;
                        hit->material = mi->material_index;
                        hit->location = loc;     /* location of the ACCEPTED hit (the reference's field also
                                                    takes rejected candidates, Q6; nothing shipped reads it) */
                        hit->prim = index; hit->inst = mesh_idx;
                        if (any_hit) return;
                    }
                }
            }
        }
    }
}

int o_cast_ray(const o_scene* sc, o_f3 origin, o_f3 direction, int* prim, int* inst, float* t,
               o_f3* location, o_f3* normal, o_f2* uv)
{
    hit_t h;
    cast_ray(sc, origin, direction, &h, 0, NULL);
    if (prim) *prim = h.prim;
    if (inst) *inst = h.inst;
    if (t) *t = h.min;
    if (location) *location = h.location;
    if (normal) *normal = h.normal;
    if (uv) *uv = h.uv;
    return h.min != FLT_MAX;
}

/* ---- test support for the parity report (no counterpart in the reference): why does a pixel differ? ----
 * o_candidate: the accept predicate of raycast.cu:87-109 for ONE (instance, triangle) pair -- 1 and the world distance when the
 * candidate passes ray_intersect + point_inside + the front-face rule, else 0.
 * o_cast_ray_exhaustive: the closest candidate over ALL instances and triangles with no BVH at all, i.e. what cast_ray returns
 * when no box test ever culls; ties go to the lowest (instance, triangle), the product's documented tie-break. */
static int candidate(const o_instance* mi, const o_mesh* mesh, const ray_t* r_ray, o_f3 origin, int index, float* distance)
{
    const o_tri* tri = &mesh->triangles[index];
    o_f3 isect;
    o_f2 uv;
    if (!tri_ray_intersect(tri, r_ray, &isect)) return 0;
    if (isect.x == FLT_MAX) return 0;
    if (!tri_point_inside(tri, isect, &uv)) return 0;
    o_f3 loc = f3(isect.x * mi->scale.x, isect.y * mi->scale.y, isect.z * mi->scale.z);
    loc = o_apply_lre(mi->inv_pose, loc);
    *distance = magnitude3(sub3(loc, origin));
    return dot3(r_ray->direction, tri->normal) < 0;
}

static ray_t instance_ray(const o_instance* mi, o_f3 origin, o_f3 direction)
{
    o_f3 r_direction = o_apply_euler(mi->rotation, direction);
    r_direction.x *= mi->inv_scale.x; r_direction.y *= mi->inv_scale.y; r_direction.z *= mi->inv_scale.z;
    o_f3 r_origin = o_apply_lre(mi->pose, origin);
    r_origin.x *= mi->inv_scale.x; r_origin.y *= mi->inv_scale.y; r_origin.z *= mi->inv_scale.z;
    return make_ray(r_origin, r_direction);
}

int o_candidate(const o_scene* sc, o_f3 origin, o_f3 direction, int inst, int prim, float* distance)
{
    if (inst < 0 || inst >= sc->num_instances) return 0;
    const o_instance* mi = &sc->instances[inst];
    const o_mesh* mesh = sc->meshes[mi->mesh_index];
    if (prim < 0 || prim >= mesh->num_triangles) return 0;
    ray_t r = instance_ray(mi, origin, direction);
    return candidate(mi, mesh, &r, origin, prim, distance);
}

/* o_set_exhaustive(1): every cast_ray of o_render / o_render_pixel becomes the BVH-free search below (all triangles of all
 * instances, ties to the lowest (instance, triangle)): "what the reference's accept rule yields when no box test culls".  The
 * parity suite uses it to show that a pixel on which product and reference differ is one where the reference's own culling
 *  or its traversal-order tie-break (H5) decided -- never to produce expected values. */
void o_set_exhaustive(int on) { g_exhaustive = on; }

static void cast_ray_exhaustive_hit(const o_scene* sc, o_f3 origin, o_f3 direction, hit_t* hit, int any_hit)
{
    for (int m = 0; m < sc->num_instances; m++) {
        const o_instance* mi = &sc->instances[m];
        const o_mesh* mesh = sc->meshes[mi->mesh_index];
        ray_t r = instance_ray(mi, origin, direction);
        for (int i = 0; i < mesh->num_triangles; i++) {
            const o_tri* tri = &mesh->triangles[i];
            o_f3 isect; o_f2 uv;
            if (!tri_ray_intersect(tri, &r, &isect) || !tri_point_inside(tri, isect, &uv)) continue;
            if (!(dot3(r.direction, tri->normal) < 0)) continue;
            o_f3 loc = f3(isect.x * mi->scale.x, isect.y * mi->scale.y, isect.z * mi->scale.z);
            loc = o_apply_lre(mi->inv_pose, loc);
            float distance = magnitude3(sub3(loc, origin));
            if (hit->min == FLT_MAX || distance < hit->min) {          /* m, i ascend: a later equal distance never replaces */
                hit->min = distance;
                o_f3 nrm = o_apply_euler(mi->inv_rotation, tri->normal);
                nrm.x *= mi->scale.x; nrm.y *= mi->scale.y; nrm.z *= mi->scale.z;
                hit->normal = o_normalize(nrm);
                hit->uv = uv; hit->material = mi->material_index; hit->location = loc;
                hit->prim = i; hit->inst = m;
                if (any_hit) return;
            }
        }
    }
}

int o_cast_ray_exhaustive(const o_scene* sc, o_f3 origin, o_f3 direction, int* prim, int* inst, float* t)
{
    float best = FLT_MAX; int bp = -1, bi = -1;
    for (int m = 0; m < sc->num_instances; m++) {
        const o_instance* mi = &sc->instances[m];
        const o_mesh* mesh = sc->meshes[mi->mesh_index];
        ray_t r = instance_ray(mi, origin, direction);
        for (int i = 0; i < mesh->num_triangles; i++) {
            float d;
            if (candidate(mi, mesh, &r, origin, i, &d) && (best == FLT_MAX || d < best)) { best = d; bp = i; bi = m; }
        }
    }
    *prim = bp; *inst = bi; *t = best;
    return bp >= 0;
}

/* raycast.cu:156-188: pixel -> world-space ray direction.  (px,py) are the float pixel coordinates
 * (integers in the reference; the path-tracing extension adds a sub-pixel jitter). */
static o_f3 generate_direction(const o_render_params* p, float px, float py)
{
    const float* K = p->K_inv;
    o_f3 ph = f3(px, py, 1.0f);
    o_f3 d = f3(K[0] * ph.x + K[1] * ph.y + K[2] * ph.z,                         /* utils.hpp:134-140 */
                K[3] * ph.x + K[4] * ph.y + K[5] * ph.z,
                K[6] * ph.x + K[7] * ph.y + K[8] * ph.z);
    float a = d.x, b = d.y;
    float radius = sqrtf(a * a + b * b);
    float theta = atanf(radius);
    /* raycast.cu:172: float products, double sum because of the `1.0` literal, double product, narrowed */
    float thetad = (float)(theta * (1.0 + p->D[0] * theta + p->D[1] * theta * theta + p->D[2] * theta * theta * theta
                                    + p->D[3] * theta * theta * theta * theta));
    float scale = thetad / radius;
    d.x = scale * a;
    d.y = scale * b;
    d = o_normalize(d);
    d = f3(d.x, d.z, -d.y);                                                      /* raycast.cu:182 */
    d = o_apply_euler(f3(p->inv_pose.yaw, p->inv_pose.pitch, p->inv_pose.roll), d);
    d = o_normalize(d);
    return d;
}

/* the camera ray of pixel (x, y) as o_render casts it in the non-jittered modes (raycast.cu:156-188) */
void o_pixel_ray(const o_render_params* p, int x, int y, o_f3* origin, o_f3* direction)
{
    *origin = f3(p->pose.x, p->pose.y, p->pose.z);
    *direction = generate_direction(p, (float)x, (float)y);
}

/* raycast.cu:224-245: base colour of the hit (BGR floats) */
static o_f3 base_colour(const o_scene* sc, const hit_t* h)
{
    const o_material* m = &sc->materials[h->material];
    o_f3 c = f3(1.0f, 1.0f, 1.0f);                                               /* Ray.hpp:21 */
    if (m->texture_width > 0) {
        int tex_x = (int)(h->uv.x * (float)m->texture_width);
        int tex_y = (int)((1.0 - h->uv.y) * (float)m->texture_height);           /* double product */
        tex_x = (int)fmaxf((float)(tex_x % m->texture_width), 0);
        tex_y = (int)fmaxf((float)(tex_y % m->texture_height), 0);
        const uint8_t* row = m->texture + (size_t)tex_y * m->texture_pitch;
        const uint8_t* tc = row + 3 * (size_t)tex_x;
        c.x *= tc[0] * 0.0039215f; c.y *= tc[1] * 0.0039215f; c.z *= tc[2] * 0.0039215f;
    } else {
        c.x *= m->albedo.x; c.y *= m->albedo.y; c.z *= m->albedo.z;
    }
    return c;
}

/* raycast.cu:249-290.  Light direction and 1e-4 offset from the live code (249-259); the Lambert /
 * shadow-ray body is the commented block 262-280 when `lit` is set, else illumination = 1.0 (282). */
static float illumination(const o_scene* sc, const hit_t* h, int lit, ctr_t* ctr, uint64_t* rays)
{
    float illum;
    if (lit) {
        o_f3 light_direction = o_normalize(f3((float)-0.2, 0.0f, 1.0f));
        o_f3 org = add3(h->location, scale3(light_direction, (float)1e-4));      /* raycast.cu:259 */
        float cos_illum = dot3(h->normal, light_direction);
        illum = (float)(0.4 * cos_illum);
        if (dot3(h->normal, light_direction) > 0) {
            hit_t sh;
            cast_ray(sc, org, light_direction, &sh, 1, ctr);
            if (rays) (*rays)++;
            if (sh.min == FLT_MAX) illum = (float)(1.0 * cos_illum);
        }
    } else {
        illum = 1.0f;
    }
    illum = fminf(1.0f, illum);                                                  /* raycast.cu:289-290 */
    illum = fmaxf(0.4f, illum);
    return illum;
}

static inline uint8_t to_u8(float v) { return (uint8_t)(int)v; }                 /* raycast.cu:292-294 truncation */
static inline float clamp01(float v) { return fminf(1.0f, fmaxf(0.0f, v)); }

/* Philox4x32-10 (Salmon et al. 2011), counter-based so CPU and GPU draw identical numbers. */
static void philox4x32(uint32_t c[4], uint32_t k0, uint32_t k1)
{
    for (int r = 0; r < 10; r++) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c[0], p1 = (uint64_t)0xCD9E8D57u * c[2];
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c[1] ^ k0, n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c[3] ^ k1, n3 = (uint32_t)p0;
        c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
}
static inline float u01(uint32_t x) { return (float)(x >> 8) * (1.0f / 16777216.0f); }

/* mirror direction r = d - 2 (d.n) n, renormalised the reference's way */
static inline o_f3 reflect_dir(o_f3 d, o_f3 n)
{
    float k = 2.0f * dot3(d, n);
    return o_normalize(sub3(d, scale3(n, k)));
}

/* sin and cos of phi in [0, 2 pi] WITHOUT libm, so that the CPU and the GPU produce the same bits (glibc's and CUDA's sinf/cosf
 * differ in the last ulp on a few percent of arguments, which moved single samples of path-traced pixels): quadrant reduction
 * and the fdlibm kernel polynomials, evaluated in binary64 with plain multiplies and adds in a fixed order (no fma), rounded to
 * binary32 once.  The device restatement (csrc/rtb_wave_common.cuh: sincos_2pi) performs the identical operation sequence. */
static void sincos_2pi(float phi, float* s, float* c)
{
    const double x = (double)phi;
    const int k = (int)(x * 0.63661977236758134308 + 0.5);
    const double y = x - (double)k * 1.57079632679489655800, z = y * y;
    const double sp = y + y * z * (-1.66666666666666324348e-01 + z * (8.33333333332248946124e-03 + z * (-1.98412698298579493134e-04 +
                      z * (2.75573137070700676789e-06 + z * (-2.50507602534068634195e-08 + z * 1.58969099521155010221e-10)))));
    const double cp = 1.0 - 0.5 * z + z * z * (4.16666666666666019037e-02 + z * (-1.38888888888741095749e-03 + z * (2.48015872894767294178e-05 +
                      z * (-2.75573143513906633035e-07 + z * (2.08757232129817482790e-09 + z * -1.13596475577881948265e-11)))));
    const int q = k & 3;
    const double sv = q == 0 ? sp : (q == 1 ? cp : (q == 2 ? -sp : -cp));
    const double cv = q == 0 ? cp : (q == 1 ? -sp : (q == 2 ? -cp : sp));
    *s = (float)sv; *c = (float)cv;
}

/* cosine-weighted hemisphere sample around n (Duff et al. 2017 basis) */
static o_f3 cosine_dir(o_f3 n, float u1, float u2)
{
    float r = sqrtf(u1), phi = 6.2831853f * u2, sphi, cphi;
    sincos_2pi(phi, &sphi, &cphi);
    float lx = r * cphi, ly = r * sphi, lz = sqrtf(fmaxf(0.0f, 1.0f - u1));
    float sign = copysignf(1.0f, n.z);
    float a = -1.0f / (sign + n.z);
    float b = n.x * n.y * a;
    o_f3 t = f3(1.0f + sign * n.x * n.x * a, sign * b, -sign * n.x);
    o_f3 bt = f3(b, sign + n.y * n.y * a, -n.y);
    o_f3 d = add3(add3(scale3(t, lx), scale3(bt, ly)), scale3(n, lz));
    return o_normalize(d);
}

static const float SKY[3] = { 1.0f, 0.8f, 0.6f };                                /* raycast.cu:210-213 (BGR) */

static void render_pixel(const o_scene* sc, const o_render_params* p, int x, int y, uint8_t* px,
                         int32_t* aprim, int32_t* ainst, float* at, ctr_t* ctr, uint64_t* rays)
{
    o_f3 origin = f3(p->pose.x, p->pose.y, p->pose.z);                           /* raycast.cu:156 */
    hit_t h;
    if (p->mode != O_MODE_PATH) {
        o_f3 dir = generate_direction(p, (float)x, (float)y);
        cast_ray(sc, origin, dir, &h, 0, ctr);
        (*rays)++;
        if (aprim) *aprim = h.prim;
        if (ainst) *ainst = h.inst;
        if (at) *at = h.min;
        if (p->mode == O_MODE_REFERENCE || p->mode == O_MODE_SHADOW) {
            if (h.min == FLT_MAX) {                                              /* raycast.cu:208-216 */
                px[0] = to_u8((float)(1.0 * 255)); px[1] = to_u8((float)(0.8 * 255)); px[2] = to_u8((float)(0.6 * 255));
                return;
            }
            o_f3 c = base_colour(sc, &h);
            float il = illumination(sc, &h, p->mode == O_MODE_SHADOW, ctr, rays);
            px[0] = to_u8(il * c.x * 255); px[1] = to_u8(il * c.y * 255); px[2] = to_u8(il * c.z * 255);
            return;
        }
        /* O_MODE_REFLECT: Whitted-style mirror chain, weight = material.metallic (extension). */
        o_f3 thr = f3(1.0f, 1.0f, 1.0f), acc = f3(0.0f, 0.0f, 0.0f);
        for (int depth = 0;; depth++) {
            if (h.min == FLT_MAX) {
                acc.x += thr.x * SKY[0]; acc.y += thr.y * SKY[1]; acc.z += thr.z * SKY[2];
                break;
            }
            o_f3 c = base_colour(sc, &h);
            float il = illumination(sc, &h, 1, ctr, rays);
            float m = sc->materials[h.material].metallic;
            float w = 1.0f - m;
            acc.x += thr.x * (w * (il * c.x)); acc.y += thr.y * (w * (il * c.y)); acc.z += thr.z * (w * (il * c.z));
            if (!(m > 0.0f) || depth >= p->max_depth) break;
            thr.x *= m * c.x; thr.y *= m * c.y; thr.z *= m * c.z;
            dir = reflect_dir(dir, h.normal);
            o_f3 org = add3(h.location, scale3(dir, (float)1e-4));
            cast_ray(sc, org, dir, &h, 0, ctr);
            (*rays)++;
        }
        px[0] = to_u8(clamp01(acc.x) * 255); px[1] = to_u8(clamp01(acc.y) * 255); px[2] = to_u8(clamp01(acc.z) * 255);
        return;
    }
    /* O_MODE_PATH: diffuse path tracing, sky is the only emitter (extension). */
    o_f3 sum = f3(0.0f, 0.0f, 0.0f);
    uint32_t pix = (uint32_t)(y * p->width + x);
    for (int s = 0; s < p->spp; s++) {
        uint32_t c4[4] = { pix, (uint32_t)s, 0u, 0u };
        philox4x32(c4, p->seed, 0x5eed5eedu);
        o_f3 dir = generate_direction(p, (float)x + (u01(c4[0]) - 0.5f), (float)y + (u01(c4[1]) - 0.5f));
        o_f3 org = origin;
        o_f3 thr = f3(1.0f, 1.0f, 1.0f);
        for (int bounce = 0; bounce <= p->max_depth; bounce++) {
            cast_ray(sc, org, dir, &h, 0, ctr);
            (*rays)++;
            if (s == 0 && bounce == 0) {
                /* first-hit AOVs come from the un-jittered centre ray for comparability */
            }
            if (h.min == FLT_MAX) {
                sum.x += thr.x * SKY[0]; sum.y += thr.y * SKY[1]; sum.z += thr.z * SKY[2];
                break;
            }
            o_f3 c = base_colour(sc, &h);
            thr.x *= c.x; thr.y *= c.y; thr.z *= c.z;
            uint32_t r4[4] = { pix, (uint32_t)s, (uint32_t)(bounce + 1), 0u };
            philox4x32(r4, p->seed, 0x5eed5eedu);
            dir = cosine_dir(h.normal, u01(r4[0]), u01(r4[1]));
            org = add3(h.location, scale3(dir, (float)1e-4));
        }
    }
    if (aprim || ainst || at) {
        o_f3 dir = generate_direction(p, (float)x, (float)y);
        cast_ray(sc, origin, dir, &h, 0, NULL);
        if (aprim) *aprim = h.prim;
        if (ainst) *ainst = h.inst;
        if (at) *at = h.min;
    }
    float inv = 1.0f / (float)p->spp;
    px[0] = to_u8(clamp01(sum.x * inv) * 255); px[1] = to_u8(clamp01(sum.y * inv) * 255); px[2] = to_u8(clamp01(sum.z * inv) * 255);
}

/* raycast.cu:146-154 + Camera.cu:23-36: one "thread" per pixel; rows handed out to host threads. */
typedef struct {
    const o_scene* sc; const o_render_params* p; uint8_t* img; size_t pitch;
    int32_t *aov_prim, *aov_inst; float* aov_t; int want_stats;
    int y1; int next_row;           /* next_row: atomic row counter */
    uint64_t rays, nodes, boxes, tris; pthread_mutex_t mu;
} job_t;

/* one pixel of o_render (test support: re-render single pixels, e.g. with o_set_exhaustive(1)) */
void o_render_pixel(const o_scene* sc, const o_render_params* p, int x, int y, uint8_t bgr[3], int32_t* prim, int32_t* inst, float* t)
{
    uint64_t rays = 0;
    render_pixel(sc, p, x, y, bgr, prim, inst, t, NULL, &rays);
}

static void* render_worker(void* arg)
{
    job_t* j = (job_t*)arg;
    ctr_t ctr = { 0, 0, 0 };
    uint64_t rays = 0;
    const o_render_params* p = j->p;
    for (;;) {
        int y = __atomic_fetch_add(&j->next_row, 1, __ATOMIC_RELAXED);
        if (y >= j->y1) break;
        uint8_t* row = j->img + (size_t)y * j->pitch;
        for (int x = 0; x < p->width; x++) {
            size_t i = (size_t)y * (size_t)p->width + (size_t)x;
            render_pixel(j->sc, p, x, y, row + 3 * (size_t)x, j->aov_prim ? j->aov_prim + i : NULL,
                         j->aov_inst ? j->aov_inst + i : NULL, j->aov_t ? j->aov_t + i : NULL,
                         j->want_stats ? &ctr : NULL, &rays);
        }
    }
    pthread_mutex_lock(&j->mu);
    j->rays += rays; j->nodes += ctr.nodes; j->boxes += ctr.boxes; j->tris += ctr.tris;
    pthread_mutex_unlock(&j->mu);
    return NULL;
}

void o_render(const o_scene* sc, const o_render_params* p, uint8_t* img, size_t pitch,
              int32_t* aov_prim, int32_t* aov_inst, float* aov_t, int nthreads, o_stats* stats)
{
    job_t j;
    memset(&j, 0, sizeof j);
    j.sc = sc; j.p = p; j.img = img; j.pitch = pitch;
    j.aov_prim = aov_prim; j.aov_inst = aov_inst; j.aov_t = aov_t; j.want_stats = stats != NULL;
    j.next_row = p->y0; j.y1 = p->y1 > 0 ? p->y1 : p->height;
    pthread_mutex_init(&j.mu, NULL);
    if (nthreads < 1) nthreads = 1;
    if (nthreads > 256) nthreads = 256;
    pthread_t th[256];
    for (int i = 1; i < nthreads; i++) pthread_create(&th[i], NULL, render_worker, &j);
    render_worker(&j);
    for (int i = 1; i < nthreads; i++) pthread_join(th[i], NULL);
    pthread_mutex_destroy(&j.mu);
    if (stats) { stats->rays = j.rays; stats->nodes_popped = j.nodes; stats->boxes_tested = j.boxes; stats->tris_tested = j.tris; }
}
