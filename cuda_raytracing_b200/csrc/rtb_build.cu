// rtb_build.cu -- on-device LBVH builder for sm_100a.
//
// Replaces the reference's single-threaded recursive host builder (BVHTree::fill, BVHTree.hpp:203-292,
// 15 re-scans per node) and its one-cudaMalloc-per-leaf upload (BVHTree.hpp:97-114,364-383):
//   prim boxes -> scene bounds -> Morton keys -> radix sort -> radix tree + AABB fit + SAH leaf decision -> collapse to 112-byte
//   BVH4 nodes -> triangle records.
// Meshes of >= FUSED_MIN_PRIMS primitives take the fused form: k_fit_collapse does tree, fit and collapse of everything below
// the 256-leaf chunks of the sorted order in shared memory, k_fit_top_fused + k_collapse finish the few nodes above them.
// Smaller meshes and the TLAS take the one-pass form: k_hierarchy (Karras 2012), k_fit + k_fit_top, k_collapse over the whole tree.
// The tree only culls (rtb_trace.cuh keeps the reference's hit predicate), so it is free to differ from the reference's tree.
// Streaming stages run near the HBM peak; the tree stages are latency chains (DESIGN.md 4 has the launch list).
#include "rtb_build.cuh"
#include "rtb_sort.cuh"
#include <cstdio>

namespace rtb {

// ------------------------------------------------------------------ arena
cudaError_t Arena::reserve(size_t bytes)
{
    if (bytes <= cap) return cudaSuccess;
    if (base) cudaFree(base);
    base = nullptr; cap = 0;
    cudaError_t e = cudaMalloc((void**)&base, bytes);
    if (e == cudaSuccess) cap = bytes;
    return e;
}
void Arena::release()
{
    if (base) cudaFree(base);
    base = nullptr; cap = used = 0;
    if (side) cudaStreamDestroy(side);
    if (ev_fork) cudaEventDestroy(ev_fork);
    if (ev_join) cudaEventDestroy(ev_join);
    side = nullptr; ev_fork = ev_join = nullptr;
}

// ------------------------------------------------------------------ helpers
__device__ __forceinline__ int f2ord(float f) { int i = __float_as_int(f); return i >= 0 ? i : i ^ 0x7fffffff; }
__device__ __forceinline__ float ord2f(int i) { return __int_as_float(i >= 0 ? i : i ^ 0x7fffffff); }

struct BuildCounters {
    int bounds[6];          // ordered-int encoded scene bounds (min xyz, max xyz)
    int cbounds[6];         // centroid bounds
    int n_wide;             // wide nodes allocated
    int n_leaves, max_leaf, max_level;
    float sah;
    int qcount[72];         // collapse work-queue sizes per level
    unsigned bar;           // grid barrier of the collapse kernel
    int top_count;          // nodes handed from k_fit to k_fit_top
    int n_local;            // wide nodes emitted by k_fit_collapse (they occupy indices 1 .. n_local; the top tree follows)
    int max_level2;         // deepest level reached below a pre-collapsed subtree (k_collapse, fused build)
};

// ------------------------------------------------------------------ 1. primitive boxes + scene bounds
__global__ void k_init_counters(BuildCounters* c)
{
    int t = threadIdx.x;
    if (t < 3) { c->bounds[t] = f2ord(FLT_MAX); c->bounds[t + 3] = f2ord(-FLT_MAX); c->cbounds[t] = f2ord(FLT_MAX); c->cbounds[t + 3] = f2ord(-FLT_MAX); }
    if (t == 0) { c->n_wide = 1; c->n_leaves = 0; c->max_leaf = 0; c->max_level = 0; c->sah = 0.f; }
    if (t < 72) c->qcount[t] = 0;
    if (t == 72) { c->bar = 0u; c->top_count = 0; c->n_local = 0; c->max_level2 = 0; }
}

// Scene and centroid bounds of n boxes: BB_PER boxes per thread (all loads issued before the first use), reduced per warp
// by shuffles, per block through shared memory, one atomic set per block.
// TRIS: the boxes are not read but MADE here, from the reference's 72-byte AoS triangles (TrianglePrimitive.hpp:8-11; bounds as
// BVHTree::grow_to_include, 175-190), and written out -- one pass over the triangles instead of a box pass plus a second
// pass that re-reads 32 bytes per primitive just to reduce them.
constexpr int BB_PER = 4;
template <bool TRIS>
__global__ void __launch_bounds__(256) k_scene_bounds(const rtb_triangle* __restrict__ tris, float4* __restrict__ bmin, float4* __restrict__ bmax,
                                                      int n, BuildCounters* c)
{
    __shared__ float red[8][12];
    float acc[12];                                  // lo xyz, hi xyz, centroid lo xyz, centroid hi xyz
#pragma unroll
    for (int k = 0; k < 3; k++) { acc[k] = FLT_MAX; acc[3 + k] = -FLT_MAX; acc[6 + k] = FLT_MAX; acc[9 + k] = -FLT_MAX; }
    const int base = blockIdx.x * (256 * BB_PER) + threadIdx.x;
    float4 a[BB_PER], b[BB_PER];
    if (TRIS) {
        float v[BB_PER][9];
#pragma unroll
        for (int j = 0; j < BB_PER; j++) {
            const int i = base + j * 256;
            if (i < n) {
                const float* t = (const float*)(tris + i);
#pragma unroll
                for (int k = 0; k < 9; k++) v[j][k] = __ldg(t + k);
            }
        }
#pragma unroll
        for (int j = 0; j < BB_PER; j++) {
            const int i = base + j * 256;
            if (i < n) {
                a[j] = make_float4(fminf(fminf(v[j][0], v[j][3]), v[j][6]), fminf(fminf(v[j][1], v[j][4]), v[j][7]), fminf(fminf(v[j][2], v[j][5]), v[j][8]), 0.f);
                b[j] = make_float4(fmaxf(fmaxf(v[j][0], v[j][3]), v[j][6]), fmaxf(fmaxf(v[j][1], v[j][4]), v[j][7]), fmaxf(fmaxf(v[j][2], v[j][5]), v[j][8]), 0.f);
                bmin[i] = a[j]; bmax[i] = b[j];
            }
        }
    } else {
#pragma unroll
        for (int j = 0; j < BB_PER; j++) { const int i = base + j * 256; if (i < n) { a[j] = bmin[i]; b[j] = bmax[i]; } }
    }
#pragma unroll
    for (int j = 0; j < BB_PER; j++) {
        if (base + j * 256 < n) {
            const float av[3] = { a[j].x, a[j].y, a[j].z }, bv[3] = { b[j].x, b[j].y, b[j].z };
#pragma unroll
            for (int k = 0; k < 3; k++) {
                const float cc = 0.5f * (av[k] + bv[k]);
                acc[k] = fminf(acc[k], av[k]); acc[3 + k] = fmaxf(acc[3 + k], bv[k]);
                acc[6 + k] = fminf(acc[6 + k], cc); acc[9 + k] = fmaxf(acc[9 + k], cc);
            }
        }
    }
#pragma unroll
    for (int k = 0; k < 12; k++) {
        const bool is_min = (k < 3) || (k >= 6 && k < 9);
        for (int o = 16; o > 0; o >>= 1) {
            const float other = __shfl_xor_sync(0xffffffffu, acc[k], o);
            acc[k] = is_min ? fminf(acc[k], other) : fmaxf(acc[k], other);
        }
        if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5][k] = acc[k];
    }
    __syncthreads();
    if (threadIdx.x < 12) {
        const int k = threadIdx.x;
        const bool is_min = (k < 3) || (k >= 6 && k < 9);
        float r = red[0][k];
        for (int w = 1; w < 8; w++) r = is_min ? fminf(r, red[w][k]) : fmaxf(r, red[w][k]);
        int* dst = k < 6 ? &c->bounds[k] : &c->cbounds[k - 6];
        if (is_min) atomicMin(dst, f2ord(r)); else atomicMax(dst, f2ord(r));
    }
}

// ------------------------------------------------------------------ 2. Morton keys
__device__ __forceinline__ unsigned expand10(unsigned v)
{
    v = (v * 0x00010001u) & 0xFF0000FFu; v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u; v = (v * 0x00000005u) & 0x49249249u;
    return v;
}
__device__ __forceinline__ unsigned long long expand21(unsigned long long v)
{
    v &= 0x1fffffull;
    v = (v | v << 32) & 0x1f00000000ffffull; v = (v | v << 16) & 0x1f0000ff0000ffull;
    v = (v | v << 8) & 0x100f00f00f00f00full; v = (v | v << 4) & 0x10c30c30c30c30c3ull;
    v = (v | v << 2) & 0x1249249249249249ull;
    return v;
}

template <typename K>
__global__ void k_morton(const float4* __restrict__ bmin, const float4* __restrict__ bmax, int n, const BuildCounters* __restrict__ c,
                         K* __restrict__ keys, unsigned* __restrict__ vals)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 a = bmin[i], b = bmax[i];
    float cx = 0.5f * (a.x + b.x), cy = 0.5f * (a.y + b.y), cz = 0.5f * (a.z + b.z);
    float lx = ord2f(c->cbounds[0]), ly = ord2f(c->cbounds[1]), lz = ord2f(c->cbounds[2]);
    float ex = ord2f(c->cbounds[3]) - lx, ey = ord2f(c->cbounds[4]) - ly, ez = ord2f(c->cbounds[5]) - lz;
    float nx = ex > 0.f ? (cx - lx) / ex : 0.f, ny = ey > 0.f ? (cy - ly) / ey : 0.f, nz = ez > 0.f ? (cz - lz) / ez : 0.f;
    if (sizeof(K) == 4) {
        // all 32 bits: 11 bits per axis interleaved, the last z bit dropped (x and y keep 2048 cells, z 1024) -- still a prefix of
        // the Morton code, so the order is a Morton order; four radix passes over 4-byte keys
        unsigned long long ix = min(2047u, (unsigned)(nx * 2048.f)), iy = min(2047u, (unsigned)(ny * 2048.f)), iz = min(2047u, (unsigned)(nz * 2048.f));
        keys[i] = (K)(((expand21(ix) << 2) | (expand21(iy) << 1) | expand21(iz)) >> 1);
    } else {
        // 13 bits per axis = 39-bit keys: FIVE 8-bit radix passes (48-bit keys took six, +117 us per 10 M keys); 8192 cells per axis
        // are still finer than the primitives of any mesh that fits the device (2^39 cells for <= 2^31 primitives), and equal
        // keys only make the tree locally unordered, never wrong (Karras' index tie-break)
        unsigned long long ix = min(8191u, (unsigned)(nx * 8192.f)), iy = min(8191u, (unsigned)(ny * 8192.f)),
                           iz = min(8191u, (unsigned)(nz * 8192.f));
        keys[i] = (K)((expand21(ix) << 2) | (expand21(iy) << 1) | expand21(iz));
    }
    vals[i] = (unsigned)i;
}

// ------------------------------------------------------------------ 3. Karras 2012 radix tree
// node ids: internal 0..n-2, leaf j -> n-1+j.
template <typename K>
__device__ __forceinline__ int delta(const K* __restrict__ keys, int n, int i, int j)
{
    if (j < 0 || j >= n) return -1;
    K a = keys[i], b = keys[j];
    if (a == b) return (int)(sizeof(K) * 8) + __clz(i ^ j);
    return sizeof(K) == 4 ? __clz((unsigned)(a ^ b)) : __clzll((long long)(a ^ b));
}

template <typename K>
__global__ void k_hierarchy(const K* __restrict__ keys, int n, int* __restrict__ left, int* __restrict__ right,
                            int* __restrict__ parent, int2* __restrict__ range)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    int d = (delta(keys, n, i, i + 1) - delta(keys, n, i, i - 1)) >= 0 ? 1 : -1;
    int dmin = delta(keys, n, i, i - d);
    int lmax = 2;
    while (delta(keys, n, i, i + lmax * d) > dmin) lmax <<= 1;
    int l = 0;
    for (int t = lmax >> 1; t >= 1; t >>= 1)
        if (delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
    int j = i + l * d;
    int dnode = delta(keys, n, i, j);
    int s = 0, t = l;
    do {
        t = (t + 1) >> 1;
        if (delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
    } while (t > 1);
    int gamma = i + s * d + min(d, 0);
    int lo = min(i, j), hi = max(i, j);
    int lc = lo == gamma ? (n - 1 + gamma) : gamma;
    int rc = hi == gamma + 1 ? (n - 1 + gamma + 1) : gamma + 1;
    left[i] = lc; right[i] = rc;
    parent[lc] = i; parent[rc] = i;
    range[i] = make_int2(lo, hi);
    if (i == 0) parent[0] = -1;
}

// ------------------------------------------------------------------ 4. bottom-up fit + SAH leaf decision
// nb_min[id].w = half surface area, nb_max[id].w = SAH cost of the subtree; topo[id].w = 1: emit the subtree as one leaf.
__device__ __forceinline__ float half_area(float4 a, float4 b)
{
    float x = b.x - a.x, y = b.y - a.y, z = b.z - a.z;
    return x * (y + z) + y * z;
}

// One thread per leaf climbs towards the root; the second thread to arrive at a node fits it (Karras 2012).  A block owns
// FIT_CHUNK consecutive leaves of the sorted order, and every node whose leaf range lies inside the block's chunk -- all but
// a few thousandths of the nodes -- is handled without leaving the SM: arrival counters and child boxes in shared memory, a
// block-level fence instead of a device-wide one (the box still goes to HBM once, for k_collapse).  A thread that reaches a
// node spanning two chunks does not climb on -- those chains are long (device fence, L2 atomic, child boxes back through L2
// at every level) and one such thread would keep its whole block resident -- it appends the node to a list and leaves;
// k_fit_top finishes the few nodes above the chunks.
#ifndef RTB_FIT_CHUNK
#define RTB_FIT_CHUNK 256
#endif
constexpr int FIT_CHUNK = RTB_FIT_CHUNK;
__global__ void __launch_bounds__(FIT_CHUNK) k_fit(const float4* __restrict__ bmin, const float4* __restrict__ bmax, const unsigned* __restrict__ order, int n,
                      const int* __restrict__ left, const int* __restrict__ right, const int* __restrict__ parent,
                      const int2* __restrict__ range, float4* nb_min, float4* nb_max, int* visit,
                      int4* __restrict__ topo, int max_leaf, float c_node, float c_tri, int* __restrict__ top_list,
                      BuildCounters* c)
{
    __shared__ float4 s_min[2 * FIT_CHUNK], s_max[2 * FIT_CHUNK];    // [0, CHUNK): internal node id - chunk_lo; [CHUNK, 2 CHUNK): leaf - chunk_lo
    __shared__ int s_visit[FIT_CHUNK];
    const int chunk_lo = blockIdx.x * FIT_CHUNK, chunk_hi = min(chunk_lo + FIT_CHUNK, n) - 1;
    s_visit[threadIdx.x] = 0;
    __syncthreads();
    const int j = chunk_lo + threadIdx.x;
    if (j >= n) return;
    unsigned p = order[j];
    float4 a = bmin[p], b = bmax[p];
    float ha = half_area(a, b);
    a.w = ha; b.w = ha * c_tri;
    int id = n - 1 + j;
    nb_min[id] = a; nb_max[id] = b;
    s_min[FIT_CHUNK + threadIdx.x] = a; s_max[FIT_CHUNK + threadIdx.x] = b;
    int cur = parent[id];
    while (cur >= 0) {
        const int2 rg = range[cur];
        const bool local = rg.x >= chunk_lo && rg.y <= chunk_hi;
        if (!local) { top_list[atomicAdd(&c->top_count, 1)] = cur; return; }      // both children's threads do this; k_fit_top pairs them up
        const int lc = left[cur], rc = right[cur];
        __threadfence_block();
        if (atomicAdd(&s_visit[cur - chunk_lo], 1) == 0) return;                  // first arrival: the sibling finishes the node
        __threadfence_block();
        const int li = lc >= n - 1 ? FIT_CHUNK + (lc - (n - 1)) - chunk_lo : lc - chunk_lo;
        const int ri = rc >= n - 1 ? FIT_CHUNK + (rc - (n - 1)) - chunk_lo : rc - chunk_lo;
        const float4 la = s_min[li], lb = s_max[li], ra = s_min[ri], rb = s_max[ri];
        float4 mn = make_float4(fminf(la.x, ra.x), fminf(la.y, ra.y), fminf(la.z, ra.z), 0.f);
        float4 mx = make_float4(fmaxf(lb.x, rb.x), fmaxf(lb.y, rb.y), fmaxf(lb.z, rb.z), 0.f);
        float h = half_area(mn, mx);
        int cnt = rg.y - rg.x + 1;
        float c_split = c_node * h + lb.w + rb.w;
        float c_leaf = cnt <= max_leaf ? c_tri * h * (float)cnt : FLT_MAX;
        bool leaf = cnt <= max_leaf && !(c_leaf > c_split);   // never a leaf above max_leaf, whatever the costs (inf/NaN boxes)
        mn.w = h; mx.w = leaf ? c_leaf : c_split;
        nb_min[cur] = mn; nb_max[cur] = mx;
        s_min[cur - chunk_lo] = mn; s_max[cur - chunk_lo] = mx;
        topo[cur] = make_int4(lc, rc, __float_as_int(h), leaf ? 1 : 0);
        cur = parent[cur];
    }
}

// k_fit with the collapse of everything below the chunk boundary fused in (build option "build_fused", the default for
// meshes of >= FUSED_MIN_PRIMS primitives).  After the fit the whole binary subtree of the chunk -- boxes, areas, SAH
// decisions -- sits in shared memory, so the block opens it into BVH4 nodes right there: no box, topology record or queue entry
// of a chunk-local node ever goes to HBM (k_fit + k_collapse wrote and re-read ~3 GB of them per 10 M triangles), and there is
// no grid barrier.  What differs from the one-pass collapse: a MAXIMAL chunk-local subtree root (a node inside the chunk whose
// parent spans two chunks) always starts a wide node of its own -- the top collapse (k_collapse over the few nodes above the
// chunks) sees it as a finished child, topo[].w = 2 -- so a few thousandths of the wide nodes have fewer than four children;
// and a chunk's wide nodes are contiguous in memory (breadth-first within the chunk) at [1 + base, 1 + base + W), base from
// one atomicAdd per block; node 0 stays the root, the top tree's other nodes follow all the local ones.
constexpr int LC_EMPTY = 0xffff;
#ifndef RTB_FUSED_MINB
#define RTB_FUSED_MINB 8
#endif
constexpr int FUSED_MIN_PRIMS = 16 * FIT_CHUNK;

// common prefix length of the augmented keys (key, index) of sorted positions i and i + 1 -- Karras' delta(i, i + 1)
template <typename K>
__device__ __forceinline__ int cpl_pair(K a, K b, int i)
{
    if (a == b) return (int)(sizeof(K) * 8) + __clz(i ^ (i + 1));
    return sizeof(K) == 4 ? __clz((unsigned)(a ^ b)) : __clzll((long long)(a ^ b));
}

// In the fused build the radix tree itself is made here too, bottom-up (Apetrei 2014), so k_hierarchy and its four arrays are not
// needed: a subtree with leaf range [l, r] is the LEFT child of the node that splits between r and r + 1 if
// cpl(r) > cpl(l - 1) (cpl = common prefix length of neighbouring keys, -1 outside the array), else the RIGHT child of the node
// that splits between l - 1 and l.  An internal node is named by its split position.  Which nodes lie inside the chunk is known
// up front: node i does iff cpl(i) exceeds the smallest cpl to its left and to its right within the chunk, the two cpl values
// across the chunk's ends included (tools/apetrei_model.py checks both rules against a top-down construction).  The climb
// therefore touches shared memory only.
template <typename K>
__global__ void __launch_bounds__(FIT_CHUNK, RTB_FUSED_MINB) k_fit_collapse(const K* __restrict__ keys, const float4* __restrict__ bmin, const float4* __restrict__ bmax,
                      const unsigned* __restrict__ order, int n, int2* __restrict__ range, float4* nb_min, float4* nb_max,
                      int4* __restrict__ topo, int max_leaf, float c_node, float c_tri, int4* __restrict__ top_list,
                      float4* __restrict__ nodes, BuildCounters* c)
{
    // local codes: [0, CHUNK) internal node (split position - chunk_lo); [CHUNK, 2 CHUNK) leaf - chunk_lo
    __shared__ float4 s_min[2 * FIT_CHUNK], s_max[2 * FIT_CHUNK];
    __shared__ int s_visit[FIT_CHUNK];
    __shared__ int s_cpl[FIT_CHUNK + 1];                // s_cpl[k] = cpl(chunk_lo - 1 + k)
    __shared__ int s_wpre[FIT_CHUNK / 32], s_wsuf[FIT_CHUNK / 32];
    __shared__ unsigned char s_local[FIT_CHUNK];        // node chunk_lo + t lies inside the chunk
    __shared__ int2 s_rng[FIT_CHUNK];                   // leaf range of an internal node (.x by its left child, .y by its right child)
    __shared__ ushort2 s_lr[FIT_CHUNK];                 // its children (local codes), same
    __shared__ unsigned char s_leaf[FIT_CHUNK];         // SAH: emit the subtree as one leaf
    __shared__ unsigned short s_item[FIT_CHUNK];        // wide node k of this chunk opens binary node s_item[k]
    __shared__ unsigned short s_wide_of[FIT_CHUNK];     // inverse
    __shared__ unsigned char s_root[FIT_CHUNK], s_lvl[FIT_CHUNK];   // which maximal root a wide node hangs under, how many levels below it
    __shared__ int s_rdepth[FIT_CHUNK];                 // per root: deepest level below it
    __shared__ int s_nitems, s_base, s_leaves, s_maxleaf;
    __shared__ ushort4 s_slots[FIT_CHUNK];              // a wide node's (up to) four children, local codes or LC_EMPTY
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int chunk_lo = blockIdx.x * FIT_CHUNK, chunk_hi = min(chunk_lo + FIT_CHUNK, n) - 1;
    const int j = chunk_lo + tid;
    s_visit[tid] = 0; s_rdepth[tid] = 0;
    if (tid == 0) { s_nitems = 0; s_base = 0; s_leaves = 0; s_maxleaf = 0; }
    // ---- cpl of this chunk's neighbouring keys, and which of its nodes are local
    int cv = -1;
    if (j < n - 1) cv = cpl_pair<K>(keys[j], keys[j + 1], j);
    s_cpl[tid + 1] = cv;
    if (tid == 0) s_cpl[0] = chunk_lo > 0 ? cpl_pair<K>(keys[chunk_lo - 1], keys[chunk_lo], chunk_lo - 1) : -1;
    int pre = cv, suf = cv;                             // inclusive prefix / suffix minimum within the warp
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int a = __shfl_up_sync(0xffffffffu, pre, o), b = __shfl_down_sync(0xffffffffu, suf, o);
        if (lane >= o) pre = min(pre, a);
        if (lane + o < 32) suf = min(suf, b);
    }
    if (lane == 31) s_wpre[warp] = pre;
    if (lane == 0) s_wsuf[warp] = suf;
    if (j < n) {                                        // leaf boxes
        const unsigned p = order[j];
        float4 a = bmin[p], b = bmax[p];
        const float ha = half_area(a, b);
        a.w = ha; b.w = ha * c_tri;
        s_min[FIT_CHUNK + tid] = a; s_max[FIT_CHUNK + tid] = b;
    }
    __syncthreads();
    {
        int lmin = s_cpl[0], rmin = 0x7fffffff;
        for (int w = 0; w < FIT_CHUNK / 32; w++) { if (w < warp) lmin = min(lmin, s_wpre[w]); if (w > warp) rmin = min(rmin, s_wsuf[w]); }
        const int a = __shfl_up_sync(0xffffffffu, pre, 1), b = __shfl_down_sync(0xffffffffu, suf, 1);
        if (lane > 0) lmin = min(lmin, a);
        if (lane < 31) rmin = min(rmin, b);
        s_local[tid] = (j < chunk_hi && cv > lmin && cv > rmin) ? 1 : 0;
    }
    __syncthreads();
    // ---- the climb: one thread per leaf; the second child to arrive at a node fits it and goes on (Karras 2012).  (A
    // level-synchronous form with the climbers compacted in a queue executes a third of the instructions and is slower, 716 vs
    // 660 us per 10 M triangles: the block's critical path, not the issue rate, is what the kernel waits for -- profiles/r04f.)
    if (j < n) {
        int code = FIT_CHUNK + tid, l = j, r = j;                   // the subtree this thread has finished: local code, leaf range
        for (;;) {
            const bool rp = s_cpl[r - chunk_lo + 1] > s_cpl[l - chunk_lo];      // cpl(r) > cpl(l - 1): left child of node r
            const int p = rp ? r : l - 1, pt = p - chunk_lo;
            if (pt < 0 || !s_local[pt]) {
                // the parent spans two chunks: this subtree is a maximal local root.  What the top phase reads of it: its box, its
                // range, what it is; and one work item for k_fit_top_fused (child, range, parent | left-child flag)
                const int id = code < FIT_CHUNK ? chunk_lo + code : n - 1 + j;
                nb_min[id] = s_min[code]; nb_max[id] = s_max[code];
                if (code < FIT_CHUNK) {
                    range[id] = make_int2(l, r);
                    if (s_leaf[code]) topo[id] = make_int4(0, 0, __float_as_int(s_min[code].w), 1);
                    else {                                          // a wide node of its own: the first items of the chunk's queue
                        const int k = atomicAdd(&s_nitems, 1);
                        s_item[k] = (unsigned short)code; s_wide_of[code] = (unsigned short)k;
                        s_root[k] = (unsigned char)k; s_lvl[k] = 0;
                    }
                }
                top_list[atomicAdd(&c->top_count, 1)] = make_int4(id, l, r, p | (rp ? (int)0x80000000 : 0));
                break;
            }
            if (rp) { s_lr[pt].x = (unsigned short)code; s_rng[pt].x = l; }
            else { s_lr[pt].y = (unsigned short)code; s_rng[pt].y = r; }
            __threadfence_block();
            if (atomicAdd(&s_visit[pt], 1) == 0) break;             // first arrival: the sibling finishes the node
            __threadfence_block();
            const int chx = ((volatile unsigned short*)&s_lr[pt])[0], chy = ((volatile unsigned short*)&s_lr[pt])[1];
            l = ((volatile int*)&s_rng[pt])[0]; r = ((volatile int*)&s_rng[pt])[1];
            const float4 la = s_min[chx], lb = s_max[chx], ra = s_min[chy], rb = s_max[chy];
            float4 mn = make_float4(fminf(la.x, ra.x), fminf(la.y, ra.y), fminf(la.z, ra.z), 0.f);
            float4 mx = make_float4(fmaxf(lb.x, rb.x), fmaxf(lb.y, rb.y), fmaxf(lb.z, rb.z), 0.f);
            const float h = half_area(mn, mx);
            const int cnt = r - l + 1;
            const float c_split = c_node * h + lb.w + rb.w;
            const float c_leaf = cnt <= max_leaf ? c_tri * h * (float)cnt : FLT_MAX;
            const bool leaf = cnt <= max_leaf && !(c_leaf > c_split);   // never a leaf above max_leaf, whatever the costs (inf/NaN boxes)
            mn.w = h; mx.w = leaf ? c_leaf : c_split;
            s_min[pt] = mn; s_max[pt] = mx; s_leaf[pt] = leaf ? 1 : 0;
            code = pt;
        }
    }
    __syncthreads();
    // ---- collapse, level by level, one thread per wide node; everything it reads is in shared memory
    const int n_roots = s_nitems;
    __syncthreads();                                                // everyone has read the count before anyone appends to the queue
    int my_leaves = 0, my_maxleaf = 0;
    int lv_begin = 0, lv_end = n_roots;
    while (lv_begin < lv_end) {                                     // block-uniform
        const int it = lv_begin + tid;
        if (it < lv_end) {
            const int bnode = s_item[it];
            int sl[4] = { LC_EMPTY, LC_EMPTY, LC_EMPTY, LC_EMPTY }; float ar[4] = { -1.f, -1.f, -1.f, -1.f };
            auto put = [&](int k, int code) {
                const float a = (code < FIT_CHUNK && !s_leaf[code]) ? fmaxf(s_min[code].w, 0.f) : -1.f;    // < 0: cannot be opened
#pragma unroll
                for (int q = 0; q < 4; q++) if (q == k) { sl[q] = code; ar[q] = a; }
            };
            const ushort2 ch = s_lr[bnode];
            put(0, ch.x); put(1, ch.y);
            int ns = 2;
            while (ns < 4) {                                        // greedy: open the slot with the largest half area
                int best = -1, bcode = 0; float barea = -1.f;
#pragma unroll
                for (int k = 0; k < 4; k++) if (k < ns && ar[k] > barea) { barea = ar[k]; best = k; bcode = sl[k]; }
                if (best < 0) break;
                const ushort2 g = s_lr[bcode];
                put(best, g.x); put(ns++, g.y);
            }
            int kc = 0;
#pragma unroll
            for (int k = 0; k < 4; k++) kc += ar[k] >= 0.f ? 1 : 0;
            int pos = kc ? atomicAdd(&s_nitems, kc) : 0;            // the children of one node are contiguous
            const int root = s_root[it], lvl = s_lvl[it] + 1;
            if (kc) atomicMax(&s_rdepth[root], lvl);
#pragma unroll
            for (int k = 0; k < 4; k++) {
                if (ar[k] >= 0.f) {
                    s_item[pos] = (unsigned short)sl[k]; s_wide_of[sl[k]] = (unsigned short)pos;
                    s_root[pos] = (unsigned char)root; s_lvl[pos] = (unsigned char)lvl;
                    pos++;
                }
            }
            s_slots[it] = make_ushort4((unsigned short)sl[0], (unsigned short)sl[1], (unsigned short)sl[2], (unsigned short)sl[3]);
        }
        __syncthreads();
        lv_begin = lv_end; lv_end = s_nitems;
        __syncthreads();
    }
    const int n_wide = s_nitems;
    if (n_wide == 0) return;                                        // block-uniform
    if (tid == 0) s_base = 1 + atomicAdd(&c->n_local, n_wide);
    __syncthreads();
    const int base = s_base;
    // emission: one thread per 16-byte ROW of a node (7 rows: lo/hi x, lo/hi y, lo/hi z of the four children, their refs), so a
    // warp writes 512 contiguous bytes and nobody holds a whole node in registers
    for (int r = tid; r < n_wide * NODE_F4; r += FIT_CHUNK) {
        const int node = r / NODE_F4, row = r - node * NODE_F4;
        const ushort4 sv = s_slots[node];
        const int sl[4] = { sv.x, sv.y, sv.z, sv.w };
        float4 out;
        if (row < 6) {
            const float4* src = (row & 1) ? s_max : s_min;
            const int axis = row >> 1;
            float v[4];
#pragma unroll
            for (int k = 0; k < 4; k++) v[k] = sl[k] != LC_EMPTY ? ((const float*)(src + sl[k]))[axis] : __int_as_float(0x7f800000);
            out = make_float4(v[0], v[1], v[2], v[3]);
        } else {
            int ref[4];
#pragma unroll
            for (int k = 0; k < 4; k++) {
                const int code = sl[k];
                if (code == LC_EMPTY) ref[k] = EMPTY_REF;
                else if (code >= FIT_CHUNK) { ref[k] = ~((chunk_lo + code - FIT_CHUNK) << 3); my_leaves++; my_maxleaf = max(my_maxleaf, 1); }
                else if (s_leaf[code]) {
                    const int2 rg = s_rng[code];
                    const int count = rg.y - rg.x + 1;
                    ref[k] = ~((rg.x << 3) | (count - 1)); my_leaves++; my_maxleaf = max(my_maxleaf, count);
                } else ref[k] = base + s_wide_of[code];
            }
            out = make_float4(__int_as_float(ref[0]), __int_as_float(ref[1]), __int_as_float(ref[2]), __int_as_float(ref[3]));
        }
        nodes[(size_t)base * NODE_F4 + r] = out;
    }
    if (tid < n_roots)                                              // finished child of the top tree: (wide index, levels below it, area, 2)
        topo[chunk_lo + s_item[tid]] = make_int4(base + tid, s_rdepth[tid], __float_as_int(s_min[s_item[tid]].w), 2);
    if (my_leaves) { atomicAdd(&s_leaves, my_leaves); atomicMax(&s_maxleaf, my_maxleaf); }
    __syncthreads();
    if (tid == 0 && s_leaves) { atomicAdd(&c->n_leaves, s_leaves); atomicMax(&c->max_leaf, s_maxleaf); }
}

// The nodes above the chunks: one thread per list entry, the classic climb (every node appears twice in the list, once per
// child; the second arrival fits it and goes on).  The boxes below were written by k_fit, a kernel boundary ago.
__global__ void __launch_bounds__(256) k_fit_top(const int* __restrict__ top_list, const BuildCounters* __restrict__ c, int n,
                          const int* __restrict__ left, const int* __restrict__ right, const int* __restrict__ parent,
                          const int2* __restrict__ range, float4* nb_min, float4* nb_max, int* visit,
                          int4* __restrict__ topo, int max_leaf, float c_node, float c_tri, int allow_leaf)
{
    // fused build (allow_leaf = 0): a node above the chunks is never emitted as one leaf -- the subtrees below it were collapsed
    // into wide nodes before anybody looked at it, and would be left behind unreferenced
    const int count = c->top_count;
    for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < count; e += gridDim.x * blockDim.x) {
        int cur = top_list[e];
        while (cur >= 0) {
            __threadfence();
            if (atomicAdd(&visit[cur], 1) == 0) break;
            const int lc = left[cur], rc = right[cur];
            const int2 rg = range[cur];
            const float4 la = __ldcg(nb_min + lc), lb = __ldcg(nb_max + lc), ra = __ldcg(nb_min + rc), rb = __ldcg(nb_max + rc);
            float4 mn = make_float4(fminf(la.x, ra.x), fminf(la.y, ra.y), fminf(la.z, ra.z), 0.f);
            float4 mx = make_float4(fmaxf(lb.x, rb.x), fmaxf(lb.y, rb.y), fmaxf(lb.z, rb.z), 0.f);
            float h = half_area(mn, mx);
            int cnt = rg.y - rg.x + 1;
            float c_split = c_node * h + lb.w + rb.w;
            float c_leaf = cnt <= max_leaf ? c_tri * h * (float)cnt : FLT_MAX;
            bool leaf = allow_leaf && cnt <= max_leaf && !(c_leaf > c_split);
            mn.w = h; mx.w = leaf ? c_leaf : c_split;
            nb_min[cur] = mn; nb_max[cur] = mx;
                topo[cur] = make_int4(lc, rc, __float_as_int(h), leaf ? 1 : 0);
            cur = parent[cur];
        }
    }
}

// Fused build: the nodes above the chunks, made AND fitted bottom-up from k_fit_collapse's work items (same parent rule, cpl
// from the keys in HBM).  A child deposits its id and the far end of its range in its parent's record, then counts itself in;
// the second arrival fits the parent and climbs on.  No node above the chunks is emitted as one leaf (see k_fit_collapse).
// The thread that completes [0, n - 1] has the root: it seeds the collapse queue with it.
template <typename K>
__global__ void __launch_bounds__(64) k_fit_top_fused(const K* __restrict__ keys, const int4* __restrict__ top_list, BuildCounters* c, int n,
                          int* left, int* right, int2* range, float4* nb_min, float4* nb_max, int* visit,
                          int4* __restrict__ topo, float c_node, int2* __restrict__ q0)
{
    const int count = c->top_count;
    for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < count; e += gridDim.x * blockDim.x) {
        const int4 ent = top_list[e];
        int cid = ent.x, l = ent.y, r = ent.z, p = ent.w & 0x7fffffff;
        bool rp = ent.w < 0;
        float4 mn = nb_min[cid], mx = nb_max[cid];
        for (;;) {
            if (rp) { left[p] = cid; ((int*)(range + p))[0] = l; }
            else { right[p] = cid; ((int*)(range + p))[1] = r; }
            // release: everything this thread wrote (its box, its half of the record) is visible to whoever sees the count; the
            // second arrival reads through L2 (__ldcg).  Lighter than __threadfence(), which also invalidates this SM's L1
            int seen;
            asm volatile("atom.add.release.gpu.global.s32 %0, [%1], 1;" : "=r"(seen) : "l"(visit + p) : "memory");
            if (seen == 0) break;
            const int sib = rp ? __ldcg(right + p) : __ldcg(left + p);
            const int far = rp ? __ldcg((const int*)(range + p) + 1) : __ldcg((const int*)(range + p));
            if (rp) r = far; else l = far;
            // the sibling's box and the four keys that decide the next parent: one round trip, not two
            const float4 sa = __ldcg(nb_min + sib), sb = __ldcg(nb_max + sib);
            K k0 = 0, k1 = 0, k2 = 0, k3 = 0;
            if (l > 0) { k0 = keys[l - 1]; k1 = keys[l]; }
            if (r < n - 1) { k2 = keys[r]; k3 = keys[r + 1]; }
            const float cost = mx.w + sb.w;
            mn = make_float4(fminf(mn.x, sa.x), fminf(mn.y, sa.y), fminf(mn.z, sa.z), 0.f);
            mx = make_float4(fmaxf(mx.x, sb.x), fmaxf(mx.y, sb.y), fmaxf(mx.z, sb.z), 0.f);
            const float h = half_area(mn, mx);
            mn.w = h; mx.w = c_node * h + cost;
            nb_min[p] = mn; nb_max[p] = mx;
            topo[p] = make_int4(rp ? cid : sib, rp ? sib : cid, __float_as_int(h), 0);
            cid = p;
            if (l == 0 && r == n - 1) { q0[0] = make_int2(p, 0); c->sah = mx.w; break; }
            const int cl_ = l > 0 ? cpl_pair<K>(k0, k1, l - 1) : -1;
            const int cr_ = r < n - 1 ? cpl_pair<K>(k2, k3, r) : -1;
            rp = cr_ > cl_;
            p = rp ? r : l - 1;
        }
    }
}

// ------------------------------------------------------------------ 5. top-down collapse to BVH4
// One work item = (binary node to open, wide node index to fill).  Greedy surface-area opening: keep
// replacing the open slot with the largest half-area by its two children until 4 slots are used.
__device__ __forceinline__ void slot_range(int id, int n, const int2* __restrict__ range, int& first, int& count)
{
    if (id >= n - 1) { first = id - (n - 1); count = 1; }
    else { int2 r = range[id]; first = r.x; count = r.y - r.x + 1; }
}

// ONE persistent launch for the whole collapse (it was one launch per level, 70 of them: the host does not know the depth).
// Levels are separated by a grid barrier (every CTA is resident: the grid is sized from the occupancy query); a wide
// node's index is (number of nodes of all earlier levels) + (its position in this level's queue), so one level's nodes --
// and the children of one node -- are contiguous in memory, and the only atomic is one per warp on the next level's count.
__device__ __forceinline__ void grid_barrier(unsigned* bar, unsigned target)
{
    // Release on the way in (the block's queue entries and counts, made visible to thread 0 by the CTA barrier, are visible device-wide
    // before the count); nothing on the way out: what a level reads of the previous one -- queue entries, queue sizes -- is read
    // through L2 (__ldcg / volatile), everything else the kernel reads was written by earlier kernels.  Two __threadfence() here
    // cost an L1 invalidation per level and CTA.
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned seen;
        asm volatile("atom.add.release.gpu.global.u32 %0, [%1], 1;" : "=r"(seen) : "l"(bar) : "memory");
        while (*(volatile unsigned*)bar < target) { }
    }
    __syncthreads();
}

#ifndef RTB_COLLAPSE_MINB
#define RTB_COLLAPSE_MINB 7
#endif
#ifndef RTB_COLLAPSE_CAP_FUSED
#define RTB_COLLAPSE_CAP_FUSED 2
#endif
#ifndef RTB_COLLAPSE_CAP
#define RTB_COLLAPSE_CAP 8
#endif
__global__ void __launch_bounds__(128, RTB_COLLAPSE_MINB) k_collapse(int2* __restrict__ q0, int2* __restrict__ q1, int n,
                                                  const int* __restrict__ left, const int* __restrict__ right,
                                                  const int2* __restrict__ range, const float4* __restrict__ nb_min,
                                                  const float4* __restrict__ nb_max, const int4* __restrict__ topo,
                                                  float4* __restrict__ nodes, BuildCounters* c)
{
    const int lane = threadIdx.x & 31;
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, n_warps = (gridDim.x * blockDim.x) >> 5;
    int my_leaves = 0, my_maxleaf = 0, my_deep = 0, level = 0;
    int start = 0;                                                      // wide nodes of the levels before the next one
    // fused build: k_fit_collapse already emitted n_local nodes at [1, 1 + n_local); this kernel's i-th node (breadth-first) goes
    // to index 0 (the root) or n_local + i.  One-pass build: n_local = 0, index i.
    const int n_local = *(volatile int*)&c->n_local;
    for (; level < 70; level++) {
        const int items = *(volatile int*)&c->qcount[level];
        if (items == 0) break;
        start += items;
        const int2* q_in = (level & 1) ? q1 : q0;
        int2* q_out = (level & 1) ? q0 : q1;
        for (int it0 = warp * 32; it0 < items; it0 += n_warps * 32) {
            const int it = it0 + lane;
            const bool active = it < items;
            int2 item = make_int2(0, 0);
            int slot[4] = { 0, 0, 0, 0 }; int ns = 0;
            // per slot: what the greedy opening needs of its binary node -- (left, right, half area, leaf flag), ONE 16-byte record
            // written by k_fit (it was four scattered fetches); area < 0 = cannot be opened
            int cl[4] = { 0, 0, 0, 0 }, cr[4] = { 0, 0, 0, 0 }; float area[4] = { -1.f, -1.f, -1.f, -1.f };
            bool inner[4] = { false, false, false, false };
            int kc = 0;
            if (active) {
                item = __ldcg(q_in + it);
                // slot k (a run-time index: written through an unrolled select so the arrays stay in registers)
                auto fetch = [&](int k, int id) {
                    int l = 0, r = 0; float a = -1.f;
                    if (id < n - 1) {
                        const int4 t = __ldg(topo + id); l = t.x; r = t.y; a = t.w ? -1.f : fmaxf(__int_as_float(t.z), 0.f);
                        if (t.w == 2) r = -1 - t.y;                    // collapsed by k_fit_collapse: l = its wide node, r < 0 carries the levels below it
                    }
#pragma unroll
                    for (int q = 0; q < 4; q++) if (q == k) { slot[q] = id; cl[q] = l; cr[q] = r; area[q] = a; }
                };
                const int4 t0 = __ldg(topo + item.x);
                if (t0.w) { fetch(0, item.x); ns = 1; }                      // whole tree is one leaf (tiny meshes)
                else { fetch(0, t0.x); fetch(1, t0.y); ns = 2; }
                while (ns < 4) {
                    int best = -1, l = 0, r = 0; float barea = -1.f;
#pragma unroll
                    for (int k = 0; k < 4; k++) if (k < ns && area[k] > barea) { barea = area[k]; best = k; l = cl[k]; r = cr[k]; }
                    if (best < 0) break;
                    fetch(best, l);
                    fetch(ns++, r);
                }
#pragma unroll
                for (int k = 0; k < 4; k++) { inner[k] = k < ns && area[k] >= 0.f; kc += inner[k] ? 1 : 0; }
            }
            // next level's queue positions for this warp's interior children: exclusive scan + one atomic
            int pre = kc;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { int v = __shfl_up_sync(0xffffffffu, pre, o); if (lane >= o) pre += v; }
            const int total = __shfl_sync(0xffffffffu, pre, 31);
            int base = 0;
            if (lane == 0 && total) base = atomicAdd(&c->qcount[level + 1], total);
            base = __shfl_sync(0xffffffffu, base, 0) + pre - kc;
            if (!active) continue;
            float lo[3][4], hi[3][4]; int ref[4];
#pragma unroll
            for (int k = 0; k < 4; k++) {
                if (k < ns) {
                    int id = slot[k];
                    float4 a = nb_min[id], b = nb_max[id];
                    lo[0][k] = a.x; lo[1][k] = a.y; lo[2][k] = a.z; hi[0][k] = b.x; hi[1][k] = b.y; hi[2][k] = b.z;
                    if (!inner[k] && cr[k] < 0) {                                  // a subtree k_fit_collapse finished
                        ref[k] = cl[k];
                        my_deep = max(my_deep, level + 1 - 1 - cr[k]);
                    } else if (!inner[k]) {
                        int first, count; slot_range(id, n, range, first, count);
                        ref[k] = ~((first << 3) | (count - 1));
                        my_leaves++; my_maxleaf = max(my_maxleaf, count);
                    } else {
                        const int qi = base++;
                        q_out[qi] = make_int2(id, n_local + start + qi);
                        ref[k] = n_local + start + qi;
                    }
                } else {
                    // empty slot: lo = hi = +inf can never pass a slab test (an inverted box would: min/max un-invert it)
                    lo[0][k] = lo[1][k] = lo[2][k] = __int_as_float(0x7f800000);
                    hi[0][k] = hi[1][k] = hi[2][k] = __int_as_float(0x7f800000);
                    ref[k] = EMPTY_REF;
                }
            }
            float4* nd = nodes + (size_t)item.y * NODE_F4;
            nd[0] = make_float4(lo[0][0], lo[0][1], lo[0][2], lo[0][3]); nd[1] = make_float4(hi[0][0], hi[0][1], hi[0][2], hi[0][3]);
            nd[2] = make_float4(lo[1][0], lo[1][1], lo[1][2], lo[1][3]); nd[3] = make_float4(hi[1][0], hi[1][1], hi[1][2], hi[1][3]);
            nd[4] = make_float4(lo[2][0], lo[2][1], lo[2][2], lo[2][3]); nd[5] = make_float4(hi[2][0], hi[2][1], hi[2][2], hi[2][3]);
            nd[6] = make_float4(__int_as_float(ref[0]), __int_as_float(ref[1]), __int_as_float(ref[2]), __int_as_float(ref[3]));
        }
        grid_barrier(&c->bar, (unsigned)gridDim.x * (unsigned)(level + 1));
    }
    if (my_leaves) { atomicAdd(&c->n_leaves, my_leaves); atomicMax(&c->max_leaf, my_maxleaf); }
    if (my_deep) atomicMax(&c->max_level2, my_deep);
    if (blockIdx.x == 0 && threadIdx.x == 0) { c->max_level = level - 1; c->n_wide = start + n_local; }
}

// n == 0 / n == 1 trees: a root with no child / one single-primitive leaf
__global__ void k_trivial_root(float4* nodes, const float4* bmin, const float4* bmax, int n)
{
    const float pinf = __int_as_float(0x7f800000);
    float4 lo = make_float4(pinf, pinf, pinf, pinf), hi = lo;
    float4 lx = lo, ly = lo, lz = lo, hx = hi, hy = hi, hz = hi;
    int ref0 = EMPTY_REF;
    if (n == 1) { lx.x = bmin[0].x; ly.x = bmin[0].y; lz.x = bmin[0].z; hx.x = bmax[0].x; hy.x = bmax[0].y; hz.x = bmax[0].z; ref0 = ~0; }
    nodes[0] = lx; nodes[1] = hx; nodes[2] = ly; nodes[3] = hy; nodes[4] = lz; nodes[5] = hz;
    nodes[6] = make_float4(__int_as_float(ref0), __int_as_float(EMPTY_REF), __int_as_float(EMPTY_REF), __int_as_float(EMPTY_REF));
}

__global__ void k_iota(unsigned* v, int n) { int i = blockIdx.x * blockDim.x + threadIdx.x; if (i < n) v[i] = i; }
__global__ void k_root_sah(const float4* nb_max, BuildCounters* c) { c->sah = nb_max[0].w; }

// ------------------------------------------------------------------ 6. triangle records
// 64-byte records in leaf order; every derived value uses the unfused ops of point_inside
// (TrianglePrimitive.hpp:154-162) so the tracer's predicate is bit-identical to the reference's.
// WIDE: the 72-byte record is fetched with nine 8-byte loads instead of eighteen 4-byte ones (the base pointer must be 8-byte
// aligned: checked on the host) -- ncu showed the kernel waiting on the load/store queue (lg_throttle), not on HBM.
template <bool WIDE>
__global__ void k_emit_tris(const rtb_triangle* __restrict__ tris, int n, const unsigned* __restrict__ order,
                            float4* __restrict__ rec, float2* __restrict__ uvs)
{
    int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    unsigned p = order[j];
    float t[18];
    if (WIDE) {
        const float2* t2 = (const float2*)(tris + p);
#pragma unroll
        for (int k = 0; k < 9; k++) { const float2 w = __ldg(t2 + k); t[2 * k] = w.x; t[2 * k + 1] = w.y; }
    } else {
        const float* t1 = (const float*)(tris + p);
#pragma unroll
        for (int k = 0; k < 18; k++) t[k] = __ldg(t1 + k);
    }
    f3 A = mk3(t[0], t[1], t[2]), B = mk3(t[3], t[4], t[5]), C = mk3(t[6], t[7], t[8]);
    f3 nrm = mk3(t[9], t[10], t[11]);
    f3 e0 = xsub(C, A), e1 = xsub(B, A);
    float4* r = rec + (size_t)j * TRI_F4;
    r[0] = make_float4(nrm.x, nrm.y, nrm.z, __int_as_float((int)p));
    r[1] = make_float4(A.x, A.y, A.z, xdot(e0, e0));
    r[2] = make_float4(e0.x, e0.y, e0.z, xdot(e0, e1));
    r[3] = make_float4(e1.x, e1.y, e1.z, xdot(e1, e1));
    float2* u = uvs + (size_t)j * 3;                        // in leaf order, like the records: coalesced here, and the shader has the slot
    u[0] = make_float2(t[12], t[13]); u[1] = make_float2(t[14], t[15]); u[2] = make_float2(t[16], t[17]);
}

// ------------------------------------------------------------------ host side
#define CK(x) do { cudaError_t _e = (x); if (_e != cudaSuccess) return _e; } while (0)

static cudaError_t emit_records(const rtb_triangle* tris_dev, int n, const unsigned* order, float4* tri_rec, float2* uvs, cudaStream_t st);

// `fork`: triangle records wanted (tris / tri_rec / uvs set).  In the fused build they are written on the arena's side stream, from
// the moment k_fit_collapse has been enqueued: k_emit_tris needs the sorted order only, and it streams 1.6 GB while the two kernels
// that finish the top of the tree are chains of dependent L2 round trips on a handful of warps.  *forked tells the caller to join.
struct EmitJob { const rtb_triangle* tris; float4* tri_rec; float2* uvs; bool forked; };

template <typename K>
static cudaError_t build_tree(const float4* bmin, const float4* bmax, int n, int max_leaf, int key_bits, float sah_node_cost, BvhResult& out,
                              Arena& ws, BuildCounters* ctr, cudaStream_t st, bool fused, EmitJob* job)
{
    K* keys[2]; unsigned* vals[2];
    keys[0] = ws.take<K>(n); keys[1] = ws.take<K>(n);
    vals[0] = ws.take<unsigned>(n); vals[1] = ws.take<unsigned>(n);
    const int nblocks = (n + RS_TILE - 1) / RS_TILE;
    unsigned* hist = ws.take<unsigned>((size_t)256 * nblocks + 256);
    int* left = ws.take<int>(n); int* right = ws.take<int>(n);
    int* parent = ws.take<int>(2 * (size_t)n); int2* range = ws.take<int2>(n);
    int* visit = ws.take<int>(n);
    float4* nb_min = ws.take<float4>(2 * (size_t)n); float4* nb_max = ws.take<float4>(2 * (size_t)n);
    int4* topo = ws.take<int4>(n);
    int2* q0 = ws.take<int2>(n); int2* q1 = ws.take<int2>(n);
    if (!q1) return cudaErrorMemoryAllocation;

    const int g = (n + 255) / 256;
    k_morton<K><<<g, 256, 0, st>>>(bmin, bmax, n, ctr, keys[0], vals[0]);
    int cur = radix_sort_pairs<K>(keys, vals, n, key_bits, hist, st);
    CK(cudaMemcpyAsync(out.order, vals[cur], sizeof(unsigned) * (size_t)n, cudaMemcpyDeviceToDevice, st));
    CK(cudaMemsetAsync(visit, 0, sizeof(int) * (size_t)n, st));
    int one = 1;
    if (fused) {
        // radix tree + fit + BVH4 nodes of everything below the chunk boundaries in one kernel; then the few nodes above them
        int4* top_list = ws.take<int4>(n);
        if (!top_list) return cudaErrorMemoryAllocation;
        // 8 CTAs x 27 KB of shared memory per SM (per device and cheap: set on every build)
        CK(cudaFuncSetAttribute(k_fit_collapse<K>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
        k_fit_collapse<K><<<(n + FIT_CHUNK - 1) / FIT_CHUNK, FIT_CHUNK, 0, st>>>(keys[cur], bmin, bmax, out.order, n, range, nb_min, nb_max, topo, max_leaf,
                                                                                   sah_node_cost, 1.0f, top_list, out.nodes, ctr);
        // one thread per work item where possible (their number is only known on the device; n / 8 threads cover every mesh seen so
        // far, the loop in the kernel covers the rest): a thread that still has a second item to start after a long climb keeps
        // everybody who needs that item's subtree waiting.  Small CTAs, so that the slots of finished climbers are reused early.
        const int top_blocks = std::max(148 * 16, std::min((n / 8 + 63) / 64, 148 * 256));
        const bool fork = job && job->tris;
        if (fork) {
            if (!ws.side) {
                CK(cudaStreamCreateWithFlags(&ws.side, cudaStreamNonBlocking));
                CK(cudaEventCreateWithFlags(&ws.ev_fork, cudaEventDisableTiming));
                CK(cudaEventCreateWithFlags(&ws.ev_join, cudaEventDisableTiming));
            }
            CK(cudaEventRecord(ws.ev_fork, st));                        // = k_fit_collapse done (the event must not cover the top fit)
        }
        k_fit_top_fused<K><<<top_blocks, 64, 0, st>>>(keys[cur], top_list, ctr, n, left, right, range, nb_min, nb_max, visit, topo, sah_node_cost, q0);
        if (fork) {                                                     // enqueued after the top fit, so its few climbers get their slots first
            CK(cudaStreamWaitEvent(ws.side, ws.ev_fork, 0));
            CK(emit_records(job->tris, n, out.order, job->tri_rec, job->uvs, ws.side));
            CK(cudaEventRecord(ws.ev_join, ws.side));
            job->forked = true;
        }
    } else {
        k_hierarchy<K><<<g, 256, 0, st>>>(keys[cur], n, left, right, parent, range);
        k_fit<<<(n + FIT_CHUNK - 1) / FIT_CHUNK, FIT_CHUNK, 0, st>>>(bmin, bmax, out.order, n, left, right, parent, range, nb_min, nb_max, visit, topo, max_leaf, sah_node_cost, 1.0f,
                                                                       (int*)q0, ctr);
        k_fit_top<<<148 * 4, 256, 0, st>>>((const int*)q0, ctr, n, left, right, parent, range, nb_min, nb_max, visit, topo, max_leaf, sah_node_cost, 1.0f, 1);
        k_root_sah<<<1, 1, 0, st>>>(nb_max, ctr);
        // level 0 work item: open binary root into wide node 0
        int2 root_item = make_int2(0, 0);
        CK(cudaMemcpyAsync(q0, &root_item, sizeof(int2), cudaMemcpyHostToDevice, st));
    }
    CK(cudaMemcpyAsync(&ctr->qcount[0], &one, sizeof(int), cudaMemcpyHostToDevice, st));
    int resident = 0;                                                   // CTAs of k_collapse that fit THIS device at once
    {
        int dev = 0, sms = 0, per_sm = 0;
        CK(cudaGetDevice(&dev));
        CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_collapse, 128, 0));
        resident = sms * std::max(1, std::min(per_sm, fused ? RTB_COLLAPSE_CAP_FUSED : RTB_COLLAPSE_CAP));
    }
    // cooperative launch: the driver starts the grid only when every CTA can be resident at once, so the grid barrier cannot
    // deadlock against another persistent kernel (a second context building on another stream, a render in flight)
    float4* nodes_out = out.nodes;
    void* args[] = { &q0, &q1, &n, &left, &right, &range, &nb_min, &nb_max, &topo, &nodes_out, &ctr };
    CK(cudaLaunchCooperativeKernel((const void*)k_collapse, dim3(std::min(resident, (n + 127) / 128)), dim3(128), args, 0, st));
    return cudaGetLastError();
}

static cudaError_t emit_records(const rtb_triangle* tris_dev, int n, const unsigned* order, float4* tri_rec, float2* uvs, cudaStream_t st)
{
    if (((uintptr_t)tris_dev & 7u) == 0) k_emit_tris<true><<<(n + 255) / 256, 256, 0, st>>>(tris_dev, n, order, tri_rec, uvs);
    else k_emit_tris<false><<<(n + 255) / 256, 256, 0, st>>>(tris_dev, n, order, tri_rec, uvs);
    return cudaGetLastError();
}

static cudaError_t read_counters(BvhResult& out, const BuildCounters& h, int n)
{
    auto dec = [](int i) { int j = i >= 0 ? i : i ^ 0x7fffffff; float f; memcpy(&f, &j, 4); return f; };
    for (int k = 0; k < 6; k++) out.bounds[k] = dec(h.bounds[k]);
    if (n <= 1) { out.num_nodes = 1; out.num_leaves = n; out.max_leaf = n; out.depth = 1; out.sah = 0.f; }
    else {
        out.num_nodes = h.n_wide; out.num_leaves = h.n_leaves; out.max_leaf = h.max_leaf; out.depth = std::max(h.max_level, h.max_level2) + 1; out.sah = h.sah;
        if (h.qcount[70] != 0) return cudaErrorLaunchFailure;      // tree deeper than the level loop: refuse, never truncate
    }
    return cudaSuccess;
}

size_t build_counters_bytes() { return sizeof(BuildCounters); }

cudaError_t build_bvh_finish(BvhResult& out, const void* host_counters)
{
    return read_counters(out, *(const BuildCounters*)host_counters, out.n);
}

cudaError_t build_bvh(const rtb_triangle* tris_dev, float4* bmin, float4* bmax, int n, int max_leaf, int key_bits, float sah_node_cost, BvhResult& out,
                      Arena& ws, cudaStream_t st, float4* tri_rec, float2* uvs, cudaEvent_t done, void* deferred_counters, int build_fused)
{
    const bool fused = build_fused != 0 && n >= FUSED_MIN_PRIMS;
    // outputs (grow-only)
    int need_nodes = n > 1 ? n : 1;
    if (out.node_cap < need_nodes) {
        if (out.nodes) cudaFree(out.nodes);
        out.nodes = nullptr; out.node_cap = 0;
        CK(cudaMalloc((void**)&out.nodes, sizeof(float4) * NODE_F4 * (size_t)need_nodes));
        out.node_cap = need_nodes;
    }
    if (out.order_cap < (n > 0 ? n : 1)) {
        if (out.order) cudaFree(out.order);
        out.order = nullptr; out.order_cap = 0;
        CK(cudaMalloc((void**)&out.order, sizeof(unsigned) * (size_t)(n > 0 ? n : 1)));
        out.order_cap = n > 0 ? n : 1;
    }
    out.n = n;
    if (max_leaf < 1) max_leaf = 1;
    if (max_leaf > MAX_LEAF) max_leaf = MAX_LEAF;
    const bool wide_keys = key_bits > 32;
    size_t bytes = 4096 + (size_t)(n + 64) * (size_t)((wide_keys ? 16 : 8) + 8 + 4 * 6 + 8 + 64 + 2 + 16 + 16 + 16) + (size_t)256 * 4 * ((n + RS_TILE - 1) / RS_TILE + 1) + 64 * 256;
    CK(ws.reserve(bytes));
    ws.reset();
    BuildCounters* ctr = ws.take<BuildCounters>(1);
    EmitJob job = { (tris_dev && tri_rec && uvs && n > 0) ? tris_dev : nullptr, tri_rec, uvs, false };
    k_init_counters<<<1, 96, 0, st>>>(ctr);
    if (n > 0) {
        const int gb = (n + 256 * BB_PER - 1) / (256 * BB_PER);
        if (tris_dev) k_scene_bounds<true><<<gb, 256, 0, st>>>(tris_dev, bmin, bmax, n, ctr);
        else k_scene_bounds<false><<<gb, 256, 0, st>>>(nullptr, bmin, bmax, n, ctr);
    }
    if (n <= 1) {
        k_trivial_root<<<1, 1, 0, st>>>(out.nodes, bmin, bmax, n);
        if (n == 1) k_iota<<<1, 32, 0, st>>>(out.order, 1);
    } else if (wide_keys) {
        CK(build_tree<unsigned long long>(bmin, bmax, n, max_leaf, 40, sah_node_cost, out, ws, ctr, st, fused, &job));
    } else {
        CK(build_tree<unsigned>(bmin, bmax, n, max_leaf, 32, sah_node_cost, out, ws, ctr, st, fused, &job));
    }
    // triangle records in leaf order: enqueued BEFORE the counters are read back, so the device does not idle while the host
    // wakes up from the synchronisation below (one host round trip per build instead of two)
    if (job.forked) CK(cudaStreamWaitEvent(st, ws.ev_join, 0));        // the records were written on the side stream: join before anything that follows
    else if (tris_dev && tri_rec && uvs && n > 0) CK(emit_records(tris_dev, n, out.order, tri_rec, uvs, st));
    if (deferred_counters) {
        // asynchronous build: the counters travel to pinned host memory on the stream, `done` is recorded after them and the
        // caller reads them with build_bvh_finish() once that event has completed -- no host synchronisation here
        CK(cudaMemcpyAsync(deferred_counters, ctr, sizeof(BuildCounters), cudaMemcpyDeviceToHost, st));
        if (done) CK(cudaEventRecord(done, st));
        return cudaSuccess;
    }
    if (done) CK(cudaEventRecord(done, st));
    BuildCounters h;
    CK(cudaMemcpyAsync(&h, ctr, sizeof h, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    CK(read_counters(out, h, n));
    return cudaSuccess;
}

}  // namespace rtb
