// rtb_build.cu -- on-device LBVH builder for sm_100a.
//
// Replaces the reference's single-threaded recursive host builder (BVHTree::fill, BVHTree.hpp:203-292,
// 15 re-scans per node) and its one-cudaMalloc-per-leaf upload (BVHTree.hpp:97-114,364-383):
//   prim boxes -> scene bounds -> Morton keys -> radix sort -> Karras 2012 radix tree -> bottom-up
//   AABB fit + SAH leaf decision -> top-down collapse to 128-byte BVH4 nodes -> triangle records.
// The tree only culls (rtb_trace.cuh keeps the reference's hit predicate), so it is free to differ
// from the reference's tree.  All stages are HBM-bound streaming kernels.
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
};

// ------------------------------------------------------------------ 1. primitive boxes + scene bounds
__global__ void k_init_counters(BuildCounters* c)
{
    int t = threadIdx.x;
    if (t < 3) { c->bounds[t] = f2ord(FLT_MAX); c->bounds[t + 3] = f2ord(-FLT_MAX); c->cbounds[t] = f2ord(FLT_MAX); c->cbounds[t + 3] = f2ord(-FLT_MAX); }
    if (t == 0) { c->n_wide = 1; c->n_leaves = 0; c->max_leaf = 0; c->max_level = 0; c->sah = 0.f; }
    if (t < 72) c->qcount[t] = 0;
    if (t == 72) { c->bar = 0u; c->top_count = 0; }
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
        unsigned ix = min(1023u, (unsigned)(nx * 1024.f)), iy = min(1023u, (unsigned)(ny * 1024.f)), iz = min(1023u, (unsigned)(nz * 1024.f));
        keys[i] = (K)((expand10(ix) << 2) | (expand10(iy) << 1) | expand10(iz));
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

// The nodes above the chunks: one thread per list entry, the classic climb (every node appears twice in the list, once per
// child; the second arrival fits it and goes on).  The boxes below were written by k_fit, a kernel boundary ago.
__global__ void __launch_bounds__(256) k_fit_top(const int* __restrict__ top_list, const BuildCounters* __restrict__ c, int n,
                          const int* __restrict__ left, const int* __restrict__ right, const int* __restrict__ parent,
                          const int2* __restrict__ range, float4* nb_min, float4* nb_max, int* visit,
                          int4* __restrict__ topo, int max_leaf, float c_node, float c_tri)
{
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
            bool leaf = cnt <= max_leaf && !(c_leaf > c_split);
            mn.w = h; mx.w = leaf ? c_leaf : c_split;
            nb_min[cur] = mn; nb_max[cur] = mx;
                topo[cur] = make_int4(lc, rc, __float_as_int(h), leaf ? 1 : 0);
            cur = parent[cur];
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
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        atomicAdd(bar, 1u);
        while (*(volatile unsigned*)bar < target) { }
        __threadfence();
    }
    __syncthreads();
}

#ifndef RTB_COLLAPSE_MINB
#define RTB_COLLAPSE_MINB 7
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
    int my_leaves = 0, my_maxleaf = 0, level = 0;
    int start = 0;                                                      // wide nodes of the levels before the next one
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
                    if (id < n - 1) { const int4 t = __ldg(topo + id); l = t.x; r = t.y; a = t.w ? -1.f : fmaxf(__int_as_float(t.z), 0.f); }
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
                    if (!inner[k]) {
                        int first, count; slot_range(id, n, range, first, count);
                        ref[k] = ~((first << 3) | (count - 1));
                        my_leaves++; my_maxleaf = max(my_maxleaf, count);
                    } else {
                        const int qi = base++;
                        q_out[qi] = make_int2(id, start + qi);
                        ref[k] = start + qi;
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
    if (blockIdx.x == 0 && threadIdx.x == 0) { c->max_level = level - 1; c->n_wide = start; }
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

template <typename K>
static cudaError_t build_tree(const float4* bmin, const float4* bmax, int n, int max_leaf, int key_bits, float sah_node_cost, BvhResult& out,
                              Arena& ws, BuildCounters* ctr, cudaStream_t st)
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
    k_hierarchy<K><<<g, 256, 0, st>>>(keys[cur], n, left, right, parent, range);
    CK(cudaMemsetAsync(visit, 0, sizeof(int) * (size_t)n, st));
    k_fit<<<(n + FIT_CHUNK - 1) / FIT_CHUNK, FIT_CHUNK, 0, st>>>(bmin, bmax, out.order, n, left, right, parent, range, nb_min, nb_max, visit, topo, max_leaf, sah_node_cost, 1.0f,
                                                                   (int*)q0, ctr);
    k_fit_top<<<148 * 4, 256, 0, st>>>((const int*)q0, ctr, n, left, right, parent, range, nb_min, nb_max, visit, topo, max_leaf, sah_node_cost, 1.0f);
    k_root_sah<<<1, 1, 0, st>>>(nb_max, ctr);
    // level 0 work item: open binary root into wide node 0
    int2 root_item = make_int2(0, 0);
    CK(cudaMemcpyAsync(q0, &root_item, sizeof(int2), cudaMemcpyHostToDevice, st));
    int one = 1;
    CK(cudaMemcpyAsync(&ctr->qcount[0], &one, sizeof(int), cudaMemcpyHostToDevice, st));
    int resident = 0;                                                   // CTAs of k_collapse that fit THIS device at once
    {
        int dev = 0, sms = 0, per_sm = 0;
        CK(cudaGetDevice(&dev));
        CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_collapse, 128, 0));
        resident = sms * std::max(1, std::min(per_sm, RTB_COLLAPSE_CAP));
    }
    // cooperative launch: the driver starts the grid only when every CTA can be resident at once, so the grid barrier cannot
    // deadlock against another persistent kernel (a second context building on another stream, a render in flight)
    float4* nodes_out = out.nodes;
    void* args[] = { &q0, &q1, &n, &left, &right, &range, &nb_min, &nb_max, &topo, &nodes_out, &ctr };
    CK(cudaLaunchCooperativeKernel((const void*)k_collapse, dim3(std::min(resident, (n + 127) / 128)), dim3(128), args, 0, st));
    return cudaGetLastError();
}

static cudaError_t read_counters(BvhResult& out, const BuildCounters& h, int n)
{
    auto dec = [](int i) { int j = i >= 0 ? i : i ^ 0x7fffffff; float f; memcpy(&f, &j, 4); return f; };
    for (int k = 0; k < 6; k++) out.bounds[k] = dec(h.bounds[k]);
    if (n <= 1) { out.num_nodes = 1; out.num_leaves = n; out.max_leaf = n; out.depth = 1; out.sah = 0.f; }
    else {
        out.num_nodes = h.n_wide; out.num_leaves = h.n_leaves; out.max_leaf = h.max_leaf; out.depth = h.max_level + 1; out.sah = h.sah;
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
                      Arena& ws, cudaStream_t st, float4* tri_rec, float2* uvs, cudaEvent_t done, void* deferred_counters)
{
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
    size_t bytes = 4096 + (size_t)(n + 64) * (size_t)((wide_keys ? 16 : 8) + 8 + 4 * 6 + 8 + 64 + 2 + 16 + 16) + (size_t)256 * 4 * ((n + RS_TILE - 1) / RS_TILE + 1) + 64 * 256;
    CK(ws.reserve(bytes));
    ws.reset();
    BuildCounters* ctr = ws.take<BuildCounters>(1);
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
        CK(build_tree<unsigned long long>(bmin, bmax, n, max_leaf, 40, sah_node_cost, out, ws, ctr, st));
    } else {
        CK(build_tree<unsigned>(bmin, bmax, n, max_leaf, 32, sah_node_cost, out, ws, ctr, st));
    }
    // triangle records in leaf order: enqueued BEFORE the counters are read back, so the device does not idle while the host
    // wakes up from the synchronisation below (one host round trip per build instead of two)
    if (tris_dev && tri_rec && uvs && n > 0) {
        if (((uintptr_t)tris_dev & 7u) == 0) k_emit_tris<true><<<(n + 255) / 256, 256, 0, st>>>(tris_dev, n, out.order, tri_rec, uvs);
        else k_emit_tris<false><<<(n + 255) / 256, 256, 0, st>>>(tris_dev, n, out.order, tri_rec, uvs);
    }
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
