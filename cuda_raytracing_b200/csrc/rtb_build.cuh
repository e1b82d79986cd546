// rtb_build.cuh -- host interface of the on-device LBVH builder (rtb_build.cu).
#pragma once
#include "rtb_internal.cuh"
#include <vector>

namespace rtb {

// Grow-only device arena so per-frame rebuilds (C5) do not call cudaMalloc.
struct Arena {
    char* base = nullptr;
    size_t cap = 0, used = 0;
    cudaError_t reserve(size_t bytes);
    template <typename T> T* take(size_t count)
    {
        size_t off = (used + 255) & ~(size_t)255;
        used = off + count * sizeof(T);
        return used <= cap ? (T*)(base + off) : nullptr;
    }
    void reset() { used = 0; }
    void release();
    // A second stream of the builder that owns this arena (created on first use, destroyed by release()): the triangle records are
    // written there while the latency-bound top of the tree is made on the caller's stream (rtb_build.cu build_bvh).
    cudaStream_t side = nullptr;
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
};

struct BvhResult {
    float4* nodes = nullptr;        // device, capacity node_cap wide nodes
    unsigned* order = nullptr;      // device, sorted slot -> primitive index
    int node_cap = 0, order_cap = 0;
    int n = 0, num_nodes = 0, num_leaves = 0, max_leaf = 0, depth = 0;
    float bounds[6] = { 0, 0, 0, 0, 0, 0 };
    float sah = 0.f;
};

// Builds a wide BVH over n primitive boxes (device float4 arrays, w ignored).  With `tris_dev` the boxes are first made from
// those n triangles (written to bmin/bmax, fused with the scene-bounds reduction); without, bmin/bmax are the input.
// With `tri_rec`/`uvs` as well, the 64-byte triangle records and the uvs are written in leaf order (k_emit_tris).
// `done` (optional) is recorded after the last kernel.  Asynchronous on `st` except for one small read-back of the build
// counters at the end.
// build_fused: fit + collapse of everything below the 256-leaf chunks in one kernel (k_fit_collapse); node 0 is the root, the
// other nodes are NOT in breadth-first order then (chunk-local subtrees are contiguous, the top tree's nodes come last).
// sah_node_cost: cost of one wide-node child box relative to one triangle test in the SAH leaf decision (option "sah_node").
cudaError_t build_bvh(const rtb_triangle* tris_dev, float4* bmin, float4* bmax, int n, int max_leaf, int key_bits, float sah_node_cost, BvhResult& out,
                      Arena& ws, cudaStream_t st, float4* tri_rec = nullptr, float2* uvs = nullptr, cudaEvent_t done = nullptr,
                      void* deferred_counters = nullptr, int build_fused = 1);
// Asynchronous form: with `deferred_counters` (pinned host memory of build_counters_bytes()) build_bvh does not synchronise; the
// counters are copied there on the stream BEFORE `done` is recorded, and build_bvh_finish() turns them into out.num_nodes,
// depth, bounds ... once the caller has seen `done` complete.
size_t build_counters_bytes();
cudaError_t build_bvh_finish(BvhResult& out, const void* host_counters);


}  // namespace rtb
