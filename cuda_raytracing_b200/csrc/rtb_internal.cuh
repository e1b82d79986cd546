// rtb_internal.cuh -- device-side data layout shared by the builder, the tracer and the C-ABI.
//
// HBM layout (all records 16-byte aligned float4, SoA inside a record):
//   wide node    80 B = 5 x float4: an 8-wide node whose child boxes are quantised to 8 bits on the node's own box
//               (compressed wide BVH, Ylitie et al. 2017; the layout and the traversal are this library's own):
//               n0 = origin.xyz, bits(ex | ey << 8 | ez << 16 | imask << 24)
//                    cell size of axis a = 2^(e_a - 127 - 15): the stored exponent is that of cell * 2^15, which is what the
//                    decoder multiplies with (rtb_trace.cuh); imask bit k = child k is an inner node
//               n1 = child_base, tri_base, meta[0..3], meta[4..7]
//                    inner child k is wide node child_base + popc(imask & ((1 << k) - 1)) -- the inner children of a node
//                    are contiguous, in slot order; meta[k] of a leaf child = 0x80 | (count - 1) << 5 | offset: its
//                    triangles are records tri_base + offset .. + count - 1 (count <= 4, so offset <= 28); 0 otherwise
//               n2 = lo.x[0..3] lo.x[4..7] hi.x[0..3] hi.x[4..7]   one byte per child and plane:
//               n3 = the same for y                                 plane = origin + q * cell, snapped outwards and
//               n4 = the same for z                                 verified in double precision by the builder
//               Children sit in the slot whose octant (bit a set = on the + side of axis a) best matches their position, so a
//               ray visits the slots in the order of descending (slot ^ octinv) -- no per-visit sort (rtb_trace_state.cuh).
//               At an 80-byte stride field k of node n sits in 16-byte bank group (5 n + k) mod 8: lanes on different nodes
//               spread over the L1 banks (the reason the 4-wide node of round 1 was 112 and not 128 bytes, tools/l1_probe.cu).
//   triangle     64 B = 4 x float4, stored node by node (the leaf children of one wide node are contiguous), never straddles a line:
//               q0 = (n.xyz, bits(original index))          n  = stored face normal
//               q1 = (A.xyz, dot00)   A = vertices[0]        dot00 = (C-A).(C-A)
//               q2 = (C-A,   dot01)                          dot01 = (C-A).(B-A)
//               q3 = (B-A,   dot11)                          dot11 = (B-A).(B-A)
//               i.e. exactly the quantities TrianglePrimitive::ray_intersect / point_inside
//               (TrianglePrimitive.hpp:62-79,151-185) recompute per test, precomputed with the same
//               unfused float ops so the per-ray predicate is bit-identical to the reference's.
//   instance     96 B = 6 x float4 : q_pose | pose.xyz,mesh | inv_scale,material | q_inv | inv_pose.xyz | scale
//               q_* = euler2quat (transforms.hpp:148-163) evaluated once on the host instead of
//               12 sinf/cosf per ray per instance (raycast.cu:33,40).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <float.h>
#include "../../include/rtb200.h"
#include "exact.cuh"

namespace rtb {

constexpr int NODE_F4 = 5;          // float4 per wide node (80 bytes, see above)
constexpr int TRI_F4 = 4;           // float4 per triangle record
constexpr int INST_F4 = 6;          // float4 per instance record
constexpr int MAX_LEAF = 4;         // triangles per leaf child (2-bit count); 8 leaf children x 4 = 32 offsets fit the 5-bit offset
constexpr int STACK_CAP = 24;       // traversal stack entries (32 bytes each) per ray: one per tree level + one per instance entered;
                                    // scenes whose trees need more are refused at upload (rtb_api.cu: stack_need)
constexpr int EMPTY_REF = (int)0x80000000;   // TraceState::cur of a lane without a ray
constexpr int BAND_ROWS = 8;        // image rows per multi-GPU band

struct MeshDev {
    const float4* nodes;            // wide nodes, root = 0
    const float4* tris;             // triangle records in leaf order
    const float2* uvs;              // 3 x float2 per triangle in leaf order, like `tris` (TrianglePrimitive::uv_coords)
    int num_triangles;
    float extent;                   // largest |coordinate| of the mesh bounds: scales the rounding slack of the slab test (setup_slabs)
};

struct SceneDev {
    const float4* tlas_nodes;       // wide nodes over instance world boxes (leaf = one instance slot)
    const int* tlas_index;          // leaf slot -> instance index
    const float4* inst;             // INST_F4 float4 per instance
    const MeshDev* meshes;
    const rtb_material* materials;
    int num_instances;
    int num_materials;
    float tlas_extent;              // largest |coordinate| of any instance world box: slab-test slack in world space, pruning slack
    // top tree levels staged into shared memory by the persistent kernels (rtb_tma.cuh); 0 counts = not staged
    const float4* top_blas;         // node array of the BLAS whose first top_blas_n nodes are staged
    int top_tlas_n, top_blas_n;
    int spec_thr;                   // speculative traversal: leave the interior loop once at most this many lanes hold no postponed leaf
};

struct FrameParams {
    int width, height;
    float K_inv[9];
    float D[4];
    float ox, oy, oz;               // camera pose xyz (raycast.cu:156)
    float4 q_cam;                   // euler2quat(inv_camera_pose yaw,pitch,roll) (raycast.cu:185), host-evaluated
    int mode, max_depth, spp;
    uint32_t seed;
    int tile_rank, tile_count;
    float3 light_dir;               // normalize(-0.2,0,1) (raycast.cu:249-250), host-evaluated with Q_rsqrt
};

struct Counters {                   // device-side frame counters
    unsigned long long rays, nodes, tris;
    int overflow;
    int pad;
};

struct Hit {
    float min;                      // HitInfo::min: world-space Euclidean distance (raycast.cu:104,110)
    int slot;                       // triangle record index (leaf order) within the mesh
    int prim;                       // original triangle index
    int inst;                       // instance index
    float u, v;                     // barycentrics of point_inside (TrianglePrimitive.hpp:165-166)
    f3 loc;                         // world-space hit location (raycast.cu:98-102)
};

}  // namespace rtb
