// rtb_internal.cuh -- device-side data layout shared by the builder, the tracer and the C-ABI.
//
// HBM layout (all records 16-byte aligned float4, SoA inside a record):
//   wide node   112 B = 7 x float4 : lo.x[4] hi.x[4] lo.y[4] hi.y[4] lo.z[4] hi.z[4] ref[4]  (no pad: see NODE_F4 below)
//               empty slots hold (+inf,+inf) boxes and ref = EMPTY_REF.
//               ref >= 0  : index of a wide node;  ref < 0 : leaf, ~ref = (first << 3) | (count-1)
//   triangle     64 B = 4 x float4, stored in leaf (Morton) order, never straddles a line:
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

// float4 per wide node.  7, not 8: with a 128-byte stride every lane of a warp reads field k of ITS node from the same four
// L1 banks (32-way conflict, 32 cycles per LDG.128: tools/l1_probe.cu); a 112-byte stride puts field k of node n in bank
// group (k - n) mod 8, so lanes on different nodes spread over all banks.
constexpr int NODE_F4 = 7;
constexpr int TRI_F4 = 4;           // float4 per triangle record
constexpr int INST_F4 = 6;          // float4 per instance record
constexpr int MAX_LEAF = 8;         // triangles per leaf (3-bit count)
constexpr int STACK_CAP = 96;       // traversal stack entries per ray: worst case 3 per tree level (rtb_api.cu make_scene refuses deeper scenes).
                                    // Per-thread local memory: only the entries a ray reaches are ever touched, so the size costs nothing;
                                    // 64 refused C3 once the 1 M-triangle BLAS came out 16 levels deep (3 * 5 + 1 + 3 * 16 + 2 = 66)
constexpr int EMPTY_REF = (int)0x80000000;   // child slot without a subtree
constexpr int BAND_ROWS = 8;        // image rows per multi-GPU band

struct MeshDev {
    const float4* nodes;            // wide nodes, root = 0
    const float4* tris;             // triangle records in leaf order
    const float2* uvs;              // 3 x float2 per triangle in leaf order, like `tris` (TrianglePrimitive::uv_coords)
    int num_triangles;
    float extent;                   // largest |coordinate| of the mesh bounds
};

struct SceneDev {
    const float4* tlas_nodes;       // wide nodes over instance world boxes (leaf = one instance slot)
    const int* tlas_index;          // leaf slot -> instance index
    const float4* inst;             // INST_F4 float4 per instance
    const MeshDev* meshes;
    const rtb_material* materials;
    int num_instances;
    int num_materials;
    float tlas_extent;              // largest |coordinate| of any instance world box: scales the pruning slack (TraceState::begin)
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
