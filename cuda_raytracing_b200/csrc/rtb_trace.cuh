// rtb_trace.cuh -- two-level wide-BVH traversal with the reference's closest-hit / any-hit semantics.
//
// Replaces cast_ray (raycast.cu:21-142): the linear instance loop becomes a TLAS descent, the
// 48-byte AoS BVH2 nodes + per-leaf index lists become 128-byte BVH4 nodes + in-order triangle
// records, and the 12 sinf/cosf per (ray, instance) become one precomputed quaternion.  What is kept
// bit-for-bit is everything that decides the answer: the world->mesh ray transform (raycast.cu:33-45),
// the plane hit and barycentric test (TrianglePrimitive.hpp:62-79,151-185), the world-space distance
// (raycast.cu:98-104) and the accept rule (raycast.cu:107-109).  Box tests only cull and are
// conservative, so the BVH shape cannot change the result; exact distance ties are broken by
// (instance, original triangle index) instead of by the reference's traversal order.
#pragma once
#include "rtb_internal.cuh"

namespace rtb {

struct InstRay {                 // ray in the current space (world or instance)
    f3 o, d;
    // Slab-test constants (culling only).  Per axis a: tnear = near_a * idn_a - oin_a, tfar = far_a * idf_a - oif_a,
    // where near/far is the box plane the ray meets first/last (chosen by the sign of d_a through the load
    // offsets below, so no min/max per child).  Rounding slack is folded in: oin_a = o*id + 4 eps |o*id|,
    // idf/oif carry a (1 + 1e-6) factor and the same absolute slack with the opposite sign.  The test can
    // therefore only accept more than the exact-arithmetic test, never less, and it stays tight for rays with
    // a zero direction component (1/d clamped to 1e20).
    float idnx, idny, idnz, oinx, oiny, oinz;
    float idfx, idfy, idfz, oifx, oify, oifz;
    int nx, ny, nz;              // float4 index of the NEAR plane vector inside a node: x 0|1, y 2|3, z 4|5
};

XF float safe_rcp(float d)
{
    float a = fabsf(d) < 1e-20f ? copysignf(1e-20f, d) : d;
    return 1.0f / a;
}

XF void setup_slabs(InstRay& r)
{
    const float ix = safe_rcp(r.d.x), iy = safe_rcp(r.d.y), iz = safe_rcp(r.d.z);
    const float ox = r.o.x * ix, oy = r.o.y * iy, oz = r.o.z * iz;
    const float k = 1.000001f, e = 2.4e-7f;
    r.idnx = ix; r.idny = iy; r.idnz = iz;
    r.oinx = ox + e * fabsf(ox); r.oiny = oy + e * fabsf(oy); r.oinz = oz + e * fabsf(oz);
    r.idfx = ix * k; r.idfy = iy * k; r.idfz = iz * k;
    r.oifx = ox * k - e * fabsf(ox); r.oify = oy * k - e * fabsf(oy); r.oifz = oz * k - e * fabsf(oz);
    r.nx = ix < 0.0f ? 1 : 0; r.ny = iy < 0.0f ? 3 : 2; r.nz = iz < 0.0f ? 5 : 4;
}

// 4 slab tests against one wide node; keys[i] = (bits(tnear) & ~3) | i for hits, 0xffffffff for misses
XF void test_node(const float4* __restrict__ node, const InstRay& r, float tmax, unsigned keys[4], int4& refs)
{
    const float4 nxv = __ldg(node + r.nx), fxv = __ldg(node + (r.nx ^ 1));
    const float4 nyv = __ldg(node + r.ny), fyv = __ldg(node + (r.ny ^ 1));
    const float4 nzv = __ldg(node + r.nz), fzv = __ldg(node + (r.nz ^ 1));
    const float4 rf = __ldg(node + 6);
    refs = make_int4(__float_as_int(rf.x), __float_as_int(rf.y), __float_as_int(rf.z), __float_as_int(rf.w));
    const float ax[4] = { nxv.x, nxv.y, nxv.z, nxv.w }, bx[4] = { fxv.x, fxv.y, fxv.z, fxv.w };
    const float ay[4] = { nyv.x, nyv.y, nyv.z, nyv.w }, by[4] = { fyv.x, fyv.y, fyv.z, fyv.w };
    const float az[4] = { nzv.x, nzv.y, nzv.z, nzv.w }, bz[4] = { fzv.x, fzv.y, fzv.z, fzv.w };
    const int rfi[4] = { refs.x, refs.y, refs.z, refs.w };
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const float tnx = fmaf(ax[i], r.idnx, -r.oinx), tny = fmaf(ay[i], r.idny, -r.oiny), tnz = fmaf(az[i], r.idnz, -r.oinz);
        const float tfx = fmaf(bx[i], r.idfx, -r.oifx), tfy = fmaf(by[i], r.idfy, -r.oify), tfz = fmaf(bz[i], r.idfz, -r.oifz);
        const float tn = fmaxf(fmaxf(tnx, tny), fmaxf(tnz, 0.0f));
        const float tf = fminf(fminf(tfx, tfy), fminf(tfz, tmax));
        const bool hit = (tn <= tf) && (rfi[i] != EMPTY_REF);
        keys[i] = hit ? ((__float_as_uint(tn) & ~3u) | (unsigned)i) : 0xffffffffu;
    }
}

#define RTB_CSWAP(a, b) { unsigned _lo = min(a, b), _hi = max(a, b); a = _lo; b = _hi; }

struct InstXf {                  // instance transform, loaded when a TLAS leaf is entered
    float4 q_pose, q_inv;
    f3 pose_t, inv_scale, inv_t, scale;
    int mesh, material;
};

XF void load_instance(const float4* __restrict__ inst, int i, InstXf& x)
{
    const float4* p = inst + (size_t)i * INST_F4;
    float4 a = __ldg(p + 0), b = __ldg(p + 1), c = __ldg(p + 2), d = __ldg(p + 3), e = __ldg(p + 4), f = __ldg(p + 5);
    x.q_pose = a;
    x.pose_t = mk3(b.x, b.y, b.z); x.mesh = __float_as_int(b.w);
    x.inv_scale = mk3(c.x, c.y, c.z); x.material = __float_as_int(c.w);
    x.q_inv = d;
    x.inv_t = mk3(e.x, e.y, e.z);
    x.scale = mk3(f.x, f.y, f.z);
}

// Closest hit (ANY=false) or any hit (ANY=true; returns at the first accepted candidate, the
// commented lighting_pass early-out of raycast.cu:128-132).
//
// Control flow is strictly structured (while / while, no goto) so the warp reconverges at the top of the
// outer loop: lanes first walk interior nodes until each holds a leaf (or is finished), then all lanes
// process their leaves together.  `cur` is the node being visited: >= 0 interior, < 0 leaf, EMPTY_REF done.
template <bool ANY, bool COUNT>
__device__ __forceinline__ void trace(const SceneDev& sc, f3 org, f3 dir, Hit& hit, Counters* ctr)
{
    hit.min = FLT_MAX; hit.slot = -1; hit.prim = -1; hit.inst = -1; hit.u = 0.f; hit.v = 0.f;
    hit.loc = mk3(0.f, 0.f, 0.f);

    uint2 stack[STACK_CAP];
    int sp = 0;
    unsigned n_nodes = 0, n_tris = 0;
    bool overflow = false;
    unsigned guard = 0;

    InstRay r;                                  // ray in the current space; the world ray is (org, dir)
    r.o = org; r.d = dir;
    setup_slabs(r);
    // parametric bound: Euclidean distance = t * |d_world| for every instance (scale cancels), so a
    // subtree entered at t can only hold hits farther than t*|d|.  Slack covers the reference's
    // inverse-pose round-trip error.
    const float inv_len = 1.0002f * rsqrtf(dir.x * dir.x + dir.y * dir.y + dir.z * dir.z);
    float tmax = FLT_MAX;

    const bool single = sc.num_instances == 1;
    const float4* nodes = sc.tlas_nodes;
    const float4* tris = nullptr;
    bool in_blas = false;
    int blas_base = 0;
    int cur_inst = -1;
    InstXf xf;
    int cur = 0;                                // TLAS root

    // raycast.cu:33-45: r_dir = apply_euler(rotation, dir) * inv_scale ; r_org = apply_lre(pose, org) * inv_scale
    auto enter_instance = [&](int inst) {
        cur_inst = inst;
        load_instance(sc.inst, inst, xf);
        r.d = xmul3(xapply_quat(xf.q_pose, dir), xf.inv_scale);
        r.o = xmul3(xapply_quat(xf.q_pose, xsub(org, xf.pose_t)), xf.inv_scale);
        const MeshDev m = sc.meshes[xf.mesh];
        setup_slabs(r);
        nodes = m.nodes; tris = m.tris;
        in_blas = true; blas_base = sp;
        cur = 0;                                // BLAS root
    };
    auto pop = [&]() -> int {
        for (;;) {
            if (in_blas && sp == blas_base) {   // instance exhausted: back to world space
                if (single) return EMPTY_REF;
                in_blas = false; r.o = org; r.d = dir; setup_slabs(r); nodes = sc.tlas_nodes;
            }
            if (sp == 0) return EMPTY_REF;
            const uint2 e = stack[--sp];
            if (__uint_as_float(e.x) < tmax) return (int)e.y;
        }
    };

    if (single) enter_instance(0);

    while (cur != EMPTY_REF) {
        while (cur >= 0) {                      // interior nodes
            if (++guard > (1u << 22)) { overflow = true; cur = EMPTY_REF; break; }   // a frame can never hang the device
            unsigned k[4]; int4 refs;
            if (COUNT) n_nodes++;
            test_node(nodes + (size_t)cur * NODE_F4, r, tmax, k, refs);
            RTB_CSWAP(k[0], k[1]); RTB_CSWAP(k[2], k[3]); RTB_CSWAP(k[0], k[2]); RTB_CSWAP(k[1], k[3]); RTB_CSWAP(k[1], k[2]);
            if (k[0] == 0xffffffffu) { cur = pop(); continue; }
            const int rr[4] = { refs.x, refs.y, refs.z, refs.w };
            if (sp + 3 > STACK_CAP) { overflow = true; cur = EMPTY_REF; break; }
#pragma unroll
            for (int i = 3; i >= 1; i--) {      // far first, so the nearer children are popped first
                if (k[i] != 0xffffffffu) {
                    const int s = k[i] & 3;
                    const int cref = s == 0 ? rr[0] : (s == 1 ? rr[1] : (s == 2 ? rr[2] : rr[3]));
                    stack[sp++] = make_uint2(k[i] & ~3u, (unsigned)cref);
                }
            }
            const int s0 = k[0] & 3;
            cur = s0 == 0 ? rr[0] : (s0 == 1 ? rr[1] : (s0 == 2 ? rr[2] : rr[3]));
        }
        if (cur == EMPTY_REF) break;
        // leaf
        const unsigned bits = ~(unsigned)cur;
        const int first = (int)(bits >> 3), count = (int)(bits & 7u) + 1;
        if (!in_blas) {                         // TLAS leaf: exactly one instance slot
            enter_instance(__ldg(sc.tlas_index + first));
            continue;
        }
        bool found = false;
        for (int i = 0; i < count; i++) {
            const float4* q = tris + (size_t)(first + i) * TRI_F4;
            const float4 q0 = __ldg(q + 0), q1 = __ldg(q + 1), q2 = __ldg(q + 2), q3 = __ldg(q + 3);
            if (COUNT) n_tris++;
            const f3 n = mk3(q0.x, q0.y, q0.z);
            // TrianglePrimitive.hpp:64-68  denom = dot(ray.direction, normal); |denom| < 1e-6 (double compare) rejects.
            // raycast.cu:107-109 same_dir is the same expression and must be < 0.
            const float denom = xdot(r.d, n);
            if (!(denom < 0.0f) || fabsf(denom) <= 1e-6f) continue;
            const f3 A = mk3(q1.x, q1.y, q1.z);
            const float t = xd(xdot(xsub(A, r.o), n), denom);                 // TrianglePrimitive.hpp:70
            if (t < 0.0f) continue;                                           // :72
            const f3 p = xadd(r.o, xscale(r.d, t));                           // :76
            if (p.x == FLT_MAX) continue;                                     // raycast.cu:91
            const f3 e0 = mk3(q2.x, q2.y, q2.z), e1 = mk3(q3.x, q3.y, q3.z);
            const f3 v2 = xsub(p, A);                                         // TrianglePrimitive.hpp:154-156
            const float dot00 = q1.w, dot01 = q2.w, dot11 = q3.w;
            const float dot02 = xdot(e0, v2), dot12 = xdot(e1, v2);
            const float invDenom = xd(1.0f, xs(xm(dot00, dot11), xm(dot01, dot01)));   // :164
            const float u = xm(xs(xm(dot11, dot02), xm(dot01, dot12)), invDenom);
            const float v = xm(xs(xm(dot00, dot12), xm(dot01, dot02)), invDenom);
            if (!((u >= 0.0f) && (v >= 0.0f) && (xa(u, v) <= 1.0f))) continue;          // :169
            // raycast.cu:98-104 world location through inv_pose, Euclidean distance to the WORLD origin
            f3 loc = xmul3(p, xf.scale);
            loc = xapply_quat(xf.q_inv, xsub(loc, xf.inv_t));
            const float distance = xmagnitude(xsub(loc, org));
            const int prim = __float_as_int(q0.w);
            const bool closer = (hit.min == FLT_MAX) || (distance < hit.min) ||
                                (distance == hit.min && (cur_inst < hit.inst || (cur_inst == hit.inst && prim < hit.prim)));
            if (closer) {
                hit.min = distance; hit.slot = first + i; hit.prim = prim; hit.inst = cur_inst;
                hit.u = u; hit.v = v; hit.loc = loc;
                tmax = fmaf(distance, inv_len, 1e-3f * inv_len);
                found = true;
            }
        }
        if (ANY && found) break;
        cur = pop();
    }
    if (COUNT && ctr) {
        atomicAdd(&ctr->nodes, (unsigned long long)n_nodes);
        atomicAdd(&ctr->tris, (unsigned long long)n_tris);
    }
    if (overflow && ctr) ctr->overflow = 1;
}

}  // namespace rtb
