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
    float idx, idy, idz;         // reciprocal direction for slab tests (culling only)
    float oix, oiy, oiz;         // o * inv_d
};

XF float safe_rcp(float d)
{
    float a = fabsf(d) < 1e-20f ? copysignf(1e-20f, d) : d;
    return 1.0f / a;
}

XF void setup_slabs(InstRay& r)
{
    r.idx = safe_rcp(r.d.x); r.idy = safe_rcp(r.d.y); r.idz = safe_rcp(r.d.z);
    r.oix = r.o.x * r.idx; r.oiy = r.o.y * r.idy; r.oiz = r.o.z * r.idz;
}

// 4 slab tests against one wide node; keys[i] = (bits(tnear) & ~3) | i for hits, 0xffffffff for misses
XF void test_node(const float4* __restrict__ node, const InstRay& r, float tmax, unsigned keys[4], int4& refs)
{
    const float4 lox = __ldg(node + 0), hix = __ldg(node + 1);
    const float4 loy = __ldg(node + 2), hiy = __ldg(node + 3);
    const float4 loz = __ldg(node + 4), hiz = __ldg(node + 5);
    const float4 rf = __ldg(node + 6);
    refs = make_int4(__float_as_int(rf.x), __float_as_int(rf.y), __float_as_int(rf.z), __float_as_int(rf.w));
    const float lx[4] = { lox.x, lox.y, lox.z, lox.w }, hx[4] = { hix.x, hix.y, hix.z, hix.w };
    const float ly[4] = { loy.x, loy.y, loy.z, loy.w }, hy[4] = { hiy.x, hiy.y, hiy.z, hiy.w };
    const float lz[4] = { loz.x, loz.y, loz.z, loz.w }, hz[4] = { hiz.x, hiz.y, hiz.z, hiz.w };
#pragma unroll
    for (int i = 0; i < 4; i++) {
        float t0x = fmaf(lx[i], r.idx, -r.oix), t1x = fmaf(hx[i], r.idx, -r.oix);
        float t0y = fmaf(ly[i], r.idy, -r.oiy), t1y = fmaf(hy[i], r.idy, -r.oiy);
        float t0z = fmaf(lz[i], r.idz, -r.oiz), t1z = fmaf(hz[i], r.idz, -r.oiz);
        float tn = fmaxf(fmaxf(fminf(t0x, t1x), fminf(t0y, t1y)), fmaxf(fminf(t0z, t1z), 0.0f));
        float tf = fminf(fminf(fmaxf(t0x, t1x), fmaxf(t0y, t1y)), fmaxf(t0z, t1z));
        // conservative: widen the far side by a few ulps and the near side likewise (robust traversal)
        const int rfi = i == 0 ? refs.x : (i == 1 ? refs.y : (i == 2 ? refs.z : refs.w));
        bool hit = (tn * 0.9999995f <= tf * 1.0000005f) && (tn * 0.9999995f < tmax) && (rfi != EMPTY_REF);
        keys[i] = hit ? ((__float_as_uint(tn * 0.9999995f) & ~3u) | (unsigned)i) : 0xffffffffu;
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
template <bool ANY, bool COUNT>
__device__ __noinline__ void trace(const SceneDev& sc, f3 org, f3 dir, Hit& hit, Counters* ctr)
{
    hit.min = FLT_MAX; hit.slot = -1; hit.prim = -1; hit.inst = -1; hit.u = 0.f; hit.v = 0.f;
    hit.loc = mk3(0.f, 0.f, 0.f);

    uint2 stack[STACK_CAP];
    int sp = 0;
    unsigned n_nodes = 0, n_tris = 0;
    bool overflow = false;
    unsigned guard = 0;

    InstRay wr;                                 // world ray
    wr.o = org; wr.d = dir;
    setup_slabs(wr);
    InstRay r = wr;
    // parametric bound: Euclidean distance = t * |d_world| for every instance (scale cancels), so a
    // subtree entered at t can only hold hits farther than t*|d|.  Slack covers the reference's
    // inverse-pose round-trip error.
    const float inv_len = 1.0002f * rsqrtf(dir.x * dir.x + dir.y * dir.y + dir.z * dir.z);
    float tmax = FLT_MAX;

    const float4* nodes = sc.tlas_nodes;
    bool in_blas = false;
    int blas_base = 0;
    InstXf xf;
    const float4* tris = nullptr;
    int cur_inst = -1;

    if (sc.num_instances == 1) {                // single instance: no TLAS level
        cur_inst = 0;
        load_instance(sc.inst, 0, xf);
        goto enter_instance;
    }
    stack[sp++] = make_uint2(0u, 0u);

    while (true) {
        if (++guard > (1u << 22)) { overflow = true; break; }   // a frame can never hang the device
        if (in_blas && sp == blas_base) {       // instance exhausted: back to world space
            if (sc.num_instances == 1) break;
            in_blas = false; r = wr; nodes = sc.tlas_nodes;
        }
        if (sp == 0) break;
        {
            uint2 e = stack[--sp];
            if (__uint_as_float(e.x) >= tmax) continue;
            int ref = (int)e.y;
            if (ref >= 0) {
                unsigned k[4]; int4 refs;
                if (COUNT) n_nodes++;
                test_node(nodes + (size_t)ref * NODE_F4, r, tmax, k, refs);
                RTB_CSWAP(k[0], k[1]); RTB_CSWAP(k[2], k[3]); RTB_CSWAP(k[0], k[2]); RTB_CSWAP(k[1], k[3]); RTB_CSWAP(k[1], k[2]);
                const int rr[4] = { refs.x, refs.y, refs.z, refs.w };
                if (sp + 4 > STACK_CAP) { overflow = true; break; }
#pragma unroll
                for (int i = 3; i >= 0; i--) {  // far first, so the nearest child is popped next
                    if (k[i] != 0xffffffffu) {
                        int s = k[i] & 3;
                        int cref = s == 0 ? rr[0] : (s == 1 ? rr[1] : (s == 2 ? rr[2] : rr[3]));
                        stack[sp++] = make_uint2(k[i] & ~3u, (unsigned)cref);
                    }
                }
                continue;
            }
            // leaf
            unsigned bits = ~(unsigned)ref;
            int first = (int)(bits >> 3), count = (int)(bits & 7u) + 1;
            if (!in_blas) {                     // TLAS leaf: exactly one instance slot
                cur_inst = __ldg(sc.tlas_index + first);
                load_instance(sc.inst, cur_inst, xf);
                goto enter_instance;
            }
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
                bool closer = (hit.min == FLT_MAX) || (distance < hit.min) ||
                              (distance == hit.min && (cur_inst < hit.inst || (cur_inst == hit.inst && prim < hit.prim)));
                if (closer) {
                    hit.min = distance; hit.slot = first + i; hit.prim = prim; hit.inst = cur_inst;
                    hit.u = u; hit.v = v; hit.loc = loc;
                    tmax = fmaf(distance, inv_len, 1e-3f * inv_len);
                    if (ANY) goto done;
                }
            }
            continue;
        }
    enter_instance:
        {
            // raycast.cu:33-45: r_dir = apply_euler(rotation, dir) * inv_scale ; r_org = apply_lre(pose, org) * inv_scale
            r.d = xmul3(xapply_quat(xf.q_pose, wr.d), xf.inv_scale);
            r.o = xmul3(xapply_quat(xf.q_pose, xsub(wr.o, xf.pose_t)), xf.inv_scale);
            setup_slabs(r);
            const MeshDev m = sc.meshes[xf.mesh];
            nodes = m.nodes; tris = m.tris;
            in_blas = true; blas_base = sp;
            if (sp + 1 > STACK_CAP) { overflow = true; break; }
            stack[sp++] = make_uint2(0u, 0u);
        }
    }
done:
    if (COUNT && ctr) {
        atomicAdd(&ctr->nodes, (unsigned long long)n_nodes);
        atomicAdd(&ctr->tris, (unsigned long long)n_tris);
    }
    if (overflow && ctr) ctr->overflow = 1;
}

}  // namespace rtb
