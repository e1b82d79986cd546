// rtb_trace_state.cuh -- per-lane traversal state machine (included by rtb_trace.cuh).
//
// One ray per lane.  `cur` is the node being visited: >= 0 interior, < 0 leaf, EMPTY_REF = no ray / finished.
// advance() = "walk interior nodes until this lane holds a leaf, process that leaf, pop the next entry".
// Control flow is strictly structured (while / while, no goto) so the warp reconverges after every advance():
// lanes first walk interior nodes until each holds a leaf (or is finished), then all lanes process their
// leaves together.  Two drivers use it:
//   * trace<>()            one ray to completion (direct kernel, modes 0/1);
//   * k_trace_queue<>      persistent warps that refill finished lanes from a ray queue (wavefront, rtb_wavefront.cu).
#pragma once

namespace rtb {

struct TraceState {
    f3 org, dir;                 // world ray
    InstRay r;                   // ray in the current space
    float inv_len, tmax;
    Hit hit;
    const float4* nodes;
    const float4* tris;
    // only what the hit block needs of the instance transform stays live (raycast.cu:98-104)
    float4 q_inv;
    f3 inv_t, scale;
    int cur, sp, blas_base, cur_inst;
    bool in_blas;
    unsigned n_nodes, n_tris, guard;
    bool overflow;
    uint2 stack[STACK_CAP];

    XF void clear() { cur = EMPTY_REF; sp = 0; n_nodes = n_tris = guard = 0; overflow = false; in_blas = false; blas_base = 0; cur_inst = -1; }

    // raycast.cu:33-45: r_dir = apply_euler(rotation, dir) * inv_scale ; r_org = apply_lre(pose, org) * inv_scale
    XF void enter_instance(const SceneDev& sc, int inst)
    {
        InstXf xf;
        cur_inst = inst;
        load_instance(sc.inst, inst, xf);
        r.d = xmul3(xapply_quat(xf.q_pose, dir), xf.inv_scale);
        r.o = xmul3(xapply_quat(xf.q_pose, xsub(org, xf.pose_t)), xf.inv_scale);
        setup_slabs(r);
        const MeshDev m = sc.meshes[xf.mesh];
        nodes = m.nodes; tris = m.tris;
        q_inv = xf.q_inv; inv_t = xf.inv_t; scale = xf.scale;
        in_blas = true; blas_base = sp;
        cur = 0;                                // BLAS root
    }

    XF void begin(const SceneDev& sc, f3 o, f3 d)
    {
        org = o; dir = d;
        hit.min = FLT_MAX; hit.slot = -1; hit.prim = -1; hit.inst = -1; hit.u = 0.f; hit.v = 0.f; hit.loc = mk3(0.f, 0.f, 0.f);
        sp = 0; in_blas = false; blas_base = 0; cur_inst = -1;
        // parametric bound: Euclidean distance = t * |d_world| for every instance (scale cancels), so a subtree entered
        // at t can only hold hits farther than t*|d|.  Slack covers the reference's inverse-pose round-trip error.
        inv_len = 1.0002f * rsqrtf(d.x * d.x + d.y * d.y + d.z * d.z);
        tmax = FLT_MAX;
        if (sc.num_instances == 1) {            // single instance: no TLAS level
            enter_instance(sc, 0);
        } else {
            r.o = o; r.d = d;
            setup_slabs(r);
            nodes = sc.tlas_nodes; tris = nullptr;
            cur = 0;                            // TLAS root
        }
    }

    XF int pop(const SceneDev& sc)
    {
        for (;;) {
            if (in_blas && sp == blas_base) {   // instance exhausted: back to world space
                if (sc.num_instances == 1) return EMPTY_REF;
                in_blas = false; r.o = org; r.d = dir; setup_slabs(r); nodes = sc.tlas_nodes;
            }
            if (sp == 0) return EMPTY_REF;
            const uint2 e = stack[--sp];
            if (__uint_as_float(e.x) < tmax) return (int)e.y;
        }
    }

    template <bool ANY, bool COUNT>
    XF void advance(const SceneDev& sc)
    {
        while (cur >= 0) {                      // interior nodes
            if (++guard > (1u << 22)) { overflow = true; cur = EMPTY_REF; break; }   // a frame can never hang the device
            unsigned k[4]; int4 refs;
            if (COUNT) n_nodes++;
            test_node(nodes + (size_t)cur * NODE_F4, r, tmax, k, refs);
            RTB_CSWAP(k[0], k[1]); RTB_CSWAP(k[2], k[3]); RTB_CSWAP(k[0], k[2]); RTB_CSWAP(k[1], k[3]); RTB_CSWAP(k[1], k[2]);
            if (k[0] == 0xffffffffu) { cur = pop(sc); continue; }
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
        if (cur == EMPTY_REF) return;
        // leaf
        const unsigned bits = ~(unsigned)cur;
        const int first = (int)(bits >> 3), count = (int)(bits & 7u) + 1;
        if (!in_blas) {                         // TLAS leaf: exactly one instance slot
            enter_instance(sc, __ldg(sc.tlas_index + first));
            return;
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
            f3 loc = xmul3(p, scale);
            loc = xapply_quat(q_inv, xsub(loc, inv_t));
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
        if (ANY && found) { cur = EMPTY_REF; return; }
        cur = pop(sc);
    }

    template <bool COUNT>
    XF void flush_counters(Counters* ctr)
    {
        if (COUNT && ctr) {
            atomicAdd(&ctr->nodes, (unsigned long long)n_nodes);
            atomicAdd(&ctr->tris, (unsigned long long)n_tris);
        }
        if (overflow && ctr) ctr->overflow = 1;
        n_nodes = n_tris = 0;
    }
};

// Closest hit (ANY=false) or any hit (ANY=true; stops at the first accepted candidate, the commented
// lighting_pass early-out of raycast.cu:128-132) for one ray.
template <bool ANY, bool COUNT>
__device__ __forceinline__ void trace(const SceneDev& sc, f3 org, f3 dir, Hit& hit, Counters* ctr)
{
    TraceState S;
    S.clear();
    S.begin(sc, org, dir);
    while (S.cur != EMPTY_REF) S.template advance<ANY, COUNT>(sc);
    S.template flush_counters<COUNT>(ctr);
    hit = S.hit;
}

}  // namespace rtb
