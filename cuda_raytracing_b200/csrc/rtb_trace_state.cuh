// rtb_trace_state.cuh -- per-lane traversal state machine (included by rtb_trace.cuh).
//
// One ray per lane.  The lane walks compressed 8-wide nodes (rtb_internal.cuh).  What it still has to look at of a node is a
// GROUP: (first inner child, node index, 8-bit mask of the children its box test let through, their entry parameters as
// bf16).  The next child is the one whose slot maximises slot ^ octinv -- the builder put every child in the slot whose
// octant matches its position, so this is front to back without a sort (one lookup in a 2 KB shared-memory table) -- and it
// is skipped when its entry parameter is no longer below tmax.  When the lane descends and the group still has children
// it goes to the stack: ONE 32-byte entry per level instead of one entry per child.
//
// `cur` says what the lane does next: WALK (pick the next child), EMPTY_REF (no ray / finished) or an EVENT: TRIS (two
// triangle leaves are waiting), ENTER (a TLAS leaf = instance was picked), LEAVE (the sentinel pushed when an instance was
// entered came off the stack: back to world space).  advance() = "walk until this lane holds an event, then handle it".
// Control flow is strictly structured (while / if, no goto) so the warp reconverges after every advance(): lanes first
// walk until each holds an event, then all lanes handle their events together.  Every space change (enter / leave an
// instance) is an event, so it never serialises the walk.
//
// The state is scalars only -- the traversal stack is a separate array handed in by pointer -- so that the whole state
// lives in registers (an array member with a dynamic index pins the entire struct to local memory).  What is only needed
// when a candidate passes the inside test (instance inverse pose, scale) is re-read from the 96-byte instance record at
// that point instead of being carried in registers through the traversal.
//
// Two drivers use it:
//   * trace<>()            one ray to completion (direct kernel, modes 0/1);
//   * persistent kernels   that refill finished lanes (rtb_wavefront.cu, rtb_mega.cu).
#pragma once

namespace rtb {

constexpr int WALK = 0;
constexpr int EV_TRIS = 1, EV_ENTER = 2, EV_LEAVE = 3;
constexpr unsigned NO_LEAF = 0xffffffffu;

// Traversal stack of 32-byte group entries in local memory: a = (first inner child, node, cand | imask << 8, -), b = entry
// parameters; the sentinel has node = 0xffffffff.
struct Stack {
    uint4 a[STACK_CAP], b[STACK_CAP];
};
#define RTB_DECL_STACK(name) Stack name

// Child order table: pick[octinv << 8 | mask] = the set bit k of mask with the largest k ^ octinv (0 for an empty mask).
// 2 KB of static shared memory per CTA; every kernel that traces calls init_pick_lut() once, before its first ray.
XF unsigned char* pick_lut()
{
    __shared__ unsigned char lut[8 * 256];
    return lut;
}
XF void init_pick_lut()
{
    unsigned char* L = pick_lut();
    for (int i = threadIdx.x; i < 8 * 256; i += blockDim.x) {
        const int o = i >> 8, m = i & 255;
        int best = 0, bp = -1;
        for (int k = 0; k < 8; k++) if (((m >> k) & 1) && (k ^ o) > bp) { bp = k ^ o; best = k; }
        L[i] = (unsigned char)best;
    }
    __syncthreads();
}

// Hit record kept in registers (direct kernel).  The persistent kernels use a sink that writes accepted candidates
// straight to the pixel's hit record in HBM and keeps only what the accept rule compares (rtb_mega.cu, rtb_wavefront.cu).
struct RegSink {
    Hit hit;
    XF void reset() { hit.min = FLT_MAX; hit.slot = -1; hit.prim = -1; hit.inst = -1; hit.u = 0.f; hit.v = 0.f; hit.loc = mk3(0.f, 0.f, 0.f); }
    XF float min() const { return hit.min; }
    XF int inst() const { return hit.inst; }
    XF int prim() const { return hit.prim; }
    XF bool any_hit() const { return false; }
    XF void accept(float distance, int slot, int prim, int inst, float u, float v, f3 loc)
    {
        hit.min = distance; hit.slot = slot; hit.prim = prim; hit.inst = inst; hit.u = u; hit.v = v; hit.loc = loc;
    }
};

struct TraceState {
    f3 org, dir;                 // world ray
    InstRay r;                   // ray in the current space
    float inv_len, tmax, tslack;
    const float4* nodes;
    const float4* tris;
    unsigned g_cb, g_pn, g_hm, g_t01, g_t23, g_t45, g_t67;     // current group
    int cur, sp, cur_inst;
    bool in_blas;
    unsigned n_nodes, n_tris, guard;
    bool overflow;

    XF void clear() { cur = EMPTY_REF; sp = 0; n_nodes = n_tris = guard = 0; overflow = false; in_blas = false; cur_inst = -1; g_hm = 0; }

    // the group "child 0 of nothing": picking it visits node 0 of the current node array
    XF void root_group() { g_cb = 0u; g_pn = 0u; g_hm = 1u | (1u << 8); g_t01 = g_t23 = g_t45 = g_t67 = 0u; }

    // raycast.cu:33-45: r_dir = apply_euler(rotation, dir) * inv_scale ; r_org = apply_lre(pose, org) * inv_scale
    XF void enter_instance(const SceneDev& sc, int inst, Stack& stk, bool sentinel)
    {
        const float4* p = sc.inst + (size_t)inst * INST_F4;
        const float4 q_pose = __ldg(p + 0), b = __ldg(p + 1), c = __ldg(p + 2);
        const f3 inv_scale = mk3(c.x, c.y, c.z);
        cur_inst = inst;
        r.d = xmul3(xapply_quat(q_pose, dir), inv_scale);
        r.o = xmul3(xapply_quat(q_pose, xsub(org, mk3(b.x, b.y, b.z))), inv_scale);
        const MeshDev* m = sc.meshes + __float_as_int(b.w);
        setup_slabs(r, m->extent);
        nodes = m->nodes; tris = m->tris;
        in_blas = true;
        if (sentinel) { stk.a[sp] = make_uint4(0u, 0xffffffffu, 0u, 0u); sp++; }
        root_group();
        cur = WALK;
    }

    XF void begin(const SceneDev& sc, f3 o, f3 d, Stack& stk)
    {
        org = o; dir = d;
        sp = 0; in_blas = false; cur_inst = -1;
        // parametric bound: Euclidean distance = t * |d_world| for every instance (scale cancels), so a subtree entered
        // at t can only hold hits farther than t*|d|.  Slack covers the reference's inverse-pose round-trip error,
        // which grows with the magnitude of the coordinates involved, not with the distance.
        const float il = rsqrtf(d.x * d.x + d.y * d.y + d.z * d.z);
        inv_len = 1.0002f * il;
        tslack = (1e-3f + 1.2e-6f * (sc.tlas_extent + fmaxf(fmaxf(fabsf(o.x), fabsf(o.y)), fabsf(o.z)))) * il;
        tmax = FLT_MAX;
        if (sc.num_instances == 1) {            // single instance: no TLAS level, no sentinel
            enter_instance(sc, 0, stk, false);
        } else {
            r.o = o; r.d = d;
            setup_slabs(r, sc.tlas_extent);
            nodes = sc.tlas_nodes; tris = nullptr;
            root_group();
            cur = WALK;
        }
    }

    // node i of the current array (SMEM: from the staged copy when it is among the top levels, rtb_tma.cuh)
    template <bool SMEM>
    XF const float4* node_ptr(unsigned i, const TopLevels& T) const
    {
        const float4* np = nodes + (size_t)i * NODE_F4;
        if (SMEM) {
            if (!in_blas) { if ((int)i < T.n_tlas) np = T.s_tlas + i * NODE_F4; }
            else if ((int)i < T.n_blas && nodes == T.g_blas) np = T.s_blas + i * NODE_F4;
        }
        return np;
    }

    // leaf child k of node pn -> first << 3 | count - 1, or NO_LEAF for a slot that holds no leaf (an empty slot whose
    // inverted box let the ray through: possible only when the rounding slack exceeds the node, i.e. far away from it)
    template <bool SMEM>
    XF unsigned leaf_ref(unsigned pn, unsigned k, const TopLevels& T) const
    {
        const float4 n1 = node_ld<SMEM>(node_ptr<SMEM>(pn, T) + 1);
        const unsigned w = (k & 4u) ? __float_as_uint(n1.w) : __float_as_uint(n1.z);
        const unsigned mb = (w >> (8u * (k & 3u))) & 0xffu;
        const unsigned first = (unsigned)__float_as_int(n1.y) + (mb & 31u), count = (mb >> 5) & 3u;
        return (mb & 0x80u) ? ((first << 3) | count) : NO_LEAF;
    }

    template <bool ANY, bool COUNT, class Sink>
    XF void advance(const SceneDev& sc, Stack& stk, Sink& sink)
    {
        TopLevels none; none.n_tlas = none.n_blas = 0; none.s_tlas = none.s_blas = none.g_blas = nullptr;
        advance<ANY, COUNT, false>(sc, stk, sink, none);
    }

    // SMEM: nodes whose index is below the staged count of their array are read from shared memory (rtb_tma.cuh)
    template <bool ANY, bool COUNT, bool SMEM, class Sink>
    XF void advance(const SceneDev& sc, Stack& stk, Sink& sink, const TopLevels& T)
    {
        if (++guard > (1u << 22)) { overflow = true; cur = EMPTY_REF; }   // a frame can never hang the device
        const unsigned char* lut = pick_lut();
        // Speculative traversal: the first triangle leaf a lane picks is postponed (`pend`) and the lane keeps walking, so it
        // does useful work while the slower lanes of the warp are still descending; the walk ends for the warp when (almost)
        // every lane still walking holds a postponed leaf.  Leaves are remembered as (node << 3 | slot); their triangle range is
        // read from the node's meta word in the triangle phase, so the walk contains ONE kind of load: the node fetch.
        // All 32 lanes stay in the loop until the warp leaves it together (a full-mask ballot): lanes that took a short path
        // through the pick phase -- a culled child, a pop -- must not run ahead of the others, or the box test below ends up
        // executing a few lanes wide (measured: 5.8 lanes, profiles/r03a).
        unsigned pend = NO_LEAF, leaf = NO_LEAF;
        for (;;) {
            int child = -1;
            if (cur == WALK) {
#pragma unroll 1
                for (int tries = 0; tries < 3; tries++) {
                    if ((g_hm & 0xffu) == 0u) {             // group exhausted: the next one comes off the stack
                        if (sp == 0) { cur = EMPTY_REF; break; }
                        sp--;
                        const uint4 a = stk.a[sp];
                        if (a.y == 0xffffffffu) { cur = EV_LEAVE; break; }
                        const uint4 b = stk.b[sp];
                        g_cb = a.x; g_pn = a.y; g_hm = a.z; g_t01 = b.x; g_t23 = b.y; g_t45 = b.z; g_t67 = b.w;
                    }
                    const unsigned k = lut[(r.octinv << 8) | (g_hm & 0xffu)];
                    g_hm &= ~(1u << k);
                    const unsigned tw = (k & 4u) ? ((k & 2u) ? g_t67 : g_t45) : ((k & 2u) ? g_t23 : g_t01);
                    const float tnk = __uint_as_float((k & 1u) ? (tw & 0xffff0000u) : (tw << 16));
                    if (tnk < tmax) {                       // else: a closer hit was found since the box test (cull only)
                        if ((g_hm >> (8u + k)) & 1u) { child = (int)(g_cb + __popc((g_hm >> 8) & ((1u << k) - 1u))); break; }
                        const unsigned id = (g_pn << 3) | k;    // leaf child: triangles (BLAS) or one instance (TLAS)
                        if (!in_blas) { leaf = id; cur = EV_ENTER; break; }
                        if (pend == NO_LEAF) pend = id;
                        else { leaf = id; cur = EV_TRIS; break; }
                    }
                }
            }
            __syncwarp();
            if (child >= 0) {
                if (g_hm & 0xffu) {
                    if (sp >= STACK_CAP) { overflow = true; cur = EMPTY_REF; child = -1; }
                    else {
                        stk.a[sp] = make_uint4(g_cb, g_pn, g_hm, 0u);
                        stk.b[sp] = make_uint4(g_t01, g_t23, g_t45, g_t67);
                        sp++;
                    }
                }
            }
            if (child >= 0) {
                if (COUNT) n_nodes++;
                const float4* np = node_ptr<SMEM>((unsigned)child, T);
                const float4 n0 = node_ld<SMEM>(np);
                const unsigned cb = SMEM ? *(const unsigned*)(np + 1) : __ldg((const unsigned*)(np + 1));
                NodeHits H;
                test_node8<SMEM>(np, n0, r, tmax, H);
                g_cb = cb; g_pn = (unsigned)child;
                g_hm = H.hits | ((__float_as_uint(n0.w) >> 24) << 8);
                g_t01 = H.tn01; g_t23 = H.tn23; g_t45 = H.tn45; g_t67 = H.tn67;
            }
            if (__popc(__ballot_sync(0xffffffffu, cur == WALK && pend == NO_LEAF)) <= sc.spec_thr) break;
        }
        // triangle ranges of the (up to two) leaves: meta byte of the slot in its node
        int t0 = 0, t1 = 0, u0 = 0, u1 = 0;
        if (pend != NO_LEAF) {
            const unsigned ref = leaf_ref<SMEM>(pend >> 3, pend & 7u, T);
            if (ref != NO_LEAF) { t0 = (int)(ref >> 3); t1 = t0 + (int)(ref & 7u) + 1; }
        }
        if (cur == EV_TRIS) {
            const unsigned ref = leaf_ref<SMEM>(leaf >> 3, leaf & 7u, T);
            if (ref != NO_LEAF) { u0 = (int)(ref >> 3); u1 = u0 + (int)(ref & 7u) + 1; }
            if (t0 == t1) { t0 = u0; t1 = u1; u0 = u1 = 0; }
            cur = WALK;
        }
        // triangle phase: the postponed leaf, then the second one (flattened into one loop over triangle slots)
        bool found = false;
        while (t0 < t1) {
            const int slot = t0;
            const float4* q = tris + (size_t)slot * TRI_F4;
            const float4 q0 = ld_keep(q + 0), q1 = ld_keep(q + 1);
            if (++t0 == t1) { t0 = u0; t1 = u1; u0 = u1 = 0; }
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
            const float4 q2 = ld_keep(q + 2), q3 = ld_keep(q + 3);
            const f3 e0 = mk3(q2.x, q2.y, q2.z), e1 = mk3(q3.x, q3.y, q3.z);
            const f3 v2 = xsub(p, A);                                         // TrianglePrimitive.hpp:154-156
            const float dot00 = q1.w, dot01 = q2.w, dot11 = q3.w;
            const float dot02 = xdot(e0, v2), dot12 = xdot(e1, v2);
            const float invDenom = xd(1.0f, xs(xm(dot00, dot11), xm(dot01, dot01)));   // :164
            const float u = xm(xs(xm(dot11, dot02), xm(dot01, dot12)), invDenom);
            const float v = xm(xs(xm(dot00, dot12), xm(dot01, dot02)), invDenom);
            if (!((u >= 0.0f) && (v >= 0.0f) && (xa(u, v) <= 1.0f))) continue;          // :169
            // raycast.cu:98-104 world location through inv_pose, Euclidean distance to the WORLD origin
            const float4* ip = sc.inst + (size_t)cur_inst * INST_F4;
            const float4 q_inv = ld_keep(ip + 3), it = ld_keep(ip + 4), sc4 = ld_keep(ip + 5);
            f3 loc = xmul3(p, mk3(sc4.x, sc4.y, sc4.z));
            loc = xapply_quat(q_inv, xsub(loc, mk3(it.x, it.y, it.z)));
            const float distance = xmagnitude(xsub(loc, org));
            const int prim = __float_as_int(q0.w);
            const float hmin = sink.min();
            const bool closer = (hmin == FLT_MAX) || (distance < hmin) ||
                                (distance == hmin && (cur_inst < sink.inst() || (cur_inst == sink.inst() && prim < sink.prim())));
            if (closer) {
                sink.accept(distance, slot, prim, cur_inst, u, v, loc);
                tmax = fmaf(distance, inv_len, tslack);
                found = true;
            }
        }
        if ((ANY || sink.any_hit()) && found) { cur = EMPTY_REF; return; }
        // events: leave the instance (sentinel popped), enter the instance of a TLAS leaf
        if (cur == EV_LEAVE) {                  // instance exhausted: back to world space
            in_blas = false; nodes = sc.tlas_nodes;
            r.o = org; r.d = dir;
            setup_slabs(r, sc.tlas_extent);
            g_hm = 0u;
            cur = WALK;
        } else if (cur == EV_ENTER) {           // TLAS leaf: exactly one instance slot
            const unsigned ref = leaf_ref<SMEM>(leaf >> 3, leaf & 7u, T);
            if (ref == NO_LEAF) { cur = WALK; return; }
            const int first = (int)(ref >> 3);
            if (sp + 2 > STACK_CAP) { overflow = true; cur = EMPTY_REF; return; }
            if (g_hm & 0xffu) {
                stk.a[sp] = make_uint4(g_cb, g_pn, g_hm, 0u);
                stk.b[sp] = make_uint4(g_t01, g_t23, g_t45, g_t67);
                sp++;
            }
            enter_instance(sc, __ldg(sc.tlas_index + first), stk, true);
        }
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
// Must be called by all 32 lanes of the warp (advance() holds a warp barrier); lanes without a ray pass active=false.
// The kernel must have called init_pick_lut().
template <bool ANY, bool COUNT>
__device__ __forceinline__ void trace(const SceneDev& sc, Stack& stk, bool active, f3 org, f3 dir, Hit& hit, Counters* ctr)
{
    TraceState S;
    RegSink sink;
    sink.reset();
    S.clear();
    if (active) S.begin(sc, org, dir, stk);
    while (__any_sync(0xffffffffu, S.cur != EMPTY_REF)) S.template advance<ANY, COUNT>(sc, stk, sink);
    S.template flush_counters<COUNT>(ctr);
    hit = sink.hit;
}

}  // namespace rtb
