// rtb_trace_state.cuh -- per-lane traversal state machine (included by rtb_trace.cuh).
//
// One ray per lane.  `cur` is the node being visited: >= 0 interior, < 0 an EVENT: a leaf, LEAVE_REF (the
// sentinel pushed when an instance was entered: "back to world space") or EMPTY_REF (no ray / finished).
// advance() = "walk interior nodes until this lane holds an event, then handle that event".
// Control flow is strictly structured (while / if, no goto) so the warp reconverges after every advance():
// lanes first walk interior nodes until each holds an event, then all lanes handle their events together.
// The interior loop contains nothing but box tests, the 4-key sort, pushes and a cull-only pop; every
// space change (enter / leave an instance) is an event, so it never serialises the interior loop.
//
// The state is scalars only -- the traversal stack is a separate array handed in by pointer -- so that the
// whole state lives in registers (an array member with a dynamic index pins the entire struct to local memory).
// What is only needed when a candidate passes the inside test (instance inverse pose, scale) is re-read from
// the 96-byte instance record at that point instead of being carried in registers through the traversal.
//
// Two drivers use it:
//   * trace<>()            one ray to completion (direct kernel, modes 0/1);
//   * k_trace_queue<>      persistent warps that refill finished lanes from a ray queue (wavefront, rtb_wavefront.cu).
#pragma once

namespace rtb {

constexpr int LEAVE_REF = (int)0x80000001;     // sentinel stack entry: instance exhausted, restore the world-space ray

// Traversal stack of (tnear bits, ref) entries.  Default: a per-thread local-memory array.  With -DRTB_SMEM_STACK the
// first SMEM_STACK_DEPTH entries live in shared memory (column of this thread, stride = 128 threads, conflict-free for
// lanes at the same depth) and only deeper entries spill to local memory.
constexpr int SMEM_STACK_DEPTH = 12;
struct Stack {
#ifdef RTB_SMEM_STACK
    uint2* sm;
    uint2 loc[STACK_CAP - SMEM_STACK_DEPTH];
    XF void put(int i, uint2 v) { if (i < SMEM_STACK_DEPTH) sm[i * 128] = v; else loc[i - SMEM_STACK_DEPTH] = v; }
    XF uint2 get(int i) const { return i < SMEM_STACK_DEPTH ? sm[i * 128] : loc[i - SMEM_STACK_DEPTH]; }
#else
    uint2 loc[STACK_CAP];
    XF void put(int i, uint2 v) { loc[i] = v; }
    XF uint2 get(int i) const { return loc[i]; }
#endif
};
#ifdef RTB_SMEM_STACK
#define RTB_DECL_STACK(name) __shared__ uint2 name##_sm[SMEM_STACK_DEPTH * 128]; Stack name; name.sm = name##_sm + threadIdx.x
#else
#define RTB_DECL_STACK(name) Stack name
#endif

// Hit record kept in registers (direct kernel).  The queue tracer uses a sink that writes accepted candidates
// straight to the hit queue in HBM and keeps only what the accept rule compares (rtb_wavefront.cu).
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
    int cur, sp, cur_inst;
    bool in_blas;
    unsigned n_nodes, n_tris, guard;
    bool overflow;

    XF void clear() { cur = EMPTY_REF; sp = 0; n_nodes = n_tris = guard = 0; overflow = false; in_blas = false; cur_inst = -1; }

    // raycast.cu:33-45: r_dir = apply_euler(rotation, dir) * inv_scale ; r_org = apply_lre(pose, org) * inv_scale
    XF void enter_instance(const SceneDev& sc, int inst, Stack& stk, bool sentinel)
    {
        const float4* p = sc.inst + (size_t)inst * INST_F4;
        const float4 q_pose = __ldg(p + 0), b = __ldg(p + 1), c = __ldg(p + 2);
        const f3 inv_scale = mk3(c.x, c.y, c.z);
        cur_inst = inst;
        r.d = xmul3(xapply_quat(q_pose, dir), inv_scale);
        r.o = xmul3(xapply_quat(q_pose, xsub(org, mk3(b.x, b.y, b.z))), inv_scale);
        setup_slabs(r);
        const MeshDev* m = sc.meshes + __float_as_int(b.w);
        nodes = m->nodes; tris = m->tris;
        in_blas = true;
        if (sentinel) stk.put(sp++, make_uint2(0xbf800000u /* -1.0f: never culled */, (unsigned)LEAVE_REF));
        cur = 0;                                // BLAS root
    }

    XF void begin(const SceneDev& sc, f3 o, f3 d, Stack& stk)
    {
        org = o; dir = d;
        sp = 0; in_blas = false; cur_inst = -1;
        // parametric bound: Euclidean distance = t * |d_world| for every instance (scale cancels), so a subtree entered
        // at t can only hold hits farther than t*|d|.  Slack covers the reference's inverse-pose round-trip error,
        // which grows with the magnitude of the coordinates involved (scene extent, ray origin), not with the distance.
        const float il = rsqrtf(d.x * d.x + d.y * d.y + d.z * d.z);
        inv_len = 1.0002f * il;
        tslack = (1e-3f + 1.2e-6f * (sc.tlas_extent + fmaxf(fmaxf(fabsf(o.x), fabsf(o.y)), fabsf(o.z)))) * il;
        tmax = FLT_MAX;
        if (sc.num_instances == 1) {            // single instance: no TLAS level, no sentinel
            enter_instance(sc, 0, stk, false);
        } else {
            r.o = o; r.d = d;
            setup_slabs(r);
            nodes = sc.tlas_nodes; tris = nullptr;
            cur = 0;                            // TLAS root
        }
    }

    XF int pop(Stack& stk)
    {
        while (sp > 0) {
            const uint2 e = stk.get(--sp);
            if (__uint_as_float(e.x) < tmax) return (int)e.y;
        }
        return EMPTY_REF;
    }

    // ref of the child in slot s (0..3) of a node, branch-free (three selects)
    static XF int sel4(unsigned s, const int4& r)
    {
        const int lo = (s & 1u) ? r.y : r.x, hi = (s & 1u) ? r.w : r.z;
        return (s & 2u) ? hi : lo;
    }
    static XF bool is_leaf_ref(int ref) { return (unsigned)ref > (unsigned)LEAVE_REF; }   // negative, not EMPTY / LEAVE

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
        // Speculative traversal: the first triangle leaf a lane meets is postponed (`pend`) and the lane keeps walking
        // interior nodes, so it does useful work while the slower lanes of the warp are still descending; the loop
        // ends for the warp when every lane still in it holds a postponed leaf.
        int pend = EMPTY_REF;
        while (cur >= 0) {                      // interior nodes
            unsigned k[4]; int4 refs;
            if (COUNT) n_nodes++;
            if (sp + 3 > STACK_CAP) { overflow = true; cur = EMPTY_REF; break; }
            const float4* np = nodes + (size_t)cur * NODE_F4;
            if (SMEM) {
                if (!in_blas) { if (cur < T.n_tlas) np = T.s_tlas + cur * NODE_F4; }
                else if (cur < T.n_blas && nodes == T.g_blas) np = T.s_blas + cur * NODE_F4;
            }
            test_node<SMEM>(np, r, tmax, k, refs);
#ifdef RTB_NEAREST_ONLY
            // A/B variant, measured SLOWER and therefore off (C3 8.17 -> 8.57 ms, same image; profiles/r01f_ab_experiments.txt): only
            // the nearest child is singled out; the other hit children are pushed in SLOT order -- no sorting network, no ref
            // selects for them (tools/bvh_lab.py: +3.5 % node visits for hitting rays).  Every entry carries its own entry
            // distance, so the pop still culls.
            const unsigned kmin = min(min(k[0], k[1]), min(k[2], k[3]));      // keys are distinct (slot in the low bits) unless all miss
            if (k[0] != kmin && k[0] != 0xffffffffu) stk.put(sp++, make_uint2(k[0] & ~3u, (unsigned)refs.x));
            if (k[1] != kmin && k[1] != 0xffffffffu) stk.put(sp++, make_uint2(k[1] & ~3u, (unsigned)refs.y));
            if (k[2] != kmin && k[2] != 0xffffffffu) stk.put(sp++, make_uint2(k[2] & ~3u, (unsigned)refs.z));
            if (k[3] != kmin && k[3] != 0xffffffffu) stk.put(sp++, make_uint2(k[3] & ~3u, (unsigned)refs.w));
            int nxt = sel4(kmin & 3u, refs);
            if (kmin == 0xffffffffu) nxt = pop(stk);
#else
            RTB_CSWAP(k[0], k[1]); RTB_CSWAP(k[2], k[3]); RTB_CSWAP(k[0], k[2]); RTB_CSWAP(k[1], k[3]); RTB_CSWAP(k[1], k[2]);
            const int c0 = sel4(k[0] & 3u, refs), c1 = sel4(k[1] & 3u, refs), c2 = sel4(k[2] & 3u, refs), c3 = sel4(k[3] & 3u, refs);
            // far first, so the nearer children are popped first (misses sort last: key 0xffffffff)
            if (k[3] != 0xffffffffu) stk.put(sp++, make_uint2(k[3] & ~3u, (unsigned)c3));
            if (k[2] != 0xffffffffu) stk.put(sp++, make_uint2(k[2] & ~3u, (unsigned)c2));
            if (k[1] != 0xffffffffu) stk.put(sp++, make_uint2(k[1] & ~3u, (unsigned)c1));
            int nxt = c0;
            if (k[0] == 0xffffffffu) nxt = pop(stk);
#endif
            if (in_blas && pend == EMPTY_REF && is_leaf_ref(nxt)) { pend = nxt; nxt = pop(stk); }
            cur = nxt;
            if (__popc(__ballot_sync(__activemask(), pend == EMPTY_REF)) <= sc.spec_thr) break;
        }
        // Reconvergence point: every lane of the warp calls advance() (both drivers guarantee it), and without this
        // barrier lanes leave the interior loop in small groups and run the leaf code 2 lanes wide (ncu, profiles/r01b).
        __syncwarp();
        // triangle phase: the postponed leaf, then the leaf in `cur` (flattened into one loop over triangle slots)
        const bool cur_leaf = in_blas && is_leaf_ref(cur);
        int t0 = 0, t1 = 0, u0 = 0, u1 = 0;
        if (pend != EMPTY_REF) { const unsigned b = ~(unsigned)pend; t0 = (int)(b >> 3); t1 = t0 + (int)(b & 7u) + 1; }
        if (cur_leaf) {
            const unsigned b = ~(unsigned)cur;
            const int f = (int)(b >> 3), e = f + (int)(b & 7u) + 1;
            if (pend != EMPTY_REF) { u0 = f; u1 = e; } else { t0 = f; t1 = e; }
        }
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
        if (cur_leaf) cur = pop(stk);
        // events: leave the instance (sentinel popped), enter the instance of a TLAS leaf
        if (cur == LEAVE_REF) {                 // instance exhausted: back to world space
            in_blas = false; nodes = sc.tlas_nodes;
            cur = pop(stk);
            if (cur >= 0) { r.o = org; r.d = dir; setup_slabs(r); }
        }
        if (!in_blas && is_leaf_ref(cur)) {     // TLAS leaf: exactly one instance slot
            const int first = (int)((~(unsigned)cur) >> 3);
            if (sp + 1 > STACK_CAP) { overflow = true; cur = EMPTY_REF; return; }
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
