// rtb_node8.cuh -- 8-wide BVH node with child boxes quantised to 8 bits on the node's own box (96 bytes).
//
// STATUS: groundwork for the next tracer step.  The encoder / decoder below are complete and unit-tested on
// the host (tests/test_oracle_cpu.py::test_node8_quantisation_is_conservative compiles this header with g++); the builder's
// collapse and the traversal loop do not use them yet, so nothing here is on the product path of this round.
//
// Layout (six 16-byte words = three 32-byte sectors; at a 96-byte stride consecutive nodes rotate through the L1 bank groups
// like the 112-byte 4-wide node does, rtb_internal.cuh):
//   word 0      origin.x, origin.y, origin.z, exps = ex | ey << 8 | ez << 16 | (valid-child mask) << 24
//               cell size of axis a = 2^(e_a - 127): a power of two, so q * cell is exact in binary32
//   words 1-3   plane bytes: lo.x[8] hi.x[8] | lo.y[8] hi.y[8] | lo.z[8] hi.z[8]
//   words 4-5   ref[8]: >= 0 wide node, < 0 leaf (~ref = first << 3 | count - 1), 0x80000000 empty -- as in the 4-wide node
//
// The decoder is ONE fma per plane: plane = fmaf(q, cell, origin).  The sum rounds, so snapping `lo` down / `hi` up on the
// grid is not enough to stay conservative: the encoder checks every plane under the decoder's own arithmetic and steps one
// cell further out when rounding would cut into the exact box.  The tree only culls (rtb_trace.cuh), so a box that is a
// cell too fat costs a visit, never a pixel; a box that is an ulp too thin would.
#pragma once
#include <math.h>
#include <stdint.h>
#include <string.h>

#ifdef __CUDACC__
#define RTB_HD __host__ __device__ __forceinline__
#else
#define RTB_HD static inline
#endif

namespace rtb {

struct Node8 {
    float ox, oy, oz;
    uint32_t exps;
    uint8_t q[3][2][8];             // [axis][0 = lo, 1 = hi][child]
    int32_t ref[8];
};
static_assert(sizeof(Node8) == 96, "Node8 must be six 16-byte words");

constexpr int32_t NODE8_EMPTY = (int32_t)0x80000000;

RTB_HD float node8_pow2(int biased) { uint32_t b = (uint32_t)biased << 23; float f; memcpy(&f, &b, 4); return f; }
RTB_HD float node8_cell(const Node8& n, int axis) { return node8_pow2((int)((n.exps >> (8 * axis)) & 0xffu)); }
RTB_HD float node8_origin(const Node8& n, int axis) { return axis == 0 ? n.ox : (axis == 1 ? n.oy : n.oz); }

// the decoder: the box the tracer tests for child k
RTB_HD void node8_child_box(const Node8& n, int k, float lo[3], float hi[3])
{
    for (int a = 0; a < 3; a++) {
        const float cell = node8_cell(n, a), org = node8_origin(n, a);
        lo[a] = fmaf((float)n.q[a][0][k], cell, org);
        hi[a] = fmaf((float)n.q[a][1][k], cell, org);
    }
}

// Encodes up to 8 child boxes (finite, lo <= hi).  Returns false when a box is not finite or the node's extent cannot be
// represented (the caller then keeps that subtree in exact 4-wide nodes, or refuses the mesh as the builder does for NaNs).
RTB_HD bool node8_encode(const float lo[][3], const float hi[][3], const int32_t* ref, int count, Node8& out)
{
    float org[3] = { 0.f, 0.f, 0.f };
    uint32_t exps = 0;
    for (int k = 0; k < 8; k++) out.ref[k] = k < count ? ref[k] : NODE8_EMPTY;
    for (int a = 0; a < 3; a++) {
        float mn = INFINITY, mx = -INFINITY;
        for (int k = 0; k < count; k++) {
            if (!(lo[k][a] <= hi[k][a]) || isinf(lo[k][a]) || isinf(hi[k][a])) return false;
            mn = fminf(mn, lo[k][a]); mx = fmaxf(mx, hi[k][a]);
        }
        if (count == 0) { mn = mx = 0.f; }
        org[a] = mn;
        // smallest power-of-two cell with 255 cells covering the extent; the subtraction may round, the checks below absorb it
        const float ext = mx - mn;
        int e;
        if (!(ext > 0.f)) e = 1;                                       // flat on this axis: any cell does (q = 0 everywhere)
        else {
            int x; (void)frexpf(ext / 255.0f, &x);                     // ext / 255 = m * 2^x, m in [0.5, 1)  =>  2^x >= ext / 255
            e = x + 127;
            if (e < 1) e = 1;
        }
        for (;;) {
            if (e > 254) return false;
            const float cell = node8_pow2(e);
            bool ok = true;
            for (int k = 0; k < count && ok; k++) {
                float ql = floorf((lo[k][a] - mn) / cell), qh = ceilf((hi[k][a] - mn) / cell);
                if (ql < 0.f) ql = 0.f;
                if (qh < 0.f) qh = 0.f;
                while (ql > 0.f && fmaf(ql, cell, mn) > lo[k][a]) ql -= 1.f;      // decoder arithmetic, not exact arithmetic
                if (fmaf(ql, cell, mn) > lo[k][a]) { ok = false; break; }         // ql == 0 decodes to the origin = min lo: cannot happen
                while (qh <= 255.f && fmaf(qh, cell, mn) < hi[k][a]) qh += 1.f;
                if (qh > 255.f || ql > 255.f) { ok = false; break; }              // does not fit: double the cell
                out.q[a][0][k] = (uint8_t)ql; out.q[a][1][k] = (uint8_t)qh;
            }
            if (ok) break;
            e++;
        }
        // empty slot: the valid mask (exps bits 24-31) and ref == EMPTY are what the tracer must test; the inverted planes are
        // only a second line of defence (on a flat axis, or when 255 cells are below the origin's ulp, they decode to lo == hi)
        for (int k = count; k < 8; k++) { out.q[a][0][k] = 255; out.q[a][1][k] = 0; }
        exps |= (uint32_t)e << (8 * a);
    }
    uint32_t valid = count >= 8 ? 0xffu : ((1u << count) - 1u);
    out.ox = org[0]; out.oy = org[1]; out.oz = org[2];
    out.exps = exps | (valid << 24);
    return true;
}

}  // namespace rtb
