// rtb_cwnode.cuh -- the 80-byte compressed 8-wide node (layout: rtb_internal.cuh): quantisation frame, plane encoding,
// host-side decoding.  Shared by the builder (device, rtb_build.cu), the C-ABI (host: instance boxes from a BLAS cut,
// rtb_api.cu) and the tests (tests/cwnode_check.cpp compiles this header with g++).
//
// Child boxes are snapped OUTWARDS on a grid of 255 power-of-two cells per axis anchored at the node's own lower corner and
// the snap is verified in double precision against origin + q * cell, which is exact there; the tracer evaluates the same
// planes with a rounding slack folded into its ray constants (setup_slabs, rtb_trace.cuh).  The tree only culls: a box that
// is a cell too fat costs a visit, never a pixel; a box that is an ulp too thin would.
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

constexpr int CW_EXP_BIAS = 15;     // stored exponent = biased exponent of cell * 2^15 (what the decoder multiplies with)

RTB_HD double cw_cell(int e_stored) { return ldexp(1.0, e_stored - 127 - CW_EXP_BIAS); }

// Smallest stored exponent whose 255 cells cover [lo, hi] (hi >= lo, both finite); 0 = not representable.
RTB_HD int cw_axis_exponent(float lo, float hi)
{
    const double ext = (double)hi - (double)lo;
    if (!(ext >= 0.0) || ext > 1e37) return 0;
    int e = 1;
    if (ext > 0.0) {
        int x;
        (void)frexp(ext / 255.0, &x);                      // ext / 255 = m * 2^x, m in [0.5, 1)  =>  2^x >= ext / 255
        e = x + 127 + CW_EXP_BIAS;
        if (e < 1) e = 1;
    }
    return e <= 254 ? e : 0;
}

// Plane bytes of [clo, chi] on the frame (origin, e): qlo the last grid plane <= clo, qhi the first >= chi.  Returns false
// when the box does not fit 255 cells (the caller doubles the cell) -- cannot happen for a child inside [origin, origin + ext].
RTB_HD bool cw_quantise(float origin, int e_stored, float clo, float chi, int& qlo, int& qhi)
{
    const double cell = cw_cell(e_stored), o = (double)origin;
    double a = floor(((double)clo - o) / cell), b = ceil(((double)chi - o) / cell);
    if (a < 0.0) a = 0.0;
    if (b < 0.0) b = 0.0;
    if (a > 255.0) a = 255.0;
    while (a > 0.0 && o + a * cell > (double)clo) a -= 1.0;
    while (b <= 255.0 && o + b * cell < (double)chi) b += 1.0;
    if (b > 255.0 || o + a * cell > (double)clo) return false;
    qlo = (int)a; qhi = (int)b;
    return true;
}

// ---- host-side view of one node (5 x 16 bytes as read back from the device)
struct CwNodeView {
    float origin[3];
    int e[3];
    unsigned imask;
    int child_base, tri_base;
    unsigned char meta[8];
    unsigned char q[3][2][8];       // [axis][0 = lo, 1 = hi][child]
};

static inline CwNodeView cw_view(const void* node80)
{
    CwNodeView v;
    const unsigned char* p = (const unsigned char*)node80;
    uint32_t w;
    memcpy(v.origin, p, 12);
    memcpy(&w, p + 12, 4);
    v.e[0] = (int)(w & 0xffu); v.e[1] = (int)((w >> 8) & 0xffu); v.e[2] = (int)((w >> 16) & 0xffu); v.imask = w >> 24;
    memcpy(&v.child_base, p + 16, 4); memcpy(&v.tri_base, p + 20, 4);
    memcpy(v.meta, p + 24, 8);
    for (int a = 0; a < 3; a++) { memcpy(v.q[a][0], p + 32 + 16 * a, 8); memcpy(v.q[a][1], p + 40 + 16 * a, 8); }
    return v;
}

// kind of child k: 0 empty, 1 inner node, 2 leaf
static inline int cw_child_kind(const CwNodeView& v, int k) { return ((v.imask >> k) & 1u) ? 1 : ((v.meta[k] & 0x80u) ? 2 : 0); }
static inline int cw_child_node(const CwNodeView& v, int k) { return v.child_base + __builtin_popcount(v.imask & ((1u << k) - 1u)); }
static inline void cw_child_tris(const CwNodeView& v, int k, int& first, int& count) { first = v.tri_base + (v.meta[k] & 31); count = ((v.meta[k] >> 5) & 3) + 1; }
// decoded box, rounded outwards to binary32
static inline void cw_child_box(const CwNodeView& v, int k, float lo[3], float hi[3])
{
    for (int a = 0; a < 3; a++) {
        const double cell = cw_cell(v.e[a]);
        const double l = (double)v.origin[a] + v.q[a][0][k] * cell, h = (double)v.origin[a] + v.q[a][1][k] * cell;
        float fl = (float)l, fh = (float)h;
        if ((double)fl > l) fl = nextafterf(fl, -INFINITY);
        if ((double)fh < h) fh = nextafterf(fh, INFINITY);
        lo[a] = fl; hi[a] = fh;
    }
}

}  // namespace rtb
