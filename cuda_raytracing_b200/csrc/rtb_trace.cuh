// rtb_trace.cuh -- two-level wide-BVH traversal with the reference's closest-hit / any-hit semantics.
//
// Replaces cast_ray (raycast.cu:21-142): the linear instance loop becomes a TLAS descent, the
// 48-byte AoS BVH2 nodes + per-leaf index lists become 80-byte compressed 8-wide nodes + node-ordered triangle
// records, and the 12 sinf/cosf per (ray, instance) become one precomputed quaternion.  What is kept
// bit-for-bit is everything that decides the answer: the world->mesh ray transform (raycast.cu:33-45),
// the plane hit and barycentric test (TrianglePrimitive.hpp:62-79,151-185), the world-space distance
// (raycast.cu:98-104) and the accept rule (raycast.cu:107-109).  Box tests only cull and are
// conservative, so the BVH shape cannot change the result; exact distance ties are broken by
// (instance, original triangle index) instead of by the reference's traversal order.
#pragma once
#include "rtb_internal.cuh"
#include "rtb_tma.cuh"
#include "rtb_cache.cuh"

namespace rtb {

struct InstRay {                 // ray in the current space (world or instance)
    f3 o, d;
    // Slab-test constants (culling only).  A plane byte q of axis a decodes to the parameter
    //     t = q * A + B,   A = cell_a / d_a,   B = (origin_a - o_a) / d_a,
    // evaluated as fma(f, A', D -+ S) with f = 1 + q 2^-15 built by ONE byte permute (PRMT: half rate; the int->float
    // conversion I2F.U8 issues at 1/8 of that rate on sm_100, tools/pipe_probe.cu), A' = 2^15 A (the builder stores the
    // exponent of cell 2^15, so A' is one exponent insert + one multiply), D = fma(origin, id, -oi) - A', and
    // S = s + 2^-22 |A'|: s is the per-axis rounding slack of the whole evaluation, 2^-20 (E + |o|) |id| for a space whose
    // node coordinates stay below E, the second term covers the rounding of D (<= 2^-9 cell).  Near planes use D - S, far
    // planes D + S: the test can only accept more than the exact-arithmetic test, never less.  1/d is clamped to 1e20.
    float idx, idy, idz, oix, oiy, oiz, sx, sy, sz;
    unsigned octinv;             // bit a set = d_a >= 0: a child in slot k is visited in the order of descending k ^ octinv
};

XF float safe_rcp(float d)
{
    float a = fabsf(d) < 1e-20f ? copysignf(1e-20f, d) : d;
    return 1.0f / a;
}

// extent: upper bound of |node coordinate| in this space (MeshDev::extent / SceneDev::tlas_extent); the quantisation frame of a
// node may overshoot its box by less than the box's own extent, hence the factor
XF void setup_slabs(InstRay& r, float extent)
{
    const float ix = safe_rcp(r.d.x), iy = safe_rcp(r.d.y), iz = safe_rcp(r.d.z);
    const float E = 4.0f * extent, k = 9.5367431640625e-7f;            // 2^-20
    r.idx = ix; r.idy = iy; r.idz = iz;
    r.oix = r.o.x * ix; r.oiy = r.o.y * iy; r.oiz = r.o.z * iz;
    r.sx = k * (E + fabsf(r.o.x)) * fabsf(ix); r.sy = k * (E + fabsf(r.o.y)) * fabsf(iy); r.sz = k * (E + fabsf(r.o.z)) * fabsf(iz);
    r.octinv = (ix < 0.0f ? 0u : 1u) | (iy < 0.0f ? 0u : 2u) | (iz < 0.0f ? 0u : 4u);
}

// GENERIC: the node may live in shared memory (staged top levels), so plain generic loads instead of ld.global.nc
template <bool GENERIC>
XF float4 node_ld(const float4* p) { return GENERIC ? *p : ld_keep(p); }

// 1 + q 2^-15 for byte j of w: bytes (0x00, q, 0x80, 0x3f)
template <int J>
XF float plane_f(unsigned w) { return __uint_as_float(__byte_perm(w, 0x3f800000u, 0x7604u | ((unsigned)J << 4))); }

struct NodeHits {
    unsigned hits;               // bit k = the ray meets child k's box before tmax (slot order; empty slots may be set, see advance)
    unsigned tn01, tn23, tn45, tn67;   // entry parameters of the 8 children, truncated to bf16 (a lower bound), two per word
};

// 4 children of one half (bytes 0-3 of the six plane words of that half); appends their miss bits to `miss` (child 0 of the
// half ends up highest) and returns their packed entry parameters
template <int HALF>
XF void test_half(unsigned nx, unsigned ny, unsigned nz, unsigned fx, unsigned fy, unsigned fz, float Ax, float Ay, float Az,
                  float Bnx, float Bny, float Bnz, float Bfx, float Bfy, float Bfz, float tmax, unsigned& miss, unsigned& tlo, unsigned& thi)
{
    float tn[4];
#define RTB_CHILD(J)                                                                                                                 \
    {                                                                                                                                \
        const float tnx = fmaf(plane_f<J>(nx), Ax, Bnx), tny = fmaf(plane_f<J>(ny), Ay, Bny), tnz = fmaf(plane_f<J>(nz), Az, Bnz);   \
        const float tfx = fmaf(plane_f<J>(fx), Ax, Bfx), tfy = fmaf(plane_f<J>(fy), Ay, Bfy), tfz = fmaf(plane_f<J>(fz), Az, Bfz);   \
        const float a = fmaxf(fmaxf(tnx, tny), tnz), b = fminf(fminf(tfx, tfy), tfz);                                                \
        /* miss = entry after exit, entry beyond tmax, or exit behind the origin: the sign bit of one of three differences */       \
        const unsigned m = __float_as_uint(b - a) | __float_as_uint(tmax - a) | __float_as_uint(b);                                  \
        miss = __funnelshift_l(m, miss, 1);                                                                                          \
        tn[J] = a;                                                                                                                   \
    }
    RTB_CHILD(3) RTB_CHILD(2) RTB_CHILD(1) RTB_CHILD(0)
#undef RTB_CHILD
    tlo = __byte_perm(__float_as_uint(tn[0]), __float_as_uint(tn[1]), 0x7632);    // (bf16 of child 0) | (bf16 of child 1) << 16
    thi = __byte_perm(__float_as_uint(tn[2]), __float_as_uint(tn[3]), 0x7632);
}

// 8 slab tests against one compressed wide node
template <bool GENERIC>
XF void test_node8(const float4* __restrict__ node, const float4 n0, const InstRay& r, float tmax, NodeHits& H)
{
    const float4 px = node_ld<GENERIC>(node + 2), py = node_ld<GENERIC>(node + 3), pz = node_ld<GENERIC>(node + 4);
    const unsigned hdr = __float_as_uint(n0.w);
    // A' = 2^(e - 127) / d: exponent byte moved into place, one multiply
    const float Ax = __uint_as_float((hdr << 23) & 0x7f800000u) * r.idx, Ay = __uint_as_float((hdr << 15) & 0x7f800000u) * r.idy,
                Az = __uint_as_float((hdr << 7) & 0x7f800000u) * r.idz;
    const float Dx = fmaf(n0.x, r.idx, -r.oix) - Ax, Dy = fmaf(n0.y, r.idy, -r.oiy) - Ay, Dz = fmaf(n0.z, r.idz, -r.oiz) - Az;
    const float k = 2.384185791015625e-7f;                               // 2^-22
    const float Sx = fmaf(fabsf(Ax), k, r.sx), Sy = fmaf(fabsf(Ay), k, r.sy), Sz = fmaf(fabsf(Az), k, r.sz);
    const float Bnx = Dx - Sx, Bny = Dy - Sy, Bnz = Dz - Sz, Bfx = Dx + Sx, Bfy = Dy + Sy, Bfz = Dz + Sz;
    // near / far plane words per axis: lo planes are near for d >= 0 (bit of octinv set)
    const unsigned mx = (r.octinv & 1u) ? 0xffffffffu : 0u, my = (r.octinv & 2u) ? 0xffffffffu : 0u, mz = (r.octinv & 4u) ? 0xffffffffu : 0u;
    const unsigned lx0 = __float_as_uint(px.x), lx1 = __float_as_uint(px.y), hx0 = __float_as_uint(px.z), hx1 = __float_as_uint(px.w);
    const unsigned ly0 = __float_as_uint(py.x), ly1 = __float_as_uint(py.y), hy0 = __float_as_uint(py.z), hy1 = __float_as_uint(py.w);
    const unsigned lz0 = __float_as_uint(pz.x), lz1 = __float_as_uint(pz.y), hz0 = __float_as_uint(pz.z), hz1 = __float_as_uint(pz.w);
#define RTB_MUX(m, a, b) (((a) & (m)) | ((b) & ~(m)))
    unsigned miss = 0;
    test_half<1>(RTB_MUX(mx, lx1, hx1), RTB_MUX(my, ly1, hy1), RTB_MUX(mz, lz1, hz1), RTB_MUX(mx, hx1, lx1), RTB_MUX(my, hy1, ly1), RTB_MUX(mz, hz1, lz1),
                 Ax, Ay, Az, Bnx, Bny, Bnz, Bfx, Bfy, Bfz, tmax, miss, H.tn45, H.tn67);
    test_half<0>(RTB_MUX(mx, lx0, hx0), RTB_MUX(my, ly0, hy0), RTB_MUX(mz, lz0, hz0), RTB_MUX(mx, hx0, lx0), RTB_MUX(my, hy0, ly0), RTB_MUX(mz, hz0, lz0),
                 Ax, Ay, Az, Bnx, Bny, Bnz, Bfx, Bfy, Bfz, tmax, miss, H.tn01, H.tn23);
#undef RTB_MUX
    // the funnel shifts pushed the 8 sign bits in from child 7 down to child 0: child k is bit k
    H.hits = ~miss & 0xffu;
}

struct InstXf {                  // instance transform, loaded when a TLAS leaf is entered
    float4 q_pose, q_inv;
    f3 pose_t, inv_scale, inv_t, scale;
    int mesh, material;
};

XF void load_instance(const float4* __restrict__ inst, int i, InstXf& x)
{
    const float4* p = inst + (size_t)i * INST_F4;
    float4 a = ld_keep(p + 0), b = ld_keep(p + 1), c = ld_keep(p + 2), d = ld_keep(p + 3), e = ld_keep(p + 4), f = ld_keep(p + 5);
    x.q_pose = a;
    x.pose_t = mk3(b.x, b.y, b.z); x.mesh = __float_as_int(b.w);
    x.inv_scale = mk3(c.x, c.y, c.z); x.material = __float_as_int(c.w);
    x.q_inv = d;
    x.inv_t = mk3(e.x, e.y, e.z);
    x.scale = mk3(f.x, f.y, f.z);
}

}  // namespace rtb

#include "rtb_trace_state.cuh"
