// rtb_trace.cuh -- two-level wide-BVH traversal with the reference's closest-hit / any-hit semantics.
//
// Replaces cast_ray (raycast.cu:21-142): the linear instance loop becomes a TLAS descent, the
// 48-byte AoS BVH2 nodes + per-leaf index lists become 112-byte BVH4 nodes + in-order triangle
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
// GENERIC: the node may live in shared memory (staged top levels), so plain generic loads instead of ld.global.nc
template <bool GENERIC>
XF float4 node_ld(const float4* p) { return GENERIC ? *p : ld_keep(p); }

template <bool GENERIC>
XF void test_node(const float4* __restrict__ node, const InstRay& r, float tmax, unsigned keys[4], int4& refs)
{
    const float4 nxv = node_ld<GENERIC>(node + r.nx), fxv = node_ld<GENERIC>(node + (r.nx ^ 1));
    const float4 nyv = node_ld<GENERIC>(node + r.ny), fyv = node_ld<GENERIC>(node + (r.ny ^ 1));
    const float4 nzv = node_ld<GENERIC>(node + r.nz), fzv = node_ld<GENERIC>(node + (r.nz ^ 1));
    const float4 rf = node_ld<GENERIC>(node + 6);
    refs = make_int4(__float_as_int(rf.x), __float_as_int(rf.y), __float_as_int(rf.z), __float_as_int(rf.w));
    const float ax[4] = { nxv.x, nxv.y, nxv.z, nxv.w }, bx[4] = { fxv.x, fxv.y, fxv.z, fxv.w };
    const float ay[4] = { nyv.x, nyv.y, nyv.z, nyv.w }, by[4] = { fyv.x, fyv.y, fyv.z, fyv.w };
    const float az[4] = { nzv.x, nzv.y, nzv.z, nzv.w }, bz[4] = { fzv.x, fzv.y, fzv.z, fzv.w };
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const float tnx = fmaf(ax[i], r.idnx, -r.oinx), tny = fmaf(ay[i], r.idny, -r.oiny), tnz = fmaf(az[i], r.idnz, -r.oinz);
        const float tfx = fmaf(bx[i], r.idfx, -r.oifx), tfy = fmaf(by[i], r.idfy, -r.oify), tfz = fmaf(bz[i], r.idfz, -r.oifz);
        const float tn = fmaxf(fmaxf(tnx, tny), fmaxf(tnz, 0.0f));
        const float tf = fminf(fminf(tfx, tfy), fminf(tfz, tmax));
        const bool hit = tn <= tf;        // empty slots hold (+inf, +inf) boxes: tn = +inf or tf = -inf, never a hit
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
