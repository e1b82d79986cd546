// rtb_shade.cuh -- camera ray generation and material shading with the reference's arithmetic.
#pragma once
#include "rtb_trace.cuh"

namespace rtb {

// raycast.cu:156-188  pixel -> world-space ray direction.  The inverse-camera rotation quaternion
// (raycast.cu:185 -> transforms.hpp:148-163) is constant per frame and arrives precomputed in P.q_cam.
XF f3 generate_direction(const FrameParams& P, float px, float py)
{
    const float* K = P.K_inv;
    float dx = xa(xa(xm(K[0], px), xm(K[1], py)), xm(K[2], 1.0f));     // utils.hpp:134-140
    float dy = xa(xa(xm(K[3], px), xm(K[4], py)), xm(K[5], 1.0f));
    float dz = xa(xa(xm(K[6], px), xm(K[7], py)), xm(K[8], 1.0f));
    const float a = dx, b = dy;
    const float radius = __fsqrt_rn(xa(xm(a, a), xm(b, b)));
    const float theta = xatanf(radius);                              // raycast.cu:170, host-libm bits (exact.cuh)
    // raycast.cu:172: float products, summed in double because of the `1.0` literal
    const double s = (((1.0 + (double)xm(P.D[0], theta)) + (double)xm(xm(P.D[1], theta), theta)) +
                      (double)xm(xm(xm(P.D[2], theta), theta), theta)) +
                     (double)xm(xm(xm(xm(P.D[3], theta), theta), theta), theta);
    const float thetad = __double2float_rn(__dmul_rn((double)theta, s));
    const float scale = xd(thetad, radius);
    dx = xm(scale, a);
    dy = xm(scale, b);
    f3 d = xnormalize(mk3(dx, dy, dz));
    d = mk3(d.x, d.z, -d.y);                                           // raycast.cu:182
    d = xapply_quat(P.q_cam, d);                                       // raycast.cu:185
    return xnormalize(d);                                              // raycast.cu:188
}

// raycast.cu:224-245 base colour (BGR floats in [0,1]); uv from TrianglePrimitive.hpp:171-181
XF f3 base_colour(const SceneDev& sc, const Hit& h, int material, int mesh)
{
    const rtb_material* m = sc.materials + material;
    const int tw = m->texture_width;
    if (tw > 0) {
        const int th = m->texture_height;
        const float2* uvp = sc.meshes[mesh].uvs + (size_t)h.slot * 3;
        const float2 uv0 = __ldg(uvp + 0), uv1 = __ldg(uvp + 1), uv2 = __ldg(uvp + 2);
        const float w = xs(xs(1.0f, h.u), h.v);
        const float ux = xa(xa(xm(w, uv0.x), xm(h.v, uv1.x)), xm(h.u, uv2.x));
        const float uy = xa(xa(xm(w, uv0.y), xm(h.v, uv1.y)), xm(h.u, uv2.y));
        int tex_x = __float2int_rz(xm(ux, (float)tw));
        int tex_y = __double2int_rz(__dmul_rn(1.0 - (double)uy, (double)(float)th));
        tex_x = (int)fmaxf((float)(tex_x % tw), 0.0f);
        tex_y = (int)fmaxf((float)(tex_y % th), 0.0f);
        const unsigned char* tc = m->texture + (size_t)tex_y * m->texture_pitch + 3 * (size_t)tex_x;
        return mk3(xm((float)tc[0], 0.0039215f), xm((float)tc[1], 0.0039215f), xm((float)tc[2], 0.0039215f));
    }
    return mk3(m->albedo[0], m->albedo[1], m->albedo[2]);
}

// raycast.cu:115-122 world-space normal of the accepted hit
XF f3 world_normal(const SceneDev& sc, const Hit& h, const InstXf& xf)
{
    const float4 q0 = __ldg(sc.meshes[xf.mesh].tris + (size_t)h.slot * TRI_F4);
    f3 n = xapply_quat(xf.q_inv, mk3(q0.x, q0.y, q0.z));
    return xnormalize(xmul3(n, xf.scale));
}

XF unsigned char to_u8(float v) { return (unsigned char)__float2int_rz(v); }   // raycast.cu:292-294
XF float clamp01(float v) { return fminf(1.0f, fmaxf(0.0f, v)); }

// Lambert + shadow term of the commented block raycast.cu:262-280 with the live light setup 249-259.
// Returns the clamped illumination (289-290); counts the any-hit ray if one was cast.  Warp-collective: every
// lane calls it, lanes without a hit pass lit=false.
template <bool COUNT>
XF float lit_illumination(const SceneDev& sc, Stack& stk, const FrameParams& P, bool lit, const Hit& h, f3 nrm, Counters* ctr, unsigned& rays)
{
    const f3 L = mk3(P.light_dir.x, P.light_dir.y, P.light_dir.z);
    const float cos_illum = lit ? xdot(nrm, L) : 0.0f;
    float illum = __double2float_rn(__dmul_rn(0.4, (double)cos_illum));
    const bool cast = lit && cos_illum > 0.0f;
    const f3 so = xadd(h.loc, xscale(L, 1e-4f));                       // raycast.cu:259
    Hit sh;
    trace<true, COUNT>(sc, stk, cast, so, L, sh, ctr);
    if (cast) {
        rays++;
        if (sh.min == FLT_MAX) illum = cos_illum;
    }
    illum = fminf(1.0f, illum);
    illum = fmaxf(0.4f, illum);
    return illum;
}

}  // namespace rtb
