// rtb_wave_common.cuh -- helpers shared by the wavefront pipeline (rtb_wavefront.cu) and the fused persistent
// path kernel (rtb_mega.cu): sky colour, Philox4x32-10, mirror / cosine-hemisphere directions, pixel mapping.
#pragma once
#include "rtb_render.cuh"
#include "rtb_shade.cuh"

namespace rtb {

__device__ constexpr float c_sky[3] = { 1.0f, 0.8f, 0.6f };     // raycast.cu:210-213 (BGR)

// Philox4x32-10, identical to oracle/rt_oracle.c
XF void philox4x32(unsigned c[4], unsigned k0, unsigned k1)
{
#pragma unroll
    for (int r = 0; r < 10; r++) {
        unsigned long long p0 = (unsigned long long)0xD2511F53u * c[0], p1 = (unsigned long long)0xCD9E8D57u * c[2];
        unsigned n0 = (unsigned)(p1 >> 32) ^ c[1] ^ k0, n1 = (unsigned)p1;
        unsigned n2 = (unsigned)(p0 >> 32) ^ c[3] ^ k1, n3 = (unsigned)p0;
        c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
}
XF float u01(unsigned x) { return xm((float)(x >> 8), 1.0f / 16777216.0f); }

XF f3 reflect_dir(f3 d, f3 n)
{
    const float k = xm(2.0f, xdot(d, n));
    return xnormalize(xsub(d, xscale(n, k)));
}

XF f3 cosine_dir(f3 n, float u1, float u2)
{
    const float r = __fsqrt_rn(u1), phi = xm(6.2831853f, u2);
    const float lx = xm(r, cosf(phi)), ly = xm(r, sinf(phi)), lz = __fsqrt_rn(fmaxf(0.0f, xs(1.0f, u1)));
    const float sign = copysignf(1.0f, n.z);
    const float a = xd(-1.0f, xa(sign, n.z));
    const float b = xm(xm(n.x, n.y), a);
    const f3 t = mk3(xa(1.0f, xm(xm(xm(sign, n.x), n.x), a)), xm(sign, b), xm(-sign, n.x));
    const f3 bt = mk3(b, xa(sign, xm(xm(n.y, n.y), a)), -n.y);
    return xnormalize(xadd(xadd(xscale(t, lx), xscale(bt, ly)), xscale(n, lz)));
}

// index within this rank's pixel set -> pixel coordinates (8x4 warp tiles, BAND_ROWS-row bands)
XF bool local_to_pixel(const FrameParams& P, int idx, int tiles_x, int& x, int& y)
{
    const int tile = idx >> 5, l = idx & 31;
    const int tx = tile % tiles_x, ty = tile / tiles_x;
    x = tx * 8 + (l & 7);
    const int yl = ty * 4 + (l >> 3);
    const int band = (yl / BAND_ROWS) * P.tile_count + P.tile_rank;
    y = band * BAND_ROWS + (yl % BAND_ROWS);
    return x < P.width && y < P.height;
}

}  // namespace rtb
