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

// sin and cos of phi in [0, 2 pi] without libm: the operation sequence of oracle/rt_oracle.c: sincos_2pi, in unfused binary64
// (CUDA's sinf/cosf and glibc's differ in the last ulp on a few percent of arguments; a path-traced pixel is only reproducible
// when both sides draw the same directions bit for bit)
XF void sincos_2pi(float phi, float& s, float& c)
{
    const double x = (double)phi;
    const int k = (int)__dadd_rn(__dmul_rn(x, 0.63661977236758134308), 0.5);
    const double y = __dsub_rn(x, __dmul_rn((double)k, 1.57079632679489655800)), z = __dmul_rn(y, y);
#define RTB_H(a, b) __dadd_rn(a, __dmul_rn(z, b))                     /* a + z * b */
    const double ps = RTB_H(-1.66666666666666324348e-01, RTB_H(8.33333333332248946124e-03, RTB_H(-1.98412698298579493134e-04,
                      RTB_H(2.75573137070700676789e-06, RTB_H(-2.50507602534068634195e-08, 1.58969099521155010221e-10)))));
    const double pc = RTB_H(4.16666666666666019037e-02, RTB_H(-1.38888888888741095749e-03, RTB_H(2.48015872894767294178e-05,
                      RTB_H(-2.75573143513906633035e-07, RTB_H(2.08757232129817482790e-09, -1.13596475577881948265e-11)))));
#undef RTB_H
    const double sp = __dadd_rn(y, __dmul_rn(__dmul_rn(y, z), ps));
    const double cp = __dadd_rn(__dsub_rn(1.0, __dmul_rn(0.5, z)), __dmul_rn(__dmul_rn(z, z), pc));
    const int q = k & 3;
    const double sv = q == 0 ? sp : (q == 1 ? cp : (q == 2 ? -sp : -cp));
    const double cv = q == 0 ? cp : (q == 1 ? -sp : (q == 2 ? -cp : sp));
    s = __double2float_rn(sv); c = __double2float_rn(cv);
}

XF f3 cosine_dir(f3 n, float u1, float u2)
{
    const float r = __fsqrt_rn(u1), phi = xm(6.2831853f, u2);
    float sphi, cphi;
    sincos_2pi(phi, sphi, cphi);
    const float lx = xm(r, cphi), ly = xm(r, sphi), lz = __fsqrt_rn(fmaxf(0.0f, xs(1.0f, u1)));
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

// ------------------------------------------------------------------ frame emission with 16-byte stores
// The persistent kernels finish pixels one at a time, in any order, so they do not write the frame: a finished pixel goes to a
// rank-local staging array as ONE packed word (b | g << 8 | r << 16; the staged pipeline leaves its radiance sums instead) and
// this pass, one thread per 16 pixels of a row, turns them into three aligned 16-byte stores (48 bytes = 16 BGR pixels).  At
// N > 1 the frame may be rank 0's memory mapped over NVLink: the link then carries 16-byte stores, not single bytes (the
// reference stores three bytes per pixel, raycast.cu:292-294).
template <bool PACKED>
__global__ void __launch_bounds__(256) k_emit_frame(FrameParams P, int my_rows, int chunks, float scale, const unsigned* __restrict__ pix,
                                                    const float4* __restrict__ acc, unsigned char* __restrict__ img, size_t pitch)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= my_rows * chunks) return;
    const int r = t / chunks, c = t - r * chunks;
    const int y = ((r / BAND_ROWS) * P.tile_count + P.tile_rank) * BAND_ROWS + (r % BAND_ROWS);
    if (y >= P.height) return;
    const int x0 = 16 * c, n = min(16, P.width - x0);
    const size_t row = (size_t)y * P.width + x0;
    unsigned px[16];
    if (PACKED && n == 16 && (P.width & 3) == 0) {
        const uint4* q = (const uint4*)(pix + row);
#pragma unroll
        for (int i = 0; i < 4; i++) { const uint4 v = __ldcg(q + i); px[4 * i] = v.x; px[4 * i + 1] = v.y; px[4 * i + 2] = v.z; px[4 * i + 3] = v.w; }
    } else {
#pragma unroll
        for (int i = 0; i < 16; i++) {
            px[i] = 0u;
            if (i < n) {
                if (PACKED) px[i] = __ldcg(pix + row + i);
                else {
                    const float4 a = __ldcg(acc + row + i);
                    px[i] = (unsigned)to_u8(xm(clamp01(xm(a.x, scale)), 255.0f)) | ((unsigned)to_u8(xm(clamp01(xm(a.y, scale)), 255.0f)) << 8) |
                            ((unsigned)to_u8(xm(clamp01(xm(a.z, scale)), 255.0f)) << 16);
                }
            }
        }
    }
    unsigned char* dst = img + (size_t)y * pitch + 3 * (size_t)x0;
    if (n == 16 && (((size_t)img | pitch) & 15) == 0) {
        unsigned w[12];
#pragma unroll
        for (int g = 0; g < 4; g++) {                       // 4 pixels -> 3 words
            const unsigned p0 = px[4 * g], p1 = px[4 * g + 1], p2 = px[4 * g + 2], p3 = px[4 * g + 3];
            w[3 * g] = (p0 & 0xffffffu) | (p1 << 24);
            w[3 * g + 1] = ((p1 >> 8) & 0xffffu) | (p2 << 16);
            w[3 * g + 2] = ((p2 >> 16) & 0xffu) | (p3 << 8);
        }
        uint4* d4 = (uint4*)dst;
        d4[0] = make_uint4(w[0], w[1], w[2], w[3]); d4[1] = make_uint4(w[4], w[5], w[6], w[7]); d4[2] = make_uint4(w[8], w[9], w[10], w[11]);
    } else {
        for (int i = 0; i < n; i++) { dst[3 * i] = (unsigned char)px[i]; dst[3 * i + 1] = (unsigned char)(px[i] >> 8); dst[3 * i + 2] = (unsigned char)(px[i] >> 16); }
    }
}

template <bool PACKED>
inline cudaError_t launch_emit_frame(const FrameParams& P, float scale, const unsigned* pix, const float4* acc, unsigned char* img, size_t pitch, cudaStream_t st)
{
    const int bands = (P.height + BAND_ROWS - 1) / BAND_ROWS;
    const int my_bands = (bands - P.tile_rank + P.tile_count - 1) / P.tile_count;
    if (my_bands <= 0) return cudaSuccess;
    const int my_rows = my_bands * BAND_ROWS, chunks = (P.width + 15) / 16;
    k_emit_frame<PACKED><<<(my_rows * chunks + 255) / 256, 256, 0, st>>>(P, my_rows, chunks, scale, pix, acc, img, pitch);
    return cudaGetLastError();
}

}  // namespace rtb
