// exact.cuh -- reference-exact arithmetic for the predicates that decide pixels.
//
// The reference's host path evaluates IEEE binary32 expressions left to right with no fused
// multiply-add.  Everything that decides WHICH triangle a pixel reports (ray generation, the
// world->mesh ray transform, plane hit, barycentric inside test, world-space distance) is written
// here with __fmul_rn/__fadd_rn/__fsub_rn/__fdiv_rn/__fsqrt_rn, which nvcc never contracts, in the
// reference's operand order.  Each helper cites the reference lines it reproduces
// (/root/reference/CudaRaytracer/...).  Conservative culling (box tests) lives elsewhere and is free
// to use FMA.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace rtb {

#define XF __device__ __forceinline__

XF float xm(float a, float b) { return __fmul_rn(a, b); }
XF float xa(float a, float b) { return __fadd_rn(a, b); }
XF float xs(float a, float b) { return __fsub_rn(a, b); }
XF float xd(float a, float b) { return __fdiv_rn(a, b); }

struct f3 { float x, y, z; };
XF f3 mk3(float x, float y, float z) { f3 r; r.x = x; r.y = y; r.z = z; return r; }

// utils.hpp:53-55  a.x*b.x + a.y*b.y + a.z*b.z
XF float xdot(f3 a, f3 b) { return xa(xa(xm(a.x, b.x), xm(a.y, b.y)), xm(a.z, b.z)); }
XF f3 xsub(f3 a, f3 b) { return mk3(xs(a.x, b.x), xs(a.y, b.y), xs(a.z, b.z)); }
XF f3 xadd(f3 a, f3 b) { return mk3(xa(a.x, b.x), xa(a.y, b.y), xa(a.z, b.z)); }
XF f3 xscale(f3 a, float b) { return mk3(xm(a.x, b), xm(a.y, b), xm(a.z, b)); }
XF f3 xmul3(f3 a, f3 b) { return mk3(xm(a.x, b.x), xm(a.y, b.y), xm(a.z, b.z)); }
// utils.hpp:49-51
XF f3 xcross(f3 a, f3 b)
{
    return mk3(xs(xm(a.y, b.z), xm(a.z, b.y)), xs(xm(a.z, b.x), xm(a.x, b.z)), xs(xm(a.x, b.y), xm(a.y, b.x)));
}

// utils.hpp:12-27, `long` read as 32 bits
XF float q_rsqrt(float number)
{
    float x2 = xm(number, 0.5f);
    int i = __float_as_int(number);
    i = 0x5f3759df - (i >> 1);
    float y = __int_as_float(i);
    return xm(y, xs(1.5f, xm(xm(x2, y), y)));
}
// utils.hpp:37-47
XF f3 xnormalize(f3 v)
{
    float inv = q_rsqrt(xa(xa(xm(v.x, v.x), xm(v.y, v.y)), xm(v.z, v.z)));
    return mk3(xm(v.x, inv), xm(v.y, inv), xm(v.z, inv));
}
// utils.hpp:29-31
XF float xmagnitude(f3 v) { return __fsqrt_rn(xa(xa(xm(v.x, v.x), xm(v.y, v.y)), xm(v.z, v.z))); }

// transforms.hpp:165-176 apply_quat
XF f3 xapply_quat(float4 q, f3 v)
{
    float a = xs(xs(xm(-v.x, q.y), xm(v.y, q.z)), xm(v.z, q.w));
    float b = xs(xa(xm(v.x, q.x), xm(v.y, q.w)), xm(v.z, q.z));
    float c = xs(xa(xm(v.y, q.x), xm(v.z, q.y)), xm(v.x, q.w));
    float d = xs(xa(xm(v.z, q.x), xm(v.x, q.z)), xm(v.y, q.y));
    return mk3(xa(xs(xs(xm(q.x, b), xm(q.y, a)), xm(q.z, d)), xm(q.w, c)),
               xa(xs(xs(xm(q.x, c), xm(q.z, a)), xm(q.w, b)), xm(q.y, d)),
               xa(xs(xs(xm(q.x, d), xm(q.w, a)), xm(q.y, c)), xm(q.z, b)));
}

// atanf with the bits of the host libm the reference's host path calls at raycast.cu:170 (`atan(radius)`).
// glibc 2.39 (this image and the GPU box) implements atanf with the FreeBSD/fdlibm single-precision
// algorithm (argument reduction to |x| < 7/16 via atan(0.5), atan(1), atan(1.5), atan(inf) + an 11-term
// odd/even polynomial).  Restated here with unfused float ops; checked bit-identical to glibc's atanf on
// 5e7 random arguments in [0,4] (tests/test_oracle_cpu.py keeps a sampled fixture of that check).  CUDA's
// own atanf differs from it by 1 ulp on ~6 % of arguments, which is enough to move a grazing-angle hit
// distance by more than the 1e-5 tolerance.
XF float xatanf(float x)
{
    const float hi[4] = { 4.6364760399e-01f, 7.8539812565e-01f, 9.8279368877e-01f, 1.5707962513e+00f };
    const float lo[4] = { 5.0121582440e-09f, 3.7748947079e-08f, 3.4473217170e-08f, 7.5497894159e-08f };
    const int hx = __float_as_int(x), ix = hx & 0x7fffffff;
    if (ix >= 0x4c000000) {                         // |x| >= 2^25 or NaN
        if (ix > 0x7f800000) return xa(x, x);
        return hx > 0 ? xa(hi[3], lo[3]) : xs(-hi[3], lo[3]);
    }
    int id;
    if (ix < 0x3ee00000) {                          // |x| < 0.4375
        if (ix < 0x31000000) return x;              // |x| < 2^-29
        id = -1;
    } else {
        x = fabsf(x);
        if (ix < 0x3f980000) {
            if (ix < 0x3f300000) { id = 0; x = xd(xs(xm(2.0f, x), 1.0f), xa(2.0f, x)); }
            else { id = 1; x = xd(xs(x, 1.0f), xa(x, 1.0f)); }
        } else {
            if (ix < 0x401c0000) { id = 2; x = xd(xs(x, 1.5f), xa(1.0f, xm(1.5f, x))); }
            else { id = 3; x = xd(-1.0f, x); }
        }
    }
    const float z = xm(x, x), w = xm(z, z);
    const float s1 = xm(z, xa(3.3333334327e-01f, xm(w, xa(1.4285714924e-01f, xm(w, xa(9.0908870101e-02f,
                     xm(w, xa(6.6610731184e-02f, xm(w, xa(4.9768779427e-02f, xm(w, 1.6285819933e-02f)))))))))));
    const float s2 = xm(w, xa(-2.0000000298e-01f, xm(w, xa(-1.1111110449e-01f, xm(w, xa(-7.6918758452e-02f,
                     xm(w, xa(-5.8335702866e-02f, xm(w, -3.6531571299e-02f)))))))));
    if (id < 0) return xs(x, xm(x, xa(s1, s2)));
    const float h = id == 0 ? hi[0] : (id == 1 ? hi[1] : (id == 2 ? hi[2] : hi[3]));
    const float l = id == 0 ? lo[0] : (id == 1 ? lo[1] : (id == 2 ? lo[2] : lo[3]));
    const float r = xs(h, xs(xs(xm(x, xa(s1, s2)), l), x));
    return hx < 0 ? -r : r;
}

}  // namespace rtb
