// Host check of the compressed wide node's arithmetic (csrc/rtb_cwnode.cuh + the decoder of csrc/rtb_trace.cuh), built and
// run by tests/test_oracle_cpu.py::test_cwnode_quantisation_and_slab_slack.
//  1. cw_quantise: every child box lies inside its decoded box (exact arithmetic), no plane more than one cell out.
//  2. the tracer's plane evaluation  t = fma(1 + q 2^-15, A', D -+ S)  restated here operation by operation in binary32:
//     for random rays the computed near value never exceeds and the computed far value never falls below the exact parameter
//     (o, d, origin, q, cell taken as exact reals; long double), i.e. the box test can only accept more than the exact test.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <random>
#include "../cuda_raytracing_b200/csrc/rtb_cwnode.cuh"
using namespace rtb;

static float as_float(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
static float safe_rcp(float d) { float a = std::fabs(d) < 1e-20f ? std::copysign(1e-20f, d) : d; return 1.0f / a; }

int main()
{
    std::mt19937_64 rng(12345);
    std::uniform_real_distribution<double> U(0.0, 1.0);
    long boxes = 0, planes = 0, worst_cells = 0;
    double worst_rel = 0;
    for (int it = 0; it < 200000; it++) {
        // a node box of random magnitude and position, children inside it (some flat, some points)
        const double mag = std::pow(10.0, -6.0 + 12.0 * U(rng));
        float nlo[3], nhi[3];
        for (int a = 0; a < 3; a++) {
            const double c = (U(rng) - 0.5) * 2.0 * mag, e = U(rng) < 0.1 ? 0.0 : U(rng) * mag * std::pow(10.0, -4.0 * U(rng));
            nlo[a] = (float)c; nhi[a] = (float)(c + e);
            if (nhi[a] < nlo[a]) nhi[a] = nlo[a];
        }
        int e[3];
        for (int a = 0; a < 3; a++) { e[a] = cw_axis_exponent(nlo[a], nhi[a]); if (!e[a]) { printf("exponent failed\n"); return 1; } }
        const float extent = (float)std::max({ std::fabs((double)nlo[0]), std::fabs((double)nhi[0]), std::fabs((double)nlo[1]), std::fabs((double)nhi[1]),
                                               std::fabs((double)nlo[2]), std::fabs((double)nhi[2]) });
        // a ray: origin anywhere up to 1000 extents away, any direction incl. axis-parallel ones
        float o[3], d[3];
        for (int a = 0; a < 3; a++) {
            o[a] = (float)((U(rng) - 0.5) * 2.0 * extent * std::pow(10.0, 3.0 * U(rng) - 1.0));
            d[a] = U(rng) < 0.05 ? 0.0f : (float)((U(rng) - 0.5) * std::pow(10.0, -3.0 * U(rng)));
        }
        for (int child = 0; child < 8; child++) {
            float clo[3], chi[3];
            for (int a = 0; a < 3; a++) {
                const double x = (double)nlo[a] + U(rng) * ((double)nhi[a] - nlo[a]), y = (double)nlo[a] + U(rng) * ((double)nhi[a] - nlo[a]);
                clo[a] = (float)std::min(x, y); chi[a] = (float)std::max(x, y);
                if (clo[a] < nlo[a]) clo[a] = nlo[a];
                if (chi[a] > nhi[a]) chi[a] = nhi[a];
                if (U(rng) < 0.1) chi[a] = clo[a];
            }
            boxes++;
            for (int a = 0; a < 3; a++) {
                int ql, qh;
                if (!cw_quantise(nlo[a], e[a], clo[a], chi[a], ql, qh)) { printf("quantise failed\n"); return 1; }
                const long double cell = std::ldexp((long double)1.0, e[a] - 127 - CW_EXP_BIAS);
                const long double L = (long double)nlo[a] + ql * cell, H = (long double)nlo[a] + qh * cell;
                if (L > clo[a] || H < chi[a] || ql < 0 || qh > 255) { printf("box not contained\n"); return 1; }
                const long cells = (long)std::ceil((double)std::max((clo[a] - L) / cell, (H - chi[a]) / cell));
                if (cells > worst_cells) worst_cells = cells;
                // ---- the tracer's evaluation of both planes (rtb_trace.cuh: setup_slabs + test_node8), binary32 step by step
                const float id = safe_rcp(d[a]);
                const float oi = o[a] * id;
                const float s = 9.5367431640625e-7f * (4.0f * extent + std::fabs(o[a])) * std::fabs(id);
                const float A = as_float((uint32_t)e[a] << 23) * id;
                const float D = std::fmaf(nlo[a], id, -oi) - A;
                const float S = std::fmaf(std::fabs(A), 2.384185791015625e-7f, s);
                const float Bn = D - S, Bf = D + S;
                const int qn = id >= 0.0f ? ql : qh, qf = id >= 0.0f ? qh : ql;
                const float fn = as_float(0x3f800000u | ((uint32_t)qn << 8)), ff = as_float(0x3f800000u | ((uint32_t)qf << 8));
                const float tn = std::fmaf(fn, A, Bn), tf = std::fmaf(ff, A, Bf);
                // exact parameters of the same planes for the ray (o, d): 1/d exact unless clamped
                const long double dd = std::fabs(d[a]) < 1e-20f ? (long double)std::copysign(1e-20f, d[a]) : (long double)d[a];
                const long double en = ((long double)nlo[a] + qn * cell - o[a]) / dd, ef = ((long double)nlo[a] + qf * cell - o[a]) / dd;
                if (!std::isfinite(tn) || !std::isfinite(tf)) { printf("non-finite plane parameter\n"); return 1; }
                if ((long double)tn > en || (long double)tf < ef) {
                    printf("NOT conservative: axis %d near %.9g exact %.9Lg far %.9g exact %.9Lg (o %g d %g origin %g e %d q %d %d)\n", a, tn, en, tf, ef,
                           o[a], d[a], nlo[a], e[a], qn, qf);
                    return 1;
                }
                const double scale = (double)(std::fabs(en) + std::fabs(ef)) + 1e-300;
                worst_rel = std::max(worst_rel, (double)std::max(en - tn, tf - ef) / scale);
                planes += 2;
            }
        }
    }
    printf("ok: %ld boxes, %ld planes, worst snap %ld cells, worst slack %.3g of |t|\n", boxes, planes, worst_cells, worst_rel);
    return worst_cells <= 1 ? 0 : 1;
}
