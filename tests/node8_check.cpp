// Host check of cuda_raytracing_b200/csrc/rtb_node8.cuh (compiled by tests/test_oracle_cpu.py with g++): encode random
// child boxes of many magnitudes / shapes, decode them with the tracer's arithmetic, and require (1) containment -- the
// decoded box holds the exact one on every axis, (2) tightness -- no plane is more than two cells out and the cell is
// within a factor 4 of extent / 255, (3) empty slots carry the EMPTY ref and a clear bit in the valid mask.
#include "../cuda_raytracing_b200/csrc/rtb_node8.cuh"
#include <cstdio>
#include <cstdlib>
#include <random>

int main()
{
    std::mt19937 rng(12345);
    std::uniform_real_distribution<float> U(0.f, 1.f);
    long boxes = 0, refused = 0;
    double worst_cells = 0;
    for (int trial = 0; trial < 200000; trial++) {
        const int count = 1 + (int)(rng() % 8);
        const float scale = std::ldexp(1.0f, (int)(rng() % 60) - 30);            // node extent 2^-30 .. 2^29
        const float centre = (U(rng) - 0.5f) * std::ldexp(1.0f, (int)(rng() % 40) - 10);
        const int shape = (int)(rng() % 4);                                       // 0 generic, 1 flat on an axis, 2 point boxes, 3 nested tiny
        float lo[8][3], hi[8][3]; int32_t ref[8];
        for (int k = 0; k < count; k++) {
            for (int a = 0; a < 3; a++) {
                float c = centre + (U(rng) - 0.5f) * scale, h = U(rng) * scale * (shape == 3 ? 1e-4f : 0.3f);
                if (shape == 2 || (shape == 1 && a == (trial % 3))) h = 0.f;
                if (shape == 1 && a == (trial % 3)) c = centre;
                lo[k][a] = c - h; hi[k][a] = c + h;
                if (lo[k][a] > hi[k][a]) { float t = lo[k][a]; lo[k][a] = hi[k][a]; hi[k][a] = t; }
            }
            ref[k] = k;
        }
        rtb::Node8 n;
        if (!rtb::node8_encode(lo, hi, ref, count, n)) { refused++; continue; }
        for (int a = 0; a < 3; a++) {
            float mn = INFINITY, mx = -INFINITY;
            for (int k = 0; k < count; k++) { mn = fminf(mn, lo[k][a]); mx = fmaxf(mx, hi[k][a]); }
            const float cell = rtb::node8_cell(n, a), ext = mx - mn;
            if (ext > 0.f && cell > 1.2e-38f * 4 && !(cell <= 4.0f * ext / 255.0f * 1.0001f) && cell > std::fabs(mn) * 2.4e-7f) {
                // a larger cell is legitimate only when the origin's own ulp forces it (|origin| >> extent)
                printf("cell too coarse: trial %d axis %d cell %g ext %g origin %g\n", trial, a, cell, ext, mn); return 1;
            }
        }
        for (int k = 0; k < 8; k++) {
            float dl[3], dh[3];
            rtb::node8_child_box(n, k, dl, dh);
            if (k >= count) {
                if (n.ref[k] != rtb::NODE8_EMPTY || ((n.exps >> (24 + k)) & 1u)) { printf("empty slot not marked\n"); return 1; }
                continue;
            }
            boxes++;
            for (int a = 0; a < 3; a++) {
                const float cell = rtb::node8_cell(n, a);
                if (!(dl[a] <= lo[k][a]) || !(dh[a] >= hi[k][a])) {
                    printf("NOT CONSERVATIVE: trial %d child %d axis %d: [%a, %a] decoded [%a, %a]\n", trial, k, a, lo[k][a], hi[k][a], dl[a], dh[a]);
                    return 1;
                }
                const double out = std::fmax((double)lo[k][a] - dl[a], (double)dh[a] - hi[k][a]) / cell;
                if (out > worst_cells) worst_cells = out;
                if (out > 2.0 + 1e-6) { printf("loose: trial %d child %d axis %d: %g cells out\n", trial, k, a, out); return 1; }
            }
        }
        if ((n.exps >> 24) != (count >= 8 ? 0xffu : ((1u << count) - 1u))) { printf("valid mask wrong\n"); return 1; }
    }
    printf("ok: %ld boxes, %ld nodes refused, worst plane %.3f cells out\n", boxes, refused, worst_cells);
    return refused > 0 ? 1 : 0;
}
