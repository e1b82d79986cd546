// rtb_wavefront.cu -- wavefront pipeline for secondary rays (RTB_MODE_REFLECT / RTB_MODE_PATH).
//
// The reference traces primary rays only (raycast.cu:204); reflections and path tracing are the
// north-star extensions whose semantics are fixed by oracle/rt_oracle.c (render_pixel).  One kernel
// launch per bounce: persistent warps pull batches of 32 rays from the bounce's queue with a global
// work counter, trace + shade them, and append surviving paths to the next queue with warp-ballot
// compaction (one atomicAdd per warp).  Bounce 0 generates camera rays in-kernel (no queue read).
// Queue record: 2 x float4 = 32 B per ray (origin|pixel, direction|unused); per-pixel throughput and
// radiance live in float4 arrays indexed by pixel.
#include "rtb_render.cuh"
#include "rtb_shade.cuh"
#include <algorithm>

struct rtb_context;

namespace rtb {

struct WaveState {
    float4* q_org[2] = { nullptr, nullptr };
    float4* q_dir[2] = { nullptr, nullptr };
    float4* thr = nullptr;      // per pixel path throughput
    float4* acc = nullptr;      // per pixel radiance (REFLECT: running sum; PATH: sum over samples)
    int* qcount = nullptr;      // [n_counters] queue sizes
    int* qhead = nullptr;       // [n_counters] work counters
    size_t cap_pixels = 0;
    int n_counters = 0;
};

__constant__ float c_sky[3] = { 1.0f, 0.8f, 0.6f };     // raycast.cu:210-213 (BGR)

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

template <bool PATH, bool COUNT>
__global__ void __launch_bounds__(128) k_wave(SceneDev sc, FrameParams P, int bounce, int sample, int counter_base, int n_local,
                                              int tiles_x, WaveState ws, unsigned char* __restrict__ img, size_t pitch,
                                              int32_t* __restrict__ aov_prim, int32_t* __restrict__ aov_inst,
                                              float* __restrict__ aov_t, Counters* ctr)
{
    const int lane = threadIdx.x & 31;
    const unsigned lt = (1u << lane) - 1u;
    const int cin = counter_base + bounce, cout = cin + 1;
    const int count = bounce == 0 ? n_local : ws.qcount[cin];
    const float4* __restrict__ qo_in = ws.q_org[bounce & 1];
    const float4* __restrict__ qd_in = ws.q_dir[bounce & 1];
    float4* __restrict__ qo_out = ws.q_org[(bounce + 1) & 1];
    float4* __restrict__ qd_out = ws.q_dir[(bounce + 1) & 1];
    unsigned rays = 0;
    for (;;) {
        int base = 0;
        if (lane == 0) base = atomicAdd(&ws.qhead[cin], 32);
        base = __shfl_sync(0xffffffffu, base, 0);
        if (base >= count) break;
        const int idx = base + lane;
        bool active = idx < count;
        int x = 0, y = 0;
        f3 org = mk3(0.f, 0.f, 0.f), dir = mk3(0.f, 0.f, 1.f);
        if (active) {
            if (bounce == 0) {
                active = local_to_pixel(P, idx, tiles_x, x, y);
                if (active) {
                    org = mk3(P.ox, P.oy, P.oz);
                    float px = (float)x, py = (float)y;
                    if (PATH) {
                        unsigned c4[4] = { (unsigned)(y * P.width + x), (unsigned)sample, 0u, 0u };
                        philox4x32(c4, P.seed, 0x5eed5eedu);
                        px = xa(px, xs(u01(c4[0]), 0.5f)); py = xa(py, xs(u01(c4[1]), 0.5f));
                    }
                    dir = generate_direction(P, px, py);
                }
            } else {
                const float4 a = qo_in[idx], b = qd_in[idx];
                org = mk3(a.x, a.y, a.z); dir = mk3(b.x, b.y, b.z);
                const int pix = __float_as_int(a.w);
                y = pix / P.width; x = pix - y * P.width;
            }
        }
        bool cont = false;
        f3 norg = org, ndir = dir;
        const int pix = y * P.width + x;
        if (active) {
            Hit h;
            trace<false, COUNT>(sc, org, dir, h, ctr);
            rays++;
            if (!PATH && bounce == 0) {
                if (aov_prim) aov_prim[pix] = h.prim;
                if (aov_inst) aov_inst[pix] = h.inst;
                if (aov_t) aov_t[pix] = h.min;
            }
            f3 thr = mk3(1.f, 1.f, 1.f), acc = mk3(0.f, 0.f, 0.f);
            if (bounce > 0) { const float4 t4 = ws.thr[pix]; thr = mk3(t4.x, t4.y, t4.z); }
            if (bounce > 0 || (PATH && sample > 0)) { const float4 a4 = ws.acc[pix]; acc = mk3(a4.x, a4.y, a4.z); }
            if (h.min == FLT_MAX) {
                acc = mk3(xa(acc.x, xm(thr.x, c_sky[0])), xa(acc.y, xm(thr.y, c_sky[1])), xa(acc.z, xm(thr.z, c_sky[2])));
            } else {
                InstXf xf;
                load_instance(sc.inst, h.inst, xf);
                const f3 c = base_colour(sc, h, xf.material, xf.mesh);
                const f3 nrm = world_normal(sc, h, xf);
                if (!PATH) {
                    const float il = lit_illumination<COUNT>(sc, P, h, nrm, ctr, rays);
                    const float m = sc.materials[xf.material].metallic;
                    const float w = xs(1.0f, m);
                    acc = mk3(xa(acc.x, xm(thr.x, xm(w, xm(il, c.x)))), xa(acc.y, xm(thr.y, xm(w, xm(il, c.y)))),
                              xa(acc.z, xm(thr.z, xm(w, xm(il, c.z)))));
                    if (m > 0.0f && bounce < P.max_depth) {
                        cont = true;
                        thr = mk3(xm(thr.x, xm(m, c.x)), xm(thr.y, xm(m, c.y)), xm(thr.z, xm(m, c.z)));
                        ndir = reflect_dir(dir, nrm);
                        norg = xadd(h.loc, xscale(ndir, 1e-4f));
                    }
                } else {
                    thr = mk3(xm(thr.x, c.x), xm(thr.y, c.y), xm(thr.z, c.z));
                    if (bounce < P.max_depth) {
                        cont = true;
                        unsigned r4[4] = { (unsigned)pix, (unsigned)sample, (unsigned)(bounce + 1), 0u };
                        philox4x32(r4, P.seed, 0x5eed5eedu);
                        ndir = cosine_dir(nrm, u01(r4[0]), u01(r4[1]));
                        norg = xadd(h.loc, xscale(ndir, 1e-4f));
                    }
                }
            }
            if (cont) {
                ws.thr[pix] = make_float4(thr.x, thr.y, thr.z, 0.f);
                ws.acc[pix] = make_float4(acc.x, acc.y, acc.z, 0.f);
            } else if (PATH) {
                ws.acc[pix] = make_float4(acc.x, acc.y, acc.z, 0.f);          // resolved after the last sample
            } else {
                unsigned char* p = img + (size_t)y * pitch + 3 * (size_t)x;
                p[0] = to_u8(xm(clamp01(acc.x), 255.0f)); p[1] = to_u8(xm(clamp01(acc.y), 255.0f)); p[2] = to_u8(xm(clamp01(acc.z), 255.0f));
            }
        }
        // warp-ballot compaction of surviving paths into the next bounce's queue
        const unsigned mask = __ballot_sync(0xffffffffu, cont);
        if (mask) {
            int wbase = 0;
            const int leader = __ffs(mask) - 1;
            if (lane == leader) wbase = atomicAdd(&ws.qcount[cout], __popc(mask));
            wbase = __shfl_sync(0xffffffffu, wbase, leader);
            if (cont) {
                const int pos = wbase + __popc(mask & lt);
                qo_out[pos] = make_float4(norg.x, norg.y, norg.z, __int_as_float(pix));
                qd_out[pos] = make_float4(ndir.x, ndir.y, ndir.z, 0.f);
            }
        }
    }
    for (int o = 16; o > 0; o >>= 1) rays += __shfl_xor_sync(0xffffffffu, rays, o);
    if (lane == 0 && rays) atomicAdd(&ctr->rays, (unsigned long long)rays);
}

// PATH: mean over samples -> 8-bit pixel
__global__ void k_resolve(FrameParams P, int n_local, int tiles_x, const float4* __restrict__ acc, unsigned char* __restrict__ img, size_t pitch)
{
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n_local) return;
    int x, y;
    if (!local_to_pixel(P, idx, tiles_x, x, y)) return;
    const float4 a = acc[y * P.width + x];
    const float inv = xd(1.0f, (float)P.spp);
    unsigned char* p = img + (size_t)y * pitch + 3 * (size_t)x;
    p[0] = to_u8(xm(clamp01(xm(a.x, inv)), 255.0f)); p[1] = to_u8(xm(clamp01(xm(a.y, inv)), 255.0f)); p[2] = to_u8(xm(clamp01(xm(a.z, inv)), 255.0f));
}

#define WCK(x) do { cudaError_t _e = (x); if (_e != cudaSuccess) return _e; } while (0)

static void wave_free(WaveState* w)
{
    for (int i = 0; i < 2; i++) { if (w->q_org[i]) cudaFree(w->q_org[i]); if (w->q_dir[i]) cudaFree(w->q_dir[i]); w->q_org[i] = w->q_dir[i] = nullptr; }
    if (w->thr) cudaFree(w->thr);
    if (w->acc) cudaFree(w->acc);
    if (w->qcount) cudaFree(w->qcount);
    if (w->qhead) cudaFree(w->qhead);
    w->thr = w->acc = nullptr; w->qcount = w->qhead = nullptr; w->cap_pixels = 0; w->n_counters = 0;
}

int g_wave_launches = 0;

cudaError_t launch_wavefront(rtb_context*, void** state, const SceneDev& sc, const FrameParams& P, unsigned char* img, size_t pitch,
                             const rtb_aovs& aov, Counters* ctr, bool count, cudaStream_t st)
{
    WaveState* w = (WaveState*)*state;
    if (!w) { w = new WaveState(); *state = w; }
    const bool path = P.mode == RTB_MODE_PATH;
    const int bands = (P.height + BAND_ROWS - 1) / BAND_ROWS;
    const int my_bands = (bands - P.tile_rank + P.tile_count - 1) / P.tile_count;
    if (my_bands <= 0) return cudaSuccess;
    const int tiles_x = (P.width + 7) / 8;
    const int n_local = tiles_x * 8 * my_bands * BAND_ROWS;           // padded pixel slots of this rank
    const size_t npix = (size_t)P.width * P.height;
    const int passes = path ? P.spp : 1;
    const int per_pass = P.max_depth + 3;
    const int n_counters = passes * per_pass;
    if (w->cap_pixels < npix || w->n_counters < n_counters) {
        wave_free(w);
        for (int i = 0; i < 2; i++) {
            WCK(cudaMalloc((void**)&w->q_org[i], sizeof(float4) * (size_t)n_local));
            WCK(cudaMalloc((void**)&w->q_dir[i], sizeof(float4) * (size_t)n_local));
        }
        WCK(cudaMalloc((void**)&w->thr, sizeof(float4) * npix));
        WCK(cudaMalloc((void**)&w->acc, sizeof(float4) * npix));
        WCK(cudaMalloc((void**)&w->qcount, sizeof(int) * (size_t)n_counters));
        WCK(cudaMalloc((void**)&w->qhead, sizeof(int) * (size_t)n_counters));
        w->cap_pixels = npix; w->n_counters = n_counters;
    }
    WCK(cudaMemsetAsync(w->qcount, 0, sizeof(int) * (size_t)n_counters, st));
    WCK(cudaMemsetAsync(w->qhead, 0, sizeof(int) * (size_t)n_counters, st));
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int grid = std::min(sms * 8, (n_local + 127) / 128);
    g_wave_launches = 0;
    if (path && (aov.prim || aov.inst || aov.t)) {
        FrameParams Q = P; Q.mode = RTB_MODE_REFERENCE;                 // first-hit AOVs from the un-jittered centre ray
        WCK(launch_render(sc, Q, img, pitch, aov, ctr, count, st));
        g_wave_launches++;
    }
    for (int s = 0; s < passes; s++) {
        for (int b = 0; b <= P.max_depth; b++) {
            if (path) {
                if (count) k_wave<true, true><<<grid, 128, 0, st>>>(sc, P, b, s, s * per_pass, n_local, tiles_x, *w, img, pitch, aov.prim, aov.inst, aov.t, ctr);
                else k_wave<true, false><<<grid, 128, 0, st>>>(sc, P, b, s, s * per_pass, n_local, tiles_x, *w, img, pitch, aov.prim, aov.inst, aov.t, ctr);
            } else {
                if (count) k_wave<false, true><<<grid, 128, 0, st>>>(sc, P, b, s, s * per_pass, n_local, tiles_x, *w, img, pitch, aov.prim, aov.inst, aov.t, ctr);
                else k_wave<false, false><<<grid, 128, 0, st>>>(sc, P, b, s, s * per_pass, n_local, tiles_x, *w, img, pitch, aov.prim, aov.inst, aov.t, ctr);
            }
            g_wave_launches++;
        }
    }
    if (path) { k_resolve<<<(n_local + 255) / 256, 256, 0, st>>>(P, n_local, tiles_x, w->acc, img, pitch); g_wave_launches++; }
    return cudaGetLastError();
}

void wave_release(void* state)
{
    if (!state) return;
    WaveState* w = (WaveState*)state;
    wave_free(w);
    delete w;
}

}  // namespace rtb
