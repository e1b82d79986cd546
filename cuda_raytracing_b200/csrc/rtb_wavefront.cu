// rtb_wavefront.cu -- wavefront pipeline for secondary and shadow rays (RTB_MODE_REFLECT / RTB_MODE_PATH).
//
// The reference traces primary rays only (raycast.cu:204); reflections and path tracing are the north-star
// extensions whose semantics are fixed by oracle/rt_oracle.c (render_pixel).  Per bounce, four stages, each a
// kernel whose warps all run the same code:
//   k_gen          camera rays of this rank's pixels -> ray queue (bounce 0 only)
//   k_trace_queue  closest hit: persistent warps, one ray per lane; a lane that finishes its ray writes the hit
//                  record and is REFILLED from the queue (warp-ballot count of idle lanes, one atomicAdd per warp),
//                  so lanes do not idle while the slowest ray of the warp finishes
//   k_shade        hit record -> material, radiance bookkeeping; appends the shadow ray and the continuation ray to
//                  their queues with warp-ballot compaction
//   k_trace_queue<ANY>  shadow rays (any-hit), adds the lit / shadowed Lambert term to the pixel
// and k_emit_frame (rtb_wave_common.cuh) converts the per-pixel radiance to the 8-bit BGR frame.
// Queue records: ray 2 x float4 (origin|pixel, direction), hit 2 x float4 (min,u,v,inst | location,slot),
// shadow 3 x float4 (origin|pixel, throughput|w, colour|cos).  Per-pixel throughput / radiance: float4 arrays.
#include "rtb_wave_common.cuh"
#include <algorithm>

struct rtb_context;

namespace rtb {


struct WaveState {
    float4* q_org[2] = { nullptr, nullptr };
    float4* q_dir[2] = { nullptr, nullptr };
    float4* hit0 = nullptr; float4* hit1 = nullptr;
    float4* s_org = nullptr; float4* s_p0 = nullptr; float4* s_p1 = nullptr;
    float4* thr = nullptr;          // per pixel path throughput
    float4* acc = nullptr;          // per pixel radiance (REFLECT: running sum; PATH: sum over samples)
    int* counters = nullptr;        // 4 ints per (pass, bounce): ray count, ray head, shadow count, shadow head
    size_t cap_rays = 0, cap_pixels = 0;
    int n_counters = 0;
    void* mega = nullptr;           // state of the fused persistent kernel (rtb_mega.cu)
};

cudaError_t launch_mega(const Tune& tune, LaunchInfo& info, void** state, const SceneDev& sc, const FrameParams& P, unsigned char* img, size_t pitch,
                        const rtb_aovs& aov, Counters* ctr, bool count, cudaStream_t st);
void mega_release(void* state);

// ------------------------------------------------------------------ stage 0: camera rays
template <bool PATH>
__global__ void __launch_bounds__(256) k_gen(FrameParams P, int sample, int n_local, int tiles_x, float4* __restrict__ q_org,
                                             float4* __restrict__ q_dir, int* __restrict__ ctr4)
{
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx == 0) { ctr4[0] = n_local; }
    if (idx >= n_local) return;
    int x, y;
    if (!local_to_pixel(P, idx, tiles_x, x, y)) { q_org[idx] = make_float4(0.f, 0.f, 0.f, __int_as_float(-1)); return; }
    float px = (float)x, py = (float)y;
    if (PATH) {
        unsigned c4[4] = { (unsigned)(y * P.width + x), (unsigned)sample, 0u, 0u };
        philox4x32(c4, P.seed, 0x5eed5eedu);
        px = xa(px, xs(u01(c4[0]), 0.5f)); py = xa(py, xs(u01(c4[1]), 0.5f));
    }
    const f3 d = generate_direction(P, px, py);
    q_org[idx] = make_float4(P.ox, P.oy, P.oz, __int_as_float(y * P.width + x));
    q_dir[idx] = make_float4(d.x, d.y, d.z, 0.f);
}

// ------------------------------------------------------------------ stage 1 / 3: queue tracer with lane refill
// Hit sink of the queue tracer: an accepted candidate goes straight to this ray's hit record in HBM (accepts are
// rare, ~1-2 per ray), so only what the accept rule compares stays in registers during the traversal.
template <bool ANY>
struct QueueSink {
    float hmin; int hinst, hprim, ray;
    float4* __restrict__ h0; float4* __restrict__ h1;
    XF void reset(int r) { hmin = FLT_MAX; hinst = -1; hprim = -1; ray = r; }
    XF float min() const { return hmin; }
    XF int inst() const { return hinst; }
    XF int prim() const { return hprim; }
    XF bool any_hit() const { return false; }
    XF void accept(float distance, int slot, int prim, int inst, float u, float v, f3 loc)
    {
        hmin = distance; hinst = inst; hprim = prim;
        if (!ANY) {
            h0[ray] = make_float4(distance, u, v, __int_as_float(inst));
            h1[ray] = make_float4(loc.x, loc.y, loc.z, __int_as_float(slot));
        }
    }
};

template <bool ANY, bool COUNT, int MINB>
__global__ void __launch_bounds__(128, MINB) k_trace_queue(int refill_idle, SceneDev sc, const float4* __restrict__ q_org, const float4* __restrict__ q_dir,
                                                     float3 cdir, const int* __restrict__ count_ptr, int* __restrict__ head,
                                                     float4* __restrict__ hit0, float4* __restrict__ hit1,
                                                     const float4* __restrict__ sp0, const float4* __restrict__ sp1,
                                                     float4* __restrict__ acc, Counters* ctr)
{
    const int lane = threadIdx.x & 31;
    const unsigned lt = (1u << lane) - 1u;
    const int count = *count_ptr;
    RTB_DECL_STACK(stk);
    TraceState S;
    QueueSink<ANY> sink;
    sink.h0 = hit0; sink.h1 = hit1;
    sink.reset(-1);
    S.clear();
    int pix = -1;
    bool more = true;                           // warp-uniform: the queue may still hold rays
    unsigned rays = 0;
    for (;;) {
        const bool idle = S.cur == EMPTY_REF;
        if (idle && sink.ray >= 0) {            // retire the finished ray
            if (!ANY) {
                if (sink.hmin == FLT_MAX) hit0[sink.ray] = make_float4(FLT_MAX, 0.f, 0.f, __int_as_float(-1));
            } else {
                // commented block raycast.cu:262-290: illumination = 0.4 cos, or cos when the light is visible; clamp [0.4, 1]
                const float4 p0 = sp0[sink.ray], p1 = sp1[sink.ray];
                float il = sink.hmin == FLT_MAX ? p1.w : __double2float_rn(__dmul_rn(0.4, (double)p1.w));
                il = fmaxf(0.4f, fminf(1.0f, il));
                float4 a = acc[pix];
                a.x = xa(a.x, xm(p0.x, xm(p0.w, xm(il, p1.x))));
                a.y = xa(a.y, xm(p0.y, xm(p0.w, xm(il, p1.y))));
                a.z = xa(a.z, xm(p0.z, xm(p0.w, xm(il, p1.z))));
                acc[pix] = a;
            }
            sink.ray = -1;
        }
        const unsigned im = __ballot_sync(0xffffffffu, idle);
        if (more && im) {                       // refill idle lanes: one atomic per warp
            const int n = __popc(im), leader = __ffs(im) - 1;
            int base = 0;
            if (lane == leader) base = atomicAdd(head, n);
            base = __shfl_sync(0xffffffffu, base, leader);
            if (idle) {
                const int i = base + __popc(im & lt);
                if (i < count) {
                    const float4 o = q_org[i];
                    pix = __float_as_int(o.w);
                    if (pix >= 0) {
                        f3 d;
                        if (ANY) d = mk3(cdir.x, cdir.y, cdir.z);
                        else { const float4 dd = q_dir[i]; d = mk3(dd.x, dd.y, dd.z); }
                        sink.reset(i);
                        S.begin(sc, mk3(o.x, o.y, o.z), d, stk);
                        rays++;
                    }
                }
            }
            if (base + n >= count) more = false;
        }
        if (__ballot_sync(0xffffffffu, S.cur != EMPTY_REF) == 0) {
            if (!more) break;
            continue;
        }
        const int keep = more ? 32 - refill_idle : 0;   // leave the loop to refill once refill_idle lanes are idle
        do {
            S.template advance<ANY, COUNT>(sc, stk, sink);
        } while (__popc(__ballot_sync(0xffffffffu, S.cur != EMPTY_REF)) > keep);
    }
    S.template flush_counters<COUNT>(ctr);
    for (int o = 16; o > 0; o >>= 1) rays += __shfl_xor_sync(0xffffffffu, rays, o);
    if (lane == 0 && rays) atomicAdd(&ctr->rays, (unsigned long long)rays);
}

// ------------------------------------------------------------------ stage 2: shading + queue emission
template <bool PATH>
__global__ void __launch_bounds__(128) k_shade(SceneDev sc, FrameParams P, int bounce, int sample, WaveState ws, int* __restrict__ ctr4,
                                               int32_t* __restrict__ aov_prim, int32_t* __restrict__ aov_inst, float* __restrict__ aov_t)
{
    const int lane = threadIdx.x & 31;
    const unsigned lt = (1u << lane) - 1u;
    const int count = ctr4[0];
    int* next_count = ctr4 + 4;                 // ray count of bounce + 1
    int* shadow_count = ctr4 + 2;
    const float4* __restrict__ qo = ws.q_org[bounce & 1];
    const float4* __restrict__ qd = ws.q_dir[bounce & 1];
    float4* __restrict__ qo_out = ws.q_org[(bounce + 1) & 1];
    float4* __restrict__ qd_out = ws.q_dir[(bounce + 1) & 1];
    const int stride = gridDim.x * blockDim.x;
    const int rounds = (count + stride - 1) / stride;
    for (int k = 0; k < rounds; k++) {          // every lane runs every round: the ballots below need the full warp
        const int i = k * stride + blockIdx.x * blockDim.x + threadIdx.x;
        bool cont = false, shadow = false;
        f3 norg = mk3(0.f, 0.f, 0.f), ndir = norg, sorg = norg, thr = mk3(1.f, 1.f, 1.f), c = norg;
        float w = 0.f, cosl = 0.f;
        int pix = -1;
        if (i < count) {
            const float4 o = qo[i];
            pix = __float_as_int(o.w);
        }
        if (pix >= 0) {
            const float4 dd = qd[i], h0 = ws.hit0[i], h1 = ws.hit1[i];
            const f3 dir = mk3(dd.x, dd.y, dd.z);
            Hit h;
            h.min = h0.x; h.u = h0.y; h.v = h0.z; h.inst = __float_as_int(h0.w);
            h.loc = mk3(h1.x, h1.y, h1.z); h.slot = __float_as_int(h1.w); h.prim = -1;
            f3 acc = mk3(0.f, 0.f, 0.f);
            if (bounce > 0) { const float4 t4 = ws.thr[pix]; thr = mk3(t4.x, t4.y, t4.z); }
            if (bounce > 0 || (PATH && sample > 0)) { const float4 a4 = ws.acc[pix]; acc = mk3(a4.x, a4.y, a4.z); }
            if (h.min == FLT_MAX) {
                acc = mk3(xa(acc.x, xm(thr.x, c_sky[0])), xa(acc.y, xm(thr.y, c_sky[1])), xa(acc.z, xm(thr.z, c_sky[2])));
            } else {
                InstXf xf;
                load_instance(sc.inst, h.inst, xf);
                const float4 q0 = __ldg(sc.meshes[xf.mesh].tris + (size_t)h.slot * TRI_F4);
                h.prim = __float_as_int(q0.w);
                c = base_colour(sc, h, xf.material, xf.mesh);
                const f3 nrm = xnormalize(xmul3(xapply_quat(xf.q_inv, mk3(q0.x, q0.y, q0.z)), xf.scale));   // raycast.cu:115-122
                if (!PATH) {
                    const f3 L = mk3(P.light_dir.x, P.light_dir.y, P.light_dir.z);
                    const float m = sc.materials[xf.material].metallic;
                    w = xs(1.0f, m);
                    cosl = xdot(nrm, L);
                    if (cosl > 0.0f) {          // raycast.cu:268: the Lambert term waits for the shadow ray
                        shadow = true;
                        sorg = xadd(h.loc, xscale(L, 1e-4f));                                    // raycast.cu:259
                    } else {
                        const float il = fmaxf(0.4f, fminf(1.0f, __double2float_rn(__dmul_rn(0.4, (double)cosl))));
                        acc = mk3(xa(acc.x, xm(thr.x, xm(w, xm(il, c.x)))), xa(acc.y, xm(thr.y, xm(w, xm(il, c.y)))),
                                  xa(acc.z, xm(thr.z, xm(w, xm(il, c.z)))));
                    }
                    if (m > 0.0f && bounce < P.max_depth) {
                        cont = true;
                        ndir = reflect_dir(dir, nrm);
                        norg = xadd(h.loc, xscale(ndir, 1e-4f));
                    }
                    if (cont) ws.thr[pix] = make_float4(xm(thr.x, xm(m, c.x)), xm(thr.y, xm(m, c.y)), xm(thr.z, xm(m, c.z)), 0.f);
                } else {
                    thr = mk3(xm(thr.x, c.x), xm(thr.y, c.y), xm(thr.z, c.z));
                    if (bounce < P.max_depth) {
                        cont = true;
                        unsigned r4[4] = { (unsigned)pix, (unsigned)sample, (unsigned)(bounce + 1), 0u };
                        philox4x32(r4, P.seed, 0x5eed5eedu);
                        ndir = cosine_dir(nrm, u01(r4[0]), u01(r4[1]));
                        norg = xadd(h.loc, xscale(ndir, 1e-4f));
                        ws.thr[pix] = make_float4(thr.x, thr.y, thr.z, 0.f);
                    }
                }
            }
            ws.acc[pix] = make_float4(acc.x, acc.y, acc.z, 0.f);
            if (!PATH && bounce == 0) {
                if (aov_prim) aov_prim[pix] = h.prim;
                if (aov_inst) aov_inst[pix] = h.inst;
                if (aov_t) aov_t[pix] = h.min;
            }
        }
        // warp-ballot compaction: continuation rays -> next bounce queue, shadow rays -> shadow queue
        const unsigned cm = __ballot_sync(0xffffffffu, cont);
        if (cm) {
            const int leader = __ffs(cm) - 1;
            int base = 0;
            if (lane == leader) base = atomicAdd(next_count, __popc(cm));
            base = __shfl_sync(0xffffffffu, base, leader);
            if (cont) {
                const int pos = base + __popc(cm & lt);
                qo_out[pos] = make_float4(norg.x, norg.y, norg.z, __int_as_float(pix));
                qd_out[pos] = make_float4(ndir.x, ndir.y, ndir.z, 0.f);
            }
        }
        const unsigned sm = __ballot_sync(0xffffffffu, shadow);
        if (sm) {
            const int leader = __ffs(sm) - 1;
            int base = 0;
            if (lane == leader) base = atomicAdd(shadow_count, __popc(sm));
            base = __shfl_sync(0xffffffffu, base, leader);
            if (shadow) {
                const int pos = base + __popc(sm & lt);
                ws.s_org[pos] = make_float4(sorg.x, sorg.y, sorg.z, __int_as_float(pix));
                ws.s_p0[pos] = make_float4(thr.x, thr.y, thr.z, w);
                ws.s_p1[pos] = make_float4(c.x, c.y, c.z, cosl);
            }
        }
    }
}

#define WCK(x) do { cudaError_t _e = (x); if (_e != cudaSuccess) return _e; } while (0)

static void wave_free(WaveState* w)
{
    float4** bufs[] = { &w->q_org[0], &w->q_org[1], &w->q_dir[0], &w->q_dir[1], &w->hit0, &w->hit1, &w->s_org, &w->s_p0, &w->s_p1, &w->thr, &w->acc };
    for (auto b : bufs) { if (*b) cudaFree(*b); *b = nullptr; }
    if (w->counters) cudaFree(w->counters);
    w->counters = nullptr; w->cap_rays = w->cap_pixels = 0; w->n_counters = 0;
}

cudaError_t launch_wavefront(const Tune& tune, LaunchInfo& info, void** state, const SceneDev& sc, const FrameParams& P, unsigned char* img, size_t pitch,
                             const rtb_aovs& aov, Counters* ctr, bool count, cudaStream_t st)
{
    WaveState* w = (WaveState*)*state;
    if (!w) { w = new WaveState(); *state = w; }
    if (tune.fused) return launch_mega(tune, info, &w->mega, sc, P, img, pitch, aov, ctr, count, st);
    const bool path = P.mode == RTB_MODE_PATH;
    const int bands = (P.height + BAND_ROWS - 1) / BAND_ROWS;
    const int my_bands = (bands - P.tile_rank + P.tile_count - 1) / P.tile_count;
    if (my_bands <= 0) return cudaSuccess;
    const int tiles_x = (P.width + 7) / 8;
    const int n_local = tiles_x * 8 * my_bands * BAND_ROWS;           // padded pixel slots of this rank
    const size_t npix = (size_t)P.width * P.height;
    const int passes = path ? P.spp : 1;
    const int per_pass = P.max_depth + 3;
    const int n_counters = 4 * passes * per_pass;
    if (w->cap_pixels < npix || w->cap_rays < (size_t)n_local || w->n_counters < n_counters) {
        wave_free(w);
        float4** rays[] = { &w->q_org[0], &w->q_org[1], &w->q_dir[0], &w->q_dir[1], &w->hit0, &w->hit1, &w->s_org, &w->s_p0, &w->s_p1 };
        for (auto b : rays) WCK(cudaMalloc((void**)b, sizeof(float4) * (size_t)n_local));
        WCK(cudaMalloc((void**)&w->thr, sizeof(float4) * npix));
        WCK(cudaMalloc((void**)&w->acc, sizeof(float4) * npix));
        WCK(cudaMalloc((void**)&w->counters, sizeof(int) * (size_t)n_counters));
        w->cap_pixels = npix; w->cap_rays = (size_t)n_local; w->n_counters = n_counters;
    }
    WCK(cudaMemsetAsync(w->counters, 0, sizeof(int) * (size_t)n_counters, st));
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int mb = tune.trace_minb <= 4 ? 4 : (tune.trace_minb == 5 ? 5 : (tune.trace_minb == 6 ? 6 : 8));
    const int ri = tune.refill_idle;
    const int grid_t = sms * (count ? 4 : mb), grid_s = grid_t;
    const int grid_shade = std::min(sms * 8, (n_local + 127) / 128);
    const float3 L = P.light_dir;
    info.launches = 0; info.variant = 0;
    if (path && (aov.prim || aov.inst || aov.t)) {
        FrameParams Q = P; Q.mode = RTB_MODE_REFERENCE;                 // first-hit AOVs from the un-jittered centre ray
        WCK(launch_render(tune, sc, Q, img, pitch, aov, ctr, count, st));
        info.launches++;
    }
    for (int s = 0; s < passes; s++) {
        int* c0 = w->counters + 4 * s * per_pass;
        if (path) k_gen<true><<<(n_local + 255) / 256, 256, 0, st>>>(P, s, n_local, tiles_x, w->q_org[0], w->q_dir[0], c0);
        else k_gen<false><<<(n_local + 255) / 256, 256, 0, st>>>(P, s, n_local, tiles_x, w->q_org[0], w->q_dir[0], c0);
        info.launches++;
        for (int b = 0; b <= P.max_depth; b++) {
            int* c4 = c0 + 4 * b;
#define RTB_TQ(ANY, COUNT, MINB, GRID, ...) k_trace_queue<ANY, COUNT, MINB><<<GRID, 128, 0, st>>>(ri, sc, __VA_ARGS__)
#define RTB_TQ_MB(ANY, GRID, ...) do { if (count) RTB_TQ(ANY, true, 4, GRID, __VA_ARGS__); else if (mb == 4) RTB_TQ(ANY, false, 4, GRID, __VA_ARGS__); \
            else if (mb == 5) RTB_TQ(ANY, false, 5, GRID, __VA_ARGS__); else if (mb == 6) RTB_TQ(ANY, false, 6, GRID, __VA_ARGS__); \
            else RTB_TQ(ANY, false, 8, GRID, __VA_ARGS__); } while (0)
            RTB_TQ_MB(false, grid_t, w->q_org[b & 1], w->q_dir[b & 1], L, c4 + 0, c4 + 1, w->hit0, w->hit1, nullptr, nullptr, nullptr, ctr);
            if (path) k_shade<true><<<grid_shade, 128, 0, st>>>(sc, P, b, s, *w, c4, aov.prim, aov.inst, aov.t);
            else k_shade<false><<<grid_shade, 128, 0, st>>>(sc, P, b, s, *w, c4, aov.prim, aov.inst, aov.t);
            info.launches += 2;
            if (!path) {
                RTB_TQ_MB(true, grid_s, w->s_org, nullptr, L, c4 + 2, c4 + 3, nullptr, nullptr, w->s_p0, w->s_p1, w->acc, ctr);
                info.launches++;
            }
        }
    }
    // radiance -> 8-bit BGR frame (REFLECT: sum of the chain; PATH: mean over samples), 16-byte stores
    info.launches++;
    return launch_emit_frame<false>(P, path ? 1.0f / (float)P.spp : 1.0f, nullptr, w->acc, img, pitch, st);
}

void wave_release(void* state)
{
    if (!state) return;
    WaveState* w = (WaveState*)state;
    wave_free(w);
    mega_release(w->mega);
    delete w;
}

}  // namespace rtb
