// rtb_mega.cu -- fused persistent path kernel for RTB_MODE_REFLECT / RTB_MODE_PATH (the default; the staged
// wavefront pipeline of rtb_wavefront.cu stays selectable with rtb_set_option("fused", 0)).
//
// Why: in the staged pipeline every bounce is 3 launches (closest-hit queue, shade, shadow queue) and every
// queue-tracer launch ends with a tail in which a few long rays keep the whole GPU waiting (ncu launch list
// profiles/r01d: the five bounce >= 2 launches of C3 take 1.6 ms for < 10 % of the rays).  Here ONE persistent
// kernel renders the frame: a lane owns a pixel's whole chain -- camera ray, closest hit, shading, shadow ray,
// next bounce, ... -- and when the chain ends it writes the pixel and claims the next pixel from a global
// counter (one atomicAdd per warp, ballot-ranked).  A ray that a lane produces while shading is by construction
// produced by an idle lane, so it is traced by that same lane: no queue hop, no HBM ray record, no second launch.
// Shading runs batched, when at least `refill_idle` lanes of the warp have finished their rays, so it executes
// nearly full-width.  The only tail left is the one at the end of the frame.
//
// Semantics (oracle/rt_oracle.c render_pixel) and arithmetic are those of k_shade / k_trace_queue<ANY>; the
// per-pixel order of radiance additions is the chain order, so frames are bit-identical to the staged pipeline.
// Per-pixel scratch in HBM (indexed by the rank-local pixel slot): radiance, throughput, the hit record the
// traversal writes on every accepted candidate, and the shadow ray's pending Lambert term + continuation ray.
#include "rtb_wave_common.cuh"
#include <algorithm>

namespace rtb {

struct MegaBufs {
    float4* acc = nullptr;          // radiance so far (REFLECT: chain sum; PATH: sum over samples)
    float4* thr = nullptr;          // path throughput
    float4* hit0 = nullptr;         // (min, u, v, inst)
    float4* hit1 = nullptr;         // (location, slot)
    float4* sp0 = nullptr;          // pending Lambert term: (throughput, w = 1 - metallic)
    float4* sp1 = nullptr;          //                       (base colour, cos)
    float4* cont_o = nullptr;       // continuation ray origin, .w = 1 when the chain continues after the shadow ray
    float4* cont_d = nullptr;       // continuation ray direction
};

struct MegaState {
    MegaBufs b;
    int* head = nullptr;
    unsigned* cost = nullptr;       // per coarse block, accumulated by the frame being rendered
    int* order = nullptr;           // claim order of the blocks for the frame being rendered
    size_t cap = 0;
    int nb_cap = 0;
    // the cost map is only reused when the next frame has the same pixel set
    int hist_w = 0, hist_h = 0, hist_rank = -1, hist_count = 0, hist_mode = -1;
};

// Accepted candidates go straight to the pixel's hit record; shadow rays (any-hit) record nothing.
struct MegaSink {
    float hmin; int hinst, hprim, idx; bool shadow;
    float4* __restrict__ h0; float4* __restrict__ h1;
    XF void reset(int i, bool sh) { hmin = FLT_MAX; hinst = -1; hprim = -1; idx = i; shadow = sh; }
    XF float min() const { return hmin; }
    XF int inst() const { return hinst; }
    XF int prim() const { return hprim; }
    XF bool any_hit() const { return shadow; }
    XF void accept(float distance, int slot, int prim, int inst, float u, float v, f3 loc)
    {
        hmin = distance; hinst = inst; hprim = prim;
        if (!shadow) {
            h0[idx] = make_float4(distance, u, v, __int_as_float(inst));
            h1[idx] = make_float4(loc.x, loc.y, loc.z, __int_as_float(slot));
        }
    }
};

XF f3 camera_ray(const FrameParams& P, int x, int y, int sample, bool path)
{
    float px = (float)x, py = (float)y;
    if (path) {
        unsigned c4[4] = { (unsigned)(y * P.width + x), (unsigned)sample, 0u, 0u };
        philox4x32(c4, P.seed, 0x5eed5eedu);
        px = xa(px, xs(u01(c4[0]), 0.5f)); py = xa(py, xs(u01(c4[1]), 0.5f));
    }
    return generate_direction(P, px, py);
}

// Claim order.  The rank-local pixel set is cut into coarse blocks of 8 x 8 warp tiles (64 x 32 pixels, 2048 slots);
// a claimed slot index is (position of the block in `order`) * 2048 + (tile in block) * 32 + (pixel in tile).  `order`
// lists the blocks by DESCENDING cost of the previous frame (longest-processing-time-first: the expensive pixels start
// early, the cheap ones fill the end of the frame, so the tail of the one persistent launch is short); the cost of a
// block is the number of traversal rounds its pixels took, accumulated by this kernel for the next frame.
struct MegaGeo {
    int tiles_x, tile_rows, nbx, nb, n_slots;
    const int* order;               // nb entries or nullptr (identity)
    unsigned* cost;                 // nb entries, accumulated this frame
};
constexpr int BLK_SLOTS = 2048;

XF bool slot_to_pixel(const FrameParams& P, const MegaGeo& G, int idx, int& x, int& y, int& blk)
{
    const int k = idx >> 11, o = idx & (BLK_SLOTS - 1);
    blk = G.order ? __ldg(G.order + k) : k;
    const int bx = blk % G.nbx, by = blk / G.nbx;
    const int t = o >> 5, l = o & 31;
    const int tx = bx * 8 + (t & 7), ty = by * 8 + (t >> 3);
    x = tx * 8 + (l & 7);
    const int yl = ty * 4 + (l >> 3);
    const int band = (yl / BAND_ROWS) * P.tile_count + P.tile_rank;
    y = band * BAND_ROWS + (yl % BAND_ROWS);
    return tx < G.tiles_x && ty < G.tile_rows && x < P.width && y < P.height;
}

// order[] = blocks sorted by descending log-bucketed cost (counting sort, one CTA), cost[] cleared for the new frame
__global__ void __launch_bounds__(1024) k_order(unsigned* __restrict__ cost, int nb, int* __restrict__ order)
{
    __shared__ unsigned hist[256];
    __shared__ unsigned offs[256];
    for (int i = threadIdx.x; i < 256; i += blockDim.x) hist[i] = 0;
    __syncthreads();
    for (int b = threadIdx.x; b < nb; b += blockDim.x) {
        const int key = 255 - min(255, (int)(log2f((float)cost[b] + 1.0f) * 8.0f));
        atomicAdd(&hist[key], 1u);
    }
    __syncthreads();
    if (threadIdx.x == 0) { unsigned run = 0; for (int i = 0; i < 256; i++) { offs[i] = run; run += hist[i]; } }
    __syncthreads();
    for (int b = threadIdx.x; b < nb; b += blockDim.x) {
        const int key = 255 - min(255, (int)(log2f((float)cost[b] + 1.0f) * 8.0f));
        order[atomicAdd(&offs[key], 1u)] = b;
    }
    __syncthreads();
    for (int b = threadIdx.x; b < nb; b += blockDim.x) cost[b] = 0u;
}

// meta word of the ray a lane is tracing: bit 0 = shadow ray, bits 1-7 = bounce, bits 8-31 = sample
template <bool PATH, bool COUNT, int MINB, bool SMEM>
__global__ void __launch_bounds__(128, MINB) k_mega(int refill_idle, int tail_idle, SceneDev sc, FrameParams P, MegaGeo G,
                                                    int* __restrict__ head, MegaBufs B, unsigned char* __restrict__ img, size_t pitch,
                                                    int32_t* __restrict__ aov_prim, int32_t* __restrict__ aov_inst,
                                                    float* __restrict__ aov_t, Counters* ctr)
{
    const int lane = threadIdx.x & 31;
    const unsigned lt = (1u << lane) - 1u;
    const f3 L = mk3(P.light_dir.x, P.light_dir.y, P.light_dir.z);
    extern __shared__ __align__(128) float4 s_top[];
    TopLevels T;
    T.n_tlas = T.n_blas = 0; T.s_tlas = T.s_blas = T.g_blas = nullptr;
    if (SMEM) stage_top_levels(sc, s_top, T);
    RTB_DECL_STACK(stk);
    TraceState S;
    MegaSink sink;
    sink.h0 = B.hit0; sink.h1 = B.hit1;
    sink.reset(-1, false);
    S.clear();
    int idx = -1;                               // rank-local pixel slot this lane owns, -1 = none
    unsigned meta = 0;
    bool more = true;                           // warp-uniform: unclaimed pixels may remain
    unsigned rays = 0, steps = 0;                // steps: traversal rounds spent on the current pixel (cost map)
    for (;;) {
        const bool idle = S.cur == EMPTY_REF;
        // ---------------- retire the finished ray of every idle lane: shade, continue the chain or end the pixel
        if (idle && idx >= 0) {
            int x, y, blk;
            slot_to_pixel(P, G, idx, x, y, blk);
            const int pix = y * P.width + x;
            const int bounce = (int)((meta >> 1) & 127u), sample = (int)(meta >> 8);
            bool next = false, ended = false;
            f3 no = mk3(0.f, 0.f, 0.f), nd = no, acc = no;
            unsigned nmeta = 0;
            if (meta & 1u) {
                // shadow ray done.  commented block raycast.cu:262-290: illumination = 0.4 cos, or cos when the light is visible
                const float4 p0 = B.sp0[idx], p1 = B.sp1[idx], a = B.acc[idx], co = B.cont_o[idx];
                float il = sink.hmin == FLT_MAX ? p1.w : __double2float_rn(__dmul_rn(0.4, (double)p1.w));
                il = fmaxf(0.4f, fminf(1.0f, il));
                acc = mk3(xa(a.x, xm(p0.x, xm(p0.w, xm(il, p1.x)))), xa(a.y, xm(p0.y, xm(p0.w, xm(il, p1.y)))),
                          xa(a.z, xm(p0.z, xm(p0.w, xm(il, p1.z)))));
                if (co.w != 0.0f) {
                    const float4 cd = B.cont_d[idx];
                    next = true; no = mk3(co.x, co.y, co.z); nd = mk3(cd.x, cd.y, cd.z);
                    nmeta = (meta & ~255u) | ((unsigned)(bounce + 1) << 1);
                } else ended = true;
            } else {
                f3 thr = mk3(1.f, 1.f, 1.f);
                if (bounce > 0) { const float4 t4 = B.thr[idx]; thr = mk3(t4.x, t4.y, t4.z); }
                if (bounce > 0 || (PATH && sample > 0)) { const float4 a4 = B.acc[idx]; acc = mk3(a4.x, a4.y, a4.z); }
                Hit h;
                h.min = sink.hmin; h.prim = -1; h.inst = -1;
                if (h.min == FLT_MAX) {
                    acc = mk3(xa(acc.x, xm(thr.x, c_sky[0])), xa(acc.y, xm(thr.y, c_sky[1])), xa(acc.z, xm(thr.z, c_sky[2])));
                    ended = true;
                } else {
                    const float4 h0 = B.hit0[idx], h1 = B.hit1[idx];
                    h.u = h0.y; h.v = h0.z; h.inst = __float_as_int(h0.w);
                    h.loc = mk3(h1.x, h1.y, h1.z); h.slot = __float_as_int(h1.w);
                    InstXf xf;
                    load_instance(sc.inst, h.inst, xf);
                    const float4 q0 = __ldg(sc.meshes[xf.mesh].tris + (size_t)h.slot * TRI_F4);
                    h.prim = __float_as_int(q0.w);
                    const f3 c = base_colour(sc, h, xf.material, xf.mesh);
                    const f3 nrm = xnormalize(xmul3(xapply_quat(xf.q_inv, mk3(q0.x, q0.y, q0.z)), xf.scale));   // raycast.cu:115-122
                    if (!PATH) {
                        const float m = sc.materials[xf.material].metallic;
                        const float w = xs(1.0f, m);
                        const float cosl = xdot(nrm, L);
                        const bool cont = m > 0.0f && bounce < P.max_depth;
                        f3 cdir = mk3(0.f, 0.f, 0.f), corg = cdir;
                        if (cont) {
                            cdir = reflect_dir(S.dir, nrm);
                            corg = xadd(h.loc, xscale(cdir, 1e-4f));
                            B.thr[idx] = make_float4(xm(thr.x, xm(m, c.x)), xm(thr.y, xm(m, c.y)), xm(thr.z, xm(m, c.z)), 0.f);
                        }
                        if (cosl > 0.0f) {          // raycast.cu:268: the Lambert term waits for the shadow ray
                            B.sp0[idx] = make_float4(thr.x, thr.y, thr.z, w);
                            B.sp1[idx] = make_float4(c.x, c.y, c.z, cosl);
                            B.cont_o[idx] = make_float4(corg.x, corg.y, corg.z, cont ? 1.0f : 0.0f);
                            if (cont) B.cont_d[idx] = make_float4(cdir.x, cdir.y, cdir.z, 0.f);
                            next = true; no = xadd(h.loc, xscale(L, 1e-4f)); nd = L;              // raycast.cu:259
                            nmeta = meta | 1u;
                        } else {
                            const float il = fmaxf(0.4f, fminf(1.0f, __double2float_rn(__dmul_rn(0.4, (double)cosl))));
                            acc = mk3(xa(acc.x, xm(thr.x, xm(w, xm(il, c.x)))), xa(acc.y, xm(thr.y, xm(w, xm(il, c.y)))),
                                      xa(acc.z, xm(thr.z, xm(w, xm(il, c.z)))));
                            if (cont) { next = true; no = corg; nd = cdir; nmeta = (meta & ~255u) | ((unsigned)(bounce + 1) << 1); }
                            else ended = true;
                        }
                    } else {
                        thr = mk3(xm(thr.x, c.x), xm(thr.y, c.y), xm(thr.z, c.z));
                        if (bounce < P.max_depth) {
                            unsigned r4[4] = { (unsigned)pix, (unsigned)sample, (unsigned)(bounce + 1), 0u };
                            philox4x32(r4, P.seed, 0x5eed5eedu);
                            nd = cosine_dir(nrm, u01(r4[0]), u01(r4[1]));
                            no = xadd(h.loc, xscale(nd, 1e-4f));
                            B.thr[idx] = make_float4(thr.x, thr.y, thr.z, 0.f);
                            next = true; nmeta = (meta & ~255u) | ((unsigned)(bounce + 1) << 1);
                        } else ended = true;
                    }
                }
                if (!PATH && bounce == 0) {
                    if (aov_prim) aov_prim[pix] = h.prim;
                    if (aov_inst) aov_inst[pix] = h.inst;
                    if (aov_t) aov_t[pix] = h.min;
                }
            }
            if (ended && PATH && sample + 1 < P.spp) {          // next sample of the same pixel
                ended = false; next = true;
                no = mk3(P.ox, P.oy, P.oz); nd = camera_ray(P, x, y, sample + 1, true);
                nmeta = (unsigned)(sample + 1) << 8;
            }
            if (next) {
                B.acc[idx] = make_float4(acc.x, acc.y, acc.z, 0.f);
                meta = nmeta;
                sink.reset(idx, (nmeta & 1u) != 0u);
                S.begin(sc, no, nd, stk);
                rays++;
            }
            if (ended) {                                        // radiance -> 8-bit BGR pixel (PATH: mean over samples)
                const float scale = PATH ? 1.0f / (float)P.spp : 1.0f;
                unsigned char* p = img + (size_t)y * pitch + 3 * (size_t)x;
                p[0] = to_u8(xm(clamp01(xm(acc.x, scale)), 255.0f));
                p[1] = to_u8(xm(clamp01(xm(acc.y, scale)), 255.0f));
                p[2] = to_u8(xm(clamp01(xm(acc.z, scale)), 255.0f));
                if ((idx & 7) == 0) atomicAdd(G.cost + blk, steps);       // 1 pixel in 8 reports: the map is a heuristic, the atomics are not free
                idx = -1;
            }
        }
        // ---------------- lanes without a pixel claim the next ones: one atomic per warp
        const unsigned fm = __ballot_sync(0xffffffffu, idx < 0);
        if (more && fm) {
            const int n = __popc(fm), leader = __ffs(fm) - 1;
            int base = 0;
            if (lane == leader) base = atomicAdd(head, n);
            base = __shfl_sync(0xffffffffu, base, leader);
            if (idx < 0) {
                const int i = base + __popc(fm & lt);
                int x, y, blk;
                if (i < G.n_slots && slot_to_pixel(P, G, i, x, y, blk)) {
                    idx = i; meta = 0; steps = 0;
                    sink.reset(idx, false);
                    S.begin(sc, mk3(P.ox, P.oy, P.oz), camera_ray(P, x, y, 0, PATH), stk);
                    rays++;
                }
            }
            if (base + n >= G.n_slots) more = false;
        }
        const unsigned am = __ballot_sync(0xffffffffu, S.cur != EMPTY_REF);
        if (am == 0) {
            if (!more) break;
            continue;
        }
        // trace until `refill_idle` of the lanes that can still get work are idle; lanes without a pixel are dead once
        // the counter is exhausted
        // Tail of the frame (no unclaimed pixels left): retire as soon as `tail_idle` live lanes finish, so that the chains
        // of a warp advance independently instead of in lock-step generations (each as long as its slowest ray).
        const int dead = more ? 0 : __popc(__ballot_sync(0xffffffffu, idx < 0));
        const int keep = more ? 32 - refill_idle : max(0, 32 - dead - tail_idle);
        do {
            steps += S.cur != EMPTY_REF;
            S.template advance<false, COUNT, SMEM>(sc, stk, sink, T);
        } while (__popc(__ballot_sync(0xffffffffu, S.cur != EMPTY_REF)) > keep);
    }
    S.template flush_counters<COUNT>(ctr);
    for (int o = 16; o > 0; o >>= 1) rays += __shfl_xor_sync(0xffffffffu, rays, o);
    if (lane == 0 && rays) atomicAdd(&ctr->rays, (unsigned long long)rays);
}

#define MCK(x) do { cudaError_t _e = (x); if (_e != cudaSuccess) return _e; } while (0)

static void mega_free(MegaState* m)
{
    float4** bufs[] = { &m->b.acc, &m->b.thr, &m->b.hit0, &m->b.hit1, &m->b.sp0, &m->b.sp1, &m->b.cont_o, &m->b.cont_d };
    for (auto p : bufs) { if (*p) cudaFree(*p); *p = nullptr; }
    if (m->head) cudaFree(m->head);
    if (m->cost) cudaFree(m->cost);
    if (m->order) cudaFree(m->order);
    m->head = nullptr; m->cost = nullptr; m->order = nullptr; m->cap = 0; m->nb_cap = 0; m->hist_rank = -1;
}

void mega_release(void* state)
{
    if (!state) return;
    MegaState* m = (MegaState*)state;
    mega_free(m);
    delete m;
}

extern int g_wave_launches;
static int g_wave_launches_pre = 0;

cudaError_t launch_mega(void** state, const SceneDev& sc, const FrameParams& P, unsigned char* img, size_t pitch,
                        const rtb_aovs& aov, Counters* ctr, bool count, cudaStream_t st)
{
    MegaState* m = (MegaState*)*state;
    if (!m) { m = new MegaState(); *state = m; }
    const bool path = P.mode == RTB_MODE_PATH;
    const int bands = (P.height + BAND_ROWS - 1) / BAND_ROWS;
    const int my_bands = (bands - P.tile_rank + P.tile_count - 1) / P.tile_count;
    if (my_bands <= 0) return cudaSuccess;
    MegaGeo G;
    G.tiles_x = (P.width + 7) / 8;
    G.tile_rows = my_bands * (BAND_ROWS / 4);
    G.nbx = (G.tiles_x + 7) / 8;
    G.nb = G.nbx * ((G.tile_rows + 7) / 8);
    G.n_slots = G.nb * BLK_SLOTS;
    if (P.max_depth > 126 || P.spp >= (1 << 24) || G.nb > (1 << 20)) return cudaErrorInvalidValue;
    if (m->cap < (size_t)G.n_slots || m->nb_cap < G.nb) {
        mega_free(m);
        float4** bufs[] = { &m->b.acc, &m->b.thr, &m->b.hit0, &m->b.hit1, &m->b.sp0, &m->b.sp1, &m->b.cont_o, &m->b.cont_d };
        for (auto p : bufs) MCK(cudaMalloc((void**)p, sizeof(float4) * (size_t)G.n_slots));
        MCK(cudaMalloc((void**)&m->head, sizeof(int)));
        MCK(cudaMalloc((void**)&m->cost, sizeof(unsigned) * (size_t)G.nb));
        MCK(cudaMalloc((void**)&m->order, sizeof(int) * (size_t)G.nb));
        m->cap = (size_t)G.n_slots; m->nb_cap = G.nb;
    }
    const bool have_hist = g_tune.cost_order && m->hist_w == P.width && m->hist_h == P.height && m->hist_rank == P.tile_rank &&
                           m->hist_count == P.tile_count && m->hist_mode == P.mode;
    if (have_hist) { k_order<<<1, 1024, 0, st>>>(m->cost, G.nb, m->order); g_wave_launches_pre = 1; }
    else { MCK(cudaMemsetAsync(m->cost, 0, sizeof(unsigned) * (size_t)G.nb, st)); g_wave_launches_pre = 0; }
    G.order = have_hist ? m->order : nullptr;
    G.cost = m->cost;
    m->hist_w = P.width; m->hist_h = P.height; m->hist_rank = P.tile_rank; m->hist_count = P.tile_count; m->hist_mode = P.mode;
    MCK(cudaMemsetAsync(m->head, 0, sizeof(int), st));
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    g_wave_launches = g_wave_launches_pre;
    if (path && (aov.prim || aov.inst || aov.t)) {
        FrameParams Q = P; Q.mode = RTB_MODE_REFERENCE;                 // first-hit AOVs from the un-jittered centre ray
        MCK(launch_render(sc, Q, img, pitch, aov, ctr, count, st));
        g_wave_launches++;
    }
    const int mb = g_tune.trace_minb <= 4 ? 4 : (g_tune.trace_minb == 5 ? 5 : (g_tune.trace_minb == 6 ? 6 : 8));
    const int ri = g_tune.refill_idle;
    const int grid = std::min(sms * (count ? 4 : mb), (G.n_slots + 127) / 128);
    const bool smem = sc.top_tlas_n + sc.top_blas_n > 0;
    const size_t shb = smem ? (size_t)(sc.top_tlas_n + sc.top_blas_n) * 128 + 16 : 0;
#define RTB_MEGA(PATH, COUNT, MINB) do { if (smem) k_mega<PATH, COUNT, MINB, true><<<grid, 128, shb, st>>>(ri, g_tune.tail_idle, sc, P, G, m->head, m->b, img, pitch, \
                                                                                     aov.prim, aov.inst, aov.t, ctr); \
                                         else k_mega<PATH, COUNT, MINB, false><<<grid, 128, 0, st>>>(ri, g_tune.tail_idle, sc, P, G, m->head, m->b, img, pitch, \
                                                                                     aov.prim, aov.inst, aov.t, ctr); } while (0)
#define RTB_MEGA_MB(PATH) do { if (count) RTB_MEGA(PATH, true, 4); else if (mb == 4) RTB_MEGA(PATH, false, 4); else if (mb == 5) RTB_MEGA(PATH, false, 5); \
                               else if (mb == 6) RTB_MEGA(PATH, false, 6); else RTB_MEGA(PATH, false, 8); } while (0)
    if (path) RTB_MEGA_MB(true); else RTB_MEGA_MB(false);
    g_wave_launches++;
    return cudaGetLastError();
}

}  // namespace rtb
