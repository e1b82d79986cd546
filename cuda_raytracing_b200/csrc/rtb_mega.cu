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
#include <mutex>

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
    unsigned* pix = nullptr;        // finished pixels, b | g << 8 | r << 16, indexed y * width + x (k_emit_frame writes the frame)
};

struct MegaState {
    MegaBufs b;
    int* head = nullptr;
    unsigned* cost = nullptr;       // per coarse block, accumulated by the frame being rendered
    int* order = nullptr;           // claim order of the blocks for the frame being rendered
    float4* q_org = nullptr; float4* q_dir = nullptr;   // ray ring of the queue variant (k_megaq)
    unsigned* q_ticket = nullptr; unsigned* q_ctl = nullptr;
    size_t q_cap = 0;
    size_t cap = 0, pix_cap = 0;
    int nb_cap = 0;
    // the cost map is only reused when the next frame has the same pixel set
    int hist_w = 0, hist_h = 0, hist_rank = -1, hist_count = 0, hist_mode = -1, hist_bs = -1;
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
            st_stream(h0 + idx, make_float4(distance, u, v, __int_as_float(inst)));
            st_stream(h1 + idx, make_float4(loc.x, loc.y, loc.z, __int_as_float(slot)));
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

// Claim order.  The rank-local pixel set is cut into blocks of 2^bs x 2^bs warp tiles (bs = 3: 64 x 32 pixels, 2048 slots; bs = 0:
// one 8 x 4 tile); a claimed slot index is (position of the block in `order`) * slots per block + (tile in block) * 32 + (pixel in tile).  `order`
// lists the blocks by DESCENDING cost of the previous frame (longest-processing-time-first: the expensive pixels start
// early, the cheap ones fill the end of the frame, so the tail of the one persistent launch is short); the cost of a
// block is the number of traversal rounds its pixels took, accumulated by this kernel for the next frame.
struct MegaGeo {
    int tiles_x, tile_rows, nbx, nb, n_slots;
    int bs;                         // a claim block is 2^bs x 2^bs warp tiles (option "cost_block")
    const int* order;               // nb entries or nullptr (identity)
    unsigned* cost;                 // nb entries, accumulated this frame
};

XF bool slot_to_pixel(const FrameParams& P, const MegaGeo& G, int idx, int& x, int& y, int& blk)
{
    const int sh = 5 + 2 * G.bs;
    const int k = idx >> sh, o = idx & ((1 << sh) - 1);
    blk = G.order ? __ldg(G.order + k) : k;
    const int bx = blk % G.nbx, by = blk / G.nbx;
    const int t = o >> 5, l = o & 31;
    const int tx = (bx << G.bs) + (t & ((1 << G.bs) - 1)), ty = (by << G.bs) + (t >> G.bs);
    x = tx * 8 + (l & 7);
    const int yl = ty * 4 + (l >> 3);
    const int band = (yl / BAND_ROWS) * P.tile_count + P.tile_rank;
    y = band * BAND_ROWS + (yl % BAND_ROWS);
    return tx < G.tiles_x && ty < G.tile_rows && x < P.width && y < P.height;
}

// order[] = blocks sorted by descending log-bucketed cost (counting sort, one CTA), cost[] cleared for the new frame.
// rows != 0: the key of a block is the cost of its whole ROW of blocks (nbx blocks per row).  An orbiting camera shifts the image
// sideways by more than a block per frame, so last frame's cost of a block says little about this frame's; the cost profile over
// the image rows (sky above, geometry below) survives the shift, and what the order has to achieve is only that the LAST pixels
// claimed are cheap ones: the launch ends one longest-remaining-chain after the last claim.
__global__ void __launch_bounds__(1024) k_order(unsigned* __restrict__ cost, int nb, int nbx, int rows, int* __restrict__ order)
{
    __shared__ unsigned hist[256];
    __shared__ unsigned offs[256];
    __shared__ unsigned rowc[1024];
    const int nby = (nb + nbx - 1) / nbx;
    const bool by_row = rows == 1 && nby <= 1024;
    for (int i = threadIdx.x; i < 256; i += blockDim.x) hist[i] = 0;
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) rowc[i] = 0;
    __syncthreads();
    if (by_row) {
        for (int b = threadIdx.x; b < nb; b += blockDim.x) atomicAdd(&rowc[b / nbx], cost[b]);
        __syncthreads();
    }
    auto block_cost = [&](int b) -> unsigned {
        if (by_row) return rowc[b / nbx];
        if (rows == 0 || rows == 2) return cost[b];
        unsigned c = 0;                                     // rows == 3: the largest cost within +-2 blocks of the same row
        const int x = b % nbx, r0 = b - x;
        for (int d = -2; d <= 2; d++) if (x + d >= 0 && x + d < nbx && r0 + x + d < nb) c = max(c, cost[r0 + x + d]);
        return c;
    };
    for (int b = threadIdx.x; b < nb; b += blockDim.x) {
        const unsigned c = block_cost(b);
        const int key = 255 - min(255, (int)(log2f((float)c + 1.0f) * 8.0f));
        atomicAdd(&hist[key], 1u);
    }
    __syncthreads();
    if (threadIdx.x == 0) { unsigned run = 0; for (int i = 0; i < 256; i++) { offs[i] = run; run += hist[i]; } }
    __syncthreads();
    for (int b = threadIdx.x; b < nb; b += blockDim.x) {
        const unsigned c = block_cost(b);
        const int key = 255 - min(255, (int)(log2f((float)c + 1.0f) * 8.0f));
        order[atomicAdd(&offs[key], 1u)] = b;
    }
    __syncthreads();
    for (int b = threadIdx.x; b < nb; b += blockDim.x) cost[b] = 0u;
}

// meta word of the ray a lane is tracing: bit 0 = shadow ray, bits 1-7 = bounce, bits 8-31 = sample
// What retiring a finished ray decided: the next ray of the pixel's chain (`next`) or the end of the pixel (`ended`).
struct Retired {
    bool next, ended;
    f3 o, d, acc;
    unsigned meta;
    int x, y, blk;
};

// Shade the finished ray of pixel slot `idx` (k_shade's arithmetic): closest-hit rays add their unshadowed terms and
// decide about a shadow ray / continuation, shadow rays add the pending Lambert term and release the continuation.
// Scratch another SM may have written is read with ld.cg (L2), never through L1.
template <bool PATH>
XF Retired retire_ray(const SceneDev& sc, const FrameParams& P, const MegaGeo& G, const MegaBufs& B, int idx, unsigned meta,
                      float hmin, f3 ray_dir, f3 L, int32_t* __restrict__ aov_prim, int32_t* __restrict__ aov_inst,
                      float* __restrict__ aov_t)
{
    Retired R;
    int x, y, blk;
    slot_to_pixel(P, G, idx, x, y, blk);
    R.x = x; R.y = y; R.blk = blk;
    const int pix = y * P.width + x;
    const int bounce = (int)((meta >> 1) & 127u), sample = (int)(meta >> 8);
    bool next = false, ended = false;
    f3 no = mk3(0.f, 0.f, 0.f), nd = no, acc = no;
    unsigned nmeta = 0;
    if (meta & 1u) {
        // shadow ray done.  commented block raycast.cu:262-290: illumination = 0.4 cos, or cos when the light is visible
        const float4 p0 = ld_stream(B.sp0 + idx), p1 = ld_stream(B.sp1 + idx), a = ld_stream(B.acc + idx), co = ld_stream(B.cont_o + idx);
        float il = hmin == FLT_MAX ? p1.w : __double2float_rn(__dmul_rn(0.4, (double)p1.w));
        il = fmaxf(0.4f, fminf(1.0f, il));
        acc = mk3(xa(a.x, xm(p0.x, xm(p0.w, xm(il, p1.x)))), xa(a.y, xm(p0.y, xm(p0.w, xm(il, p1.y)))),
                  xa(a.z, xm(p0.z, xm(p0.w, xm(il, p1.z)))));
        if (co.w != 0.0f) {
            const float4 cd = ld_stream(B.cont_d + idx);
            next = true; no = mk3(co.x, co.y, co.z); nd = mk3(cd.x, cd.y, cd.z);
            nmeta = (meta & ~255u) | ((unsigned)(bounce + 1) << 1);
        } else ended = true;
    } else {
        f3 thr = mk3(1.f, 1.f, 1.f);
        if (bounce > 0) { const float4 t4 = ld_stream(B.thr + idx); thr = mk3(t4.x, t4.y, t4.z); }
        if (bounce > 0 || (PATH && sample > 0)) { const float4 a4 = ld_stream(B.acc + idx); acc = mk3(a4.x, a4.y, a4.z); }
        Hit h;
        h.min = hmin; h.prim = -1; h.inst = -1;
        if (h.min == FLT_MAX) {
            acc = mk3(xa(acc.x, xm(thr.x, c_sky[0])), xa(acc.y, xm(thr.y, c_sky[1])), xa(acc.z, xm(thr.z, c_sky[2])));
            ended = true;
        } else {
            const float4 h0 = ld_stream(B.hit0 + idx), h1 = ld_stream(B.hit1 + idx);
            h.u = h0.y; h.v = h0.z; h.inst = __float_as_int(h0.w);
            h.loc = mk3(h1.x, h1.y, h1.z); h.slot = __float_as_int(h1.w);
            InstXf xf;
            load_instance(sc.inst, h.inst, xf);
            const float4 q0 = ld_keep(sc.meshes[xf.mesh].tris + (size_t)h.slot * TRI_F4);
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
                    cdir = reflect_dir(ray_dir, nrm);
                    corg = xadd(h.loc, xscale(cdir, 1e-4f));
                    st_stream(B.thr + idx, make_float4(xm(thr.x, xm(m, c.x)), xm(thr.y, xm(m, c.y)), xm(thr.z, xm(m, c.z)), 0.f));
                }
                if (cosl > 0.0f) {          // raycast.cu:268: the Lambert term waits for the shadow ray
                    st_stream(B.sp0 + idx, make_float4(thr.x, thr.y, thr.z, w));
                    st_stream(B.sp1 + idx, make_float4(c.x, c.y, c.z, cosl));
                    st_stream(B.cont_o + idx, make_float4(corg.x, corg.y, corg.z, cont ? 1.0f : 0.0f));
                    if (cont) st_stream(B.cont_d + idx, make_float4(cdir.x, cdir.y, cdir.z, 0.f));
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
                    st_stream(B.thr + idx, make_float4(thr.x, thr.y, thr.z, 0.f));
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
    R.next = next; R.ended = ended; R.o = no; R.d = nd; R.acc = acc; R.meta = nmeta;
    return R;
}

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
            const Retired R = retire_ray<PATH>(sc, P, G, B, idx, meta, sink.hmin, S.dir, L, aov_prim, aov_inst, aov_t);
            const bool next = R.next, ended = R.ended;
            const f3 no = R.o, nd = R.d, acc = R.acc;
            const unsigned nmeta = R.meta;
            const int x = R.x, y = R.y, blk = R.blk;
            if (next) {
                st_stream(B.acc + idx, make_float4(acc.x, acc.y, acc.z, 0.f));
                meta = nmeta;
                sink.reset(idx, (nmeta & 1u) != 0u);
                S.begin(sc, no, nd, stk);
                rays++;
            }
            if (ended) {                                        // radiance -> 8-bit BGR pixel (PATH: mean over samples)
                const float scale = PATH ? 1.0f / (float)P.spp : 1.0f;
                B.pix[(size_t)y * P.width + x] = (unsigned)to_u8(xm(clamp01(xm(acc.x, scale)), 255.0f)) | ((unsigned)to_u8(xm(clamp01(xm(acc.y, scale)), 255.0f)) << 8) |
                                                 ((unsigned)to_u8(xm(clamp01(xm(acc.z, scale)), 255.0f)) << 16);
                // 1 pixel in 8 reports when a block is many tiles: the map is a heuristic, same-address atomics are not free
                if (G.bs == 0 || (idx & 7) == 0) atomicAdd(G.cost + blk, steps);
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


// ------------------------------------------------------------------ fused kernel with in-kernel ray queues
// Same chain semantics as k_mega, but a ray produced while shading is not traced by the lane that produced it: it is
// appended to a ring in HBM (warp-ballot compaction, one atomicAdd per warp) and warps always start 32 rays at once --
// the next 32 queued rays when a ring holds a warp's worth, else a fresh 8x4 pixel tile.  Deep bounces, which leave a
// tile with 1-4 live lanes in k_mega, run in full warps again, and it is still ONE launch per frame.
//   * QR_RINGS rings; a CTA produces into ring (blockIdx % QR_RINGS) and consumes from it first, then from the next
//     few, so the head/tail words are shared by ~10 CTAs instead of all 1184 (one global ring livelocks on its CAS).
//   * producers reserve slots with atomicAdd(tail), write the 32-byte entry, fence, publish ticket[slot] = position+1;
//     consumers reserve with atomicCAS(head) only below the tail they read, spin on the ticket, fence, read with ld.cg.
//     Positions run on from frame to frame (a frame ends with head == tail in every ring), so a ticket can never be
//     mistaken for one of an earlier frame and the ticket array is cleared once, when it is allocated -- not per frame.
//   * a full ring never blocks: the producing lanes keep their rays and trace them themselves (k_mega's behaviour).
//     "Full" leaves QR_MARGIN entries free: a consumer holds a claim of <= 32 entries between its CAS and its loads, a ring
//     is consumed by the CTAs of 4 consecutive ring numbers (<= 4 * ceil(grid / QR_RINGS) CTAs * 4 warps = 160 warps at
//     the largest grid = 5120 entries), and a producer must not lap an entry that is claimed but not yet read.
//   * once every tile is claimed nothing is enqueued any more (the frame tail is latency-bound: a queue hop only delays
//     the chain); a warp leaves when no tile is unclaimed and every ring is empty.  A ray enqueued concurrently with that
//     check is never stranded: its producer passes through the fetch path right after the enqueue and finds it.
#ifndef RTB_QR_RINGS
#define RTB_QR_RINGS 128
#endif
constexpr int QR_RINGS = RTB_QR_RINGS;
constexpr int QR_CTL = 64;                      // unsigned words per ring control block: [0] head, [32] tail (separate lines)
constexpr unsigned QR_MARGIN = 8192;            // entries a producer leaves free (see above); rings hold >= 2 * QR_MARGIN entries
struct QRing {
    float4* org;                    // [QR_RINGS][cap] (origin, pixel slot)
    float4* dir;                    // [QR_RINGS][cap] (direction, meta)
    unsigned* ticket;               // [QR_RINGS][cap]
    unsigned* ctl;                  // [QR_RINGS][QR_CTL]
    unsigned cap;                   // entries per ring, power of two
};

XF unsigned ld_volatile_u32(const unsigned* p) { return *(const volatile unsigned*)p; }

// leader-only: try to reserve up to 32 entries of ring r; returns the count (0 = nothing / lost the race), base in `h`
XF int ring_try_claim(const QRing& Q, int r, bool any, unsigned& h)
{
    unsigned* c = Q.ctl + (size_t)r * QR_CTL;
    h = ld_volatile_u32(c);
    const int avail = (int)(ld_volatile_u32(c + 32) - h);
    if (avail >= 32 || (any && avail > 0)) {
        const int n = min(32, avail);
        if (atomicCAS(c, h, h + (unsigned)n) == h) return n;
    }
    return 0;
}

template <bool PATH, bool COUNT, int MINB>
__global__ void __launch_bounds__(128, MINB) k_megaq(int keep_thr, int ring_split, SceneDev sc, FrameParams P, MegaGeo G, int* __restrict__ head, QRing Q, MegaBufs B,
                                                     unsigned char* __restrict__ img, size_t pitch, int32_t* __restrict__ aov_prim,
                                                     int32_t* __restrict__ aov_inst, float* __restrict__ aov_t, Counters* ctr)
{
    const int lane = threadIdx.x & 31;
    const unsigned lt = (1u << lane) - 1u;
    const int r0 = blockIdx.x % QR_RINGS;
    const f3 L = mk3(P.light_dir.x, P.light_dir.y, P.light_dir.z);
    TopLevels T;
    T.n_tlas = T.n_blas = 0; T.s_tlas = T.s_blas = T.g_blas = nullptr;
    RTB_DECL_STACK(stk);
    TraceState S;
    MegaSink sink;
    sink.h0 = B.hit0; sink.h1 = B.hit1;
    sink.reset(-1, false);
    S.clear();
    int idx = -1;
    unsigned meta = 0;
    bool more = true;                           // warp-uniform: unclaimed pixel tiles may remain
    unsigned rays = 0;
    for (;;) {
        // ---------------- every lane is idle here: retire the generation that just finished
        bool next = false, ended = false;
        f3 no = mk3(0.f, 0.f, 0.f), nd = no;
        unsigned nmeta = 0;
        if (idx >= 0) {
            const Retired R = retire_ray<PATH>(sc, P, G, B, idx, meta, sink.hmin, S.dir, L, aov_prim, aov_inst, aov_t);
            next = R.next; ended = R.ended; no = R.o; nd = R.d; nmeta = R.meta;
            if ((idx & 7) == 0) atomicAdd(G.cost + R.blk, 1u);                   // cost map: rays per pixel (chain depth), 1 pixel in 8 reports
            if (next) st_stream(B.acc + idx, make_float4(R.acc.x, R.acc.y, R.acc.z, 0.f));
            if (ended) {                                        // radiance -> 8-bit BGR pixel (PATH: mean over samples)
                const float scale = PATH ? 1.0f / (float)P.spp : 1.0f;
                B.pix[(size_t)R.y * P.width + R.x] = (unsigned)to_u8(xm(clamp01(xm(R.acc.x, scale)), 255.0f)) | ((unsigned)to_u8(xm(clamp01(xm(R.acc.y, scale)), 255.0f)) << 8) |
                                                     ((unsigned)to_u8(xm(clamp01(xm(R.acc.z, scale)), 255.0f)) << 16);
            }
        }
        const unsigned nm = __ballot_sync(0xffffffffu, next);
        bool kept = false;                                      // warp-uniform: the rays stayed in their lanes (ring full)
        if (nm) {                                               // append the continuing rays: one atomic per warp and ray kind
            // Rays of a kind go to the same ring (ring_split): shadow rays (short: any-hit, most leave their own instance and
            // meet nothing) to the odd ring of the CTA's pair, closest-hit continuation rays to the even one, so that the 32 rays
            // a warp later starts together are of similar length -- a warp runs until its slowest ray is done.
            const unsigned sm = ring_split ? __ballot_sync(0xffffffffu, next && (nmeta & 1u)) : 0u;   // shadow rays
            const unsigned cm = nm & ~sm;
            const int rs = ring_split ? (r0 | 1) : r0, rc = ring_split ? (r0 & ~1) : r0;
            const int leader = __ffs(nm) - 1;
            unsigned base_c = 0xffffffffu, base_s = 0xffffffffu;
            // a generation that stays wide enough keeps its lanes; so does every generation once the tiles are gone (the
            // frame tail is latency-bound: a queue hop only delays the chain)
            if (lane == leader && __popc(nm) < keep_thr && *(volatile int*)head < G.n_slots) {
                unsigned* cc = Q.ctl + (size_t)rc * QR_CTL;
                unsigned* cs = Q.ctl + (size_t)rs * QR_CTL;
                const unsigned used_c = ld_volatile_u32(cc + 32) - ld_volatile_u32(cc), used_s = ld_volatile_u32(cs + 32) - ld_volatile_u32(cs);
                if (used_c + QR_MARGIN <= Q.cap && used_s + QR_MARGIN <= Q.cap) {
                    base_c = cm ? atomicAdd(cc + 32, (unsigned)__popc(cm)) : 0u;
                    base_s = sm ? atomicAdd(cs + 32, (unsigned)__popc(sm)) : 0u;
                }
            }
            base_c = __shfl_sync(0xffffffffu, base_c, leader); base_s = __shfl_sync(0xffffffffu, base_s, leader);
            kept = base_c == 0xffffffffu;
            if (next) {
                if (!kept) {
                    const bool sh = (sm >> lane) & 1u;
                    const unsigned pos = sh ? base_s + __popc(sm & lt) : base_c + __popc(cm & lt), slot = pos & (Q.cap - 1u);
                    const size_t e = (size_t)(sh ? rs : rc) * Q.cap + slot;
                    st_stream(Q.org + e, make_float4(no.x, no.y, no.z, __int_as_float(idx)));
                    st_stream(Q.dir + e, make_float4(nd.x, nd.y, nd.z, __uint_as_float(nmeta)));
                    __threadfence();
                    *(volatile unsigned*)(Q.ticket + e) = pos + 1u;
                } else {
                    meta = nmeta;
                    sink.reset(idx, (nmeta & 1u) != 0u);
                    S.begin(sc, no, nd, stk);
                    rays++;
                }
            }
        }
        if (!(kept && next)) idx = -1;
        if (!kept) {
            // ---------------- get the next 32 rays: queued rays when a ring holds a warp's worth, else a new tile; once the
            // tiles are gone take whatever any ring still holds, and leave when every ring is empty (a ray enqueued after that
            // is picked up by its own producer, which always passes through here after an enqueue)
            int got = 0, qr = 0, qn = 0, pbase = 0;             // got: 1 = ring qr entries [qbase, qbase + qn), 2 = pixel slots, 3 = done
            unsigned qbase = 0;
            for (int tries = 0; !got; tries++) {
                if (more) {
                    if (lane == 0) {
                        for (int k = 0; k < 4 && !got; k++) {   // own ring, then three neighbours
                            const int r = (r0 + k) % QR_RINGS;
                            qn = ring_try_claim(Q, r, false, qbase);
                            if (qn) { got = 1; qr = r; }
                        }
                        if (!got) { pbase = atomicAdd(head, 32); got = 2; }
                    }
                } else {
                    bool has = false; int mine = 0;             // every lane looks at QR_RINGS / 32 rings
                    for (int k = 0; k < QR_RINGS / 32 && !has; k++) {
                        mine = (r0 + lane + 32 * k) % QR_RINGS;
                        const unsigned* c = Q.ctl + (size_t)mine * QR_CTL;
                        has = (int)(ld_volatile_u32(c + 32) - ld_volatile_u32(c)) > 0;
                    }
                    const unsigned hm = __ballot_sync(0xffffffffu, has);
                    const int any_r = hm ? __shfl_sync(0xffffffffu, mine, __ffs(hm) - 1) : -1;
                    if (lane == 0) {
                        const bool stuck = any_r >= 0 && tries > (1 << 16);
                        if (stuck) ctr->overflow = 1;           // rays left behind: the frame is reported as failed
                        if (any_r < 0 || stuck) got = 3;
                        else { qn = ring_try_claim(Q, any_r, true, qbase); if (qn) { got = 1; qr = any_r; } }
                    }
                }
                got = __shfl_sync(0xffffffffu, got, 0);
            }
            if (got == 3) break;
            if (got == 1) {
                qbase = __shfl_sync(0xffffffffu, qbase, 0); qn = __shfl_sync(0xffffffffu, qn, 0); qr = __shfl_sync(0xffffffffu, qr, 0);
                if (lane < qn) {
                    const unsigned pos = qbase + (unsigned)lane, slot = pos & (Q.cap - 1u);
                    const size_t e = (size_t)qr * Q.cap + slot;
                    unsigned wait = 0;
                    bool published = true;
                    while (ld_volatile_u32(Q.ticket + e) != pos + 1u) { if (++wait > (1u << 26)) { ctr->overflow = 1; published = false; break; } }
                    if (published) {                            // an entry that never appears is dropped (frame reported as failed), not read
                        __threadfence();
                        const float4 o = ld_stream(Q.org + e), d = ld_stream(Q.dir + e);
                        idx = __float_as_int(o.w); meta = __float_as_uint(d.w);
                        sink.reset(idx, (meta & 1u) != 0u);
                        S.begin(sc, mk3(o.x, o.y, o.z), mk3(d.x, d.y, d.z), stk);
                        rays++;
                    }
                }
            } else {
                pbase = __shfl_sync(0xffffffffu, pbase, 0);
                const int i = pbase + lane;
                int x, y, blk;
                const bool valid = i < G.n_slots && slot_to_pixel(P, G, i, x, y, blk);
                if (valid) {
                    idx = i; meta = 0;
                    sink.reset(idx, false);
                    S.begin(sc, mk3(P.ox, P.oy, P.oz), camera_ray(P, x, y, 0, PATH), stk);
                    rays++;
                }
                if (pbase + 32 >= G.n_slots) more = false;
            }
        }
        while (__any_sync(0xffffffffu, S.cur != EMPTY_REF)) S.template advance<false, COUNT, false>(sc, stk, sink, T);
    }
    S.template flush_counters<COUNT>(ctr);
    for (int o = 16; o > 0; o >>= 1) rays += __shfl_xor_sync(0xffffffffu, rays, o);
    if (lane == 0 && rays) atomicAdd(&ctr->rays, (unsigned long long)rays);
}

// Start of a frame: the pixel-slot counter back to 0; ring positions run on (no ticket reset).  A frame that ended normally
// left head == tail in every ring; one that was cut short (overflow flag) may not have: whatever it left is dropped here.
__global__ void k_frame_begin(int* __restrict__ head, unsigned* __restrict__ ctl, int rings)
{
    if (threadIdx.x == 0) *head = 0;
    if (ctl && (int)threadIdx.x < rings) { unsigned* c = ctl + (size_t)threadIdx.x * QR_CTL; c[0] = c[32]; }
}

#define MCK(x) do { cudaError_t _e = (x); if (_e != cudaSuccess) return _e; } while (0)

static void mega_free(MegaState* m)
{
    float4** bufs[] = { &m->b.acc, &m->b.thr, &m->b.hit0, &m->b.hit1, &m->b.sp0, &m->b.sp1, &m->b.cont_o, &m->b.cont_d };
    for (auto p : bufs) { if (*p) cudaFree(*p); *p = nullptr; }
    if (m->b.pix) cudaFree(m->b.pix);
    m->b.pix = nullptr; m->pix_cap = 0;
    if (m->head) cudaFree(m->head);
    if (m->cost) cudaFree(m->cost);
    if (m->order) cudaFree(m->order);
    if (m->q_org) cudaFree(m->q_org);
    if (m->q_dir) cudaFree(m->q_dir);
    if (m->q_ticket) cudaFree(m->q_ticket);
    if (m->q_ctl) cudaFree(m->q_ctl);
    m->q_org = m->q_dir = nullptr; m->q_ticket = m->q_ctl = nullptr; m->q_cap = 0;
    m->head = nullptr; m->cost = nullptr; m->order = nullptr; m->cap = 0; m->nb_cap = 0; m->hist_rank = -1;
}

void mega_release(void* state)
{
    if (!state) return;
    MegaState* m = (MegaState*)state;
    mega_free(m);
    delete m;
}

cudaError_t launch_mega(const Tune& tune, LaunchInfo& info, void** state, const SceneDev& sc, const FrameParams& P, unsigned char* img, size_t pitch,
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
    G.bs = std::max(0, std::min(tune.cost_block, 3));
    const int bt = 1 << G.bs;
    G.nbx = (G.tiles_x + bt - 1) / bt;
    G.nb = G.nbx * ((G.tile_rows + bt - 1) / bt);
    G.n_slots = G.nb * (32 << (2 * G.bs));
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
    if (m->pix_cap < (size_t)P.width * P.height) {
        if (m->b.pix) cudaFree(m->b.pix);
        m->b.pix = nullptr; m->pix_cap = 0;
        MCK(cudaMalloc((void**)&m->b.pix, sizeof(unsigned) * (size_t)P.width * P.height));
        m->pix_cap = (size_t)P.width * P.height;
    }
    // The queue variant wins when the frame is long enough to amortise its tail (measured on C3: 4.1 M slots 4.92 vs 5.12 ms,
    // 2.1 M slots 3.10 vs 2.86 ms) and always for path tracing, whose chains run for spp samples (C4 at 1/8: 9.3 vs 10.3 ms).
    const int variant = tune.fused >= 0 ? tune.fused : (path || G.n_slots >= 3000000 ? 2 : 1);
    const bool have_hist = tune.cost_order && m->hist_w == P.width && m->hist_h == P.height && m->hist_rank == P.tile_rank &&
                           m->hist_count == P.tile_count && m->hist_mode == P.mode && m->hist_bs == G.bs;
    info.launches = 0; info.variant = variant;
    if (have_hist) { k_order<<<1, 1024, 0, st>>>(m->cost, G.nb, G.nbx, tune.cost_order - 1, m->order); info.launches = 1; }
    else MCK(cudaMemsetAsync(m->cost, 0, sizeof(unsigned) * (size_t)G.nb, st));
    G.order = have_hist ? m->order : nullptr;
    G.cost = m->cost;
    m->hist_w = P.width;
    m->hist_h = P.height; m->hist_rank = P.tile_rank; m->hist_count = P.tile_count; m->hist_mode = P.mode; m->hist_bs = G.bs;
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (path && (aov.prim || aov.inst || aov.t)) {
        FrameParams Q = P; Q.mode = RTB_MODE_REFERENCE;                 // first-hit AOVs from the un-jittered centre ray
        MCK(launch_render(tune, sc, Q, img, pitch, aov, ctr, count, st));
        info.launches++;
    }
    const int mb = tune.trace_minb <= 4 ? 4 : (tune.trace_minb == 5 ? 5 : (tune.trace_minb == 6 ? 6 : 8));
    const int ri = tune.refill_idle;
    const int grid = std::min(sms * (count ? 4 : mb), (G.n_slots + 127) / 128);
    const bool smem = sc.top_tlas_n + sc.top_blas_n > 0;
    const size_t shb = smem ? (size_t)(sc.top_tlas_n + sc.top_blas_n) * NODE_F4 * 16 + 16 : 0;
    if (variant == 2) {                                                 // in-kernel ray queue variant
        size_t qcap = 2 * QR_MARGIN;                                    // entries per ring; a full ring falls back to in-lane tracing
        while (qcap * QR_RINGS < 2 * (size_t)G.n_slots) qcap <<= 1;
        if (m->q_cap < qcap) {
            if (m->q_org) cudaFree(m->q_org);
            if (m->q_dir) cudaFree(m->q_dir);
            if (m->q_ticket) cudaFree(m->q_ticket);
            if (m->q_ctl) cudaFree(m->q_ctl);
            m->q_org = m->q_dir = nullptr; m->q_ticket = m->q_ctl = nullptr; m->q_cap = 0;
            MCK(cudaMalloc((void**)&m->q_org, sizeof(float4) * qcap * QR_RINGS));
            MCK(cudaMalloc((void**)&m->q_dir, sizeof(float4) * qcap * QR_RINGS));
            MCK(cudaMalloc((void**)&m->q_ticket, sizeof(unsigned) * qcap * QR_RINGS));
            MCK(cudaMalloc((void**)&m->q_ctl, sizeof(unsigned) * (QR_RINGS * QR_CTL + 32)));
            m->q_cap = qcap;
            MCK(cudaMemsetAsync(m->q_ticket, 0, sizeof(unsigned) * m->q_cap * QR_RINGS, st));      // once per allocation
            MCK(cudaMemsetAsync(m->q_ctl, 0, sizeof(unsigned) * (QR_RINGS * QR_CTL + 32), st));
        }
        k_frame_begin<<<1, 128, 0, st>>>(m->head, m->q_ctl, QR_RINGS);
        info.launches++;
        QRing Q;
        Q.org = m->q_org; Q.dir = m->q_dir; Q.ticket = m->q_ticket; Q.ctl = m->q_ctl;
        Q.cap = (unsigned)m->q_cap;
        MegaGeo GQ = G;                                                 // same claim order and cost map as k_mega; cost = rays retired per block
#define RTB_MEGAQ(PATH, COUNT, MINB) k_megaq<PATH, COUNT, MINB><<<grid, 128, 0, st>>>(tune.queue_keep, (PATH) ? 0 : tune.ring_split, sc, P, GQ, m->head, Q, m->b, img, pitch, aov.prim, aov.inst, aov.t, ctr)
#define RTB_MEGAQ_MB(PATH) do { if (count) RTB_MEGAQ(PATH, true, 4); else if (mb <= 5) RTB_MEGAQ(PATH, false, 5); \
                                else if (mb == 6) RTB_MEGAQ(PATH, false, 6); else RTB_MEGAQ(PATH, false, 8); } while (0)
        if (path) RTB_MEGAQ_MB(true); else RTB_MEGAQ_MB(false);
        info.launches += 2;
        MCK(cudaGetLastError());
        return launch_emit_frame<true>(P, 1.0f, m->b.pix, nullptr, img, pitch, st);
    }
    k_frame_begin<<<1, 128, 0, st>>>(m->head, nullptr, 0);
    info.launches++;
#define RTB_MEGA(PATH, COUNT, MINB) do { if (smem) k_mega<PATH, COUNT, MINB, true><<<grid, 128, shb, st>>>(ri, tune.tail_idle, sc, P, G, m->head, m->b, img, pitch, \
                                                                                     aov.prim, aov.inst, aov.t, ctr); \
                                         else k_mega<PATH, COUNT, MINB, false><<<grid, 128, 0, st>>>(ri, tune.tail_idle, sc, P, G, m->head, m->b, img, pitch, \
                                                                                     aov.prim, aov.inst, aov.t, ctr); } while (0)
#define RTB_MEGA_MB(PATH) do { if (count) RTB_MEGA(PATH, true, 4); else if (mb == 4) RTB_MEGA(PATH, false, 4); else if (mb == 5) RTB_MEGA(PATH, false, 5); \
                               else if (mb == 6) RTB_MEGA(PATH, false, 6); else RTB_MEGA(PATH, false, 8); } while (0)
    if (path) RTB_MEGA_MB(true); else RTB_MEGA_MB(false);
    info.launches += 2;
    MCK(cudaGetLastError());
    return launch_emit_frame<true>(P, 1.0f, m->b.pix, nullptr, img, pitch, st);
}

// ------------------------------------------------------------------ the same kernel with its scratch in a static device array
// (rtb_render_noalloc, include/rtb200.h): for callers that must not allocate -- CUDA graph capture, serving loops, harnesses that
// forbid cudaMalloc.  Everything k_mega keeps per pixel slot (eight float4 arrays), the packed pixels k_emit_frame reads, the
// per-block cost words and the slot counter are carved out of g_static_scratch below, which every GPU that loads the library holds
// until the process exits.  All calls share it: calls that overlap on one GPU race (see the header).
constexpr int STATIC_MAX_SLOTS = RTB_NOALLOC_MAX_SLOTS;           // pixel slots (the frame's own bands padded to claim blocks)
constexpr int STATIC_MAX_PIXELS = RTB_NOALLOC_MAX_SLOTS;          // width * height of the whole frame (pix is indexed y * width + x)
constexpr size_t STATIC_OFF_PIX = sizeof(float4) * 8 * (size_t)STATIC_MAX_SLOTS;
constexpr size_t STATIC_OFF_COST = STATIC_OFF_PIX + sizeof(unsigned) * (size_t)STATIC_MAX_PIXELS;
constexpr size_t STATIC_OFF_HEAD = STATIC_OFF_COST + sizeof(unsigned) * (size_t)(STATIC_MAX_SLOTS / 32);    // a block has >= 32 slots
constexpr size_t STATIC_BYTES = STATIC_OFF_HEAD + 256;
static_assert(STATIC_OFF_PIX % 256 == 0 && STATIC_OFF_COST % 256 == 0 && STATIC_OFF_HEAD % 256 == 0, "sub-arrays stay 256-byte aligned");
static __device__ __align__(256) unsigned char g_static_scratch[STATIC_BYTES];

// In host code `g_static_scratch` is only a stand-in for the symbol, not a device address: the address a kernel can use comes from
// cudaGetSymbolAddress, once per device, with that device current.
constexpr int STATIC_MAX_DEVICES = 64;
static unsigned char* g_static_base[STATIC_MAX_DEVICES] = {};
static cudaError_t g_static_status[STATIC_MAX_DEVICES] = {};
static bool g_static_done[STATIC_MAX_DEVICES] = {};
static std::mutex g_static_mutex;

size_t mega_static_bytes() { return STATIC_BYTES; }

cudaError_t mega_static_setup(int device)
{
    if (device < 0 || device >= STATIC_MAX_DEVICES) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(g_static_mutex);
    if (g_static_done[device]) return g_static_status[device];
    int cur = -1;
    cudaError_t e = cudaGetDevice(&cur);
    if (e == cudaSuccess && cur != device) e = cudaErrorInvalidDevice;      // the caller makes the device current first
    void* ptr = nullptr;
    if (e == cudaSuccess) e = cudaGetSymbolAddress(&ptr, g_static_scratch);
    if (e == cudaSuccess && (!ptr || ((uintptr_t)ptr & 255u))) e = cudaErrorInvalidSymbol;
    g_static_base[device] = e == cudaSuccess ? (unsigned char*)ptr : nullptr;
    g_static_status[device] = e;
    g_static_done[device] = true;
    return e;
}

// Slot / pixel need of a frame for the static variant; false when it does not fit (slots and pixels are reported either way).
bool mega_static_fits(const Tune& tune, const FrameParams& P, long long* slots, long long* pixels)
{
    const int bands = (P.height + BAND_ROWS - 1) / BAND_ROWS;
    const int my_bands = std::max(0, (bands - P.tile_rank + P.tile_count - 1) / P.tile_count);
    const int bs = std::max(0, std::min(tune.cost_block, 3)), bt = 1 << bs;
    const long long tiles_x = (P.width + 7) / 8, tile_rows = (long long)my_bands * (BAND_ROWS / 4);
    const long long nb = ((tiles_x + bt - 1) / bt) * ((tile_rows + bt - 1) / bt);
    *slots = nb * (32ll << (2 * bs));
    *pixels = (long long)P.width * P.height;
    return *slots <= STATIC_MAX_SLOTS && *pixels <= STATIC_MAX_PIXELS;
}

// launch_mega's k_mega branch (lane-owned chains) with MegaState pointing into the static array: no allocation, no cost history.
cudaError_t launch_mega_static(const Tune& tune, LaunchInfo& info, int device, const SceneDev& sc, const FrameParams& P, unsigned char* img, size_t pitch,
                               const rtb_aovs& aov, Counters* ctr, bool count, cudaStream_t st)
{
    if (device < 0 || device >= STATIC_MAX_DEVICES || !g_static_done[device]) return cudaErrorNotReady;
    if (g_static_status[device] != cudaSuccess) return g_static_status[device];
    long long slots = 0, pixels = 0;
    if (!mega_static_fits(tune, P, &slots, &pixels)) return cudaErrorInvalidValue;
    const bool path = P.mode == RTB_MODE_PATH;
    const int bands = (P.height + BAND_ROWS - 1) / BAND_ROWS;
    const int my_bands = (bands - P.tile_rank + P.tile_count - 1) / P.tile_count;
    info.launches = 0; info.variant = 1;
    if (my_bands <= 0) return cudaSuccess;
    MegaGeo G;
    G.tiles_x = (P.width + 7) / 8;
    G.tile_rows = my_bands * (BAND_ROWS / 4);
    G.bs = std::max(0, std::min(tune.cost_block, 3));
    const int bt = 1 << G.bs;
    G.nbx = (G.tiles_x + bt - 1) / bt;
    G.nb = G.nbx * ((G.tile_rows + bt - 1) / bt);
    G.n_slots = G.nb * (32 << (2 * G.bs));
    if (P.max_depth > 126 || P.spp >= (1 << 24)) return cudaErrorInvalidValue;
    unsigned char* base = g_static_base[device];
    MegaState ms; MegaState* m = &ms;
    float4** bufs[] = { &m->b.acc, &m->b.thr, &m->b.hit0, &m->b.hit1, &m->b.sp0, &m->b.sp1, &m->b.cont_o, &m->b.cont_d };
    for (int i = 0; i < 8; i++) *bufs[i] = (float4*)(base + sizeof(float4) * (size_t)STATIC_MAX_SLOTS * i);
    m->b.pix = (unsigned*)(base + STATIC_OFF_PIX);
    m->cost = (unsigned*)(base + STATIC_OFF_COST);
    m->head = (int*)(base + STATIC_OFF_HEAD);
    // what must start at zero: the per-block cost words (here) and the slot counter (k_frame_begin); every other word of the
    // scratch is written by the frame before it is read, so the last call's contents do not matter
    MCK(cudaMemsetAsync(m->cost, 0, sizeof(unsigned) * (size_t)G.nb, st));
    G.order = nullptr;
    G.cost = m->cost;
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    if (path && (aov.prim || aov.inst || aov.t)) {
        FrameParams Q = P; Q.mode = RTB_MODE_REFERENCE;                 // first-hit AOVs from the un-jittered centre ray, as in launch_mega
        MCK(launch_render(tune, sc, Q, img, pitch, aov, ctr, count, st));
        info.launches++;
    }
    const int mb = tune.trace_minb <= 4 ? 4 : (tune.trace_minb == 5 ? 5 : (tune.trace_minb == 6 ? 6 : 8));
    const int ri = tune.refill_idle;
    const int grid = std::min(sms * (count ? 4 : mb), (G.n_slots + 127) / 128);
    const bool smem = false;                                            // no staged top levels: SceneDev comes without them here
    const size_t shb = 0;
    k_frame_begin<<<1, 128, 0, st>>>(m->head, nullptr, 0);
    info.launches++;
    if (path) RTB_MEGA_MB(true); else RTB_MEGA_MB(false);
    info.launches += 2;
    MCK(cudaGetLastError());
    return launch_emit_frame<true>(P, 1.0f, m->b.pix, nullptr, img, pitch, st);
}

}  // namespace rtb
