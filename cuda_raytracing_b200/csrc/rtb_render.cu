// rtb_render.cu -- frame kernels: camera rays -> two-level traversal -> shading -> framebuffer.
//
// Replaces the reference's `render` kernel (raycast.cu:146-297).  Pixels are mapped so that one warp
// owns an 8x4 pixel tile (coherent rays share node cache lines); a block owns 16x8 pixels.  Multi-GPU:
// a rank renders only the BAND_ROWS-row bands b with b % tile_count == tile_rank.
#include "rtb_render.cuh"
#include "rtb_shade.cuh"

namespace rtb {

XF void write_pixel(unsigned char* img, size_t pitch, int x, int y, unsigned char b, unsigned char g, unsigned char r)
{
    unsigned char* p = img + (size_t)y * pitch + 3 * (size_t)x;
    p[0] = b; p[1] = g; p[2] = r;
}

// modes RTB_MODE_REFERENCE / RTB_MODE_SHADOW: everything for one pixel in one thread
template <bool LIT, bool COUNT, int MINB>
__global__ void __launch_bounds__(128, MINB) k_render_direct(SceneDev sc, FrameParams P, unsigned char* __restrict__ img, size_t pitch,
                                                      int32_t* __restrict__ aov_prim, int32_t* __restrict__ aov_inst,
                                                      float* __restrict__ aov_t, Counters* ctr)
{
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int x = blockIdx.x * 16 + (warp & 1) * 8 + (lane & 7);
    const int band = blockIdx.y * P.tile_count + P.tile_rank;
    const int y = band * BAND_ROWS + (warp >> 1) * 4 + (lane >> 3);
    unsigned rays = 0;
    const bool valid = x < P.width && y < P.height;                                // the whole warp stays: trace() is warp-collective
    const f3 origin = mk3(P.ox, P.oy, P.oz);
    f3 dir = mk3(0.f, 0.f, 1.f);
    if (valid) dir = generate_direction(P, (float)x, (float)y);
    RTB_DECL_STACK(stk);
    Hit h;
    trace<false, COUNT>(sc, stk, valid, origin, dir, h, ctr);
    const bool is_hit = valid && h.min != FLT_MAX;
    InstXf xf;
    f3 c = mk3(0.f, 0.f, 0.f), nrm = mk3(0.f, 0.f, 1.f);
    if (is_hit) {
        load_instance(sc.inst, h.inst, xf);
        c = base_colour(sc, h, xf.material, xf.mesh);
        if (LIT) nrm = world_normal(sc, h, xf);
    }
    float illum = 1.0f;                                                            // raycast.cu:282
    if (LIT) illum = lit_illumination<COUNT>(sc, stk, P, is_hit, h, nrm, ctr, rays);
    unsigned bgr = 255u | (204u << 8) | (153u << 16);                             // raycast.cu:208-216 sky
    if (is_hit) {
        illum = fmaxf(0.4f, fminf(1.0f, illum));
        bgr = (unsigned)to_u8(xm(xm(illum, c.x), 255.0f)) | ((unsigned)to_u8(xm(xm(illum, c.y), 255.0f)) << 8) |
              ((unsigned)to_u8(xm(xm(illum, c.z), 255.0f)) << 16);                 // raycast.cu:292-294
    }
    if (valid) {
        rays++;
        const size_t pi = (size_t)y * P.width + x;
        if (aov_prim) aov_prim[pi] = h.prim;
        if (aov_inst) aov_inst[pi] = h.inst;
        if (aov_t) aov_t[pi] = h.min;
    }
    // Framebuffer: a warp row is 8 pixels = 24 bytes = six 32-bit words.  Lanes 0-5 of each row gather the two pixels
    // their word overlaps by shuffle and issue one aligned 4-byte store (the reference: 3 single-byte stores per pixel).
    {
        const int w = lane & 7, row0 = lane & ~7;
        const int pa = w + (w >= 3 ? 1 : 0), sh = 8 * (w % 3);
        const unsigned ca = __shfl_sync(0xffffffffu, bgr, row0 + (pa & 7)), cb = __shfl_sync(0xffffffffu, bgr, row0 + ((pa + 1) & 7));
        const int x0 = blockIdx.x * 16 + (warp & 1) * 8;
        const bool vec = x0 + 8 <= P.width && (((size_t)img | pitch) & 3) == 0;   // warp-uniform
        if (vec) {
            if (w < 6 && y < P.height) {
                const unsigned word = (ca >> sh) | (sh == 0 ? (cb << 24) : (cb << (24 - sh)));
                *(unsigned*)(img + (size_t)y * pitch + 3 * (size_t)x0 + 4 * w) = word;
            }
        } else if (valid) {
            write_pixel(img, pitch, x, y, (unsigned char)bgr, (unsigned char)(bgr >> 8), (unsigned char)(bgr >> 16));
        }
    }
    for (int o = 16; o > 0; o >>= 1) rays += __shfl_xor_sync(0xffffffffu, rays, o);
    if (lane == 0 && rays) atomicAdd(&ctr->rays, (unsigned long long)rays);
}

cudaError_t launch_render(const Tune& tune, const SceneDev& sc, const FrameParams& P, unsigned char* img, size_t pitch, const rtb_aovs& aov,
                          Counters* ctr, bool count, cudaStream_t st)
{
    const int bands = (P.height + BAND_ROWS - 1) / BAND_ROWS;
    const int my_bands = (bands - P.tile_rank + P.tile_count - 1) / P.tile_count;
    if (my_bands <= 0 || P.width <= 0) return cudaSuccess;
    dim3 grid((P.width + 15) / 16, my_bands);
    if (P.mode != RTB_MODE_REFERENCE && P.mode != RTB_MODE_SHADOW) return cudaErrorNotSupported;
    const bool lit = P.mode == RTB_MODE_SHADOW;
#define RTB_DIRECT(LIT, COUNT, MINB) k_render_direct<LIT, COUNT, MINB><<<grid, 128, 0, st>>>(sc, P, img, pitch, aov.prim, aov.inst, aov.t, ctr)
    if (count) { if (lit) RTB_DIRECT(true, true, 4); else RTB_DIRECT(false, true, 4); }
    else {
        const int mb = tune.direct_minb;
        if (lit) { if (mb <= 4) RTB_DIRECT(true, false, 4); else if (mb == 5) RTB_DIRECT(true, false, 5); else if (mb == 6) RTB_DIRECT(true, false, 6); else RTB_DIRECT(true, false, 8); }
        else { if (mb <= 4) RTB_DIRECT(false, false, 4); else if (mb == 5) RTB_DIRECT(false, false, 5); else if (mb == 6) RTB_DIRECT(false, false, 6); else RTB_DIRECT(false, false, 8); }
    }
#undef RTB_DIRECT
    return cudaGetLastError();
}

}  // namespace rtb
