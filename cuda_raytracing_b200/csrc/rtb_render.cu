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
template <bool LIT, bool COUNT>
__global__ void __launch_bounds__(128) k_render_direct(SceneDev sc, FrameParams P, unsigned char* __restrict__ img, size_t pitch,
                                                      int32_t* __restrict__ aov_prim, int32_t* __restrict__ aov_inst,
                                                      float* __restrict__ aov_t, Counters* ctr)
{
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int x = blockIdx.x * 16 + (warp & 1) * 8 + (lane & 7);
    const int band = blockIdx.y * P.tile_count + P.tile_rank;
    const int y = band * BAND_ROWS + (warp >> 1) * 4 + (lane >> 3);
    unsigned rays = 0;
    if (x < P.width && y < P.height) {
        const f3 origin = mk3(P.ox, P.oy, P.oz);
        const f3 dir = generate_direction(P, (float)x, (float)y);
        Hit h;
        trace<false, COUNT>(sc, origin, dir, h, ctr);
        rays = 1;
        const size_t pi = (size_t)y * P.width + x;
        if (aov_prim) aov_prim[pi] = h.prim;
        if (aov_inst) aov_inst[pi] = h.inst;
        if (aov_t) aov_t[pi] = h.min;
        if (h.min == FLT_MAX) {
            write_pixel(img, pitch, x, y, 255, 204, 153);                         // raycast.cu:208-216
        } else {
            InstXf xf;
            load_instance(sc.inst, h.inst, xf);
            const f3 c = base_colour(sc, h, xf.material, xf.mesh);
            float illum = 1.0f;                                                    // raycast.cu:282
            if (LIT) illum = lit_illumination<COUNT>(sc, P, h, world_normal(sc, h, xf), ctr, rays);
            illum = fmaxf(0.4f, fminf(1.0f, illum));
            write_pixel(img, pitch, x, y, to_u8(xm(xm(illum, c.x), 255.0f)), to_u8(xm(xm(illum, c.y), 255.0f)),
                        to_u8(xm(xm(illum, c.z), 255.0f)));                       // raycast.cu:292-294
        }
    }
    for (int o = 16; o > 0; o >>= 1) rays += __shfl_xor_sync(0xffffffffu, rays, o);
    if (lane == 0 && rays) atomicAdd(&ctr->rays, (unsigned long long)rays);
}

cudaError_t launch_render(const SceneDev& sc, const FrameParams& P, unsigned char* img, size_t pitch, const rtb_aovs& aov,
                          Counters* ctr, bool count, cudaStream_t st)
{
    const int bands = (P.height + BAND_ROWS - 1) / BAND_ROWS;
    const int my_bands = (bands - P.tile_rank + P.tile_count - 1) / P.tile_count;
    if (my_bands <= 0 || P.width <= 0) return cudaSuccess;
    dim3 grid((P.width + 15) / 16, my_bands);
    if (P.mode == RTB_MODE_REFERENCE) {
        if (count) k_render_direct<false, true><<<grid, 128, 0, st>>>(sc, P, img, pitch, aov.prim, aov.inst, aov.t, ctr);
        else k_render_direct<false, false><<<grid, 128, 0, st>>>(sc, P, img, pitch, aov.prim, aov.inst, aov.t, ctr);
    } else if (P.mode == RTB_MODE_SHADOW) {
        if (count) k_render_direct<true, true><<<grid, 128, 0, st>>>(sc, P, img, pitch, aov.prim, aov.inst, aov.t, ctr);
        else k_render_direct<true, false><<<grid, 128, 0, st>>>(sc, P, img, pitch, aov.prim, aov.inst, aov.t, ctr);
    } else {
        return cudaErrorNotSupported;
    }
    return cudaGetLastError();
}

}  // namespace rtb
