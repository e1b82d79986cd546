// rtb_render.cuh -- host interface of the frame kernels (rtb_render.cu, rtb_wavefront.cu).
#pragma once
#include "rtb_internal.cuh"

namespace rtb {

// Tuning knobs (rtb_set_option "trace_minb" / "direct_minb" / "refill_idle" ...); one set per context, read at launch time.
struct Tune {
    int trace_minb = 8;     // __launch_bounds__(128, N) variant of the queue tracer: resident blocks per SM the registers allow
    int direct_minb = 8;    // same for the direct kernel
    int fused = -1;         // modes 2/3: 2 = one persistent kernel + in-kernel ray queues (k_megaq), 1 = the same without queues (k_mega),
                            // 0 = staged wavefront pipeline, -1 = choose 1 or 2 per frame (launch_mega: queues pay off from ~3 M pixel slots)
    int smem_top = 0;       // BLAS nodes (top levels, BFS prefix) staged in shared memory by TMA bulk copy; 0 = off (measured slower: profiles/r01e_tune_c3_smem_top.log)
    int queue_keep = 28;    // queue variant (fused=2): a warp whose continuing rays number at least this keeps them in their lanes
    int ring_split = 1;     // queue variant: shadow rays and continuation rays go to separate rings (warps start 32 rays of one kind)
    int tail_idle = 2;      // fused kernel, after the last pixel was claimed: retire once this many live lanes are idle
    int cost_order = 2;     // persistent kernels: claim pixel blocks by descending cost of the previous frame: 1 = per block, 2 = per ROW of blocks
                            // (robust to a camera that pans: rtb_mega.cu k_order), 3 = per block, dilated by two blocks sideways, 0 = off
    int cost_block = 3;     // ... a block being 2^cost_block x 2^cost_block warp tiles of 8 x 4 pixels
    int spec_thr = 4;       // see SceneDev::spec_thr
    int refill_idle = 32;   // queue tracer: refill a warp once this many lanes are idle
};
// what a frame launched: filled by the launchers, reported through rtb_frame_stats
struct LaunchInfo {
    int launches = 0;       // kernels of this library launched for the frame
    int variant = -1;       // modes 2/3: 2 = k_megaq, 1 = k_mega, 0 = staged wavefront; -1 = the direct kernel
};

cudaError_t launch_render(const Tune& tune, const SceneDev& sc, const FrameParams& P, unsigned char* img, size_t pitch, const rtb_aovs& aov,
                          Counters* ctr, bool count, cudaStream_t st);

}  // namespace rtb
