// rtb_render.cuh -- host interface of the frame kernels (rtb_render.cu, rtb_wavefront.cu).
#pragma once
#include "rtb_internal.cuh"

namespace rtb {

// Tuning knobs (rtb_set_option "trace_minb" / "direct_minb" / "refill_idle"); process-wide, read at launch time.
struct Tune {
    int trace_minb = 8;     // __launch_bounds__(128, N) variant of the queue tracer: resident blocks per SM the registers allow
    int direct_minb = 8;    // same for the direct kernel
    int refill_idle = 28;   // queue tracer: refill a warp once this many lanes are idle
};
extern Tune g_tune;

cudaError_t launch_render(const SceneDev& sc, const FrameParams& P, unsigned char* img, size_t pitch, const rtb_aovs& aov,
                          Counters* ctr, bool count, cudaStream_t st);

}  // namespace rtb
