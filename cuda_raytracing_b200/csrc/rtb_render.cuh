// rtb_render.cuh -- host interface of the frame kernels (rtb_render.cu, rtb_wavefront.cu).
#pragma once
#include "rtb_internal.cuh"

namespace rtb {

cudaError_t launch_render(const SceneDev& sc, const FrameParams& P, unsigned char* img, size_t pitch, const rtb_aovs& aov,
                          Counters* ctr, bool count, cudaStream_t st);

}  // namespace rtb
