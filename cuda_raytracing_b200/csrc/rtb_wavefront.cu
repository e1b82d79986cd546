// rtb_wavefront.cu -- wavefront pipeline for secondary rays (RTB_MODE_REFLECT / RTB_MODE_PATH).
#include "rtb_render.cuh"
#include "rtb_shade.cuh"

struct rtb_context;

namespace rtb {

cudaError_t launch_wavefront(rtb_context*, void**, const SceneDev&, const FrameParams&, unsigned char*, size_t, const rtb_aovs&,
                             Counters*, bool, cudaStream_t)
{
    return cudaErrorNotSupported;
}

}  // namespace rtb

extern "C" void rtb_wavefront_release(rtb_context*) {}
