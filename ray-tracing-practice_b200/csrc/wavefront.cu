// wavefront.cu -- wavefront variant of the render path (generate / extend / shade over global queues).
#include "kernels.h"
namespace rtb {
cudaError_t launch_render_wavefront(const DevScene&, const FrameParams&, int, int, int, LaunchCfg, WavefrontScratch&, cudaStream_t) {
    return cudaErrorNotSupported;
}
}  // namespace rtb
