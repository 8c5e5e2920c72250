// mlp_bwd.cu -- a12 backward of the fp32 MLP (placeholder until the kernel lands).
#include "mlp_layout.cuh"
namespace dmnerf {
int launch_mlp_bwd_fp32(const float*, const float*, const float*, int64_t, int, float* const*, cudaStream_t) {
    set_error("mlp_bwd: not built yet");
    return DMNERF_EUNSUPPORTED;
}
}  // namespace dmnerf
