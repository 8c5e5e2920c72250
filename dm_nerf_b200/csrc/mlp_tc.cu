// mlp_tc.cu -- tcgen05 bf16 path (placeholder until the kernel lands).
#include "mlp_layout.cuh"
namespace dmnerf {
int64_t bf16_image_bytes(int) { return -1; }
int pack_bf16(const float* const*, int, void*, cudaStream_t) { set_error("bf16 mode: not built yet"); return DMNERF_EUNSUPPORTED; }
int launch_mlp_bf16(const MlpInput&, const void*, int, float*, cudaStream_t) { set_error("bf16 mode: not built yet"); return DMNERF_EUNSUPPORTED; }
}  // namespace dmnerf
extern "C" int dmnerf_selftest_umma(const float*, const float*, int, int, float*, void*) { return DMNERF_EUNSUPPORTED; }
