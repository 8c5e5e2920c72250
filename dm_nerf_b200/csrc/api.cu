// api.cu -- error plumbing and the extern "C" entry points of the MLP path (include/dmnerf_b200.h).
#include <stdarg.h>
#include "mlp_layout.cuh"

namespace dmnerf {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int cuda_fail(cudaError_t e, const char* what) {
    set_error("CUDA error %d (%s) at %s", static_cast<int>(e), cudaGetErrorString(e), what);
    return DMNERF_ECUDA;
}


// mlp_simt.cu
int pack_fp32(const float* const* params, int ins_num, float* packed, cudaStream_t st);
int launch_mlp_fp32(const MlpInput& in, const float* packed, int ins_num, float* raw_out, float* acts,
                    cudaStream_t st);
// mlp_bwd.cu
int launch_mlp_bwd_fp32(const float* packed, const float* acts, const float* d_raw, int64_t n_points, int ins_num,
                        float* const* grads, cudaStream_t st);
// mlp_tc.cu
int64_t bf16_image_bytes(int ins_num);
int pack_bf16(const float* const* params, int ins_num, void* packed, cudaStream_t st, bool f16 = false);
int64_t x3_image_bytes(int ins_num);
int pack_x3(const float* const* params, int ins_num, void* packed, cudaStream_t st);
int launch_mlp_x3(const MlpInput& in, const void* packed, int ins_num, float* raw_out, void* stash, cudaStream_t st,
                  bool dens = false);
int launch_mlp_bf16(const MlpInput& in, const void* packed, int ins_num, float* raw_out, void* stash, cudaStream_t st,
                    bool f16 = false, bool dens = false);
int tc_debug_flags();
// mlp_tc_bwd.cu
int64_t bf16_acts_bytes(int64_t n_points);
int64_t bf16_bwd_workspace_bytes(int64_t n_points);
int launch_mlp_bwd_bf16(const void* packed, const void* acts, const float* d_raw, int64_t n_points, int ins_num,
                        float* const* grads, void* workspace, int debug_flags, cudaStream_t st);
int launch_selftest_umma_mn(const float* a, const float* b, int n, float* d, int debug_flags, cudaStream_t st);

}  // namespace dmnerf

using namespace dmnerf;

extern "C" int dmnerf_version(void) { return 100; }

extern "C" int dmnerf_built_arch(void) { return 100; }

extern "C" int dmnerf_has_experimental(void) {
#ifdef DMNERF_EXPERIMENTAL
    return 1;
#else
    return 0;
#endif
}

extern "C" int dmnerf_last_error(char* buf, int cap) {
    const int len = static_cast<int>(strlen(g_err));
    if (buf && cap > 0) {
        strncpy(buf, g_err, cap - 1);
        buf[cap - 1] = 0;
    }
    return len;
}

static int check_ins(int ins_num, const char* who) {
    DM_REQUIRE(ins_num >= 1 && ins_num <= kMaxIns, DMNERF_EUNSUPPORTED, "%s: ins_num %d outside [1,%d]", who, ins_num,
               kMaxIns);
    return DMNERF_OK;
}

extern "C" int64_t dmnerf_packed_bytes(int ins_num, int mode) {
    if (ins_num < 1 || ins_num > kMaxIns) return -1;
    if (mode == DMNERF_MODE_FP32) return fp32_image(ins_num).total_floats * 4;
    if (mode == DMNERF_MODE_BF16 || mode == DMNERF_MODE_F16) return bf16_image_bytes(ins_num);
    if (mode == DMNERF_MODE_TC32) return x3_image_bytes(ins_num);
    return -1;
}

extern "C" int dmnerf_weights_pack(const float* const* params, int ins_num, int mode, void* packed, void* stream) {
    DM_REQUIRE(params && packed, DMNERF_EINVAL, "weights_pack: null pointer");
    if (int rc = check_ins(ins_num, "weights_pack")) return rc;
    for (int i = 0; i < kNumParams; ++i) DM_REQUIRE(params[i], DMNERF_EINVAL, "weights_pack: params[%d] is null", i);
    if (mode == DMNERF_MODE_FP32) return pack_fp32(params, ins_num, static_cast<float*>(packed), as_stream(stream));
    if (mode == DMNERF_MODE_BF16) return pack_bf16(params, ins_num, packed, as_stream(stream));
    if (mode == DMNERF_MODE_F16) return pack_bf16(params, ins_num, packed, as_stream(stream), true);
    if (mode == DMNERF_MODE_TC32) return pack_x3(params, ins_num, packed, as_stream(stream));
    DM_REQUIRE(false, DMNERF_EUNSUPPORTED, "weights_pack: unknown mode %d", mode);
}

extern "C" int64_t dmnerf_acts_bytes(int64_t n_points, int ins_num, int mode) {
    if (n_points < 0 || ins_num < 1 || ins_num > kMaxIns) return -1;
    if (mode == DMNERF_MODE_BF16) return bf16_acts_bytes(n_points);
    if (mode != DMNERF_MODE_FP32) return -1;
    const int64_t tiles = (n_points + kTileM - 1) / kTileM;
    return tiles * kActRows * kTileM * 4;
}

static int run_mlp(const MlpInput& in, const void* packed, int ins_num, int mode, float* raw_out, void* acts,
                   void* stream) {
    DM_REQUIRE(packed && raw_out, DMNERF_EINVAL, "mlp_fwd: null pointer");
    if (int rc = check_ins(ins_num, "mlp_fwd")) return rc;
    if (in.n_points <= 0) return DMNERF_OK;
    if (mode == DMNERF_MODE_FP32)
        return launch_mlp_fp32(in, static_cast<const float*>(packed), ins_num, raw_out, static_cast<float*>(acts),
                               as_stream(stream));
    if (mode == DMNERF_MODE_BF16) {
        return launch_mlp_bf16(in, packed, ins_num, raw_out, acts, as_stream(stream));
    }
    if (mode == DMNERF_MODE_F16) return launch_mlp_bf16(in, packed, ins_num, raw_out, acts, as_stream(stream), true);
    if (mode == DMNERF_MODE_TC32) return launch_mlp_x3(in, packed, ins_num, raw_out, acts, as_stream(stream));
    DM_REQUIRE(false, DMNERF_EUNSUPPORTED, "mlp_fwd: unknown mode %d", mode);
}

extern "C" int dmnerf_mlp_fwd(const float* rays_o, const float* rays_d, const float* z, int64_t n_rays, int n_samples,
                              const void* packed, int ins_num, int mode, float* raw_out, void* acts, void* stream) {
    if (n_rays == 0) return DMNERF_OK;
    DM_REQUIRE(rays_o && rays_d && z, DMNERF_EINVAL, "mlp_fwd: null ray pointer");
    DM_REQUIRE(n_rays >= 0 && n_samples >= 1 && n_samples <= 4096, DMNERF_EUNSUPPORTED,
               "mlp_fwd: n_samples %d outside [1,4096]", n_samples);
    MlpInput in{rays_o, rays_d, z, nullptr, n_rays * n_samples, n_samples};
    return run_mlp(in, packed, ins_num, mode, raw_out, acts, stream);
}

extern "C" int dmnerf_mlp_density(const float* rays_o, const float* rays_d, const float* z, int64_t n_rays, int n_samples,
                                  const void* packed, int ins_num, int mode, float* sigma_out, void* stream) {
    if (n_rays == 0) return DMNERF_OK;
    DM_REQUIRE(rays_o && rays_d && z && packed && sigma_out, DMNERF_EINVAL, "mlp_density: null pointer");
    DM_REQUIRE(n_rays >= 0 && n_samples >= 1 && n_samples <= 4096, DMNERF_EUNSUPPORTED,
               "mlp_density: n_samples %d outside [1,4096]", n_samples);
    if (int rc = check_ins(ins_num, "mlp_density")) return rc;
    DM_REQUIRE(mode == DMNERF_MODE_BF16 || mode == DMNERF_MODE_F16 || mode == DMNERF_MODE_TC32, DMNERF_EUNSUPPORTED,
               "mlp_density: mode %d (the density-only trunk exists for the tensor-core modes; take raw[...,3] of "
               "dmnerf_mlp_fwd in DMNERF_MODE_FP32)", mode);
    MlpInput in{rays_o, rays_d, z, nullptr, n_rays * n_samples, n_samples};
    if (mode == DMNERF_MODE_TC32) return launch_mlp_x3(in, packed, ins_num, sigma_out, nullptr, as_stream(stream), true);
    return launch_mlp_bf16(in, packed, ins_num, sigma_out, nullptr, as_stream(stream), mode == DMNERF_MODE_F16, true);
}

extern "C" int dmnerf_mlp_fwd_embedded(const float* x, int64_t m, const void* packed, int ins_num, int mode,
                                       float* raw_out, void* acts, void* stream) {
    if (m == 0) return DMNERF_OK;
    DM_REQUIRE(x, DMNERF_EINVAL, "mlp_fwd_embedded: null input");
    MlpInput in{nullptr, nullptr, nullptr, x, m, 1};
    return run_mlp(in, packed, ins_num, mode, raw_out, acts, stream);
}

extern "C" int64_t dmnerf_bwd_workspace_bytes(int64_t n_points, int ins_num, int mode) {
    if (n_points < 0 || ins_num < 1 || ins_num > kMaxIns) return -1;
    if (mode == DMNERF_MODE_BF16) return bf16_bwd_workspace_bytes(n_points);
    return mode == DMNERF_MODE_FP32 ? 0 : -1;
}

extern "C" int dmnerf_mlp_bwd(const void* packed, const void* acts, const float* d_raw, int64_t n_points,
                              int ins_num, int mode, float* const* grads, void* workspace, void* stream) {
    DM_REQUIRE(packed && acts && d_raw && grads, DMNERF_EINVAL, "mlp_bwd: null pointer");
    if (int rc = check_ins(ins_num, "mlp_bwd")) return rc;
    for (int i = 0; i < kNumParams; ++i) DM_REQUIRE(grads[i], DMNERF_EINVAL, "mlp_bwd: grads[%d] is null", i);
    if (n_points <= 0) return DMNERF_OK;
    if (mode == DMNERF_MODE_BF16)
        return launch_mlp_bwd_bf16(packed, acts, d_raw, n_points, ins_num, grads, workspace, tc_debug_flags(),
                                   as_stream(stream));
    DM_REQUIRE(mode != DMNERF_MODE_F16 && mode != DMNERF_MODE_TC32, DMNERF_EUNSUPPORTED,
               "mlp_bwd: mode %d is an inference-only chain (train in DMNERF_MODE_BF16 or DMNERF_MODE_FP32)", mode);
    DM_REQUIRE(mode == DMNERF_MODE_FP32, DMNERF_EUNSUPPORTED, "mlp_bwd: unknown mode %d", mode);
    return launch_mlp_bwd_fp32(static_cast<const float*>(packed), static_cast<const float*>(acts), d_raw, n_points,
                               ins_num, grads, as_stream(stream));
}

extern "C" int dmnerf_selftest_umma_mn(const float* a, const float* b, int n, float* d, void* stream) {
    DM_REQUIRE(a && b && d, DMNERF_EINVAL, "selftest_umma_mn: null pointer");
    DM_REQUIRE(n >= 16 && n <= 256 && n % 16 == 0, DMNERF_EUNSUPPORTED, "selftest_umma_mn: n=%d (need a multiple of 16, <= 256)", n);
    return launch_selftest_umma_mn(a, b, n, d, tc_debug_flags(), as_stream(stream));
}
