// common.cuh -- shared helpers for libdmnerf_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include "../../include/dmnerf_b200.h"

namespace dmnerf {

// ---- error plumbing ---------------------------------------------------------------------------
void set_error(const char* fmt, ...);
int  cuda_fail(cudaError_t e, const char* what);

#define DM_REQUIRE(cond, code, ...)                      \
    do {                                                 \
        if (!(cond)) {                                   \
            ::dmnerf::set_error(__VA_ARGS__);            \
            return (code);                               \
        }                                                \
    } while (0)

#define DM_CUDA(call)                                                        \
    do {                                                                     \
        cudaError_t e__ = (call);                                            \
        if (e__ != cudaSuccess) return ::dmnerf::cuda_fail(e__, #call);      \
    } while (0)

#define DM_LAUNCH_CHECK(name)                                                \
    do {                                                                     \
        cudaError_t e__ = cudaGetLastError();                                \
        if (e__ != cudaSuccess) return ::dmnerf::cuda_fail(e__, name);       \
    } while (0)

static inline cudaStream_t as_stream(void* s) { return reinterpret_cast<cudaStream_t>(s); }

// ---- network shape (the supported contract, include/dmnerf_b200.h) ----------------------------
constexpr int kWidth    = 256;
constexpr int kPosCh    = 63;   // 3 + 3*2*10
constexpr int kDirCh    = 27;   // 3 + 3*2*4
constexpr int kPosPad   = 64;
constexpr int kDirPad   = 32;
constexpr int kPosFreqs = 10;
constexpr int kDirFreqs = 4;
constexpr int kMaxIns   = 123;  // C = ins_num + 5 <= 128

// state_dict order of the 30 tensors (15 Linear modules) (networks/dm_nerf.py:65-78)
enum ParamIdx {
    P_MLP0_W = 0,  // mlps.i.weight = 2*i, mlps.i.bias = 2*i+1
    P_RGBF_W = 16, P_RGBF_B = 17,   // rgb_feature_linear
    P_INSF_W = 18, P_INSF_B = 19,   // ins_feature_linear
    P_RGB0_W = 20, P_RGB0_B = 21,   // rgb_feature_linears.0
    P_INS0_W = 22, P_INS0_B = 23,   // ins_feature_linears.0
    P_DENS_W = 24, P_DENS_B = 25,   // density_linear
    P_INSL_W = 26, P_INSL_B = 27,   // ins_linear
    P_RGBL_W = 28, P_RGBL_B = 29,   // rgb_linear
};
constexpr int kNumParams = 30;
static_assert(kNumParams == DMNERF_NUM_PARAMS, "header out of sync");

// ---- device math helpers --------------------------------------------------------------------------
// Same formula torch's CUDA sigmoid kernel evaluates (1/(1+exp(-x)), full-precision expf, IEEE div):
// saturates to exactly 1.0f from x >= ~17.33, which is what creates the reference's argmax ties.
__device__ __forceinline__ float sigmoidf_ref(float x) { return 1.0f / (1.0f + expf(-x)); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// per-sample alpha / transmittance for one chunk of <= 32 samples of a ray (render.py:9-14), one lane per sample;
// returns w, updates carry.  Shared by the compositing kernels and the coarse tail so that both produce the same bits.
__device__ __forceinline__ void scan_chunk(float sigma, float dist, bool valid, int lane, float& carry,
                                           float& alpha, float& trans, float& w) {
    alpha = 1.0f - expf(-fmaxf(sigma, 0.0f) * dist);
    float t = valid ? (1.0f - alpha) + 1e-10f : 1.0f;
    float incl = t;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        float v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl *= v;
    }
    float excl = __shfl_up_sync(0xffffffffu, incl, 1);
    if (lane == 0) excl = 1.0f;
    trans = carry * excl;
    w = valid ? alpha * trans : 0.0f;
    carry *= __shfl_sync(0xffffffffu, incl, 31);
}

}  // namespace dmnerf
