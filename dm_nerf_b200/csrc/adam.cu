// adam.cu -- SURVEY 8(f) rank 3: the optimizer step of the training loop (train_dmsr.py:62-72 calls
// torch.optim.Adam(params, lr, betas=(0.9, 0.999)).step(), then decays the learning rate) as ONE elementwise kernel
// over the flat parameter buffer that is also the gradient all-reduce's buffer layout.
//
// Same update as torch.optim.Adam without weight decay / amsgrad:
//   m = b1 m + (1-b1) g;  v = b2 v + (1-b2) g^2;  p -= lr / (1 - b1^t) * m / (sqrt(v) / sqrt(1 - b2^t) + eps)
// The step count t and (optionally) the learning rate are read from DEVICE scalars so the launch can sit in a CUDA
// graph; g is multiplied by grad_scale first (1 / world_size folds the data-parallel mean into the update).
// HBM-bound: 4 reads + 3 writes of 4 bytes per parameter (39 MB for both DM-NeRF models).
#include "common.cuh"

namespace dmnerf {

__global__ void __launch_bounds__(256)
adam_step_kernel(float4* __restrict__ p, const float4* __restrict__ g, float4* __restrict__ m, float4* __restrict__ v,
                 int64_t n4, const float* __restrict__ lr_dev, float lr, const int* __restrict__ step_dev, float beta1,
                 float beta2, float eps, float grad_scale) {
    __shared__ float s_c[2];
    if (threadIdx.x == 0) {
        const double t = static_cast<double>(*step_dev);
        const double bc1 = 1.0 - pow(static_cast<double>(beta1), t);
        const double bc2 = 1.0 - pow(static_cast<double>(beta2), t);
        const double rate = lr_dev ? static_cast<double>(*lr_dev) : static_cast<double>(lr);
        s_c[0] = static_cast<float>(rate / bc1);
        s_c[1] = static_cast<float>(sqrt(bc2));
    }
    __syncthreads();
    const float step_size = s_c[0], bc2_sqrt = s_c[1];
    const float omb1 = 1.0f - beta1, omb2 = 1.0f - beta2;
    for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n4;
         i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
        float4 pp = p[i], mm = m[i], vv = v[i];
        const float4 gg = g[i];
        float* pa = reinterpret_cast<float*>(&pp);
        float* ma = reinterpret_cast<float*>(&mm);
        float* va = reinterpret_cast<float*>(&vv);
        const float* ga = reinterpret_cast<const float*>(&gg);
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const float gr = ga[e] * grad_scale;
            ma[e] = beta1 * ma[e] + omb1 * gr;
            va[e] = beta2 * va[e] + omb2 * gr * gr;
            const float denom = sqrtf(va[e]) / bc2_sqrt + eps;
            pa[e] -= step_size * ma[e] / denom;
        }
        p[i] = pp; m[i] = mm; v[i] = vv;
    }
}

}  // namespace dmnerf

using namespace dmnerf;

extern "C" int dmnerf_adam_step(float* p, const float* g, float* m, float* v, int64_t n, const float* lr_dev, float lr,
                                const int* step_dev, float beta1, float beta2, float eps, float grad_scale, void* stream) {
    DM_REQUIRE(p && g && m && v && step_dev, DMNERF_EINVAL, "adam_step: null pointer");
    DM_REQUIRE(n >= 0 && (n & 3) == 0, DMNERF_EINVAL, "adam_step: n = %lld must be a multiple of 4", static_cast<long long>(n));
    DM_REQUIRE(((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(m) |
                 reinterpret_cast<uintptr_t>(v)) & 15) == 0, DMNERF_EINVAL, "adam_step: buffers must be 16-byte aligned");
    if (n == 0) return DMNERF_OK;
    const int64_t n4 = n / 4;
    const int64_t want = (n4 + 255) / 256;
    const unsigned grid = static_cast<unsigned>(want < 148 * 8 ? want : 148 * 8);
    adam_step_kernel<<<grid, 256, 0, as_stream(stream)>>>(reinterpret_cast<float4*>(p), reinterpret_cast<const float4*>(g),
                                                         reinterpret_cast<float4*>(m), reinterpret_cast<float4*>(v), n4,
                                                         lr_dev, lr, step_dev, beta1, beta2, eps, grad_scale);
    DM_LAUNCH_CHECK("adam_step_kernel");
    return DMNERF_OK;
}

// ---- pixel draw of the training loops (helpers.py:99-111: np.random.choice(H*W, N, replace=False)) -------------------
// N DISTINCT pixel indices per step as the first N values of a keyed pseudo-random permutation of [0, total): a 4-round
// Feistel network over the smallest even-width power-of-two domain, cycle-walked back into range (expected < 4 walks).
// One launch of N threads instead of permuting all H*W indices (torch.randperm: five radix-sort passes over 307,200
// keys every step); the key mixes a seed with the DEVICE-side step counter, so the draw changes under CUDA-graph replay.
namespace dmnerf {

__device__ __forceinline__ uint32_t mix32(uint32_t x) {          // murmur3 finalizer
    x ^= x >> 16; x *= 0x85EBCA6Bu; x ^= x >> 13; x *= 0xC2B2AE35u; x ^= x >> 16;
    return x;
}

__global__ void draw_pixels_kernel(int64_t* __restrict__ out, int64_t n, uint32_t total, const int* __restrict__ step_dev,
                                   uint32_t seed_lo, uint32_t seed_hi, int half_bits) {
    const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t step = static_cast<uint32_t>(*step_dev);
    const uint32_t key = mix32(seed_lo ^ mix32(step + 0x9E3779B9u)) ^ seed_hi;
    const uint32_t mask = (1u << half_bits) - 1u;
    uint32_t v = static_cast<uint32_t>(i);
    for (int walk = 0; walk < 256; ++walk) {                    // (3/4)^256: never exhausted in practice
        uint32_t l = v >> half_bits, r = v & mask;
#pragma unroll
        for (int round = 0; round < 4; ++round) {
            const uint32_t f = mix32(r ^ (key + 0x632BE5ABu * (round + 1))) & mask;
            const uint32_t nl = r;
            r = l ^ f;
            l = nl;
        }
        v = (l << half_bits) | r;
        if (v < total) break;                                       // a permutation of the domain restricted to [0,total)
    }
    out[i] = static_cast<int64_t>(v < total ? v : static_cast<uint32_t>(i));
}

}  // namespace dmnerf

extern "C" int dmnerf_draw_pixels(int64_t* idx_out, int64_t n, int64_t total, const int* step_dev, uint64_t seed,
                                  void* stream) {
    DM_REQUIRE(idx_out && step_dev, DMNERF_EINVAL, "draw_pixels: null pointer");
    DM_REQUIRE(n >= 0 && total >= 1 && n <= total && total <= (int64_t(1) << 30), DMNERF_EINVAL,
               "draw_pixels: need 0 <= n <= total <= 2^30 (n=%lld total=%lld)", static_cast<long long>(n), static_cast<long long>(total));
    if (n == 0) return DMNERF_OK;
    int bits = 2;
    while ((int64_t(1) << bits) < total) bits += 2;                  // even width: two equal halves
    draw_pixels_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, as_stream(stream)>>>(
        idx_out, n, static_cast<uint32_t>(total), step_dev, static_cast<uint32_t>(seed), static_cast<uint32_t>(seed >> 32), bits / 2);
    DM_LAUNCH_CHECK("draw_pixels_kernel");
    return DMNERF_OK;
}
