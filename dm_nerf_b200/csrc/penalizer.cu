// penalizer.cu -- a7 emptiness penalizer (networks/penalizer.py:5-62), forward and backward.
// One warp per ray: per-sample coefficients (Gaussian bell, before/middle masks) go to shared
// memory, then the ray's S*C contiguous floats are streamed with coalesced loads/stores.
//   loss = Sb / ((K+1)*max(Nb,1e-8)) + Sm / max(Nm,1e-8)
//   Sb = sum_{s,k} bce_k(p) * (1-G_s)*before_s      bce_k = -log(1-p+1e-8) (k<K),  -log(p+1e-8) (k=K)
//   Sm = sum_s -log(1-p_K+1e-8) * G_s*middle_s      Nb = sum before, Nm = sum middle
// The four sums are batch-global (penalizer.py:42-43,52): block partials + 4 atomics per block.
#include "common.cuh"

namespace dmnerf {

constexpr int kPenWarps = 8;

__device__ __forceinline__ void sample_coeffs(float zs, float depth, float norm, float tol, float dw, float& coef_b,
                                              float& coef_m, float& m_b, float& m_m) {
    const float front = (depth - tol) * norm;
    const float back = (depth + tol) * norm;
    const float surf = depth * norm;
    const float along = zs * norm;
    const float gap = surf - along;
    const float two_var = 2.0f * (dw * dw);
    const float denom = 0.4f * sqrtf(6.283185307179586f);
    const float bell = expf(-(gap * gap) / two_var) / denom + 1e-8f;
    m_b = (along < front) ? 1.0f : 0.0f;
    const float m_a = (along > back) ? 1.0f : 0.0f;
    m_m = 1.0f - (m_a + m_b);
    coef_b = (1.0f - bell) * m_b;
    coef_m = bell * m_m;
}

__global__ void __launch_bounds__(kPenWarps * 32)
penalizer_fwd_kernel(const float* __restrict__ raw, const float* __restrict__ z, const float* __restrict__ depth,
                     const float* __restrict__ rays_d, int64_t n, int S, int C, float tol, float dw,
                     float* __restrict__ partials) {
    extern __shared__ float smem[];
    __shared__ float red[kPenWarps][4];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t ray = static_cast<int64_t>(blockIdx.x) * kPenWarps + warp;
    float sb = 0.f, nb = 0.f, sm = 0.f, nm = 0.f;
    if (ray < n) {
        float* cb = smem + warp * 2 * S;
        float* cm = cb + S;
        const float dx = rays_d[ray * 3], dy = rays_d[ray * 3 + 1], dz = rays_d[ray * 3 + 2];
        const float norm = sqrtf(dx * dx + dy * dy + dz * dz);
        const float dep = depth[ray];
        for (int s = lane; s < S; s += 32) {
            float a, b, mb, mm;
            sample_coeffs(z[ray * S + s], dep, norm, tol, dw, a, b, mb, mm);
            cb[s] = a; cm[s] = b; nb += mb; nm += mm;
        }
        __syncwarp();
        const float invC = 1.0f / static_cast<float>(C);
        const float* row = raw + ray * S * static_cast<int64_t>(C);
        const int total = S * C;
        // four independent elements per lane and trip: their loads are issued before any of the transcendental
        // chains starts (memory-level parallelism x4; a single dependent load per trip left the kernel latency bound)
        constexpr int U = 4;
        for (int base = 0; base < total; base += 32 * U) {
            float x[U], wb[U], wm[U];
            bool last[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int i = base + u * 32 + lane;
                wb[u] = 0.f; wm[u] = 0.f; x[u] = 0.f; last[u] = false;
                if (i < total) {
                    const int s = __float2int_rz((static_cast<float>(i) + 0.5f) * invC);
                    const int j = i - s * C;
                    // samples behind the surface band carry zero weight in both terms, and inside the band only the
                    // "empty" channel counts: those logits are never loaded (about half of the bytes for a mid-range depth)
                    if (j >= 4) {
                        last[u] = j == C - 1;
                        wb[u] = cb[s];
                        wm[u] = last[u] ? cm[s] : 0.f;
                        if (wb[u] != 0.f || wm[u] != 0.f) x[u] = __ldg(row + i);
                    }
                }
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                if (wb[u] != 0.f || wm[u] != 0.f) {
                    const float p = sigmoidf_ref(x[u]);
                    const float l_not = -logf(1.0f - p + 1e-8f);
                    if (last[u]) {
                        sb += -logf(p + 1e-8f) * wb[u];
                        sm += l_not * wm[u];
                    } else {
                        sb += l_not * wb[u];
                    }
                }
            }
        }
    }
    sb = warp_sum(sb); nb = warp_sum(nb); sm = warp_sum(sm); nm = warp_sum(nm);
    if (lane == 0) { red[warp][0] = sb; red[warp][1] = nb; red[warp][2] = sm; red[warp][3] = nm; }
    __syncthreads();
    if (threadIdx.x < 4) {
        float t = 0.f;
        for (int w = 0; w < kPenWarps; ++w) t += red[w][threadIdx.x];
        atomicAdd(partials + threadIdx.x, t);
    }
}

__global__ void penalizer_finish_kernel(const float* __restrict__ partials, int L, float* __restrict__ loss) {
    const float lb = partials[0] / (static_cast<float>(L) * fmaxf(partials[1], 1e-8f));
    const float lm = partials[2] / fmaxf(partials[3], 1e-8f);
    *loss = lb + lm;
}

__global__ void __launch_bounds__(kPenWarps * 32)
penalizer_bwd_kernel(const float* __restrict__ raw, const float* __restrict__ z, const float* __restrict__ depth,
                     const float* __restrict__ rays_d, int64_t n, int S, int C, float tol, float dw,
                     const float* __restrict__ partials, const float* __restrict__ d_loss,
                     float* __restrict__ d_raw) {
    extern __shared__ float smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t ray = static_cast<int64_t>(blockIdx.x) * kPenWarps + warp;
    if (ray >= n) return;
    float* cb = smem + warp * 2 * S;
    float* cm = cb + S;
    const float g = d_loss ? *d_loss : 1.0f;
    const int L = C - 4;
    const float kb = g / (static_cast<float>(L) * fmaxf(partials[1], 1e-8f));
    const float km = g / fmaxf(partials[3], 1e-8f);
    const float dx = rays_d[ray * 3], dy = rays_d[ray * 3 + 1], dz = rays_d[ray * 3 + 2];
    const float norm = sqrtf(dx * dx + dy * dy + dz * dz);
    const float dep = depth[ray];
    for (int s = lane; s < S; s += 32) {
        float a, b, mb, mm;
        sample_coeffs(z[ray * S + s], dep, norm, tol, dw, a, b, mb, mm);
        cb[s] = a * kb; cm[s] = b * km;
    }
    __syncwarp();
    const float invC = 1.0f / static_cast<float>(C);
    const float* row = raw + ray * S * static_cast<int64_t>(C);
    float* out = d_raw + ray * S * static_cast<int64_t>(C);
    const int total = S * C;
    constexpr int U = 4;
    for (int base = 0; base < total; base += 32 * U) {
        float x[U], wb[U], wm[U];
        bool last[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int i = base + u * 32 + lane;
            wb[u] = 0.f; wm[u] = 0.f; x[u] = 0.f; last[u] = false;
            if (i < total) {
                const int s = __float2int_rz((static_cast<float>(i) + 0.5f) * invC);
                const int j = i - s * C;
                if (j >= 4) {
                    last[u] = j == C - 1;
                    wb[u] = cb[s];
                    wm[u] = last[u] ? cm[s] : 0.f;
                    if (wb[u] != 0.f || wm[u] != 0.f) x[u] = __ldg(row + i);
                }
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int i = base + u * 32 + lane;
            if (i < total) {
                float v = 0.f;
                if (wb[u] != 0.f || wm[u] != 0.f) {
                    const float p = sigmoidf_ref(x[u]);
                    const float dp = p * (1.0f - p);
                    const float d_not = dp / (1.0f - p + 1e-8f);          // d/dx of -log(1-p+1e-8)
                    if (last[u]) v = wb[u] * (-dp / (p + 1e-8f)) + wm[u] * d_not;
                    else         v = wb[u] * d_not;
                }
                out[i] = v;
            }
        }
    }
}

}  // namespace dmnerf

using namespace dmnerf;

static int pen_check(const void* raw, const void* z, const void* depth, const void* rays_d, int s, int c, int64_t n) {
    DM_REQUIRE(n == 0 || (raw && z && depth && rays_d), DMNERF_EINVAL, "penalizer: null pointer");
    DM_REQUIRE(s >= 1 && s <= 1024 && c >= 6 && c <= 128, DMNERF_EUNSUPPORTED,
               "penalizer: n_samples %d / channels %d unsupported", s, c);
    return DMNERF_OK;
}

extern "C" int dmnerf_penalizer_fwd(const float* raw, const float* z, const float* depth, const float* rays_d,
                                    int64_t n, int s, int c, float tolerance, float deta_w, float* partials,
                                    float* loss, void* stream) {
    if (int rc = pen_check(raw, z, depth, rays_d, s, c, n)) return rc;
    DM_REQUIRE(partials && loss, DMNERF_EINVAL, "penalizer_fwd: null output");
    cudaStream_t st = as_stream(stream);
    DM_CUDA(cudaMemsetAsync(partials, 0, 4 * sizeof(float), st));
    if (n > 0) {
        const size_t smem = static_cast<size_t>(kPenWarps) * 2 * s * sizeof(float);
        DM_CUDA(cudaFuncSetAttribute(penalizer_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        const unsigned grid = static_cast<unsigned>((n + kPenWarps - 1) / kPenWarps);
        penalizer_fwd_kernel<<<grid, kPenWarps * 32, smem, st>>>(raw, z, depth, rays_d, n, s, c, tolerance, deta_w,
                                                                partials);
        DM_LAUNCH_CHECK("penalizer_fwd_kernel");
    }
    penalizer_finish_kernel<<<1, 1, 0, st>>>(partials, c - 4, loss);
    DM_LAUNCH_CHECK("penalizer_finish_kernel");
    return DMNERF_OK;
}

extern "C" int dmnerf_penalizer_bwd(const float* raw, const float* z, const float* depth, const float* rays_d,
                                    int64_t n, int s, int c, float tolerance, float deta_w, const float* partials,
                                    const float* d_loss, float* d_raw, void* stream) {
    if (int rc = pen_check(raw, z, depth, rays_d, s, c, n)) return rc;
    DM_REQUIRE(partials && d_raw, DMNERF_EINVAL, "penalizer_bwd: null pointer");
    if (n <= 0) return DMNERF_OK;
    const size_t smem = static_cast<size_t>(kPenWarps) * 2 * s * sizeof(float);
    DM_CUDA(cudaFuncSetAttribute(penalizer_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const unsigned grid = static_cast<unsigned>((n + kPenWarps - 1) / kPenWarps);
    penalizer_bwd_kernel<<<grid, kPenWarps * 32, smem, as_stream(stream)>>>(raw, z, depth, rays_d, n, s, c, tolerance,
                                                                           deta_w, partials, d_loss, d_raw);
    DM_LAUNCH_CHECK("penalizer_bwd_kernel");
    return DMNERF_OK;
}
