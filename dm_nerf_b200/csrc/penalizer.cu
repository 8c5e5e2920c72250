// penalizer.cu -- a7 emptiness penalizer (networks/penalizer.py:5-62), forward and backward.
// One warp per ray segment (1..8 warps per ray, chosen from the batch size): per-sample coefficients (Gaussian bell, before/middle masks) go to shared
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

// The transcendental chain keeps the accurate expf / logf / IEEE division in every mode: a variant on the hardware
// approximations (ex2 / lg2) was built and measured SLOWER (262 vs 229 us on 4096 x 192 x 98 logits): with the math
// gone the warps sit on their global loads (long-scoreboard stalls 12.6 vs 2.7 per issue), while the accurate chain
// hides that latency for free.
__device__ __forceinline__ float pen_sigmoid(float x) { return sigmoidf_ref(x); }
__device__ __forceinline__ float pen_nlog(float q) { return -logf(q); }
__device__ __forceinline__ float pen_div(float a, float b) { return a / b; }

// A warp owns `sp` consecutive samples of one ray (`parts` warps per ray).  Pass A, lanes = samples: coefficients, mask
// counts, the "empty" (last) channel of every weighted sample, and a compacted list of the samples in FRONT of the
// surface band -- the only ones whose other K channels count.  Pass B, lanes = (listed sample, channel 4..C-2)
// flattened: every lane that loads a logit also needs it, index split by one multiply-high.
struct PenUnit { int64_t ray; int s0, s1; };
__device__ __forceinline__ PenUnit pen_unit(int warp, int parts, int sp, int S) {
    const int64_t unit = static_cast<int64_t>(blockIdx.x) * kPenWarps + warp;
    PenUnit u;
    u.ray = unit / parts;
    u.s0 = static_cast<int>(unit - u.ray * parts) * sp;
    u.s1 = min(S, u.s0 + sp);
    return u;
}

__global__ void __launch_bounds__(kPenWarps * 32)
penalizer_fwd_kernel(const float* __restrict__ raw, const float* __restrict__ z, const float* __restrict__ depth,
                     const float* __restrict__ rays_d, int64_t n, int S, int C, float tol, float dw,
                     float* __restrict__ partials, int parts, int sp) {
    extern __shared__ float smem[];
    __shared__ float red[kPenWarps][4];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const PenUnit u = pen_unit(warp, parts, sp, S);
    float sb = 0.f, nb = 0.f, sm = 0.f, nm = 0.f;
    if (u.ray < n && u.s0 < u.s1) {
        float* cb = smem + warp * 2 * sp;
        int* list = reinterpret_cast<int*>(cb + sp);
        const float dx = rays_d[u.ray * 3], dy = rays_d[u.ray * 3 + 1], dz = rays_d[u.ray * 3 + 2];
        const float norm = sqrtf(dx * dx + dy * dy + dz * dz);
        const float dep = depth[u.ray];
        const float* row = raw + u.ray * S * static_cast<int64_t>(C);
        int nbef = 0;
        for (int sbase = u.s0; sbase < u.s1; sbase += 32) {
            const int s = sbase + lane;
            float a = 0.f, b = 0.f;
            if (s < u.s1) {
                float mb, mm;
                sample_coeffs(z[u.ray * S + s], dep, norm, tol, dw, a, b, mb, mm);
                nb += mb; nm += mm;
                cb[s - u.s0] = a;
                if (a != 0.f || b != 0.f) {
                    const float p = pen_sigmoid(__ldg(row + static_cast<int64_t>(s) * C + C - 1));
                    if (a != 0.f) sb += pen_nlog(p + 1e-8f) * a;
                    if (b != 0.f) sm += pen_nlog(1.0f - p + 1e-8f) * b;
                }
            }
            const unsigned m = __ballot_sync(0xffffffffu, a != 0.f);
            if (a != 0.f) list[nbef + __popc(m & ((1u << lane) - 1u))] = s;
            nbef += __popc(m);
        }
        __syncwarp();
        const int per = C - 5;
        const unsigned total = static_cast<unsigned>(nbef) * per, magic = per > 1 ? 0xFFFFFFFFu / per + 1u : 0u;   // per == 1 (K = 1): li = e below
        constexpr int U = 4;
        for (unsigned base = 0; base < total; base += 32 * U) {
            float x[U], w[U];
#pragma unroll
            for (int q = 0; q < U; ++q) {
                const unsigned e = base + q * 32 + lane;
                w[q] = 0.f; x[q] = 0.f;
                if (e < total) {
                    const unsigned li = per > 1 ? __umulhi(e, magic) : e;
                    const int sidx = list[li];
                    w[q] = cb[sidx - u.s0];
                    x[q] = __ldg(row + static_cast<int64_t>(sidx) * C + 4 + (e - li * per));
                }
            }
#pragma unroll
            for (int q = 0; q < U; ++q)
                if (w[q] != 0.f) sb += pen_nlog(1.0f - pen_sigmoid(x[q]) + 1e-8f) * w[q];
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

// Backward.  kAccum = false: d_raw is this op's own gradient, every element of the warp's range is written (zeros where
// nothing flows).  kAccum = true (flags bit 1): d_raw already holds the compositing gradient of the same `raw` and only
// the weighted logits are touched (+=) -- no zero stream, and no separate gradient-sum pass afterwards.
template <bool kAccum>
__global__ void __launch_bounds__(kPenWarps * 32)
penalizer_bwd_kernel(const float* __restrict__ raw, const float* __restrict__ z, const float* __restrict__ depth,
                     const float* __restrict__ rays_d, int64_t n, int S, int C, float tol, float dw,
                     const float* __restrict__ partials, const float* __restrict__ d_loss,
                     float* __restrict__ d_raw, int parts, int sp) {
    extern __shared__ float smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const PenUnit u = pen_unit(warp, parts, sp, S);
    if (u.ray >= n || u.s0 >= u.s1) return;
    float* cb = smem + warp * 2 * sp;
    int* list = reinterpret_cast<int*>(cb + sp);
    const float g = d_loss ? *d_loss : 1.0f;
    const int L = C - 4;
    const float kb = g / (static_cast<float>(L) * fmaxf(partials[1], 1e-8f));
    const float km = g / fmaxf(partials[3], 1e-8f);
    const float dx = rays_d[u.ray * 3], dy = rays_d[u.ray * 3 + 1], dz = rays_d[u.ray * 3 + 2];
    const float norm = sqrtf(dx * dx + dy * dy + dz * dz);
    const float dep = depth[u.ray];
    const float* row = raw + u.ray * S * static_cast<int64_t>(C);
    float* out = d_raw + u.ray * S * static_cast<int64_t>(C);
    if (!kAccum) {
        // zeros first (coalesced), the weighted logits are overwritten below by the same warp
        for (int i = u.s0 * C + lane; i < u.s1 * C; i += 32) out[i] = 0.f;
        __syncwarp();
    }
    int nbef = 0;
    for (int sbase = u.s0; sbase < u.s1; sbase += 32) {
        const int s = sbase + lane;
        float a = 0.f, b = 0.f;
        if (s < u.s1) {
            float mb, mm;
            sample_coeffs(z[u.ray * S + s], dep, norm, tol, dw, a, b, mb, mm);
            a *= kb; b *= km;
            cb[s - u.s0] = a;
            if (a != 0.f || b != 0.f) {
                const int64_t idx = static_cast<int64_t>(s) * C + C - 1;
                const float p = pen_sigmoid(__ldg(row + idx));
                const float dp = p * (1.0f - p);
                const float v = a * (-pen_div(dp, p + 1e-8f)) + b * pen_div(dp, 1.0f - p + 1e-8f);
                if (kAccum) out[idx] += v; else out[idx] = v;
            }
        }
        const unsigned m = __ballot_sync(0xffffffffu, a != 0.f);
        if (a != 0.f) list[nbef + __popc(m & ((1u << lane) - 1u))] = s;
        nbef += __popc(m);
    }
    __syncwarp();
    const int per = C - 5;
    const unsigned total = static_cast<unsigned>(nbef) * per, magic = per > 1 ? 0xFFFFFFFFu / per + 1u : 0u;   // per == 1 (K = 1): li = e below
    constexpr int U = 4;
    for (unsigned base = 0; base < total; base += 32 * U) {
        float x[U], w[U], acc[U];
        int64_t idx[U];
#pragma unroll
        for (int q = 0; q < U; ++q) {
            const unsigned e = base + q * 32 + lane;
            w[q] = 0.f; x[q] = 0.f; acc[q] = 0.f; idx[q] = 0;
            if (e < total) {
                const unsigned li = per > 1 ? __umulhi(e, magic) : e;
                const int sidx = list[li];
                w[q] = cb[sidx - u.s0];
                idx[q] = static_cast<int64_t>(sidx) * C + 4 + (e - li * per);
                x[q] = __ldg(row + idx[q]);
                if (kAccum) acc[q] = out[idx[q]];
            }
        }
#pragma unroll
        for (int q = 0; q < U; ++q) {
            if (w[q] != 0.f) {
                const float p = pen_sigmoid(x[q]);
                out[idx[q]] = acc[q] + w[q] * pen_div(p * (1.0f - p), 1.0f - p + 1e-8f);   // d/dx of -log(1-p+1e-8)
            }
        }
    }
}

}  // namespace dmnerf

using namespace dmnerf;

// warps per ray: enough units for several waves of ~48 warps on each of 148 SMs (a single wave plus a thin tail costs
// two wave times: the kernels are latency bound per warp), at least 16 samples per warp
static void pen_split(int64_t n, int s, int* parts, int* sp) {
    int p = static_cast<int>((148 * 48 * 4 + n - 1) / (n > 0 ? n : 1));
    p = p < 1 ? 1 : (p > 8 ? 8 : p);
    while (p > 1 && (s + p - 1) / p < 16) --p;
    *sp = (s + p - 1) / p;
    *parts = (s + *sp - 1) / *sp;
}

static int pen_check(const void* raw, const void* z, const void* depth, const void* rays_d, int s, int c, int64_t n) {
    DM_REQUIRE(n == 0 || (raw && z && depth && rays_d), DMNERF_EINVAL, "penalizer: null pointer");
    DM_REQUIRE(s >= 1 && s <= 1024 && c >= 6 && c <= 128, DMNERF_EUNSUPPORTED,
               "penalizer: n_samples %d / channels %d unsupported", s, c);
    return DMNERF_OK;
}

extern "C" int dmnerf_penalizer_fwd(const float* raw, const float* z, const float* depth, const float* rays_d,
                                    int64_t n, int s, int c, float tolerance, float deta_w, int flags, float* partials,
                                    float* loss, void* stream) {
    if (int rc = pen_check(raw, z, depth, rays_d, s, c, n)) return rc;
    DM_REQUIRE(partials && loss, DMNERF_EINVAL, "penalizer_fwd: null output");
    cudaStream_t st = as_stream(stream);
    DM_CUDA(cudaMemsetAsync(partials, 0, 4 * sizeof(float), st));
    if (n > 0) {
        int parts, sp;
        pen_split(n, s, &parts, &sp);
        const size_t smem = static_cast<size_t>(kPenWarps) * 2 * sp * sizeof(float);
        const unsigned grid = static_cast<unsigned>((n * parts + kPenWarps - 1) / kPenWarps);
        auto kern = penalizer_fwd_kernel;
        DM_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        kern<<<grid, kPenWarps * 32, smem, st>>>(raw, z, depth, rays_d, n, s, c, tolerance, deta_w, partials, parts, sp);
        DM_LAUNCH_CHECK("penalizer_fwd_kernel");
    }
    penalizer_finish_kernel<<<1, 1, 0, st>>>(partials, c - 4, loss);
    DM_LAUNCH_CHECK("penalizer_finish_kernel");
    return DMNERF_OK;
}

extern "C" int dmnerf_penalizer_bwd(const float* raw, const float* z, const float* depth, const float* rays_d,
                                    int64_t n, int s, int c, float tolerance, float deta_w, int flags,
                                    const float* partials, const float* d_loss, float* d_raw, void* stream) {
    if (int rc = pen_check(raw, z, depth, rays_d, s, c, n)) return rc;
    DM_REQUIRE(partials && d_raw, DMNERF_EINVAL, "penalizer_bwd: null pointer");
    if (n <= 0) return DMNERF_OK;
    int parts, sp;
    pen_split(n, s, &parts, &sp);
    const size_t smem = static_cast<size_t>(kPenWarps) * 2 * sp * sizeof(float);
    const unsigned grid = static_cast<unsigned>((n * parts + kPenWarps - 1) / kPenWarps);
    auto kern = (flags & DMNERF_PEN_ACCUMULATE) ? penalizer_bwd_kernel<true> : penalizer_bwd_kernel<false>;
    DM_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<grid, kPenWarps * 32, smem, as_stream(stream)>>>(raw, z, depth, rays_d, n, s, c, tolerance, deta_w, partials,
                                                          d_loss, d_raw, parts, sp);
    DM_LAUNCH_CHECK("penalizer_bwd_kernel");
    return DMNERF_OK;
}
