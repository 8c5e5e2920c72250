// sampler.cu -- a4 hierarchical sampling (networks/helpers.py:123-155), the coarse tail of
// dm_nerf (networks/render.py:66-70), sorted merges (networks/manipulator.py:190,195) and the
// stratified jitter (networks/render.py:40-47).  One warp per ray, everything in shared memory.
//
// CDF summation order (fixed and documented; oracle/dmnerf_oracle_c.c uses the same one, so indices are
// bit-exact against it): w_j = weights_j + 1e-5 (fp32); total = fp32(sum_j double(w_j)), sequential;
// pdf_j = w_j / total (fp32 IEEE division); cdf_0 = 0, cdf_{j+1} = fp32(sum_{i<=j} double(pdf_i)), one
// running double accumulator.  The running-double prefix sum IS what ATen's CPU cumsum does for float
// (acc_type<float,false> = double), and fp32(double sum) equals ATen's cascade `sum` except where that one
// is not correctly rounded -- so against the torch-CPU golden vectors most rays are bit-identical and the
// rest differ by <= 1 ulp of the cdf (a sequential fp32 order is off by several ulp).
// searchsorted(right=True): inds = #{j : cdf_j <= u}.  The lerp uses un-fused mul/add like ATen.
#include "common.cuh"
#include <math_constants.h>

namespace dmnerf {

constexpr int kSampWarps = 4;

// torch.linspace(0,1,n) in fp32: step=(1-0)/(n-1); i < n/2 ? step*i : 1 - step*(n-1-i)
__device__ __forceinline__ float linspace01(int i, int n) {
    const float step = 1.0f / static_cast<float>(n - 1);
    return (i < n / 2) ? __fmul_rn(step, static_cast<float>(i))
                       : __fsub_rn(1.0f, __fmul_rn(step, static_cast<float>(n - 1 - i)));
}

// Builds cdf[0..nb-1] in shared memory from wv[0..nb-2] (= weights + 1e-5 already stored).
// Semantics: the total and the running sums are accumulated SEQUENTIALLY in double and rounded to float once (what
// ATen's CPU cumsum does).  When the largest and smallest term are within 2^20 of each other -- always, for weights in
// [0,1] with the 1e-5 floor -- every partial sum of these <= 512 fp32 terms is exactly representable in a double
// (common quantum ulp(min) >= min * 2^-24, sum <= 512 max < 2^53 quanta), so the additions are exact, their order
// does not matter, and the sums are taken as a warp scan instead of two ~60-step dependent chains on one lane.
// Anything else (negative or wildly scaled weights) takes the sequential path.
__device__ __forceinline__ void build_cdf(float* cdf, float* wv, int nb, int lane) {
    const int m = nb - 1;
    const int per = (m + 31) >> 5, j0 = min(m, lane * per), j1 = min(m, j0 + per);
    float lo = CUDART_INF_F, hi = 0.f;
    double part = 0.0;
    for (int j = j0; j < j1; ++j) { const float v = wv[j]; lo = fminf(lo, v); hi = fmaxf(hi, v); part += static_cast<double>(v); }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, o));
        hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, o));
    }
    const bool exact = lo > 0.f && hi < lo * 1048576.0f;                   // warp-uniform
    double tot = part;
    if (exact) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, o);
    } else {
        tot = 0.0;
        for (int j = 0; j < m; ++j) tot += static_cast<double>(wv[j]);      // every lane, same order
    }
    const float total = static_cast<float>(tot);
    __syncwarp();
    for (int j = lane; j < m; j += 32) wv[j] = __fdiv_rn(wv[j], total);
    __syncwarp();
    if (exact) {
        // the normalised terms keep their ratios (one rounding each): still far inside the exactness bound
        double run = 0.0;
        for (int j = j0; j < j1; ++j) run += static_cast<double>(wv[j]);
        double incl = run;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const double up = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += up;
        }
        double c = incl - run;                                             // exclusive prefix of this lane's segment
        if (lane == 0) cdf[0] = 0.f;
        for (int j = j0; j < j1; ++j) { c += static_cast<double>(wv[j]); cdf[j + 1] = static_cast<float>(c); }
    } else if (lane == 0) {
        double c = 0.0;
        cdf[0] = 0.f;
        for (int j = 0; j < m; ++j) { c += static_cast<double>(wv[j]); cdf[j + 1] = static_cast<float>(c); }
    }
    __syncwarp();
}

__device__ __forceinline__ float invert_cdf(const float* cdf, const float* bins, int nb, float u, int& inds) {
    int lo = 0, hi = nb;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (cdf[mid] <= u) lo = mid + 1; else hi = mid;
    }
    inds = lo;
    const int below = max(lo - 1, 0);
    const int above = min(lo, nb - 1);
    const float cb = cdf[below], ca = cdf[above];
    float denom = __fsub_rn(ca, cb);
    if (denom < 1e-5f) denom = 1.0f;
    const float t = __fdiv_rn(__fsub_rn(u, cb), denom);
    const float bb = bins[below], ba = bins[above];
    return __fadd_rn(bb, __fmul_rn(t, __fsub_rn(ba, bb)));
}

// In-warp bitonic sort of P (power of two) floats in shared memory, ascending.
__device__ __forceinline__ void bitonic_sort(float* v, int P, int lane) {
    for (int k = 2; k <= P; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = lane; i < (P >> 1); i += 32) {
                const int lo = ((i / j) * (j << 1)) + (i % j);
                const int hi = lo + j;
                const bool up = (lo & k) == 0;
                const float a = v[lo], b = v[hi];
                if ((a > b) == up) { v[lo] = b; v[hi] = a; }
            }
            __syncwarp();
        }
    }
}

__device__ __forceinline__ int next_pow2(int x) { int p = 1; while (p < x) p <<= 1; return p; }

// mode 0: weights given, cdf built here.  mode 1: cdf given (strict mode).
__global__ void __launch_bounds__(kSampWarps * 32)
sample_pdf_kernel(const float* __restrict__ bins, const float* __restrict__ weights_or_cdf, int cdf_given,
                  const float* __restrict__ u, int64_t u_stride, int64_t n, int nb, int ns,
                  float* __restrict__ z_out, int64_t* __restrict__ inds_out) {
    extern __shared__ float smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t ray = static_cast<int64_t>(blockIdx.x) * kSampWarps + warp;
    if (ray >= n) return;
    float* s_bins = smem + warp * (3 * nb);
    float* s_cdf = s_bins + nb;
    float* s_w = s_cdf + nb;
    for (int j = lane; j < nb; j += 32) s_bins[j] = bins[ray * nb + j];
    if (cdf_given) {
        for (int j = lane; j < nb; j += 32) s_cdf[j] = weights_or_cdf[ray * nb + j];
        __syncwarp();
    } else {
        for (int j = lane; j < nb - 1; j += 32) s_w[j] = __fadd_rn(weights_or_cdf[ray * (nb - 1) + j], 1e-5f);
        __syncwarp();
        build_cdf(s_cdf, s_w, nb, lane);
    }
    for (int i = lane; i < ns; i += 32) {
        const float uu = u ? u[ray * u_stride + i] : linspace01(i, ns);
        int inds;
        const float zz = invert_cdf(s_cdf, s_bins, nb, uu, inds);
        z_out[ray * ns + i] = zz;
        if (inds_out) inds_out[ray * ns + i] = inds;
    }
}

// Coarse tail: mids -> pdf over weights[1:-1] -> n_imp samples -> sort(cat(z_coarse, samples)).
// The result equals torch.sort(cat) as a sorted multiset.  Both inputs are normally sorted already (z_coarse by
// construction; the samples because the inverse CDF is monotone and, in det mode, u is ascending), so the merge is
// done by rank -- out position of z_i = i + #{samples < z_i}, of sample k = k + #{z <= s_k}, one binary search per
// element -- instead of a 256-wide bitonic network.  Random u: the 128 samples go through a 128-wide bitonic sort
// first.  Sortedness is CHECKED per ray (an ulp-level inversion at a bin edge is possible in principle); a ray
// that fails takes the full bitonic sort.
__device__ __forceinline__ bool warp_sorted(const float* v, int n, int lane) {
    bool ok = true;
    for (int i = lane + 1; i < n; i += 32) ok = ok && (v[i - 1] <= v[i]);
    return __all_sync(0xffffffffu, ok);
}

// kFromSigma: `weights` holds the coarse DENSITIES [n,sc] instead; the compositing weights (render.py:9-14) are first
// rebuilt from them with the compositing kernels' own chunked scan (CH samples per step -> the same bits) and never
// leave the SM unless weights_out asks for them.  z_stride: row stride of z_coarse (0 = one row for every ray).
template <bool kFromSigma>
__global__ void __launch_bounds__(kSampWarps * 32)
fine_z_kernel(const float* __restrict__ z_coarse, int64_t z_stride, const float* __restrict__ weights,
              const float* __restrict__ rays_d, int CH, const float* __restrict__ u, int64_t u_stride, int64_t n, int sc,
              int n_imp, int P, int P2, float* __restrict__ z_fine, float* __restrict__ z_samples,
              float* __restrict__ weights_out) {
    extern __shared__ float smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t ray = static_cast<int64_t>(blockIdx.x) * kSampWarps + warp;
    if (ray >= n) return;
    const int nb = sc - 1;
    float* s_sort = smem + warp * (P + 3 * sc + P2);
    float* s_bins = s_sort + P;
    float* s_cdf = s_bins + sc;
    float* s_w = s_cdf + sc;
    float* s_samp = s_w + sc;
    float* s_zc = s_sort + (P - sc);                       // coarse depths parked at the tail of the output buffer
    for (int j = lane; j < sc; j += 32) s_zc[j] = z_coarse[ray * z_stride + j];
    __syncwarp();
    for (int j = lane; j < nb; j += 32) s_bins[j] = __fmul_rn(0.5f, __fadd_rn(s_zc[j + 1], s_zc[j]));
    if constexpr (kFromSigma) {
        const float dx = __ldg(rays_d + ray * 3 + 0), dy = __ldg(rays_d + ray * 3 + 1), dz = __ldg(rays_d + ray * 3 + 2);
        const float norm = sqrtf(dx * dx + dy * dy + dz * dz);
        float carry = 1.0f;
        for (int c0 = 0; c0 < sc; c0 += CH) {
            const int cnt = min(CH, sc - c0);
            const int s = c0 + lane;
            const bool valid = lane < cnt;
            float dist = 0.f, sg = 0.f;
            if (valid) {
                const float zc = s_zc[s];
                dist = ((s + 1 < sc) ? (s_zc[s + 1] - zc) : 1e10f) * norm;
                sg = __ldg(weights + ray * sc + s);
            }
            float alpha, trans, w;
            scan_chunk(sg, dist, valid, lane, carry, alpha, trans, w);
            if (valid) {
                if (s >= 1 && s <= nb - 1) s_w[s - 1] = __fadd_rn(w, 1e-5f);
                if (weights_out) weights_out[ray * sc + s] = w;
            }
        }
    } else {
        for (int j = lane; j < nb - 1; j += 32) s_w[j] = __fadd_rn(weights[ray * sc + 1 + j], 1e-5f);
    }
    __syncwarp();
    build_cdf(s_cdf, s_w, nb, lane);
    for (int i = lane; i < n_imp; i += 32) {
        const float uu = u ? u[ray * u_stride + i] : linspace01(i, n_imp);
        int inds;
        const float zz = invert_cdf(s_cdf, s_bins, nb, uu, inds);
        s_samp[i] = zz;
        if (z_samples) z_samples[ray * n_imp + i] = zz;
    }
    for (int i = n_imp + lane; i < P2; i += 32) s_samp[i] = CUDART_INF_F;
    __syncwarp();
    const int tot = sc + n_imp;
    // the cdf is dead from here on: the coarse depths move into its place, freeing the whole output buffer
    for (int j = lane; j < sc; j += 32) s_cdf[j] = s_zc[j];
    __syncwarp();
    const bool coarse_ok = warp_sorted(s_cdf, sc, lane);
    if (coarse_ok) {
        if (!warp_sorted(s_samp, n_imp, lane)) bitonic_sort(s_samp, P2, lane);
        for (int i = lane; i < sc; i += 32) {              // rank of a coarse depth: samples strictly below it
            const float v = s_cdf[i];
            int lo = 0, hi = n_imp;
            while (lo < hi) { const int mid = (lo + hi) >> 1; if (s_samp[mid] < v) lo = mid + 1; else hi = mid; }
            s_sort[i + lo] = v;
        }
        for (int k = lane; k < n_imp; k += 32) {           // rank of a sample: coarse depths at or below it
            const float v = s_samp[k];
            int lo = 0, hi = sc;
            while (lo < hi) { const int mid = (lo + hi) >> 1; if (s_cdf[mid] <= v) lo = mid + 1; else hi = mid; }
            s_sort[k + lo] = v;
        }
        __syncwarp();
    } else {
        for (int j = lane; j < sc; j += 32) s_sort[j] = s_cdf[j];
        for (int i = lane; i < n_imp; i += 32) s_sort[sc + i] = s_samp[i];
        for (int i = tot + lane; i < P; i += 32) s_sort[i] = CUDART_INF_F;
        __syncwarp();
        bitonic_sort(s_sort, P, lane);
    }
    for (int i = lane; i < tot; i += 32) z_fine[ray * tot + i] = s_sort[i];
}

struct CatParts {
    const float* src[8];
    int len[8];
    int parts;
};

__global__ void __launch_bounds__(kSampWarps * 32)
sort_cat_kernel(CatParts cp, int64_t n, int tot, int P, float* __restrict__ out) {
    extern __shared__ float smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t ray = static_cast<int64_t>(blockIdx.x) * kSampWarps + warp;
    if (ray >= n) return;
    float* s_sort = smem + warp * P;
    int off = 0;
    for (int p = 0; p < cp.parts; ++p) {
        const float* src = cp.src[p] + ray * cp.len[p];
        for (int i = lane; i < cp.len[p]; i += 32) s_sort[off + i] = src[i];
        off += cp.len[p];
    }
    for (int i = tot + lane; i < P; i += 32) s_sort[i] = CUDART_INF_F;
    __syncwarp();
    bitonic_sort(s_sort, P, lane);
    for (int i = lane; i < tot; i += 32) out[ray * tot + i] = s_sort[i];
}

// z' = lower + (upper-lower)*t with mids between neighbours (render.py:42-47), un-fused like ATen.
__global__ void jitter_z_kernel(const float* __restrict__ z, const float* __restrict__ t_rand, int64_t total, int S,
                                float* __restrict__ out) {
    const int64_t idx = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    const int s = static_cast<int>(idx % S);
    const float zc = z[idx];
    const float upper = (s + 1 < S) ? __fmul_rn(0.5f, __fadd_rn(z[idx + 1], zc)) : zc;
    const float lower = (s > 0) ? __fmul_rn(0.5f, __fadd_rn(zc, z[idx - 1])) : zc;
    out[idx] = __fadd_rn(lower, __fmul_rn(__fsub_rn(upper, lower), t_rand[idx]));
}

}  // namespace dmnerf

using namespace dmnerf;

static int launch_sample(const float* bins, const float* w_or_cdf, int cdf_given, const float* u, int64_t u_stride,
                         int64_t n, int nb, int ns, float* z_out, int64_t* inds_out, void* stream) {
    if (n <= 0) return DMNERF_OK;
    DM_REQUIRE(bins && w_or_cdf && z_out, DMNERF_EINVAL, "sample_pdf: null pointer");
    DM_REQUIRE(nb >= 2 && nb <= 1024 && ns >= 2 && ns <= 4096, DMNERF_EUNSUPPORTED,
               "sample_pdf: n_bins %d / n_samples %d unsupported", nb, ns);
    if (n <= 0) return DMNERF_OK;
    const size_t smem = static_cast<size_t>(kSampWarps) * 3 * nb * sizeof(float);
    const unsigned grid = static_cast<unsigned>((n + kSampWarps - 1) / kSampWarps);
    sample_pdf_kernel<<<grid, kSampWarps * 32, smem, as_stream(stream)>>>(bins, w_or_cdf, cdf_given, u, u_stride, n,
                                                                         nb, ns, z_out, inds_out);
    DM_LAUNCH_CHECK("sample_pdf_kernel");
    return DMNERF_OK;
}

extern "C" int dmnerf_sample_pdf(const float* bins, const float* weights, const float* u, int64_t u_stride, int64_t n,
                                 int n_bins, int n_samples, float* z_out, int64_t* inds_out, void* stream) {
    return launch_sample(bins, weights, 0, u, u_stride, n, n_bins, n_samples, z_out, inds_out, stream);
}

extern "C" int dmnerf_sample_cdf(const float* bins, const float* cdf, const float* u, int64_t u_stride, int64_t n,
                                 int n_bins, int n_samples, float* z_out, int64_t* inds_out, void* stream) {
    return launch_sample(bins, cdf, 1, u, u_stride, n, n_bins, n_samples, z_out, inds_out, stream);
}

static int launch_fine_z(bool from_sigma, const float* z_coarse, int64_t z_stride, const float* w_or_sigma,
                         const float* rays_d, int ch, const float* u, int64_t u_stride, int64_t n, int sc, int n_imp,
                         float* z_fine, float* z_samples, float* weights_out, void* stream, const char* who) {
    DM_REQUIRE(sc >= 4 && sc <= 512 && n_imp >= 2 && sc + n_imp <= 1024, DMNERF_EUNSUPPORTED,
               "%s: sc %d / n_importance %d unsupported", who, sc, n_imp);
    int P = 1, P2 = 1;
    while (P < sc + n_imp) P <<= 1;
    while (P2 < n_imp) P2 <<= 1;
    const size_t smem = static_cast<size_t>(kSampWarps) * (P + 3 * sc + P2) * sizeof(float);
    const unsigned grid = static_cast<unsigned>((n + kSampWarps - 1) / kSampWarps);
    if (from_sigma)
        fine_z_kernel<true><<<grid, kSampWarps * 32, smem, as_stream(stream)>>>(
            z_coarse, z_stride, w_or_sigma, rays_d, ch, u, u_stride, n, sc, n_imp, P, P2, z_fine, z_samples, weights_out);
    else
        fine_z_kernel<false><<<grid, kSampWarps * 32, smem, as_stream(stream)>>>(
            z_coarse, z_stride, w_or_sigma, nullptr, 32, u, u_stride, n, sc, n_imp, P, P2, z_fine, z_samples, nullptr);
    DM_LAUNCH_CHECK("fine_z_kernel");
    return DMNERF_OK;
}

extern "C" int dmnerf_fine_z(const float* z_coarse, const float* weights, const float* u, int64_t u_stride, int64_t n,
                             int sc, int n_imp, float* z_fine, float* z_samples, void* stream) {
    if (n <= 0) return DMNERF_OK;
    DM_REQUIRE(z_coarse && weights && z_fine, DMNERF_EINVAL, "fine_z: null pointer");
    return launch_fine_z(false, z_coarse, sc, weights, nullptr, 32, u, u_stride, n, sc, n_imp, z_fine, z_samples, nullptr,
                         stream, "fine_z");
}

extern "C" int dmnerf_coarse_tail(const float* sigma, const float* z_coarse, int64_t z_stride, const float* rays_d,
                                  const float* u, int64_t u_stride, int64_t n, int sc, int c, int n_imp, float* z_fine,
                                  float* weights_out, void* stream) {
    if (n <= 0) return DMNERF_OK;
    DM_REQUIRE(sigma && z_coarse && rays_d && z_fine, DMNERF_EINVAL, "coarse_tail: null pointer");
    DM_REQUIRE(z_stride == 0 || z_stride >= sc, DMNERF_EINVAL, "coarse_tail: z_stride %lld (0 or >= sc)",
               static_cast<long long>(z_stride));
    DM_REQUIRE(c >= 6 && c <= 128, DMNERF_EUNSUPPORTED, "coarse_tail: channels %d unsupported", c);
    // the scan step dmnerf_composite_fwd would use for raw [n,sc,c] (16 samples for wide rows on its bulk path)
    const int ch = (c > 64 && (static_cast<int64_t>(sc) * c) % 4 == 0) ? 16 : 32;
    return launch_fine_z(true, z_coarse, z_stride, sigma, rays_d, ch, u, u_stride, n, sc, n_imp, z_fine, nullptr,
                         weights_out, stream, "coarse_tail");
}

extern "C" int dmnerf_sort_cat(const float* const* srcs, const int* lens, int parts, int64_t n, float* out,
                               void* stream) {
    if (n <= 0) return DMNERF_OK;
    DM_REQUIRE(srcs && lens && out, DMNERF_EINVAL, "sort_cat: null pointer");
    DM_REQUIRE(parts >= 1 && parts <= 8, DMNERF_EUNSUPPORTED, "sort_cat: %d parts (max 8)", parts);
    CatParts cp;
    int tot = 0;
    for (int p = 0; p < parts; ++p) {
        DM_REQUIRE(srcs[p] && lens[p] > 0, DMNERF_EINVAL, "sort_cat: part %d empty", p);
        cp.src[p] = srcs[p]; cp.len[p] = lens[p]; tot += lens[p];
    }
    cp.parts = parts;
    DM_REQUIRE(tot <= 2048, DMNERF_EUNSUPPORTED, "sort_cat: %d samples per ray (max 2048)", tot);
    if (n <= 0) return DMNERF_OK;
    int P = 1;
    while (P < tot) P <<= 1;
    const size_t smem = static_cast<size_t>(kSampWarps) * P * sizeof(float);
    const unsigned grid = static_cast<unsigned>((n + kSampWarps - 1) / kSampWarps);
    sort_cat_kernel<<<grid, kSampWarps * 32, smem, as_stream(stream)>>>(cp, n, tot, P, out);
    DM_LAUNCH_CHECK("sort_cat_kernel");
    return DMNERF_OK;
}

extern "C" int dmnerf_jitter_z(const float* z, const float* t_rand, int64_t n, int s, float* z_out, void* stream) {
    if (n <= 0) return DMNERF_OK;
    DM_REQUIRE(z && t_rand && z_out, DMNERF_EINVAL, "jitter_z: null pointer");
    DM_REQUIRE(s >= 1, DMNERF_EINVAL, "jitter_z: bad sample count");
    const int64_t total = n * s;
    if (total <= 0) return DMNERF_OK;
    const unsigned grid = static_cast<unsigned>((total + 255) / 256);
    jitter_z_kernel<<<grid, 256, 0, as_stream(stream)>>>(z, t_rand, total, s, z_out);
    DM_LAUNCH_CHECK("jitter_z_kernel");
    return DMNERF_OK;
}
