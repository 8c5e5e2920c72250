// exchanger.cu -- a10 per-sample instance exchange (networks/manipulator.py:18-83), one move label
// per launch; plus pinhole ray generation (networks/helpers.py:50-61) and the per-ray label argmax.
//
// One warp per (ray, 32-sample chunk).  Both raw chunks are staged in shared memory; lanes act as
// channels for the per-ray ("pixel") label and as samples for the per-sample labels.  Labels are
// argmax(sigmoid(logits)) with the FIRST maximum winning, evaluated with the same fp32 sigmoid the
// reference's CUDA path uses, so saturation ties (logit >~ 17.33 -> 1.0f) resolve identically.
// Integer/boolean logic afterwards is exact:
//   occluded:  pix != m && lab == m  -> lab = pix                      (manipulator.py:33-36, 57-60)
//   fill    :  pix == m && lab != m                                    (:40-42)
//   code    :  [tar_lab == m] + 2*[lab == m]; 1,3 -> take target row, 2 -> zero row   (:64-81)
#include "common.cuh"
#include "tc_ptx.cuh"

namespace dmnerf {

int tc_debug_flags();
constexpr int kExWarps = 4;

// argmax over sigmoid(v[0..cnt)) spread across the warp (lane j holds channels j, j+32, ...).
__device__ __forceinline__ int warp_argmax_sigmoid(const float* v, int cnt, int lane) {
    float best = -1.0f;
    int idx = 0x7fffffff;
    for (int j = lane; j < cnt; j += 32) {
        const float s = sigmoidf_ref(v[j]);
        if (s > best) { best = s; idx = j; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ob = __shfl_xor_sync(0xffffffffu, best, o);
        const int oi = __shfl_xor_sync(0xffffffffu, idx, o);
        if (ob > best || (ob == best && oi < idx)) { best = ob; idx = oi; }
    }
    return idx;
}

// argmax_j sigmoid(row[j]), first maximum winning -- the same result as evaluating the fp32 sigmoid of every channel,
// but the sigmoid is only evaluated where it can matter.  sigmoid is monotone up to its last-ulp rounding, so only
// channels close to the largest logit can reach the largest sigmoid VALUE: with s = sigmoid(x_max) < 1 a logit more than
// 32 * 2^-24 / (1 - s) below x_max has a true sigmoid at least 16 ulp below s (slope s (1 - s) >= (1 - s) / 2) and cannot
// tie or win; with s == 1.0f (saturation, x >~ 17.33: the reference's tie case) every logit >= 16 is a candidate, no other
// can round to 1.  Typically one candidate per row instead of K+1 expf + IEEE divisions.
__device__ __forceinline__ int row_argmax_sigmoid(const float* row, int cnt) {
    // one pass for the largest logit (first index) and the runner-up; the candidate pass below only runs when the
    // runner-up can reach the largest sigmoid VALUE (ties, saturation) -- otherwise the first maximum is the answer
    float xmax = row[0], x2 = -INFINITY;
    int i1 = 0;
#pragma unroll 4
    for (int j = 1; j < cnt; ++j) {
        const float x = row[j];
        if (x > xmax) { x2 = xmax; xmax = x; i1 = j; }
        else x2 = fmaxf(x2, x);
    }
    const float smax = sigmoidf_ref(xmax);
    // xmax < -80: sigmoid(xmax) is denormal or 0 and the ulp argument above does not hold -> every channel is a
    // candidate (all-zero rows then give index 0, like the reference's argmax over zeros)
    const float thr = xmax < -80.0f ? -INFINITY
                      : (smax >= 1.0f ? fminf(16.0f, xmax) : xmax - 1.9073486e-6f / (1.0f - smax));
    if (x2 < thr) return i1;
    float best = -1.0f;
    int idx = 0;
    for (int j = 0; j < cnt; ++j) {
        const float x = row[j];
        if (x >= thr) {
            const float s = sigmoidf_ref(x);
            if (s > best) { best = s; idx = j; }
        }
    }
    return idx;
}

// kAsync: both chunks go global -> shared as ONE bulk asynchronous copy each (cp.async.bulk, the TMA engine; lane 0
// issues, the warp waits on its own mbarrier) into rows of C floats: one DRAM round trip per chunk instead of five,
// two instructions instead of ~80 per lane, and no L1 miss-queue limit on what an SM has in flight (the 16-byte LDGSTS
// variant of this path stopped at 0.47 of the HBM peak on exactly that).  Needs C odd (a lane scanning row `lane` then
// hits bank (C*lane + j) mod 32, conflict-free without the padding column that would break the 16-byte alignment)
// and 16-byte aligned ray blocks (S*C a multiple of 4).  Other shapes take the register path.
template <bool kAsync>
__global__ void __launch_bounds__(kExWarps * 32)
exchange_kernel(float* __restrict__ ori_raw, const float* __restrict__ tar_raw, const float* __restrict__ ori_acc,
                const float* __restrict__ tar_acc, int m, int first_label, int64_t n, int S, int C,
                int64_t* __restrict__ ori_label, int64_t* __restrict__ tar_label) {
    extern __shared__ float smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int chunks = (S + 31) / 32;
    // a warp walks (ray, chunk) items grid-stride: the grid is sized to what is resident at once, so a warp's shared
    // memory slot is reused instead of a new 128-thread block being scheduled for every four chunks
    const int64_t n_items = n * chunks, stride = static_cast<int64_t>(gridDim.x) * kExWarps;
    uint32_t bar_phase = 0;
    if constexpr (kAsync) {
        if (lane == 0) {
            const int cp0 = C;
            ptx::mbar_init(static_cast<uint32_t>(__cvta_generic_to_shared(smem + warp * (2 * 32 * cp0 + 36) + 2 * 32 * cp0 + 32)), 1);
            ptx::fence_barrier_init();
        }
        __syncwarp();
    }
  for (int64_t item = static_cast<int64_t>(blockIdx.x) * kExWarps + warp; item < n_items; item += stride) {
    __syncwarp();                                             // the previous item's readers are done with the slot
    const int64_t ray = item / chunks;
    const int c0 = static_cast<int>(item % chunks) * 32;
    const int cnt = min(32, S - c0);
    const int CP = kAsync ? C : (C | 1);
    const int L = C - 4;
    float* t_ori = smem + warp * (2 * 32 * CP + 36);
    float* t_tar = t_ori + 32 * CP;
    int* act = reinterpret_cast<int*>(t_tar + 32 * CP);
    const uint32_t bar = static_cast<uint32_t>(__cvta_generic_to_shared(act + 32));     // this warp's mbarrier (kAsync)
    float* g_ori = ori_raw + (ray * S + c0) * static_cast<int64_t>(C);
    const float* g_tar = tar_raw + (ray * S + c0) * static_cast<int64_t>(C);
    const int total = cnt * C;
    const unsigned magic = 0xFFFFFFFFu / C + 1u;             // i / C by one multiply-high (i < 32 * 128)
    // staging: 8 elements of each field per lane in flight (a plain one-load-per-trip loop left every warp waiting on
    // ~37 dependent DRAM round trips per chunk: the kernel ran at 0.3 of the HBM roofline)
    constexpr int U = 8;
    if constexpr (kAsync) {
        const uint32_t so = static_cast<uint32_t>(__cvta_generic_to_shared(t_ori));
        const uint32_t st = static_cast<uint32_t>(__cvta_generic_to_shared(t_tar));
        const int whole = total & ~3;                         // bulk copies move multiples of 16 bytes
        if (lane == 0) {
            ptx::fence_proxy_async_smem();                    // the previous item's generic reads of the slot come first
            ptx::mbar_arrive_expect_tx(bar, 2u * whole * 4u);
            ptx::bulk_g2s(so, g_ori, whole * 4u, bar);
            ptx::bulk_g2s(st, g_tar, whole * 4u, bar);
        }
        // ragged last chunk (S not a multiple of 32): the odd floats by hand, never reading past the field
        for (int j = whole + lane; j < total; j += 32) { t_ori[j] = g_ori[j]; t_tar[j] = __ldg(g_tar + j); }
    } else
    for (int base = 0; base < total; base += 32 * U) {
        float va[U], vb[U];
#pragma unroll
        for (int q = 0; q < U; ++q) {
            const int i = base + q * 32 + lane;
            va[q] = i < total ? g_ori[i] : 0.f;
            vb[q] = i < total ? __ldg(g_tar + i) : 0.f;
        }
#pragma unroll
        for (int q = 0; q < U; ++q) {
            const int i = base + q * 32 + lane;
            if (i < total) {
                const int r = static_cast<int>(__umulhi(static_cast<unsigned>(i), magic));
                const int col = i - r * C;
                t_ori[r * CP + col] = va[q];
                t_tar[r * CP + col] = vb[q];
            }
        }
    }
    // per-ray labels from the accumulated maps without their last ("empty") channel
    const int pix = warp_argmax_sigmoid(ori_acc + ray * L, L - 1, lane);
    const int t_pix = warp_argmax_sigmoid(tar_acc + ray * L, L - 1, lane);
    if constexpr (kAsync) { ptx::mbar_wait(bar, bar_phase); bar_phase ^= 1u; }
    __syncwarp();
    int a = 0;
    if (lane < cnt) {
        const int64_t at = ray * S + c0 + lane;
        int lab = first_label ? row_argmax_sigmoid(t_ori + lane * CP + 4, L) : static_cast<int>(ori_label[at]);
        if (pix != m && lab == m) lab = pix;
        const bool fill = (pix == m) && (lab != m);
        int t_lab = row_argmax_sigmoid(t_tar + lane * CP + 4, L);
        if (t_pix != m && t_lab == m) t_lab = t_pix;
        const int code = (t_lab == m ? 1 : 0) + (lab == m ? 2 : 0);
        if (fill || code == 1 || code == 3) a = 1;
        else if (code == 2) a = 2;
        if (ori_label) ori_label[at] = lab;
        if (tar_label) tar_label[at] = t_lab;
    }
    act[lane] = a;
    __syncwarp();
    // rows that keep their values are not touched; most chunks have none to rewrite
    int any = 0;
    if (lane < cnt) any = act[lane];
    if (__ballot_sync(0xffffffffu, any != 0) == 0u) continue;
    for (int i = lane; i < total; i += 32) {
        const int r = static_cast<int>(__umulhi(static_cast<unsigned>(i), magic));
        const int col = i - r * C;
        const int ar = act[r];
        if (ar == 1) g_ori[i] = t_tar[r * CP + col];
        else if (ar == 2) g_ori[i] = t_ori[r * CP + col] * 0.0f;
    }
  }
}

__global__ void label_argmax_kernel(const float* __restrict__ ins, int64_t n, int K, uint8_t* __restrict__ labels) {
    const int64_t ray = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (ray >= n) return;
    // the composited maps are probabilities already (render.py:25, manipulator.py:102); the reference takes a plain
    // torch.argmax of them (evaluator.py:132, manipulator.py:321,478): first maximum wins, no second sigmoid
    const float* row = ins + ray * K;
    float best = row[0];
    int idx = 0;
    for (int j = 1; j < K; ++j) {
        const float v = row[j];
        if (v > best) { best = v; idx = j; }
    }
    labels[ray] = static_cast<uint8_t>(idx);
}

struct Camera {
    float k00, k02, k11, k12, k22;
    float r[9];
    float t[3];
};

// dirs = [(i-cx)/fx, (j-cy)/fy, K22]; rays_d = sum_k dirs[k]*R[c][k] (left-to-right, un-fused); rays_o = t.
__global__ void rays_k_kernel(Camera cam, int W, int64_t pix0, const int64_t* __restrict__ pix_idx, int64_t n,
                              float* __restrict__ rays_o, float* __restrict__ rays_d) {
    const int64_t q = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (q >= n) return;
    const int64_t pix = pix_idx ? pix_idx[q] : pix0 + q;
    const float i = static_cast<float>(pix % W), j = static_cast<float>(pix / W);
    const float d0 = __fdiv_rn(__fsub_rn(i, cam.k02), cam.k00);
    const float d1 = __fdiv_rn(__fsub_rn(j, cam.k12), cam.k11);
    const float d2 = cam.k22;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        float v = __fmul_rn(d0, cam.r[c * 3 + 0]);
        v = __fadd_rn(v, __fmul_rn(d1, cam.r[c * 3 + 1]));
        v = __fadd_rn(v, __fmul_rn(d2, cam.r[c * 3 + 2]));
        rays_d[q * 3 + c] = v;
        rays_o[q * 3 + c] = cam.t[c];
    }
}

// Same rays with the camera read from DEVICE memory (K [3,3] row-major, c2w [>=3, 4] row-major): nothing about the
// view is baked into the launch, so a captured training step can be replayed for any image / pose.
__global__ void rays_k_dev_kernel(const float* __restrict__ K, const float* __restrict__ c2w, int W, int64_t pix0,
                                  const int64_t* __restrict__ pix_idx, int64_t n, float* __restrict__ rays_o,
                                  float* __restrict__ rays_d) {
    const int64_t q = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (q >= n) return;
    const int64_t pix = pix_idx ? pix_idx[q] : pix0 + q;
    const float i = static_cast<float>(pix % W), j = static_cast<float>(pix / W);
    const float d0 = __fdiv_rn(__fsub_rn(i, __ldg(K + 2)), __ldg(K + 0));
    const float d1 = __fdiv_rn(__fsub_rn(j, __ldg(K + 5)), __ldg(K + 4));
    const float d2 = __ldg(K + 8);
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        float v = __fmul_rn(d0, __ldg(c2w + c * 4 + 0));
        v = __fadd_rn(v, __fmul_rn(d1, __ldg(c2w + c * 4 + 1)));
        v = __fadd_rn(v, __fmul_rn(d2, __ldg(c2w + c * 4 + 2)));
        rays_d[q * 3 + c] = v;
        rays_o[q * 3 + c] = __ldg(c2w + c * 4 + 3);
    }
}

// Standalone positional embedding (Embedder.embed, networks/dm_nerf.py:37-38): out[n, 3+6F].
__global__ void embed_kernel(const float* __restrict__ x, int64_t n, int F, float* __restrict__ out) {
    const int64_t idx = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (idx >= n * F) return;
    const int64_t row = idx / F;
    const int f = static_cast<int>(idx - row * F);
    const int D = 3 + 6 * F;
    const float scale = static_cast<float>(1 << f);
    float* o = out + row * D;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const float v = x[row * 3 + c];
        if (f == 0) o[c] = v;
        float sn, cs;
        sincosf(v * scale, &sn, &cs);
        o[3 + 6 * f + c] = sn;
        o[3 + 6 * f + 3 + c] = cs;
    }
}

}  // namespace dmnerf

using namespace dmnerf;

extern "C" int dmnerf_embed(const float* x, int64_t n, int n_freqs, float* out, void* stream) {
    DM_REQUIRE(x && out, DMNERF_EINVAL, "embed: null pointer");
    DM_REQUIRE(n_freqs >= 1 && n_freqs <= 16, DMNERF_EUNSUPPORTED, "embed: %d frequency bands", n_freqs);
    if (n <= 0) return DMNERF_OK;
    const int64_t total = n * n_freqs;
    embed_kernel<<<static_cast<unsigned>((total + 255) / 256), 256, 0, as_stream(stream)>>>(x, n, n_freqs, out);
    DM_LAUNCH_CHECK("embed_kernel");
    return DMNERF_OK;
}

extern "C" int dmnerf_exchange(float* ori_raw, const float* tar_raw, const float* ori_acc, const float* tar_acc,
                               int move_label, int first_label, int64_t n, int s, int c, int64_t* ori_label,
                               int64_t* tar_label, void* stream) {
    if (n <= 0) return DMNERF_OK;
    DM_REQUIRE(ori_raw && tar_raw && ori_acc && tar_acc, DMNERF_EINVAL, "exchange: null pointer");
    DM_REQUIRE(ori_label || first_label, DMNERF_EINVAL, "exchange: ori_label may only be NULL when first_label != 0");
    DM_REQUIRE(s >= 1 && c >= 7 && c <= 128, DMNERF_EUNSUPPORTED, "exchange: n_samples %d / channels %d unsupported",
               s, c);
    if (n <= 0) return DMNERF_OK;
    const bool fast = (c & 1) && (static_cast<int64_t>(s) * c) % 4 == 0 &&
                      ((reinterpret_cast<uintptr_t>(ori_raw) | reinterpret_cast<uintptr_t>(tar_raw)) & 15) == 0;
    const int CP = fast ? c : (c | 1);
    const size_t smem = static_cast<size_t>(kExWarps) * (2 * 32 * CP + 36) * sizeof(float);
    const int64_t items = n * ((s + 31) / 32);
    static int sm_count = 0;
    if (sm_count == 0) {
        int dev = 0;
        DM_CUDA(cudaGetDevice(&dev));
        DM_CUDA(cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev));
    }
    const int64_t resident = static_cast<int64_t>(sm_count) * (smem ? (220 * 1024) / static_cast<int64_t>(smem) : 1);
    const int64_t blocks = (items + kExWarps - 1) / kExWarps;
    const unsigned grid = static_cast<unsigned>(blocks < resident ? blocks : resident);
    if (fast) {
        DM_CUDA(cudaFuncSetAttribute(exchange_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        exchange_kernel<true><<<grid, kExWarps * 32, smem, as_stream(stream)>>>(ori_raw, tar_raw, ori_acc, tar_acc, move_label,
                                                                               first_label, n, s, c, ori_label, tar_label);
    } else {
        DM_CUDA(cudaFuncSetAttribute(exchange_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        exchange_kernel<false><<<grid, kExWarps * 32, smem, as_stream(stream)>>>(ori_raw, tar_raw, ori_acc, tar_acc, move_label,
                                                                                first_label, n, s, c, ori_label, tar_label);
    }
    DM_LAUNCH_CHECK("exchange_kernel");
    return DMNERF_OK;
}

extern "C" int dmnerf_label_argmax(const float* ins, int64_t n, int k, uint8_t* labels, void* stream) {
    DM_REQUIRE(ins && labels, DMNERF_EINVAL, "label_argmax: null pointer");
    DM_REQUIRE(k >= 1 && k <= 255, DMNERF_EUNSUPPORTED, "label_argmax: %d classes", k);
    if (n <= 0) return DMNERF_OK;
    label_argmax_kernel<<<static_cast<unsigned>((n + 127) / 128), 128, 0, as_stream(stream)>>>(ins, n, k, labels);
    DM_LAUNCH_CHECK("label_argmax_kernel");
    return DMNERF_OK;
}

extern "C" int dmnerf_rays_k(int H, int W, const float* K, const float* c2w, int64_t pix0, const int64_t* pix_idx,
                             int64_t n, float* rays_o, float* rays_d, void* stream) {
    DM_REQUIRE(K && c2w && rays_o && rays_d, DMNERF_EINVAL, "rays_k: null pointer");
    DM_REQUIRE(H > 0 && W > 0 && (pix_idx || (pix0 >= 0 && pix0 + n <= static_cast<int64_t>(H) * W)), DMNERF_EINVAL,
               "rays_k: pixel range [%lld,+%lld) outside %dx%d", (long long)pix0, (long long)n, H, W);
    if (n <= 0) return DMNERF_OK;
    Camera cam;
    cam.k00 = K[0]; cam.k02 = K[2]; cam.k11 = K[4]; cam.k12 = K[5]; cam.k22 = K[8];
    for (int r = 0; r < 3; ++r) {
        for (int k = 0; k < 3; ++k) cam.r[r * 3 + k] = c2w[r * 4 + k];
        cam.t[r] = c2w[r * 4 + 3];
    }
    rays_k_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, as_stream(stream)>>>(cam, W, pix0, pix_idx, n,
                                                                                       rays_o, rays_d);
    DM_LAUNCH_CHECK("rays_k_kernel");
    return DMNERF_OK;
}

extern "C" int dmnerf_rays_k_dev(int H, int W, const float* K_dev, const float* c2w_dev, int64_t pix0, const int64_t* pix_idx,
                                 int64_t n, float* rays_o, float* rays_d, void* stream) {
    DM_REQUIRE(K_dev && c2w_dev && rays_o && rays_d, DMNERF_EINVAL, "rays_k_dev: null pointer");
    DM_REQUIRE(H > 0 && W > 0 && (pix_idx || (pix0 >= 0 && pix0 + n <= static_cast<int64_t>(H) * W)), DMNERF_EINVAL,
               "rays_k_dev: pixel range [%lld,+%lld) outside %dx%d", (long long)pix0, (long long)n, H, W);
    if (n <= 0) return DMNERF_OK;
    rays_k_dev_kernel<<<static_cast<unsigned>((n + 255) / 256), 256, 0, as_stream(stream)>>>(K_dev, c2w_dev, W, pix0, pix_idx, n,
                                                                                           rays_o, rays_d);
    DM_LAUNCH_CHECK("rays_k_dev_kernel");
    return DMNERF_OK;
}
