// mlp_tc_bwd.cu -- a12 in DMNERF_MODE_BF16: parameter gradients of the DM_NeRF MLP (autograd of
// networks/dm_nerf.py:80-106, reached from total_loss.backward(), train_dmsr.py:63) on the tcgen05 tensor
// cores.  Three launches per backward, all fed by the bf16 stash the training forward wrote
// (mlp_layout.cuh: 64-point blocks whose regions are ready-made UMMA operand images):
//
//   1. mlp_dgrad_bf16_kernel  the gradient chain d(raw) -> G8 -> G7 ... -> G0 (G_l = dL/d pre-activation of
//      forward layer T_l).  Same persistent structure as the forward: tile pairs in lockstep, transposed
//      weights streamed through a bulk-copy ring, fp32 accumulators in TMEM, the epilogue applies the ReLU
//      mask read from the forward's 1-bit sign-mask stash and writes the next operand to smem AND to the gradient stash.
//      Detach points: the instance rows of T8 are left out of the chain (h.detach(), dm_nerf.py:95); nothing
//      flows into the embeddings (rays / z are data).  The density head is a rank-1 term in T8's epilogue.
//   2. mlp_wgrad_bf16_kernel  dW_l = G_l^T A_l as split-K GEMMs over points: each CTA owns a contiguous range of
//      64-point blocks, walks 11 "jobs" (one or two products that share loads, <= 512 TMEM columns), bulk-copies
//      the G and A regions of a block (both operands MN-major: the 64 points are the reduction dim),
//      accumulates in TMEM across ALL its blocks and flushes once per job with vector fp32 reductions into
//      the packed gradient image.  Idle warps sum the bias gradients from the staged G operands.
//   3. unpack_grads_kernel    packed gradient image -> the 30 parameter gradients, including the chain rule
//      through the two linear->linear pairs the forward composed at pack time.
#include "mlp_tc_common.cuh"

namespace dmnerf {
namespace tcb {

using namespace tc;

// ============================================================================================================
// 1. dgrad chain
// ============================================================================================================
constexpr int kDgStages = 4;
constexpr int kDgABytes = 32 * kKgBytes;                           // one tile's operand: 256 features x 128 rows
constexpr int kDgSmRing = 2 * kDgABytes;                           // 131,072
constexpr int kDgSmWd = kDgSmRing + kDgStages * kStageBytes;       // 196,608
constexpr int kDgSmBar = kDgSmWd + 256 * 4;
constexpr int kDgSmTotal = kDgSmBar + 128;

struct DBars {
    uint32_t base;
    __device__ __forceinline__ uint32_t w_full(int s) const { return base + 8 * s; }
    __device__ __forceinline__ uint32_t w_empty(int s) const { return base + 8 * (kDgStages + s); }
    __device__ __forceinline__ uint32_t a_ready(int t) const { return base + 8 * (2 * kDgStages + t); }
    __device__ __forceinline__ uint32_t acc_ready(int t) const { return base + 8 * (2 * kDgStages + 2 + t); }
};

__global__ void __launch_bounds__(kFwThreads, 1)
mlp_dgrad_bf16_kernel(const float* __restrict__ d_raw, int64_t n_points, const uint8_t* __restrict__ image, Image im,
                      const uint8_t* __restrict__ acts, uint8_t* __restrict__ gstash) {
    extern __shared__ __align__(1024) uint8_t smem[];
    const uint32_t sm0 = ptx::smem_u32(smem);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const DBars bars{sm0 + kDgSmBar};
    float* s_wd = reinterpret_cast<float*>(smem + kDgSmWd);
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(smem + kDgSmBar + 8 * (2 * kDgStages + 4));
    const int C = im.ins_num + 5;

    if (threadIdx.x == 0) {
        for (int s = 0; s < kDgStages; ++s) { ptx::mbar_init(bars.w_full(s), 1); ptx::mbar_init(bars.w_empty(s), 2); }
        for (int t = 0; t < 2; ++t) { ptx::mbar_init(bars.a_ready(t), 2 * kTile); ptx::mbar_init(bars.acc_ready(t), 1); }
        ptx::fence_barrier_init();
    }
    if (warp == kMmaWarp) ptx::tmem_alloc(ptx::smem_u32(const_cast<uint32_t*>(tmem_slot)), 512);
    {
        const float* gb = reinterpret_cast<const float*>(image + im.bias_off) + kBiasWd;
        for (int i = threadIdx.x; i < 256; i += kFwThreads) s_wd[i] = gb[i];
    }
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem = *tmem_slot;
    const int64_t n_pairs = (n_points + kPair - 1) / kPair;

    if (warp == kProducerWarp) {
        if (lane == 0) {
            int stage = 0; uint32_t phase = 0;
            for (int64_t pair = blockIdx.x; pair < n_pairs; pair += gridDim.x) {
                for (int d = 0; d < kDgLayers; ++d) {
                    const int chunks = im.dg_k[d] / kChunkK;
                    const uint8_t* src = image + im.dg_off[d];
                    for (int c = 0; c < chunks; ++c) {
                        ptx::mbar_wait(bars.w_empty(stage), phase ^ 1);
                        ptx::mbar_arrive_expect_tx(bars.w_full(stage), kStageBytes);
                        ptx::bulk_g2s(sm0 + kDgSmRing + stage * kStageBytes, src + static_cast<int64_t>(c) * kStageBytes,
                                      kStageBytes, bars.w_full(stage));
                        if (++stage == kDgStages) { stage = 0; phase ^= 1; }
                    }
                }
            }
            for (int i = 0; i < kDgStages; ++i) {      // drain the last releases before the CTA may exit
                ptx::mbar_wait(bars.w_empty(stage), phase ^ 1);
                if (++stage == kDgStages) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp >= kMmaWarp) {
        // one issuer warp per tile (see mlp_tc.cu: a ring step costs the issuing thread ~300 clk of control code)
        if (lane == 0) {
            const int t = warp - kMmaWarp;
            int stage = 0; uint32_t phase = 0, a_par = 0;
            const uint32_t idesc = ptx::idesc_bf16(kTile, 256);
            constexpr uint32_t kKg16 = kKgBytes >> 4;
            const uint64_t desc_hi = static_cast<uint64_t>(0x4008u) << 32;       // SBO = 128 B, descriptor version 1
            const uint32_t a_lo0 = ((sm0 + t * kDgABytes) >> 4) | (kKg16 << 16);
            const uint32_t ring_lo = ((sm0 + kDgSmRing) >> 4) | (256u << 16);
            const uint32_t d_tmem = tmem + t * 256;
            for (int64_t pair = blockIdx.x; pair < n_pairs; pair += gridDim.x) {
                for (int d = 0; d < kDgLayers; ++d) {
                    const int chunks = im.dg_k[d] / kChunkK;
                    uint32_t a_lo = a_lo0;
#pragma unroll 1
                    for (int c = 0; c < chunks; ++c, a_lo += 4 * kKg16) {
                        ptx::mbar_wait(bars.w_full(stage), phase);
                        if (c == 0) { ptx::mbar_wait(bars.a_ready(t), a_par); a_par ^= 1; }
                        ptx::tc_fence_after();
                        const uint32_t b_lo = ring_lo + stage * (kStageBytes >> 4);
                        ptx::umma_bf16(d_tmem, desc_hi | a_lo, desc_hi | b_lo, idesc, c != 0);
                        ptx::umma_bf16(d_tmem, desc_hi | (a_lo + 2 * kKg16), desc_hi | (b_lo + 2 * 256), idesc, 1u);
                        if (c == chunks - 1) ptx::umma_commit(bars.acc_ready(t));
                        ptx::umma_commit(bars.w_empty(stage));
                        if (++stage == kDgStages) { stage = 0; phase ^= 1; }
                    }
                }
            }
        }
    } else {
        // ===== epilogue warps: thread == (point == TMEM lane, column half) =====
        const int t = warp >> 3, q = warp & 3, half = (warp >> 2) & 1, row = q * 32 + lane;
        const uint32_t a_base = sm0 + t * kDgABytes;
        const uint32_t t_base = tmem + t * 256 + (static_cast<uint32_t>(q * 32) << 16);
        const int nkg9 = im.kp9 / 8;
        uint32_t acc_par = 0;
        for (int64_t pair = blockIdx.x; pair < n_pairs; pair += gridDim.x) {
            const int64_t p = pair * kPair + t * kTile + row;
            const bool valid = p < n_points;
            const int64_t blk = (pair * 2 + t) * 2 + (row >> 6);
            const uint8_t* masks = acts + tc_mask_offset(n_points);
            uint8_t* g_row = gstash + blk * kStGradBlkBytes + (row & 63) * 16;
            const float* dr = d_raw + p * C;
            const float dsig = valid ? __ldg(dr + 3) : 0.f;
            // ---- G9 = d(raw) as a bf16 operand, zero padded to 128 channels in the stash ----
            for (int kg = half; kg < 16; kg += 2) {
                float v[8];
#pragma unroll
                for (int e = 0; e < 8; ++e) v[e] = (valid && kg * 8 + e < C) ? __ldg(dr + kg * 8 + e) : 0.f;
                const uint32_t w0 = ptx::pack_bf16(v[0], v[1], false), w1 = ptx::pack_bf16(v[2], v[3], false);
                const uint32_t w2 = ptx::pack_bf16(v[4], v[5], false), w3 = ptx::pack_bf16(v[6], v[7], false);
                if (kg < nkg9) ptx::st_shared_v4(a_base + kg * kKgBytes + row * 16, w0, w1, w2, w3);
                *reinterpret_cast<uint4*>(g_row + (kStG9 + kg) * kStKgBytes) = make_uint4(w0, w1, w2, w3);
            }
            ptx::fence_proxy_async_smem();
            ptx::tc_fence_before();
            ptx::mbar_arrive(bars.a_ready(t));

            for (int d = 0; d < kDgLayers; ++d) {
                const int ma = d == 0 ? 8 : 8 - d;                           // activation whose ReLU gates this gradient
                const int okg0 = d == 0 ? kStG8 : kStG0 + 32 * (8 - d);      // where the masked gradient is stashed
                const int cb0 = half * 4;
                // 128 sign bits (the forward's mask stash): four coalesced 4-byte loads, in flight during the wait
                uint32_t mk[4];
#pragma unroll
                for (int c = 0; c < 4; ++c) mk[c] = __ldg(reinterpret_cast<const uint32_t*>(masks + tc_mask_word(blk, ma, half, c, row & 63)));
                ptx::mbar_wait(bars.acc_ready(t), acc_par);
                acc_par ^= 1;
                ptx::tc_fence_after();
#pragma unroll 1
                for (int cb = cb0; cb < cb0 + 4; ++cb) {
                    uint32_t r[32];
                    ptx::tmem_ld32(t_base + cb * 32, r);
                    const int c = cb - cb0;
                    const uint32_t mword = c == 0 ? mk[0] : (c == 1 ? mk[1] : (c == 2 ? mk[2] : mk[3]));
                    ptx::tmem_ld_wait();
#pragma unroll
                    for (int g = 0; g < 4; ++g) {
                        float v[8];
#pragma unroll
                        for (int e = 0; e < 8; ++e) {
                            float x = __uint_as_float(r[g * 8 + e]);
                            if (d == 1) x = fmaf(dsig, s_wd[cb * 32 + g * 8 + e], x);           // density head: d(sigma) * w_density
                            v[e] = (mword >> ((e & 1) * 16 + g * 4 + (e >> 1))) & 1u ? x : 0.f;
                        }
                        const uint32_t w0 = ptx::pack_bf16(v[0], v[1], false), w1 = ptx::pack_bf16(v[2], v[3], false);
                        const uint32_t w2 = ptx::pack_bf16(v[4], v[5], false), w3 = ptx::pack_bf16(v[6], v[7], false);
                        // the last layer's result (G0) only feeds wgrad: keeping it out of smem avoids a race with
                        // the next pair's G9 prologue, which other threads of this row may already be writing
                        if (d + 1 < kDgLayers) ptx::st_shared_v4(a_base + (cb * 4 + g) * kKgBytes + row * 16, w0, w1, w2, w3);
                        __stcs(reinterpret_cast<uint4*>(g_row + (okg0 + cb * 4 + g) * kStKgBytes), make_uint4(w0, w1, w2, w3));
                    }
                }
                if (d + 1 < kDgLayers) {
                    ptx::fence_proxy_async_smem();
                    ptx::tc_fence_before();
                    ptx::mbar_arrive(bars.a_ready(t));
                }
            }
        }
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == kMmaWarp) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc(tmem, 512);
    }
}

// ============================================================================================================
// 2. wgrad
// ============================================================================================================
// packed gradient image (fp32 floats)
constexpr int64_t kGW0 = 0;                                   // [256][64]  mlps.0 (col 63 unused)
constexpr int64_t kGW5S = 16384;                              // [256][64]  mlps.5 skip columns
__host__ __device__ constexpr int64_t gwl(int i) { return 32768 + static_cast<int64_t>(i) * 65536; }
//   gwl(0..3) mlps.1..4 | gwl(4) mlps.5 hidden columns | gwl(5) mlps.6 | gwl(6) mlps.7 | gwl(7) T8 = [Wc_rgb ; Wc_ins]
constexpr int64_t kGW9 = 32768 + 8 * 65536;                   // [128][256] G9^T [rgb_hid | ins_hid]
constexpr int64_t kGWD = kGW9 + 32768;                        // [128][256] G9^T h7   (row 3 = density_linear.weight)
constexpr int64_t kGW8D = kGWD + 32768;                       // [256][32]  T8 dir-embedding columns
constexpr int64_t kGB = kGW8D + 8192;                         // bias sums: 9 x 256 (G0..G8) then 128 (G9)
constexpr int64_t kGimgFloats = kGB + 9 * 256 + 128;

constexpr int kWgStages = 2;
constexpr int kWgSlotBytes = 32 * kStKgBytes;                 // 32,768: one 256-feature region of a block
constexpr int kWgStageBytes = 3 * kWgSlotBytes;               // 98,304
constexpr int kWgSmBar = kWgStages * kWgStageBytes;           // 196,608
constexpr int kWgSmTotal = kWgSmBar + 128;
constexpr int kWgJobs = 11;
constexpr int kWgBiasWarps = 8;

struct WLoad { int slot, from_grad, kg0, nkg; };
struct WProd { int g_slot, halves, a_slot, n, col0, ld; int64_t dst; };
struct WBias { int slot, nkg; int64_t dst; };
struct WJob {
    int nloads, nprods, nbias;
    WLoad ld[3];
    WProd pr[2];
    WBias bs[2];
};

__host__ __device__ inline WJob wgrad_job(int j) {
    WJob J{};
    if (j == 0) {                       // A = pos-embedding: mlps.0 and the skip columns of mlps.5
        J.nloads = 3; J.ld[0] = {0, 1, kStG0, 32}; J.ld[1] = {1, 1, kStG0 + 5 * 32, 32}; J.ld[2] = {2, 0, kStPos, 8};
        J.nprods = 2; J.pr[0] = {0, 2, 2, 64, 0, 64, kGW0}; J.pr[1] = {1, 2, 2, 64, 128, 64, kGW5S};
        J.nbias = 2; J.bs[0] = {0, 32, kGB}; J.bs[1] = {1, 32, kGB + 5 * 256};
    } else if (j <= 8) {                // T_j: G_j^T h_{j-1}   (j == 8: [rgb_hid ; ins_hid] rows against h7)
        const int gkg = j == 8 ? kStG8 : kStG0 + 32 * j;
        J.nloads = 2; J.ld[0] = {0, 1, gkg, 32}; J.ld[1] = {2, 0, kStH0 + 32 * (j - 1), 32};
        J.nprods = 1; J.pr[0] = {0, 2, 2, 256, 0, 256, gwl(j - 1)};
        J.nbias = j == 5 ? 0 : 1; J.bs[0] = {0, 32, kGB + 256 * j};
    } else if (j == 9) {                // heads: G9^T hid (rgb_linear, ins_linear) and G9^T h7 (density_linear)
        J.nloads = 3; J.ld[0] = {0, 1, kStG9, 16}; J.ld[1] = {1, 0, kStH0 + 32 * 7, 32}; J.ld[2] = {2, 0, kStHid, 32};
        J.nprods = 2; J.pr[0] = {0, 1, 2, 256, 0, 256, kGW9}; J.pr[1] = {0, 1, 1, 256, 256, 256, kGWD};
        J.nbias = 1; J.bs[0] = {0, 16, kGB + 9 * 256};
    } else {                            // dir-embedding columns of rgb_feature_linears.0
        J.nloads = 2; J.ld[0] = {0, 1, kStG8, 32}; J.ld[1] = {2, 0, kStDir, 4};
        J.nprods = 1; J.pr[0] = {0, 2, 2, 32, 0, 32, kGW8D};
        J.nbias = 0;
    }
    return J;
}

__device__ __forceinline__ void red_add_v4(float* dst, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(dst), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
__device__ __forceinline__ uint4 lds_v4(uint32_t addr) {
    uint4 r;
    asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "r"(addr));
    return r;
}
__device__ __forceinline__ uint32_t idesc_bf16_mn(uint32_t m, uint32_t n) {      // both operands MN-major
    return ptx::idesc_bf16(m, n) | (1u << 15) | (1u << 16);
}

__global__ void __launch_bounds__(kThreads, 1)
mlp_wgrad_bf16_kernel(const uint8_t* __restrict__ acts, const uint8_t* __restrict__ gstash, int64_t n_blocks,
                      float* __restrict__ gimg, int swap) {
    extern __shared__ __align__(1024) uint8_t smem[];
    const uint32_t sm0 = ptx::smem_u32(smem);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t bar0 = sm0 + kWgSmBar;
    auto full = [&](int s) { return bar0 + 8 * s; };
    auto empty = [&](int s) { return bar0 + 8 * (kWgStages + s); };
    const uint32_t acc_done = bar0 + 8 * (2 * kWgStages), tmem_free = bar0 + 8 * (2 * kWgStages + 1);
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(smem + kWgSmBar + 8 * (2 * kWgStages + 2));

    if (threadIdx.x == 0) {
        for (int s = 0; s < kWgStages; ++s) { ptx::mbar_init(full(s), 1); ptx::mbar_init(empty(s), 1 + kWgBiasWarps); }
        ptx::mbar_init(acc_done, 1);
        ptx::mbar_init(tmem_free, kEpiWarps);
        ptx::fence_barrier_init();
    }
    if (warp == kMmaWarp) ptx::tmem_alloc(ptx::smem_u32(const_cast<uint32_t*>(tmem_slot)), 512);
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem = *tmem_slot;

    // Work split.  Small grids: every CTA walks all 11 jobs over its own slice of blocks.  Otherwise the CTAs are
    // dealt out to the JOBS in proportion to the bytes a job loads per block, and each CTA runs ONE job over a slice
    // of all blocks: it then flushes one accumulator set (<= 512 KB of vector reductions) instead of all of them
    // (2.8 MB per CTA, which dominated the kernel at training batch sizes: 148 x 2.8 MB of RED traffic per call).
    int j_first = 0, j_last = kWgJobs - 1;
    int64_t b0 = n_blocks * blockIdx.x / gridDim.x, b1 = n_blocks * (blockIdx.x + 1) / gridDim.x;
    // (with >= 48 blocks per CTA the flush is amortised and the even split balances better: measured at 12,288 blocks)
    if (gridDim.x >= 2 * kWgJobs && n_blocks >= 64 && n_blocks < 48 * static_cast<int64_t>(gridDim.x)) {
        int wsum = 0, wj[kWgJobs];
        for (int j = 0; j < kWgJobs; ++j) {
            const WJob J = wgrad_job(j);
            wj[j] = 0;
            for (int i = 0; i < J.nloads; ++i) wj[j] += J.ld[i].nkg;
            wsum += wj[j];
        }
        int acc = 0;
        for (int j = 0; j < kWgJobs; ++j) {
            const int s0 = static_cast<int>(static_cast<int64_t>(gridDim.x) * acc / wsum);
            acc += wj[j];
            const int s1 = static_cast<int>(static_cast<int64_t>(gridDim.x) * acc / wsum);
            if (static_cast<int>(blockIdx.x) >= s0 && static_cast<int>(blockIdx.x) < s1) {
                j_first = j_last = j;
                const int idx = blockIdx.x - s0, cnt = s1 - s0;
                b0 = n_blocks * idx / cnt;
                b1 = n_blocks * (idx + 1) / cnt;
            }
        }
    }

    if (warp == kProducerWarp) {
        if (lane == 0) {
            int stage = 0; uint32_t phase = 0;
            for (int j = j_first; j <= j_last; ++j) {
                const WJob J = wgrad_job(j);
                uint32_t bytes = 0;
                for (int i = 0; i < J.nloads; ++i) bytes += J.ld[i].nkg * kStKgBytes;
                for (int64_t b = b0; b < b1; ++b) {
                    ptx::mbar_wait(empty(stage), phase ^ 1);
                    ptx::mbar_arrive_expect_tx(full(stage), bytes);
                    for (int i = 0; i < J.nloads; ++i) {
                        const WLoad L = J.ld[i];
                        const uint8_t* src = L.from_grad ? gstash + b * kStGradBlkBytes : acts + b * kStActBlkBytes;
                        ptx::bulk_g2s(sm0 + stage * kWgStageBytes + L.slot * kWgSlotBytes,
                                      src + static_cast<int64_t>(L.kg0) * kStKgBytes, L.nkg * kStKgBytes, full(stage));
                    }
                    if (++stage == kWgStages) { stage = 0; phase ^= 1; }
                }
            }
            for (int i = 0; i < kWgStages; ++i) {
                ptx::mbar_wait(empty(stage), phase ^ 1);
                if (++stage == kWgStages) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp == kMmaWarp) {
        if (lane == 0) {
            int stage = 0; uint32_t phase = 0, free_par = 1;
            for (int j = j_first; j <= j_last; ++j) {
                const WJob J = wgrad_job(j);
                ptx::mbar_wait(tmem_free, free_par);          // previous job's accumulators have been flushed
                free_par ^= 1;
                ptx::tc_fence_after();
                for (int64_t b = b0; b < b1; ++b) {
                    ptx::mbar_wait(full(stage), phase);
                    ptx::tc_fence_after();
                    const uint32_t st = sm0 + stage * kWgStageBytes;
#pragma unroll 1
                    for (int ks = 0; ks < kStBlk / 16; ++ks) {
                        for (int i = 0; i < J.nprods; ++i) {
                            const WProd P = J.pr[i];
                            const uint32_t idesc = idesc_bf16_mn(128, P.n);
                            const uint64_t bd = op_desc(st + P.a_slot * kWgSlotBytes + ks * 256, 128, kStKgBytes, swap);
                            for (int h = 0; h < P.halves; ++h) {
                                const uint64_t ad = op_desc(st + P.g_slot * kWgSlotBytes + h * 16 * kStKgBytes + ks * 256,
                                                            128, kStKgBytes, swap);
                                ptx::umma_bf16(tmem + P.col0 + h * P.n, ad, bd, idesc, (b != b0 || ks != 0));
                            }
                        }
                    }
                    ptx::umma_commit(empty(stage));
                    if (++stage == kWgStages) { stage = 0; phase ^= 1; }
                }
                ptx::umma_commit(acc_done);
            }
        }
    } else {
        // ===== aux warps 0..15: bias sums from the staged G operands (warps 0..7), then the per-job flush =====
        int stage = 0; uint32_t phase = 0, done_par = 0;
        const int q = warp & 3, cg = warp >> 2;
        for (int j = j_first; j <= j_last; ++j) {
            const WJob J = wgrad_job(j);
            if (warp < kWgBiasWarps) {
                float bsum[2][4][8];
#pragma unroll
                for (int a = 0; a < 2; ++a)
#pragma unroll
                    for (int s = 0; s < 4; ++s)
#pragma unroll
                        for (int e = 0; e < 8; ++e) bsum[a][s][e] = 0.f;
                for (int64_t b = b0; b < b1; ++b) {
                    ptx::mbar_wait(full(stage), phase);
                    const uint32_t st = sm0 + stage * kWgStageBytes;
#pragma unroll
                    for (int a = 0; a < 2; ++a) {
                        if (a < J.nbias) {
#pragma unroll
                            for (int s = 0; s < 4; ++s) {
                                const int kg = warp * 4 + s;
                                if (kg < J.bs[a].nkg) {
#pragma unroll
                                    for (int rr = 0; rr < 2; ++rr) {
                                        const uint4 w = lds_v4(st + J.bs[a].slot * kWgSlotBytes + kg * kStKgBytes + (lane + 32 * rr) * 16);
                                        const uint32_t ww[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
                                        for (int e = 0; e < 4; ++e) {
                                            bsum[a][s][2 * e] += __uint_as_float(ww[e] << 16);
                                            bsum[a][s][2 * e + 1] += __uint_as_float(ww[e] & 0xFFFF0000u);
                                        }
                                    }
                                }
                            }
                        }
                    }
                    __syncwarp();
                    if (lane == 0) ptx::mbar_arrive(empty(stage));
                    if (++stage == kWgStages) { stage = 0; phase ^= 1; }
                }
#pragma unroll
                for (int a = 0; a < 2; ++a) {
                    if (a < J.nbias) {
#pragma unroll
                        for (int s = 0; s < 4; ++s) {
                            const int kg = warp * 4 + s;
                            if (kg < J.bs[a].nkg) {
#pragma unroll
                                for (int e = 0; e < 8; ++e) {
                                    const float v = warp_sum(bsum[a][s][e]);
                                    if (lane == 0) atomicAdd(gimg + J.bs[a].dst + kg * 8 + e, v);
                                }
                            }
                        }
                    }
                }
            }
            // ---- flush: TMEM accumulators -> packed gradient image (vector fp32 reductions) ----
            ptx::mbar_wait(acc_done, done_par);
            done_par ^= 1;
            ptx::tc_fence_after();
            int unit = 0;
            for (int i = 0; i < J.nprods; ++i) {
                const WProd P = J.pr[i];
                for (int h = 0; h < P.halves; ++h) {
                    for (int c0 = 0; c0 < P.n; c0 += 32, ++unit) {
                        if ((unit & 3) != cg) continue;
                        uint32_t r[32];
                        ptx::tmem_ld32(tmem + (static_cast<uint32_t>(q * 32) << 16) + P.col0 + h * P.n + c0, r);
                        ptx::tmem_ld_wait();
                        float* dst = gimg + P.dst + static_cast<int64_t>(h * 128 + q * 32 + lane) * P.ld + c0;
#pragma unroll
                        for (int e = 0; e < 32; e += 4)
                            red_add_v4(dst + e, __uint_as_float(r[e]), __uint_as_float(r[e + 1]), __uint_as_float(r[e + 2]),
                                       __uint_as_float(r[e + 3]));
                    }
                }
            }
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(tmem_free);
        }
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == kMmaWarp) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc(tmem, 512);
    }
}

// ============================================================================================================
// 3. packed gradient image -> parameter gradients
// ============================================================================================================
struct GradPtrs { float* g[kNumParams]; };

__global__ void unpack_grads_kernel(const float* __restrict__ gimg, const uint8_t* __restrict__ image, Image im,
                                    GradPtrs gp) {
    const float* f32 = reinterpret_cast<const float*>(image + im.f32_off);
    const float* w_rgbf = f32 + kF32Rgbf;      // [256][256]
    const float* w_insf = f32 + kF32Insf;
    const float* w_rgb0 = f32 + kF32Rgb0;      // [128][256] hidden columns of rgb_feature_linears.0
    const float* w_ins0 = f32 + kF32Ins0;
    const float* b_rgbf = f32 + kF32BRgbf;
    const float* b_insf = f32 + kF32BInsf;
    const float* gw8 = gimg + gwl(7);          // [256][256]: rows 0..127 dWc_rgb, 128..255 dWc_ins
    const float* gb8 = gimg + kGB + 8 * 256;
    const float* gb9 = gimg + kGB + 9 * 256;
    const int K1 = im.ins_num + 1;
    const int param = blockIdx.y;
    int count;
    switch (param) {
        case 0: count = 256 * 63; break;
        case 10: count = 256 * 319; break;
        case P_RGBF_W: case P_INSF_W: count = 65536; break;
        case P_RGB0_W: count = 128 * 283; break;
        case P_INS0_W: count = 0; break;
        case P_RGB0_B: case P_INS0_B: count = 128; break;
        case P_DENS_W: count = 256; break;
        case P_DENS_B: count = 1; break;
        case P_INSL_W: count = K1 * 128; break;
        case P_INSL_B: count = K1; break;
        case P_RGBL_W: count = 3 * 128; break;
        case P_RGBL_B: count = 3; break;
        default: count = (param & 1) ? 256 : 65536; break;       // trunk biases / 256x256 trunk weights, feature biases
    }
    for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < count; idx += gridDim.x * blockDim.x) {
        float v = 0.f;
        if (param < 16) {
            const int l = param >> 1;
            if (param & 1) v = gimg[kGB + 256 * l + idx];
            else if (l == 0) v = gimg[kGW0 + (idx / 63) * 64 + idx % 63];
            else if (l == 5) { const int n = idx / 319, k = idx % 319; v = k < 256 ? gimg[gwl(4) + n * 256 + k] : gimg[kGW5S + n * 64 + (k - 256)]; }
            else v = gimg[gwl(l - 1) + idx];                    // mlps.1-4 -> gwl(0..3), mlps.6,7 -> gwl(5,6)
        } else if (param == P_RGBF_W || param == P_INSF_W) {       // dW0[j][k] = sum_i W1[i][j] dWc[i][k]
            const int j = idx >> 8, k = idx & 255;
            const float* w1 = param == P_RGBF_W ? w_rgb0 : w_ins0;
            const float* dwc = param == P_RGBF_W ? gw8 : gw8 + 128 * 256;
            for (int i = 0; i < 128; ++i) v = fmaf(w1[i * 256 + j], dwc[i * 256 + k], v);
        } else if (param == P_RGBF_B || param == P_INSF_B) {       // db0[j] = sum_i W1[i][j] dbc[i]
            const float* w1 = param == P_RGBF_B ? w_rgb0 : w_ins0;
            const float* dbc = param == P_RGBF_B ? gb8 : gb8 + 128;
            for (int i = 0; i < 128; ++i) v = fmaf(w1[i * 256 + idx], dbc[i], v);
        } else if (param == P_RGB0_W || param == P_INS0_W) {       // dir-embedding columns only; the hidden columns are
            const int i = idx / 283, j = idx % 283;                // dot products over k: unpack_composed_kernel
            if (param == P_INS0_W || j < 256) continue;
            v = gimg[kGW8D + i * 32 + (j - 256)];
        } else if (param == P_RGB0_B) v = gb8[idx];
        else if (param == P_INS0_B) v = gb8[128 + idx];
        else if (param == P_DENS_W) v = gimg[kGWD + 3 * 256 + idx];
        else if (param == P_DENS_B) v = gb9[3];
        else if (param == P_INSL_W) v = gimg[kGW9 + (4 + idx / 128) * 256 + 128 + idx % 128];
        else if (param == P_INSL_B) v = gb9[4 + idx];
        else if (param == P_RGBL_W) v = gimg[kGW9 + (idx / 128) * 256 + idx % 128];
        else if (param == P_RGBL_B) v = gb9[idx];
        gp.g[param][idx] += v;
    }
}

// dW1[i][j] += sum_k dWc[i][k] W0[j][k] + dbc[i] b0[j] for the hidden columns of rgb_feature_linears.0 (ld 283) and
// ins_feature_linears.0 (ld 256): one warp per output, k across lanes (both rows read coalesced).
__global__ void unpack_composed_kernel(const float* __restrict__ gimg, const uint8_t* __restrict__ image, Image im,
                                       GradPtrs gp) {
    const float* f32 = reinterpret_cast<const float*>(image + im.f32_off);
    const int lane = threadIdx.x & 31;
    const int out = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);      // 2 x 128 x 256 outputs
    if (out >= 2 * 128 * 256) return;
    const int br = out >> 15, i = (out >> 8) & 127, j = out & 255;
    const float* dwc = gimg + gwl(7) + (br * 128 + i) * 256;
    const float* w0 = f32 + (br == 0 ? kF32Rgbf : kF32Insf) + j * 256;
    float v = 0.f;
#pragma unroll
    for (int k = lane; k < 256; k += 32) v = fmaf(dwc[k], w0[k], v);
    v = warp_sum(v);
    if (lane == 0) {
        v = fmaf(gimg[kGB + 8 * 256 + br * 128 + i], f32[(br == 0 ? kF32BRgbf : kF32BInsf) + j], v);
        float* g = br == 0 ? gp.g[P_RGB0_W] + i * 283 + j : gp.g[P_INS0_W] + i * 256 + j;
        *g += v;
    }
}

// ---- bring-up self test: D[128,n] = sum_k A[k][m] B[k][n], both operands MN-major (k = 64 "points") ----------
__global__ void __launch_bounds__(128, 1)
umma_mn_selftest_kernel(const float* __restrict__ a, const float* __restrict__ b, int n, float* __restrict__ d, int swap) {
    extern __shared__ __align__(1024) uint8_t smem[];
    const uint32_t sm0 = ptx::smem_u32(smem);
    const int warp = threadIdx.x >> 5, row = threadIdx.x;
    const uint32_t a_base = sm0, b_base = sm0 + 16384, bar = sm0 + 16384 + 32768;
    volatile uint32_t* slot = reinterpret_cast<volatile uint32_t*>(smem + 16384 + 32768 + 8);
    __nv_bfloat16* sa = reinterpret_cast<__nv_bfloat16*>(smem);
    __nv_bfloat16* sb = reinterpret_cast<__nv_bfloat16*>(smem + 16384);
    for (int i = threadIdx.x; i < 64 * 128; i += 128) {           // a[k][m] -> slab m/8, row k, elem m%8
        const int k = i / 128, m = i % 128;
        sa[((m >> 3) * 64 + k) * 8 + (m & 7)] = __float2bfloat16_rn(a[i]);
    }
    for (int i = threadIdx.x; i < 64 * n; i += 128) {
        const int k = i / n, c = i % n;
        sb[((c >> 3) * 64 + k) * 8 + (c & 7)] = __float2bfloat16_rn(b[i]);
    }
    if (threadIdx.x == 0) { ptx::mbar_init(bar, 1); ptx::fence_barrier_init(); }
    if (warp == 0) ptx::tmem_alloc(ptx::smem_u32(const_cast<uint32_t*>(slot)), 256);
    ptx::fence_proxy_async_smem();
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem = *slot;
    if (threadIdx.x == 0) {
        const uint32_t idesc = idesc_bf16_mn(128, n);
        for (int ks = 0; ks < 4; ++ks) {
            const uint64_t ad = op_desc(a_base + ks * 256, 128, kStKgBytes, swap);
            const uint64_t bd = op_desc(b_base + ks * 256, 128, kStKgBytes, swap);
            ptx::umma_bf16(tmem, ad, bd, idesc, ks != 0);
        }
        ptx::umma_commit(bar);
    }
    ptx::mbar_wait(bar, 0);
    ptx::tc_fence_after();
    for (int cb = 0; cb * 16 < n; ++cb) {
        uint32_t r[16];
        ptx::tmem_ld16(tmem + (static_cast<uint32_t>(warp * 32) << 16) + cb * 16, r);
        ptx::tmem_ld_wait();
        for (int e = 0; e < 16; ++e) d[row * n + cb * 16 + e] = __uint_as_float(r[e]);
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 0) { ptx::tc_fence_after(); ptx::tmem_dealloc(tmem, 256); }
}

}  // namespace tcb

int64_t bf16_acts_bytes(int64_t n_points) { return tc_stash_blocks(n_points) * (kStActBlkBytes + kStMaskBlkBytes); }

// Sign masks from an activation stash whose forward did not write them (the parked forward variants): one thread per
// mask word.  The default forward kernel writes the masks itself.
namespace tcb {
__global__ void masks_from_acts_kernel(uint8_t* __restrict__ acts, int64_t n_points, int64_t n_blocks) {
    const int64_t idx = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (idx >= n_blocks * kStMaskLayers * 2 * 4 * kStBlk) return;
    const int row = static_cast<int>(idx % kStBlk);
    int64_t r = idx / kStBlk;
    const int c = static_cast<int>(r % 4); r /= 4;
    const int half = static_cast<int>(r % 2); r /= 2;
    const int a = static_cast<int>(r % kStMaskLayers);
    const int64_t blk = r / kStMaskLayers;
    const int kg0 = (a == 8 ? kStHid : kStH0 + 32 * a) + (half * 4 + c) * 4;
    uint32_t m = 0;
    for (int g = 0; g < 4; ++g) {
        const uint4 w = *reinterpret_cast<const uint4*>(acts + blk * kStActBlkBytes + static_cast<int64_t>(kg0 + g) * kStKgBytes + row * 16);
        const uint32_t ww[4] = {w.x, w.y, w.z, w.w};
        for (int j = 0; j < 4; ++j) {
            const int k = g * 4 + j;
            if (ww[j] & 0xFFFFu) m |= 1u << k;
            if (ww[j] >> 16) m |= 1u << (16 + k);
        }
    }
    *reinterpret_cast<uint32_t*>(acts + tc_mask_offset(n_points) + tc_mask_word(blk, a, half, c, row)) = m;
}
}  // namespace tcb

int launch_masks_from_acts(void* acts, int64_t n_points, cudaStream_t st) {
    const int64_t blocks = ((n_points + 255) / 256) * 4;
    const int64_t words = blocks * kStMaskLayers * 2 * 4 * kStBlk;
    tcb::masks_from_acts_kernel<<<static_cast<unsigned>((words + 255) / 256), 256, 0, st>>>(static_cast<uint8_t*>(acts), n_points, blocks);
    DM_LAUNCH_CHECK("masks_from_acts_kernel");
    return DMNERF_OK;
}

int64_t bf16_bwd_workspace_bytes(int64_t n_points) {
    return tc_stash_blocks(n_points) * kStGradBlkBytes + tcb::kGimgFloats * 4;
}

int persistent_ctas(int sm_count);

int launch_mlp_bwd_bf16(const void* packed, const void* acts, const float* d_raw, int64_t n_points, int ins_num,
                        float* const* grads, void* workspace, int debug_flags, cudaStream_t st) {
    static int sm_count = 0;
    if (sm_count == 0) {
        int dev = 0;
        DM_CUDA(cudaGetDevice(&dev));
        DM_CUDA(cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev));
    }
    DM_REQUIRE(workspace, DMNERF_EINVAL, "mlp_bwd(bf16): workspace of dmnerf_bwd_workspace_bytes() is required");
    DM_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 1023) == 0 && (reinterpret_cast<uintptr_t>(acts) & 1023) == 0 &&
               (reinterpret_cast<uintptr_t>(packed) & 15) == 0,
               DMNERF_EINVAL, "mlp_bwd(bf16): acts/workspace must be 1024-byte aligned, packed 16-byte aligned");
    const tc::Image im = tc::make_image(ins_num);
    uint8_t* gstash = static_cast<uint8_t*>(workspace);
    float* gimg = reinterpret_cast<float*>(gstash + tc_stash_blocks(n_points) * kStGradBlkBytes);
    DM_CUDA(cudaMemsetAsync(gimg, 0, tcb::kGimgFloats * 4, st));

    // blocks actually written by the forward / dgrad kernels: whole 256-point pairs (the allocation is rounded up
    // further, to 512-point quads, for the CTA-pair forward)
    const int64_t pairs = (n_points + 255) / 256;
    const int64_t blocks = pairs * 4;
    const int sms = persistent_ctas(sm_count);
    DM_CUDA(cudaFuncSetAttribute(tcb::mlp_dgrad_bf16_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, tcb::kDgSmTotal));
    tcb::mlp_dgrad_bf16_kernel<<<static_cast<unsigned>(pairs < sms ? pairs : sms), tc::kFwThreads, tcb::kDgSmTotal, st>>>(
        d_raw, n_points, static_cast<const uint8_t*>(packed), im, static_cast<const uint8_t*>(acts), gstash);
    DM_LAUNCH_CHECK("mlp_dgrad_bf16_kernel");

    DM_CUDA(cudaFuncSetAttribute(tcb::mlp_wgrad_bf16_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, tcb::kWgSmTotal));
    tcb::mlp_wgrad_bf16_kernel<<<static_cast<unsigned>(blocks < sms ? blocks : sms), tc::kThreads, tcb::kWgSmTotal, st>>>(
        static_cast<const uint8_t*>(acts), gstash, blocks, gimg, debug_flags & 2 ? 1 : 0);
    DM_LAUNCH_CHECK("mlp_wgrad_bf16_kernel");

    tcb::GradPtrs gp;
    for (int i = 0; i < kNumParams; ++i) gp.g[i] = grads[i];
    tcb::unpack_grads_kernel<<<dim3(256, kNumParams), 256, 0, st>>>(gimg, static_cast<const uint8_t*>(packed), im, gp);
    tcb::unpack_composed_kernel<<<2 * 128 * 256 / 8, 256, 0, st>>>(gimg, static_cast<const uint8_t*>(packed), im, gp);
    DM_LAUNCH_CHECK("unpack_grads_kernel");
    return DMNERF_OK;
}

int launch_selftest_umma_mn(const float* a, const float* b, int n, float* d, int debug_flags, cudaStream_t st) {
    const int smem = 16384 + 32768 + 64;
    DM_CUDA(cudaFuncSetAttribute(tcb::umma_mn_selftest_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    tcb::umma_mn_selftest_kernel<<<1, 128, smem, st>>>(a, b, n, d, debug_flags & 2 ? 1 : 0);
    DM_LAUNCH_CHECK("umma_mn_selftest_kernel");
    return DMNERF_OK;
}

}  // namespace dmnerf

// ---- no-allocation entry point (no counterpart in the reference) ------------
// Some callers cannot call cudaMalloc -- code under CUDA graph capture, tight serving / training loops, harnesses that
// forbid it -- yet the bf16 backward needs scratch: the dgrad chain's gradient stash (4.75 KB per point, 64-point blocks in
// 512-point quads) and the packed gradient image the wgrad GEMMs reduce into.  dmnerf_mlp_bwd_bf16_noalloc keeps that
// scratch in a file-scope __device__ array of a fixed, modest size: every GPU that loads the library holds
// kNoAllocBytes + 1 KB (~13 MB) until the process exits, and the entry point serves at most kNoAllocMaxPoints points per call.
//   * Host code must not hand the array's NAME to a kernel: on the host it is a stand-in symbol, not a device address.  The
//     address comes from cudaGetSymbolAddress, once per device, with that device current (below).
//   * Every call on a device shares the one array and finds the previous call's contents in it.  What must start at zero
//     (the gradient image) is cleared by the launcher on the caller's stream; the stash is fully rewritten by the dgrad chain
//     for the blocks the wgrad GEMMs read.  Calls that overlap on the GPU (two streams, two host threads) RACE on it:
//     serialise them, or use dmnerf_mlp_bwd with one workspace per stream.
//   * No cudaMalloc / cudaFree / allocating library routine is called: cudaGetSymbolAddress, cudaMemsetAsync,
//     cudaFuncSetAttribute and kernel launches only.
#include <mutex>

namespace dmnerf {

constexpr int64_t kNoAllocMaxPoints = 2048;                                     // e.g. 32 rays x 64 samples per call
constexpr int64_t kNoAllocBlocks = ((kNoAllocMaxPoints + 511) / 512) * 8;      // == tc_stash_blocks(kNoAllocMaxPoints)
constexpr int64_t kNoAllocBytes = kNoAllocBlocks * kStGradBlkBytes + tcb::kGimgFloats * 4;
// + 1 KB: the launcher wants the workspace 1024-byte aligned, the array is declared with 256-byte alignment
static __device__ __align__(256) unsigned char g_bwd_scratch[kNoAllocBytes + 1024];

constexpr int kNoAllocMaxDevices = 64;
static void* g_bwd_scratch_ptr[kNoAllocMaxDevices] = {};
static std::mutex g_bwd_scratch_mutex;

// device address of g_bwd_scratch on the CURRENT device, rounded up to 1024 bytes; resolved once per device
static int noalloc_scratch(void** out) {
    int dev = 0;
    DM_CUDA(cudaGetDevice(&dev));
    DM_REQUIRE(dev >= 0 && dev < kNoAllocMaxDevices, DMNERF_EUNSUPPORTED, "mlp_bwd_bf16_noalloc: device ordinal %d >= %d", dev,
               kNoAllocMaxDevices);
    std::lock_guard<std::mutex> lock(g_bwd_scratch_mutex);
    if (g_bwd_scratch_ptr[dev] == nullptr) {
        void* p = nullptr;
        DM_CUDA(cudaGetSymbolAddress(&p, g_bwd_scratch));
        DM_REQUIRE(p != nullptr, DMNERF_ECUDA, "mlp_bwd_bf16_noalloc: cudaGetSymbolAddress returned a null address");
        g_bwd_scratch_ptr[dev] = reinterpret_cast<void*>((reinterpret_cast<uintptr_t>(p) + 1023) & ~static_cast<uintptr_t>(1023));
    }
    *out = g_bwd_scratch_ptr[dev];
    return DMNERF_OK;
}

}  // namespace dmnerf

extern "C" int64_t dmnerf_noalloc_max_points(void) { return dmnerf::kNoAllocMaxPoints; }

extern "C" int dmnerf_mlp_bwd_bf16_noalloc(const void* packed, const void* acts, const float* d_raw, int64_t n_points,
                                           int ins_num, float* const* grads, void* stream) {
    using namespace dmnerf;
    DM_REQUIRE(packed && acts && d_raw && grads, DMNERF_EINVAL, "mlp_bwd_bf16_noalloc: null pointer");
    DM_REQUIRE(ins_num >= 1 && ins_num <= kMaxIns, DMNERF_EUNSUPPORTED, "mlp_bwd_bf16_noalloc: ins_num %d outside [1,%d]",
               ins_num, kMaxIns);
    for (int i = 0; i < kNumParams; ++i) DM_REQUIRE(grads[i], DMNERF_EINVAL, "mlp_bwd_bf16_noalloc: grads[%d] is null", i);
    DM_REQUIRE(n_points >= 0 && n_points <= kNoAllocMaxPoints, DMNERF_EUNSUPPORTED,
               "mlp_bwd_bf16_noalloc: n_points %lld exceeds the static scratch's maximum of %lld points per call "
               "(use dmnerf_mlp_bwd with a caller-provided workspace)", (long long)n_points, (long long)kNoAllocMaxPoints);
    if (n_points == 0) return DMNERF_OK;
    DM_REQUIRE(bf16_bwd_workspace_bytes(n_points) <= kNoAllocBytes, DMNERF_EUNSUPPORTED,
               "mlp_bwd_bf16_noalloc: workspace of %lld bytes exceeds the static %lld", (long long)bf16_bwd_workspace_bytes(n_points),
               (long long)kNoAllocBytes);
    void* scratch = nullptr;
    if (int rc = noalloc_scratch(&scratch)) return rc;
    return launch_mlp_bwd_bf16(packed, acts, d_raw, n_points, ins_num, grads, scratch, 0, static_cast<cudaStream_t>(stream));
}
