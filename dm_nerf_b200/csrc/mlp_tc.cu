// mlp_tc.cu -- a1+a2+a9 in DMNERF_MODE_BF16: positional encoding + the whole DM_NeRF MLP as ONE
// persistent tcgen05 kernel.  bf16 operands, fp32 accumulation in TMEM, layers chained through
// shared memory (activations never touch HBM), weights streamed L2 -> smem by the TMA engine
// (1-D bulk async copies into a 3-stage mbarrier ring).
//
// Template switches of the one kernel: kStash (training: operands and ReLU sign masks mirrored to HBM), kF16 (fp16
// instead of bf16 operands: DMNERF_MODE_F16, inference), kDens (density-only coarse pass: T0..T7, sigma out).
//
// Work decomposition: one CTA per SM, each looping over pairs of 128-point tiles (X, Y).  Both
// tiles walk the layers in lockstep so every streamed weight chunk feeds 4 MMAs (2 k-steps x 2
// tiles).  Warp roles: warps 0-3 epilogue of X, warps 4-7 epilogue of Y (thread == point/TMEM lane),
// warp 8 weight producer, warp 9 MMA issuer (one elected thread) + TMEM owner.
//
// Layers as the tensor path sees them (the two linear->linear pairs of the reference are composed
// at pack time: rgb_feature_linear o rgb_feature_linears.0 and ins_feature_linear o
// ins_feature_linears.0 have no activation between them, networks/dm_nerf.py:89-99):
//   T0      K= 64  N=256  relu   A = pos-embedding (EXT region)
//   T1..T4  K=256  N=256  relu
//   T5      K=320  N=256  relu   A = [h | pos-embedding]
//   T6,T7   K=256  N=256  relu   (T7's epilogue also takes the density dot product in fp32)
//   T8      K=288  N=256  relu   A = [h7 | dir-embedding]; rows 0..127 rgb_hid, 128..255 ins_hid
//   T9      K=256  N=Np   -      A = [rgb_hid | ins_hid]; block-diagonal rgb / ins heads, Np = round16(C)
//
// Operand layout (A and B, both K-major, no swizzle): [k-group of 8][row][8 bf16] -> 8x16-byte core
// matrices are 128 contiguous bytes; LBO (next k-group) = rows*16 B, SBO (next 8 rows) = 128 B.  An
// epilogue thread owns one row, so one 16-byte store writes 8 consecutive features of the next
// layer's A operand and a warp's stores are 512 contiguous bytes (conflict-free).
#include "mlp_tc_common.cuh"
#include <cuda_fp16.h>

namespace dmnerf {

static int g_reserved_sms = 0;     // SMs the persistent MLP kernels leave free (dmnerf_reserve_sms)
static int g_tc_debug_flags = 0;   // bit0: swap LBO/SBO in the smem descriptors (bring-up aid)

namespace tc {

// ---- pack --------------------------------------------------------------------------------------------
// composed weights: Wc[i][k] = sum_j W1[i][j] * W0[j][k]  (W1 = *_feature_linears.0[:, :256], W0 = *_feature_linear)
// bc[i] = sum_j W1[i][j]*b0[j] + b1[i].  scratch: [0,32768) rgb, [32768,65536) ins, then bc[256].
__global__ void __launch_bounds__(256)
compose_kernel(Params pr, float* __restrict__ scratch) {
    // grid (16, 8, 3): z = 0 rgb / 1 ins: a 16x16 tile of Wc; z = 2: the composed biases, eight rows per block.
    // Shared-memory tiles; every weight still accumulates its 256 products in increasing j with fmaf (same bits as the
    // one-thread-per-output loop this replaces, which took 30 us per model on the training step's critical path).
    if (blockIdx.z == 2) {
        // composed biases: one warp per output row, lanes stride the 256 products (coalesced), shuffle tree at the end
        const int row = (blockIdx.y * 16 + blockIdx.x) * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
        if (row >= 256) return;
        const int bb = row >> 7, i = row & 127;
        const float* w1b = bb == 0 ? pr.p[P_RGB0_W] : pr.p[P_INS0_W];
        const float* b0 = bb == 0 ? pr.p[P_RGBF_B] : pr.p[P_INSF_B];
        const float* b1 = bb == 0 ? pr.p[P_RGB0_B] : pr.p[P_INS0_B];
        const int ldb = bb == 0 ? 283 : 256;
        float a = 0.f;
#pragma unroll
        for (int j = lane; j < 256; j += 32) a = fmaf(w1b[i * ldb + j], b0[j], a);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
        if (lane == 0) scratch[2 * 128 * 256 + row] = a + b1[i];
        return;
    }
    const int br = blockIdx.z;
    const float* w1 = br == 0 ? pr.p[P_RGB0_W] : pr.p[P_INS0_W];
    const int ld1 = br == 0 ? 283 : 256;
    const float* w0 = br == 0 ? pr.p[P_RGBF_W] : pr.p[P_INSF_W];
    __shared__ float sa[16][17], sb[16][17];
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    const int i = blockIdx.y * 16 + ty, k = blockIdx.x * 16 + tx;
    float a = 0.f;
    for (int j0 = 0; j0 < 256; j0 += 16) {
        sa[ty][tx] = w1[i * ld1 + j0 + tx];
        sb[ty][tx] = w0[(j0 + ty) * 256 + k];
        __syncthreads();
#pragma unroll
        for (int j = 0; j < 16; ++j) a = fmaf(sa[ty][j], sb[j][tx], a);
        __syncthreads();
    }
    scratch[br * 32768 + i * 256 + k] = a;
}

// one thread per bf16 element of the chunk stream: layer l, chunk c, k-group g, row n, element e
__device__ __forceinline__ uint16_t half_bits(float v, int f16) {
    return f16 ? __half_as_ushort(__float2half_rn(v)) : __bfloat16_as_ushort(__float2bfloat16_rn(v));
}
__device__ __forceinline__ float half_value(uint16_t b, int f16) {
    return f16 ? __half2float(__ushort_as_half(b)) : __bfloat162float(__ushort_as_bfloat16(b));
}

__global__ void pack_bf16_kernel(Params pr, Image im, uint8_t* __restrict__ image, int f16) {
    const float* scratch = reinterpret_cast<const float*>(image + im.scratch_off);
    const int64_t idx = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    const int64_t total = im.bias_off / 2;
    if (idx >= total) return;
    int l = 0;
    while (l + 1 < kLayers && idx * 2 >= im.chunk_off[l + 1]) ++l;
    const int N = im.n[l];
    int64_t r = idx - im.chunk_off[l] / 2;
    const int per_chunk = kChunkK * N;
    const int c = static_cast<int>(r / per_chunk);
    r -= static_cast<int64_t>(c) * per_chunk;
    const int g = static_cast<int>(r / (N * 8));
    r -= static_cast<int64_t>(g) * N * 8;
    const int n = static_cast<int>(r >> 3), e = static_cast<int>(r & 7);
    const int k = c * kChunkK + g * 8 + e;
    reinterpret_cast<uint16_t*>(image)[idx] = half_bits(logical_weight(pr, scratch, l, n, k, im.ins_num), f16);
}

// dgrad operand B[n = forward input j][k = forward output i] of dgrad layer d (mlp_tc_common.cuh)
__device__ inline float dgrad_weight(const Params& pr, const float* scratch, int d, int n, int k, int ins_num) {
    if (d == 0) {                                       // T9^T: k = raw channel, n = [rgb_hid | ins_hid] feature
        if (n < 128) return k < 3 ? pr.p[P_RGBL_W][k * 128 + n] : 0.f;
        return (k >= 4 && k < 5 + ins_num) ? pr.p[P_INSL_W][(k - 4) * 128 + (n - 128)] : 0.f;
    }
    if (d == 1) return scratch[k * 256 + n];            // composed rgb rows of T8 (k < 128); ins rows are detached
    const int l = 9 - d;                                // d = 2 -> T7 ... d = 8 -> T1
    if (l == 5) return pr.p[10][k * 319 + n];           // hidden columns of the skip layer
    return pr.p[2 * l][k * 256 + n];
}

__global__ void pack_dgrad_kernel(Params pr, Image im, uint8_t* __restrict__ image) {
    const float* scratch = reinterpret_cast<const float*>(image + im.scratch_off);
    const int64_t idx = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    const int64_t total = (im.f32_off - im.dg_off[0]) / 2;
    if (idx >= total) return;
    int d = 0;
    while (d + 1 < kDgLayers && idx * 2 >= im.dg_off[d + 1] - im.dg_off[0]) ++d;
    int64_t r = idx - (im.dg_off[d] - im.dg_off[0]) / 2;
    const int per_chunk = kChunkK * 256;
    const int c = static_cast<int>(r / per_chunk);
    r -= static_cast<int64_t>(c) * per_chunk;
    const int g = static_cast<int>(r / (256 * 8));
    r -= static_cast<int64_t>(g) * 256 * 8;
    const int n = static_cast<int>(r >> 3), e = static_cast<int>(r & 7);
    const int k = c * kChunkK + g * 8 + e;
    reinterpret_cast<__nv_bfloat16*>(image + im.dg_off[0])[idx] = __float2bfloat16_rn(dgrad_weight(pr, scratch, d, n, k, im.ins_num));
}

__global__ void pack_f32_kernel(Params pr, Image im, uint8_t* __restrict__ image) {
    float* f = reinterpret_cast<float*>(image + im.f32_off);
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= kF32Floats) return;
    float v;
    if (i < kF32Insf) v = pr.p[P_RGBF_W][i];
    else if (i < kF32Rgb0) v = pr.p[P_INSF_W][i - kF32Insf];
    else if (i < kF32Ins0) { const int r = i - kF32Rgb0; v = pr.p[P_RGB0_W][(r >> 8) * 283 + (r & 255)]; }
    else if (i < kF32BRgbf) v = pr.p[P_INS0_W][i - kF32Ins0];
    else if (i < kF32BInsf) v = pr.p[P_RGBF_B][i - kF32BRgbf];
    else v = pr.p[P_INSF_B][i - kF32BInsf];
    f[i] = v;
}

// forward bias blocks: layer l = blockIdx.x, element (g, n, e) of a [2][256][8] K-major B operand; k = 0 holds the
// bf16 head of the bias and k = 1 the bf16 of the remainder (A = ones in both columns), ~16 mantissa bits in all
__global__ void pack_fwbias_kernel(Params pr, Image im, uint8_t* __restrict__ image, int f16) {
    const float* scratch = reinterpret_cast<const float*>(image + im.scratch_off);
    const int l = blockIdx.x;
    uint16_t* dst = reinterpret_cast<uint16_t*>(image + im.fwb_off + static_cast<int64_t>(l) * kFwBiasBlkBytes);
    for (int i = threadIdx.x; i < 2 * 256 * 8; i += blockDim.x) {
        const int g = i / (256 * 8), n = (i >> 3) & 255, e = i & 7, kin = g * 8 + e;
        const float b = l < 8 ? pr.p[2 * l + 1][n] : scratch[65536 + n];
        const float hi = half_value(half_bits(b, f16), f16);
        dst[i] = half_bits(kin == 0 ? hi : (kin == 1 ? b - hi : 0.f), f16);
    }
}

__global__ void pack_bias_kernel(Params pr, Image im, uint8_t* __restrict__ image) {
    const float* scratch = reinterpret_cast<const float*>(image + im.scratch_off);
    float* b = reinterpret_cast<float*>(image + im.bias_off);
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= kBiasFloats) return;
    float v = 0.f;
    if (i < 8 * 256) v = pr.p[2 * (i >> 8) + 1][i & 255];
    else if (i < 9 * 256) v = scratch[65536 + (i - 8 * 256)];
    else if (i < kBiasWd) {
        const int n = i - kBiasT9;
        if (n < 3) v = pr.p[P_RGBL_B][n];
        else if (n >= 4 && n < 5 + im.ins_num) v = pr.p[P_INSL_B][n - 4];
    } else if (i < kBiasBd) v = pr.p[P_DENS_W][i - kBiasWd];
    else if (i == kBiasBd) v = pr.p[P_DENS_B][0];
    b[i] = v;
}


// ---- the kernel -------------------------------------------------------------------------------------------------
// kStash (training): every A operand the backward needs (embeddings, h0..h7, [rgb_hid|ins_hid]) is mirrored
// to the per-tile bf16 stash (mlp_layout.cuh) by the same 16-byte stores that build the smem operand.
// cycle counters of block 0 when debug bit 4 is set (dmnerf_debug_counters): issuer 0 total, 1 wait w_full, 2 wait
// a_ready | epilogue warp 0: 3 total, 4 wait acc_ready | producer 5 wait w_empty
__device__ long long g_dbg1[8];

// kDens (inference): DENSITY ONLY -- the trunk T0..T7 and the density dot product, raw_out = sigma [n_points].  What a
// coarse render pass needs (its colours and instance logits are never looked at: tester.py:73-76 keeps *_fine only); the
// colour / instance layers T8, T9, the direction embedding and the [n_points, C] output are skipped.
template <bool kStash, bool kF16 = false, bool kDens = false>
__global__ void __launch_bounds__(kFwThreads, 1)
mlp_fwd_bf16_kernel(MlpInput in, const uint8_t* __restrict__ image, Image im, float* __restrict__ raw_out, int flags,
                    uint8_t* __restrict__ stash) {
    extern __shared__ __align__(1024) uint8_t smem[];
    const uint32_t sm0 = ptx::smem_u32(smem);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const FwBars bars{sm0 + kFwSmBar};
    float* s_misc = reinterpret_cast<float*>(smem + kFwSmMisc);      // [0,128) T9 bias | [128,384) w_density | [384] b_density
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(smem + kFwSmBar + 8 * (2 * kFwStages + 4));
    const int C = im.ins_num + 5;

    if (threadIdx.x == 0) {
        for (int s = 0; s < kFwStages; ++s) { ptx::mbar_init(bars.w_full(s), 1); ptx::mbar_init(bars.w_empty(s), 2); }
        for (int t = 0; t < 2; ++t) { ptx::mbar_init(bars.a_ready(t), 2 * kTile); ptx::mbar_init(bars.acc_ready(t), 1); }
        ptx::fence_barrier_init();
    }
    if (warp == kMmaWarp) ptx::tmem_alloc(ptx::smem_u32(const_cast<uint32_t*>(tmem_slot)), 512);
    {
        const float* gb = reinterpret_cast<const float*>(image + im.bias_off);
        for (int i = threadIdx.x; i < 128 + 256 + 1; i += kFwThreads) s_misc[i] = gb[kBiasT9 + i];
        uint32_t* ones = reinterpret_cast<uint32_t*>(smem + kFwSmOnes);
        for (int i = threadIdx.x; i < 32; i += kFwThreads) ones[i] = (i & 3) == 0 ? (kF16 ? 0x3C003C00u : 0x3F803F80u) : 0u;
    }
    ptx::fence_proxy_async_smem();
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem = *tmem_slot;

    const int64_t n_pairs = (in.n_points + kPair - 1) / kPair;

    if (warp == kProducerWarp) {
        // ===== weight producer: walks the chunk stream of every layer for every tile pair =====
        if (lane == 0) {
            int stage = 0; uint32_t phase = 0;
            long long dbg_w = 0;
            for (int64_t pair = blockIdx.x; pair < n_pairs; pair += gridDim.x) {
                for (int l = 0; l < (kDens ? 8 : kLayers); ++l) {
                    const uint32_t bytes = kFwChunkK * im.n[l] * 2;
                    const int chunks = layer_k(l) / kFwChunkK;
                    const uint8_t* src = image + im.chunk_off[l];
                    for (int c = 0; c < chunks + (l < 9); ++c) {
                        const long long w0 = clock64();
                        ptx::mbar_wait(bars.w_empty(stage), phase ^ 1);
                        dbg_w += clock64() - w0;
                        const uint32_t nb = c < chunks ? bytes : kFwBiasBlkBytes;
                        ptx::mbar_arrive_expect_tx(bars.w_full(stage), nb);
                        ptx::bulk_g2s(sm0 + kFwSmRing + stage * kFwStageBytes,
                                      c < chunks ? src + static_cast<int64_t>(c) * bytes
                                                 : image + im.fwb_off + static_cast<int64_t>(l) * kFwBiasBlkBytes,
                                      nb, bars.w_full(stage));
                        if (++stage == kFwStages) { stage = 0; phase ^= 1; }
                    }
                }
            }
            if ((flags & 16) && blockIdx.x == 0) g_dbg1[5] = dbg_w;
            // drain: every stage's last release (tcgen05.commit arrive) must have landed before the CTA exits
            for (int i = 0; i < kFwStages; ++i) {
                ptx::mbar_wait(bars.w_empty(stage), phase ^ 1);
                if (++stage == kFwStages) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp >= kMmaWarp) {
        // ===== MMA issuers: warp 17 drives tile 0, warp 18 tile 1; both walk the same ring (w_empty counts 2) =====
        if (lane == 0) {
            const int t = warp - kMmaWarp;
            int stage = 0; uint32_t phase = 0, a_par = 0;
            long long dbg_f = 0, dbg_a = 0;
            const long long dbg_start = clock64();
            // descriptors are built by adding 16-byte offsets to precomputed low words (K-major, no swizzle:
            // LBO = k-group stride, SBO = 128 B)
            constexpr uint32_t kKg16 = kKgBytes >> 4;
            const uint64_t desc_hi = static_cast<uint64_t>(0x4008u) << 32;
            const uint32_t a_lo0 = ((sm0 + kSmA + t * kABytes) >> 4) | (kKg16 << 16);
            // ones operand: every core matrix of the [128 x 16] A tile is the same 128 bytes (LBO = SBO = 0); both
            // k-groups then carry ones in columns 0,1, and the bias block is zero in rows k = 8.. so only k = 0,1 count
            const uint64_t ones_d = (static_cast<uint64_t>(0x4000u) << 32) | ((sm0 + kFwSmOnes) >> 4);
            const uint32_t ring_lo = (sm0 + kFwSmRing) >> 4;
            const uint32_t d_tmem = tmem + t * 256;
            for (int64_t pair = blockIdx.x; pair < n_pairs; pair += gridDim.x) {
                for (int l = 0; l < (kDens ? 8 : kLayers); ++l) {
                    const uint32_t N = im.n[l];
                    const uint32_t idesc = kF16 ? ptx::idesc_f16(kTile, N) : ptx::idesc_bf16(kTile, N);
                    const int chunks = layer_k(l) / kFwChunkK, last = chunks - (l == 9);   // T0..T8: + the bias step
                    uint32_t a_lo = a_lo0 + layer_akg0(l) * kKg16;
#pragma unroll 1
                    for (int c = 0; c <= last; ++c, a_lo += 4 * kKg16) {
                        const long long w0 = clock64();
                        ptx::mbar_wait(bars.w_full(stage), phase);
                        dbg_f += clock64() - w0;
                        if (c == 0) {
                            const long long w1 = clock64();
                            ptx::mbar_wait(bars.a_ready(t), a_par);
                            dbg_a += clock64() - w1;
                            a_par ^= 1;
                        }
                        ptx::tc_fence_after();
                        const uint32_t b_lo = (ring_lo + stage * (kFwStageBytes >> 4)) | (N << 16);
                        if (c < chunks) {
                            ptx::umma_bf16(d_tmem, desc_hi | a_lo, desc_hi | b_lo, idesc, c != 0);
                            ptx::umma_bf16(d_tmem, desc_hi | (a_lo + 2 * kKg16), desc_hi | (b_lo + 2 * N), idesc, 1u);
                        } else {
                            ptx::umma_bf16(d_tmem, ones_d, desc_hi | b_lo, idesc, 1u);
                        }
                        if (c == last) ptx::umma_commit(bars.acc_ready(t));
                        ptx::umma_commit(bars.w_empty(stage));
                        if (++stage == kFwStages) { stage = 0; phase ^= 1; }
                    }
                }
            }
            if ((flags & 16) && blockIdx.x == 0 && t == 0) { g_dbg1[0] = clock64() - dbg_start; g_dbg1[1] = dbg_f; g_dbg1[2] = dbg_a; }
        }
    } else {
        // ===== epilogue / embedding: 8 warps per tile; thread == (point == TMEM lane, column half) =====
        const int t = warp >> 3, q = warp & 3, half = (warp >> 2) & 1, row = q * 32 + lane;
        const uint32_t a_base = sm0 + kSmA + t * kABytes;
        const uint32_t t_base = tmem + t * 256 + (static_cast<uint32_t>(q * 32) << 16);
        float* stage_out = reinterpret_cast<float*>(smem + kSmA + t * kABytes);
        float* dens_part = reinterpret_cast<float*>(smem + kFwSmDens) + t * kTile;
        uint32_t acc_par = 0;
        long long dbg_acc = 0;
        const long long dbg_start = clock64();
        for (int64_t pair = blockIdx.x; pair < n_pairs; pair += gridDim.x) {
            const int64_t tile0 = pair * kPair + t * kTile;
            const int64_t p = tile0 + row;
            const bool valid = p < in.n_points;
            const int64_t sblk = (pair * 2 + t) * 2 + (row >> 6);
            uint8_t* gst = kStash ? stash + sblk * kStActBlkBytes + (row & 63) * 16 : nullptr;
            uint8_t* mst = kStash ? stash + tc_mask_offset(in.n_points) : nullptr;
            float dir[3] = {0.f, 0.f, 0.f};
            if (in.x == nullptr) {
                float pos[3] = {0.f, 0.f, 0.f};
                if (valid) {
                    const int64_t ray = p / in.S;
                    const float zz = __ldg(in.z + p);
                    float d[3];
#pragma unroll
                    for (int c = 0; c < 3; ++c) {
                        d[c] = __ldg(in.rays_d + ray * 3 + c);
                        pos[c] = __fadd_rn(__ldg(in.rays_o + ray * 3 + c), __fmul_rn(d[c], zz));
                    }
                    const float inv = 1.0f / sqrtf(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
#pragma unroll
                    for (int c = 0; c < 3; ++c) dir[c] = d[c] * inv;
                }
                if (half == 0) embed_to_ext<kPosFreqs, 0, 4, kF16>(a_base, row, pos, gst, kStPos);
                else           embed_to_ext<kPosFreqs, 4, 8, kF16>(a_base, row, pos, gst, kStPos);
            } else {
                row_to_ext<kF16>(a_base, row, valid ? in.x + p * 90 : nullptr, 63, half * 4, half * 4 + 4, gst, kStPos);
            }
            ptx::fence_proxy_async_smem();
            ptx::tc_fence_before();
            ptx::mbar_arrive(bars.a_ready(t));

            float dens = 0.f;
            for (int l = 0; l < kLayers - 1; ++l) {
                const long long w0 = clock64();
                ptx::mbar_wait(bars.acc_ready(t), acc_par);
                dbg_acc += clock64() - w0;
                acc_par ^= 1;
                ptx::tc_fence_after();
#pragma unroll 1
                for (int cb = half * 4; cb < half * 4 + 4; ++cb) {
                    uint32_t r[32];
                    ptx::tmem_ld32(t_base + cb * 32, r);
                    ptx::tmem_ld_wait();
                    uint32_t mword = 0;
#pragma unroll
                    for (int g = 0; g < 4; ++g) {
                        float v[8];
#pragma unroll
                        for (int e = 0; e < 8; ++e) v[e] = __uint_as_float(r[g * 8 + e]);
                        if (l == 7) {
#pragma unroll
                            for (int e = 0; e < 8; ++e) dens = fmaf(fmaxf(v[e], 0.f), s_misc[128 + cb * 32 + g * 8 + e], dens);
                        }
                        if constexpr (kStash)
                            mword = store_kg_mask(a_base, cb * 4 + g, row, v, gst, (l == 8 ? kStHid : kStH0 + 32 * l) + cb * 4 + g, g, mword);
                        else if (!(kDens && l == 7))          // density only: nobody reads h7 as an operand
                            store_kg<kF16>(a_base, cb * 4 + g, row, v, true);
                    }
                    // one coalesced 128-byte line of sign masks per warp and 32 columns (read by the dgrad chain)
                    if constexpr (kStash) *reinterpret_cast<uint32_t*>(mst + tc_mask_word(sblk, l, half, cb & 3, row & 63)) = mword;
                }
                if constexpr (kDens) {
                    if (l == 7) {
                        // sigma = w_density . relu(h7) + b_density: the two column halves meet in shared memory
                        if (half == 1) dens_part[row] = dens;
                        ptx::named_bar_sync(1 + t, 2 * kTile);
                        if (half == 0 && valid) raw_out[p] = (dens + dens_part[row]) + s_misc[384];
                        break;                                  // T8 / T9 are not evaluated
                    }
                } else if (l == 7) {
                    // dir-embedding replaces the pos-embedding in EXT (T5 is long done with it)
                    if (half == 0) {
                        if (in.x == nullptr) embed_to_ext<kDirFreqs, 0, 4, kF16>(a_base, row, dir, gst, kStDir);
                        else row_to_ext<kF16>(a_base, row, valid ? in.x + p * 90 + 63 : nullptr, 27, 0, 4, gst, kStDir);
                    } else {
                        dens_part[row] = dens;
                    }
                }
                ptx::fence_proxy_async_smem();
                ptx::tc_fence_before();
                ptx::mbar_arrive(bars.a_ready(t));
            }
            if constexpr (kDens) continue;
            // ---- T9: heads -> fp32 raw row [rgb3 | sigma | ins K+1], staged in smem, bulk-stored (half 0) ----
            const long long w9 = clock64();
            ptx::mbar_wait(bars.acc_ready(t), acc_par);
            dbg_acc += clock64() - w9;
            acc_par ^= 1;
            ptx::tc_fence_after();
            if (half == 0) {
                dens += dens_part[row];
                for (int cb = 0; cb * 16 < im.npad; ++cb) {
                    uint32_t r[16];
                    ptx::tmem_ld16(t_base + cb * 16, r);
                    ptx::tmem_ld_wait();
#pragma unroll
                    for (int e = 0; e < 16; ++e) {
                        const int col = cb * 16 + e;
                        if (col < C) {
                            float v = __uint_as_float(r[e]) + s_misc[col];
                            if (col == 3) v = dens + s_misc[384];
                            stage_out[row * C + col] = v;
                        }
                    }
                }
                ptx::tc_fence_before();
                ptx::fence_proxy_async_smem();
                ptx::named_bar_sync(1 + t, kTile);
                const int64_t rows_left = in.n_points - tile0;
                if (rows_left >= kTile) {
                    if (row == 0) {
                        ptx::bulk_s2g(raw_out + tile0 * C, ptx::smem_u32(stage_out), kTile * C * 4);
                        ptx::bulk_commit();
                        ptx::bulk_wait_read0();
                    }
                } else if (rows_left > 0) {
                    const int total = static_cast<int>(rows_left) * C;
                    for (int i = row; i < total; i += kTile) raw_out[tile0 * C + i] = stage_out[i];
                }
                // the next write into this region is T0's epilogue of the next pair, which cannot start before all
                // 256 threads (incl. the bulk-store issuer, after its read-wait) arrived on a_ready again.
                ptx::named_bar_sync(1 + t, kTile);
            }
        }
        if ((flags & 16) && blockIdx.x == 0 && threadIdx.x == 0) { g_dbg1[3] = clock64() - dbg_start; g_dbg1[4] = dbg_acc; }
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == kMmaWarp) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc(tmem, 512);
    }
}

// ---- bring-up self test: D[128,n] = A[128,k] * B[n,k]^T through the same descriptors ------------------------
__global__ void __launch_bounds__(128, 1)
umma_selftest_kernel(const float* __restrict__ a, const float* __restrict__ b, int n, int k, float* __restrict__ d,
                     int flags) {
    extern __shared__ __align__(1024) uint8_t smem[];
    const uint32_t sm0 = ptx::smem_u32(smem);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, row = threadIdx.x;
    const uint32_t a_base = sm0, b_base = sm0 + 65536, bar = sm0 + 65536 + 131072;
    volatile uint32_t* slot = reinterpret_cast<volatile uint32_t*>(smem + 65536 + 131072 + 8);
    __nv_bfloat16* sa = reinterpret_cast<__nv_bfloat16*>(smem);
    __nv_bfloat16* sb = reinterpret_cast<__nv_bfloat16*>(smem + 65536);
    for (int i = threadIdx.x; i < 128 * k; i += 128) {
        const int r = i / k, c = i - r * k;
        sa[((c >> 3) * 128 + r) * 8 + (c & 7)] = __float2bfloat16_rn(a[i]);
    }
    for (int i = threadIdx.x; i < n * k; i += 128) {
        const int r = i / k, c = i - r * k;
        sb[((c >> 3) * n + r) * 8 + (c & 7)] = __float2bfloat16_rn(b[i]);
    }
    if (threadIdx.x == 0) { ptx::mbar_init(bar, 1); ptx::fence_barrier_init(); }
    if (warp == 0) ptx::tmem_alloc(ptx::smem_u32(const_cast<uint32_t*>(slot)), 256);
    ptx::fence_proxy_async_smem();
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem = *slot;
    if (threadIdx.x == 0) {
        const uint32_t idesc = ptx::idesc_bf16(128, n);
        for (int ks = 0; ks < k / 16; ++ks) {
            const uint64_t ad = op_desc(a_base + ks * 2 * kKgBytes, kKgBytes, 128, flags & 1);
            const uint64_t bd = op_desc(b_base + ks * 2 * (n * 16), n * 16, 128, flags & 1);
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
    (void)lane;
}

}  // namespace tc

#ifdef DMNERF_EXPERIMENTAL
int launch_mlp_bf16_v2(const MlpInput& in, const void* packed, int ins_num, float* raw_out, void* stash, cudaStream_t st);
int launch_mlp_bf16_v3(const MlpInput& in, const void* packed, int ins_num, float* raw_out, void* stash, cudaStream_t st);
int launch_mlp_bf16_v4(const MlpInput& in, const void* packed, int ins_num, float* raw_out, void* stash, cudaStream_t st);
#endif
int launch_masks_from_acts(void* acts, int64_t n_points, cudaStream_t st);
int persistent_ctas(int sm_count);

int launch_mlp_bf16(const MlpInput& in, const void* packed, int ins_num, float* raw_out, void* stash, cudaStream_t st,
                    bool f16, bool dens) {
    DM_REQUIRE(!(dens && stash), DMNERF_EUNSUPPORTED, "mlp_density: inference only (no activation stash)");
    DM_REQUIRE(!(f16 && stash), DMNERF_EUNSUPPORTED,
               "mlp_fwd(f16): the fp16 chain is inference-only (no activation stash; train in DMNERF_MODE_BF16)");
    // debug bits 2/3/5 select the parked variants of csrc/experimental/ (built only with -DDMNERF_EXPERIMENTAL; all
    // correct, none faster than this kernel -- measurements in DESIGN.md 4.1)
#ifndef DMNERF_EXPERIMENTAL
    DM_REQUIRE(!(g_tc_debug_flags & (4 | 8 | 32)), DMNERF_EUNSUPPORTED,
               "mlp_fwd(bf16): forward variants are not in this build (compile with -DDMNERF_EXPERIMENTAL)");
#else
    if (!f16 && (g_tc_debug_flags & (4 | 8 | 32))) {
        const int rc = (g_tc_debug_flags & 4)   ? launch_mlp_bf16_v2(in, packed, ins_num, raw_out, stash, st)
                       : (g_tc_debug_flags & 8) ? launch_mlp_bf16_v3(in, packed, ins_num, raw_out, stash, st)
                                                : launch_mlp_bf16_v4(in, packed, ins_num, raw_out, stash, st);
        // the variants stash the activations only; the dgrad chain's sign masks are derived from them here
        return (rc == DMNERF_OK && stash) ? launch_masks_from_acts(stash, in.n_points, st) : rc;
    }
#endif
    static int sm_count = 0;
    if (sm_count == 0) {
        int dev = 0;
        DM_CUDA(cudaGetDevice(&dev));
        DM_CUDA(cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev));
    }
    DM_REQUIRE((reinterpret_cast<uintptr_t>(packed) & 15) == 0 && (reinterpret_cast<uintptr_t>(raw_out) & 15) == 0,
               DMNERF_EINVAL, "mlp_fwd(bf16): packed image and raw_out must be 16-byte aligned");
    const int64_t pairs = (in.n_points + tc::kPair - 1) / tc::kPair;
    const int sms = persistent_ctas(sm_count);
    const unsigned grid = static_cast<unsigned>(pairs < sms ? pairs : sms);
    if (stash) {
        DM_REQUIRE((reinterpret_cast<uintptr_t>(stash) & 15) == 0, DMNERF_EINVAL, "mlp_fwd(bf16): stash must be 16-byte aligned");
        DM_CUDA(cudaFuncSetAttribute(tc::mlp_fwd_bf16_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, tc::kFwSmTotal));
        tc::mlp_fwd_bf16_kernel<true><<<grid, tc::kFwThreads, tc::kFwSmTotal, st>>>(
            in, static_cast<const uint8_t*>(packed), tc::make_image(ins_num), raw_out, g_tc_debug_flags,
            static_cast<uint8_t*>(stash));
    } else if (dens) {
        if (f16) {
            DM_CUDA(cudaFuncSetAttribute(tc::mlp_fwd_bf16_kernel<false, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, tc::kFwSmTotal));
            tc::mlp_fwd_bf16_kernel<false, true, true><<<grid, tc::kFwThreads, tc::kFwSmTotal, st>>>(
                in, static_cast<const uint8_t*>(packed), tc::make_image(ins_num), raw_out, g_tc_debug_flags, nullptr);
        } else {
            DM_CUDA(cudaFuncSetAttribute(tc::mlp_fwd_bf16_kernel<false, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, tc::kFwSmTotal));
            tc::mlp_fwd_bf16_kernel<false, false, true><<<grid, tc::kFwThreads, tc::kFwSmTotal, st>>>(
                in, static_cast<const uint8_t*>(packed), tc::make_image(ins_num), raw_out, g_tc_debug_flags, nullptr);
        }
    } else if (f16) {
        DM_CUDA(cudaFuncSetAttribute(tc::mlp_fwd_bf16_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, tc::kFwSmTotal));
        tc::mlp_fwd_bf16_kernel<false, true><<<grid, tc::kFwThreads, tc::kFwSmTotal, st>>>(
            in, static_cast<const uint8_t*>(packed), tc::make_image(ins_num), raw_out, g_tc_debug_flags, nullptr);
    } else {
        DM_CUDA(cudaFuncSetAttribute(tc::mlp_fwd_bf16_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, tc::kFwSmTotal));
        tc::mlp_fwd_bf16_kernel<false><<<grid, tc::kFwThreads, tc::kFwSmTotal, st>>>(
            in, static_cast<const uint8_t*>(packed), tc::make_image(ins_num), raw_out, g_tc_debug_flags, nullptr);
    }
    DM_LAUNCH_CHECK("mlp_fwd_bf16_kernel");
    return DMNERF_OK;
}

int tc_debug_flags() { return g_tc_debug_flags; }
// CTAs a persistent one-CTA-per-SM kernel may launch: all SMs minus the reservation (never fewer than one)
int persistent_ctas(int sm_count) { return sm_count - g_reserved_sms > 1 ? sm_count - g_reserved_sms : 1; }

int debug_counters_v1(long long* host_out8) {
    DM_CUDA(cudaMemcpyFromSymbol(host_out8, tc::g_dbg1, 8 * sizeof(long long)));
    return DMNERF_OK;
}

#ifdef DMNERF_EXPERIMENTAL
namespace tc2 { int pack_v2(const float* const* params, int ins_num, void* packed, cudaStream_t st); }
namespace tc3 { int pack_v3(const float* const* params, int ins_num, void* packed, cudaStream_t st); }
#endif

int64_t bf16_image_bytes(int ins_num) { return tc::make_image(ins_num).total_bytes; }

int pack_bf16(const float* const* params, int ins_num, void* packed, cudaStream_t st, bool f16) {
    const tc::Image im = tc::make_image(ins_num);
    tc::Params pr;
    for (int i = 0; i < kNumParams; ++i) pr.p[i] = params[i];
    uint8_t* image = static_cast<uint8_t*>(packed);
    float* scratch = reinterpret_cast<float*>(image + im.scratch_off);
    tc::compose_kernel<<<dim3(16, 8, 3), 256, 0, st>>>(pr, scratch);
    const int64_t elems = im.bias_off / 2;
    tc::pack_bf16_kernel<<<static_cast<unsigned>((elems + 255) / 256), 256, 0, st>>>(pr, im, image, f16);
    tc::pack_bias_kernel<<<(tc::kBiasFloats + 255) / 256, 256, 0, st>>>(pr, im, image);
    tc::pack_fwbias_kernel<<<9, 256, 0, st>>>(pr, im, image, f16);
    if (f16) {                      // inference-only image: no dgrad streams, no fp32 copies
        DM_LAUNCH_CHECK("pack_f16");
        return DMNERF_OK;
    }
    const int64_t dg_elems = (im.f32_off - im.dg_off[0]) / 2;
    tc::pack_dgrad_kernel<<<static_cast<unsigned>((dg_elems + 255) / 256), 256, 0, st>>>(pr, im, image);
    tc::pack_f32_kernel<<<(tc::kF32Floats + 255) / 256, 256, 0, st>>>(pr, im, image);
    DM_LAUNCH_CHECK("pack_bf16");
    // the operand streams of the parked forward variants are only written while one of them is selected (the host
    // cache key includes that bit): training repacks every step and should not pay for them
#ifdef DMNERF_EXPERIMENTAL
    if (!(g_tc_debug_flags & (4 | 8 | 32))) return DMNERF_OK;
    if (int rc = tc2::pack_v2(params, ins_num, packed, st)) return rc;
    return tc3::pack_v3(params, ins_num, packed, st);
#else
    return DMNERF_OK;
#endif
}

}  // namespace dmnerf

using namespace dmnerf;

#ifdef DMNERF_EXPERIMENTAL
namespace dmnerf {
namespace tc2 { int debug_counters_v2(long long* host_out8); }
namespace tc3 { int debug_counters_v3(long long* host_out8); }
namespace tc4 { int debug_counters_v4(long long* host_out8); }
}
#endif

// cycle counters of block 0 of the last forward launch that ran with debug bit 4 (or a parked variant's bit) set
extern "C" int dmnerf_debug_counters(long long* host_out8) {
    DM_REQUIRE(host_out8, DMNERF_EINVAL, "debug_counters: null pointer");
    DM_CUDA(cudaDeviceSynchronize());
#ifdef DMNERF_EXPERIMENTAL
    if (g_tc_debug_flags & 32) return tc4::debug_counters_v4(host_out8);
    if (g_tc_debug_flags & 8) return tc3::debug_counters_v3(host_out8);
    if (g_tc_debug_flags & 4) return tc2::debug_counters_v2(host_out8);
#endif
    return debug_counters_v1(host_out8);
}

extern "C" int dmnerf_debug_flags(int flags) {
    const int old = g_tc_debug_flags;
    if (flags >= 0) g_tc_debug_flags = flags;
    return old;
}

extern "C" int dmnerf_reserve_sms(int n) {
    const int old = g_reserved_sms;
    if (n >= 0) g_reserved_sms = n;
    return old;
}

extern "C" int dmnerf_selftest_umma(const float* a, const float* b, int n, int k, float* d, void* stream) {
    DM_REQUIRE(a && b && d, DMNERF_EINVAL, "selftest_umma: null pointer");
    DM_REQUIRE(n >= 16 && n <= 256 && n % 16 == 0 && k >= 16 && k <= 256 && k % 16 == 0, DMNERF_EUNSUPPORTED,
               "selftest_umma: n=%d k=%d (need multiples of 16, <= 256)", n, k);
    const int smem = 65536 + 131072 + 64;
    DM_CUDA(cudaFuncSetAttribute(tc::umma_selftest_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    tc::umma_selftest_kernel<<<1, 128, smem, as_stream(stream)>>>(a, b, n, k, d, g_tc_debug_flags);
    DM_LAUNCH_CHECK("umma_selftest_kernel");
    return DMNERF_OK;
}

// ---- micro-benchmark: TMEM -> register read throughput (tcgen05.ld 32x32b.x32), `warps` warps (multiple of 4)
// each reading its lane quarter of all 512 columns `iters` times.  out[0] = cycles, out[1] = bytes read per CTA.
namespace dmnerf { namespace tc {
__global__ void __launch_bounds__(512, 1)
tmem_read_bw_kernel(int iters, int batched, long long* __restrict__ out) {
    __shared__ uint32_t slot;
    const int warp = threadIdx.x >> 5;
    if (warp == 0) ptx::tmem_alloc(ptx::smem_u32(&slot), 512);
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem = slot;
    const uint32_t base = tmem + (static_cast<uint32_t>((warp & 3) * 32) << 16);
    const int nw = blockDim.x >> 5, groups = nw >> 2, grp = warp >> 2;
    const int c_lo = 512 / groups * grp, c_hi = c_lo + 512 / groups;
    uint32_t sink = 0;
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
        if (batched) {
            for (int c = c_lo; c < c_hi; c += 64) {
                uint32_t r0[32], r1[32];
                ptx::tmem_ld32(base + c, r0);
                ptx::tmem_ld32(base + c + 32, r1);
                ptx::tmem_ld_wait();
#pragma unroll
                for (int e = 0; e < 32; ++e) sink ^= r0[e] ^ r1[e];
            }
        } else {
            for (int c = c_lo; c < c_hi; c += 32) {
                uint32_t r[32];
                ptx::tmem_ld32(base + c, r);
                ptx::tmem_ld_wait();
#pragma unroll
                for (int e = 0; e < 32; ++e) sink ^= r[e];
            }
        }
    }
    __syncthreads();
    const long long t1 = clock64();
    if (threadIdx.x == 0 && blockIdx.x == 0) { out[0] = t1 - t0; out[1] = static_cast<long long>(iters) * 512 * 128 * 4; }
    if (sink == 0x12345678u) out[2] = sink;
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 0) { ptx::tc_fence_after(); ptx::tmem_dealloc(tmem, 512); }
}
}}  // namespace dmnerf::tc

extern "C" int dmnerf_selftest_tmem_read_bw(int warps, int iters, int batched, long long* out, void* stream) {
    DM_REQUIRE(out && warps >= 4 && warps <= 16 && warps % 4 == 0, DMNERF_EINVAL, "selftest_tmem_read_bw: warps must be 4, 8, 12 or 16");
    dmnerf::tc::tmem_read_bw_kernel<<<1, warps * 32, 0, as_stream(stream)>>>(iters, batched, out);
    DM_LAUNCH_CHECK("tmem_read_bw_kernel");
    return DMNERF_OK;
}

// ---- micro-benchmark: back-to-back tcgen05.mma issue rate for M=128, N=n, K=16 (bf16), one CTA.
// out[0] = cycles for `count` MMAs (committed and waited for).
namespace dmnerf { namespace tc {
__global__ void __launch_bounds__(128, 1)
umma_rate_kernel(int n, int count, long long* __restrict__ out) {
    extern __shared__ __align__(1024) uint8_t smem[];
    const uint32_t sm0 = ptx::smem_u32(smem);
    const int warp = threadIdx.x >> 5;
    const uint32_t a_base = sm0, b_base = sm0 + 65536, bar = sm0 + 65536 + 131072;
    volatile uint32_t* slot = reinterpret_cast<volatile uint32_t*>(smem + 65536 + 131072 + 8);
    for (int i = threadIdx.x; i < (65536 + 131072) / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0;
    if (threadIdx.x == 0) { ptx::mbar_init(bar, 1); ptx::fence_barrier_init(); }
    if (warp == 0) ptx::tmem_alloc(ptx::smem_u32(const_cast<uint32_t*>(slot)), 512);
    ptx::fence_proxy_async_smem();
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem = *slot;
    if (threadIdx.x == 0) {
        const uint32_t idesc = ptx::idesc_bf16(128, n);
        // descriptors precomputed and the issue loop unrolled: measures the tensor pipe, not this thread's
        // instruction latency (a naive loop with descriptor construction issues only one MMA per ~72 cycles)
        uint64_t ad[16], bd[16];
#pragma unroll
        for (int ks = 0; ks < 16; ++ks) {
            ad[ks] = ptx::smem_desc(a_base + ks * 2 * kKgBytes, kKgBytes, 128);
            bd[ks] = ptx::smem_desc(b_base + ks * 2 * (n * 16), n * 16, 128);
        }
        const long long t0 = clock64();
        for (int i = 0; i < count; i += 16) {
            const uint32_t d = tmem + ((i >> 4) & 1) * 256;
#pragma unroll
            for (int ks = 0; ks < 16; ++ks) ptx::umma_bf16(d, ad[ks], bd[ks], idesc, ks != 0);
        }
        ptx::umma_commit(bar);
        ptx::mbar_wait(bar, 0);
        out[0] = clock64() - t0;
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 0) { ptx::tc_fence_after(); ptx::tmem_dealloc(tmem, 512); }
}
}}  // namespace dmnerf::tc

extern "C" int dmnerf_selftest_umma_rate(int n, int count, long long* out, void* stream) {
    DM_REQUIRE(out && n >= 16 && n <= 256 && n % 16 == 0 && count > 0, DMNERF_EINVAL, "selftest_umma_rate: bad arguments");
    const int smem = 65536 + 131072 + 64;
    DM_CUDA(cudaFuncSetAttribute(dmnerf::tc::umma_rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    dmnerf::tc::umma_rate_kernel<<<1, 128, smem, as_stream(stream)>>>(n, count, out);
    DM_LAUNCH_CHECK("umma_rate_kernel");
    return DMNERF_OK;
}

// ---- micro-benchmark: what the issuing thread pays per ring step besides the MMAs themselves.  Groups of two
// M=128,N=16,K=16 MMAs (44 clk each in the pipe) plus, by `mode` bit: 1 a tcgen05.commit, 2 an mbarrier wait that
// succeeds at once, 4 tcgen05.fence::after_thread_sync, 8 a clock64() read.  out[0] = cycles for `count` groups.
namespace dmnerf { namespace tc {
__global__ void __launch_bounds__(128, 1)
issue_cost_kernel(int mode, int count, long long* __restrict__ out) {
    extern __shared__ __align__(1024) uint8_t smem[];
    const uint32_t sm0 = ptx::smem_u32(smem);
    const int warp = threadIdx.x >> 5;
    const uint32_t a_base = sm0, b_base = sm0 + 8192, bar = sm0 + 16384;
    volatile uint32_t* slot = reinterpret_cast<volatile uint32_t*>(smem + 16384 + 32);
    for (int i = threadIdx.x; i < 16384 / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0;
    if (threadIdx.x == 0) { ptx::mbar_init(bar, 1); ptx::mbar_init(bar + 8, 1); ptx::mbar_init(bar + 16, 1); ptx::fence_barrier_init(); }
    if (warp == 0) ptx::tmem_alloc(ptx::smem_u32(const_cast<uint32_t*>(slot)), 512);
    ptx::fence_proxy_async_smem();
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem = *slot;
    const uint32_t idesc = ptx::idesc_bf16(128, 16);
    const uint64_t ad0 = ptx::smem_desc(a_base, kKgBytes, 128), ad1 = ptx::smem_desc(a_base + 2 * kKgBytes, kKgBytes, 128);
    const uint64_t bd0 = ptx::smem_desc(b_base, 256, 128), bd1 = ptx::smem_desc(b_base + 512, 256, 128);
    if ((mode & 16) && warp == 0) {
        // whole warp walks the loop; one elected lane issues
        long long sink = 0;
        const long long t0 = clock64();
#pragma unroll 1
        for (int i = 0; i < count; ++i) {
            if (mode & 8) sink += clock64();
            if (mode & 2) ptx::mbar_wait(bar + 8, 1);
            if (mode & 4) ptx::tc_fence_after();
            if (ptx::elect_one()) {
                ptx::umma_bf16(tmem, ad0, bd0, idesc, (mode & 32) ? 1u : 0u);
                ptx::umma_bf16(tmem, ad1, bd1, idesc, 1u);
                if (mode & 64) {
                    ptx::umma_bf16(tmem + ((mode & 128) ? 256 : 0), ad0, bd0, idesc, 1u);
                    ptx::umma_bf16(tmem + ((mode & 128) ? 256 : 0), ad1, bd1, idesc, 1u);
                }
                if (mode & 1) ptx::umma_commit(bar + 16);
            }
            __syncwarp();
        }
        if (ptx::elect_one()) ptx::umma_commit(bar);
        __syncwarp();
        ptx::mbar_wait(bar, 0);
        if (threadIdx.x == 0) { out[0] = clock64() - t0; out[1] = sink; }
    } else if (!(mode & 16) && threadIdx.x == 0) {
        long long sink = 0;
        const long long t0 = clock64();
#pragma unroll 1
        for (int i = 0; i < count; ++i) {
            if (mode & 8) sink += clock64();
            if (mode & 2) ptx::mbar_wait(bar + 8, 1);          // fresh barrier: the "previous" phase counts as complete
            if (mode & 4) ptx::tc_fence_after();
            ptx::umma_bf16(tmem, ad0, bd0, idesc, 0u);
            ptx::umma_bf16(tmem, ad1, bd1, idesc, 1u);
            if (mode & 1) ptx::umma_commit(bar + 16);
        }
        ptx::umma_commit(bar);
        ptx::mbar_wait(bar, 0);
        out[0] = clock64() - t0;
        out[1] = sink;
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 0) { ptx::tc_fence_after(); ptx::tmem_dealloc(tmem, 512); }
}
}}  // namespace dmnerf::tc

extern "C" int dmnerf_selftest_issue_cost(int mode, int count, long long* out, void* stream) {
    DM_REQUIRE(out && count > 0, DMNERF_EINVAL, "selftest_issue_cost: bad arguments");
    const int smem = 16384 + 64;
    dmnerf::tc::issue_cost_kernel<<<1, 128, smem, as_stream(stream)>>>(mode, count, out);
    DM_LAUNCH_CHECK("issue_cost_kernel");
    return DMNERF_OK;
}

// ---- bring-up self test of the CTA-pair MMA: D[256,n] = A[256,k] * B[n,k]^T, cluster of 2 CTAs; CTA r holds rows
// [128r, 128r+128) of A and rows [n/2 r, n/2 (r+1)) of B (K-major, no swizzle), and receives rows [128r, ..) of D.
namespace dmnerf { namespace tc {
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128, 1)
umma2_selftest_kernel(const float* __restrict__ a, const float* __restrict__ b, int n, int k, float* __restrict__ d) {
    extern __shared__ __align__(1024) uint8_t smem[];
    const uint32_t sm0 = ptx::smem_u32(smem);
    const uint32_t rank = ptx::cluster_ctarank();
    const int warp = threadIdx.x >> 5, row = threadIdx.x;
    const int nh = n / 2;
    const uint32_t a_base = sm0, b_base = sm0 + 65536, bar = sm0 + 65536 + 65536;
    volatile uint32_t* slot = reinterpret_cast<volatile uint32_t*>(smem + 65536 + 65536 + 8);
    __nv_bfloat16* sa = reinterpret_cast<__nv_bfloat16*>(smem);
    __nv_bfloat16* sb = reinterpret_cast<__nv_bfloat16*>(smem + 65536);
    for (int i = threadIdx.x; i < 128 * k; i += 128) {
        const int r = i / k, c = i - r * k;
        sa[((c >> 3) * 128 + r) * 8 + (c & 7)] = __float2bfloat16_rn(a[(rank * 128 + r) * k + c]);
    }
    for (int i = threadIdx.x; i < nh * k; i += 128) {
        const int r = i / k, c = i - r * k;
        sb[((c >> 3) * nh + r) * 8 + (c & 7)] = __float2bfloat16_rn(b[(rank * nh + r) * k + c]);
    }
    if (threadIdx.x == 0) { ptx::mbar_init(bar, 1); ptx::fence_barrier_init(); }
    if (warp == 0) ptx::tmem_alloc2(ptx::smem_u32(const_cast<uint32_t*>(slot)), 256);
    ptx::fence_proxy_async_smem();
    ptx::tc_fence_before();
    ptx::cluster_sync();
    ptx::tc_fence_after();
    const uint32_t tmem = *slot;
    if (rank == 0 && threadIdx.x == 0) {
        const uint32_t idesc = ptx::idesc_bf16(256, n);
        for (int ks = 0; ks < k / 16; ++ks) {
            const uint64_t ad = ptx::smem_desc(a_base + ks * 2 * kKgBytes, kKgBytes, 128);
            const uint64_t bd = ptx::smem_desc(b_base + ks * 2 * (nh * 16), nh * 16, 128);
            ptx::umma2_bf16(tmem, ad, bd, idesc, ks != 0);
        }
        ptx::umma_commit2(bar, 3);
    }
    ptx::mbar_wait(bar, 0);
    ptx::tc_fence_after();
    for (int cb = 0; cb * 16 < n; ++cb) {
        uint32_t r[16];
        ptx::tmem_ld16(tmem + (static_cast<uint32_t>(warp * 32) << 16) + cb * 16, r);
        ptx::tmem_ld_wait();
        for (int e = 0; e < 16; ++e) d[(rank * 128 + row) * n + cb * 16 + e] = __uint_as_float(r[e]);
    }
    ptx::tc_fence_before();
    ptx::cluster_sync();
    if (warp == 0) { ptx::tc_fence_after(); ptx::tmem_dealloc2(tmem, 256); }
}
}}  // namespace dmnerf::tc

extern "C" int dmnerf_selftest_umma2(const float* a, const float* b, int n, int k, float* d, void* stream) {
    DM_REQUIRE(a && b && d, DMNERF_EINVAL, "selftest_umma2: null pointer");
    DM_REQUIRE(n >= 32 && n <= 256 && n % 32 == 0 && k >= 16 && k <= 256 && k % 16 == 0, DMNERF_EUNSUPPORTED,
               "selftest_umma2: n=%d k=%d (n multiple of 32 <= 256, k multiple of 16 <= 256)", n, k);
    const int smem = 65536 + 65536 + 64;
    DM_CUDA(cudaFuncSetAttribute(dmnerf::tc::umma2_selftest_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    dmnerf::tc::umma2_selftest_kernel<<<2, 128, smem, as_stream(stream)>>>(a, b, n, k, d);
    DM_LAUNCH_CHECK("umma2_selftest_kernel");
    return DMNERF_OK;
}
