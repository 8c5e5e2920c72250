// mlp_tc_x3.cu -- a1+a2+a9 in DMNERF_MODE_TC32: the fp32-accurate tensor-core chain.  Same network walk as
// mlp_tc.cu (T0..T9, the two linear->linear pairs composed at pack time) but every operand is an fp16 hi+lo
// PAIR (x = hi + lo, 22 significand bits) and every layer is three tcgen05 products into ONE fp32 accumulator:
//     D = A_hi W_hi^T + A_lo W_hi^T + A_hi W_lo^T + bias (a three-term fp16 split against a ones operand)
// (the dropped A_lo W_lo^T term is ~2^-22 relative).  This is the "3x split" SURVEY 7 hard-part 1 names: it
// holds the reference's fp32 forward (networks/dm_nerf.py:80-106) to ~1e-6 where the bf16 / fp16 chains sit at
// 2e-3 / 2e-4, at 3x the MMA count instead of the CUDA-core kernel's 34x the time.
//
// Work decomposition: one CTA per SM, ONE 128-point tile at a time (the hi and lo images of the activations fill
// the two A regions the pair kernel uses for its two tiles).  Warps 0-15 epilogue (thread == TMEM lane x 16-column
// slot of every output quarter), warp 16 weight producer, warp 17 the issuer.  The weight stream interleaves
// [chunk c hi][chunk c lo] through a 4 x 16 KB ring: a hi stage feeds 4 MMAs, a lo stage 2.  The accumulator
// alternates between the two halves of TMEM from layer to layer and the epilogue hands the output on in four
// 64-column quarters (a_ready[j]), so the next layer's products start one quarter into the epilogue (see the
// comment at the kernel).  Embeddings: sin/cos of x*2^f evaluated per octave in fp32 (no angle doubling -- its
// error grows as 2^f ulp).  Inference only (no activation stash); kDens = density-only coarse pass.
#include "mlp_tc_common.cuh"
#include <cuda_fp16.h>

namespace dmnerf {
namespace tc {

__global__ void compose_kernel(Params pr, float* __restrict__ scratch);                 // mlp_tc.cu
__global__ void pack_bias_kernel(Params pr, Image im, uint8_t* __restrict__ image);     // mlp_tc.cu

namespace x3 {

constexpr int kThreads = 19 * 32;
constexpr int kAHi = 0, kALo = kABytes;                        // two 81,920-byte operand images
constexpr int kDensKg = 36;                                    // density partials [4][128] fp32 live in k-group 36 of
                                                               //   A_hi between T7's epilogue and T9's (T8 reads k-groups 0..35)
// Image map of this mode (fields of tc::Image reused): chunk_off[l] -> [c][hi|lo][k-group][n][8] fp16, fwb_off ->
// 9 bias blocks whose rows k = 0,1,2 hold a three-term fp16 split of the bias, bias_off / scratch_off as in mlp_tc.cu.
__host__ __device__ inline Image make_image_x3(int ins_num) {
    Image im = make_image(ins_num);
    int64_t off = 0;
    for (int l = 0; l < kLayers; ++l) {
        im.chunk_off[l] = off;
        off += 2 * static_cast<int64_t>(layer_k(l)) * im.n[l] * 2;
    }
    im.bias_off = off;       off += kBiasFloats * 4;
    im.scratch_off = off;    off += (2 * 128 * 256 + 256) * 4;
    im.fwb_off = off;        off += 9 * kFwBiasBlkBytes;
    for (int d = 0; d < kDgLayers; ++d) im.dg_off[d] = 0;
    im.f32_off = 0;
    im.total_bytes = off;
    return im;
}

__device__ __forceinline__ float f16_round(float v) { return __half2float(__float2half_rn(v)); }

// one thread per fp16 element of the interleaved stream: layer l, chunk c, part (hi|lo), k-group g, row n, element e
__global__ void pack_chunks_kernel(Params pr, Image im, uint8_t* __restrict__ image) {
    const float* scratch = reinterpret_cast<const float*>(image + im.scratch_off);
    const int64_t idx = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (idx >= im.bias_off / 2) return;
    int l = 0;
    while (l + 1 < kLayers && idx * 2 >= im.chunk_off[l + 1]) ++l;
    const int N = im.n[l];
    int64_t r = idx - im.chunk_off[l] / 2;
    const int per_part = kFwChunkK * N;
    const int cp = static_cast<int>(r / per_part);
    r -= static_cast<int64_t>(cp) * per_part;
    const int c = cp >> 1, part = cp & 1;
    const int g = static_cast<int>(r / (N * 8));
    r -= static_cast<int64_t>(g) * N * 8;
    const int n = static_cast<int>(r >> 3), e = static_cast<int>(r & 7);
    const float w = logical_weight(pr, scratch, l, n, c * kFwChunkK + g * 8 + e, im.ins_num);
    const float hi = f16_round(w);
    reinterpret_cast<__half*>(image)[idx] = __float2half_rn(part == 0 ? hi : w - hi);
}

__global__ void pack_bias3_kernel(Params pr, Image im, uint8_t* __restrict__ image) {
    const float* scratch = reinterpret_cast<const float*>(image + im.scratch_off);
    const int l = blockIdx.x;
    __half* dst = reinterpret_cast<__half*>(image + im.fwb_off + static_cast<int64_t>(l) * kFwBiasBlkBytes);
    for (int i = threadIdx.x; i < 2 * 256 * 8; i += blockDim.x) {
        const int g = i / (256 * 8), n = (i >> 3) & 255, e = i & 7, kin = g * 8 + e;
        const float b = l < 8 ? pr.p[2 * l + 1][n] : scratch[65536 + n];
        const float t0 = f16_round(b), t1 = f16_round(b - t0), t2 = b - t0 - t1;
        dst[i] = __float2half_rn(kin == 0 ? t0 : (kin == 1 ? t1 : (kin == 2 ? t2 : 0.f)));
    }
}

// 8 consecutive fp32 features of one row -> hi and lo fp16 images of k-group `kg` (one 16-byte store each)
__device__ __forceinline__ void store_split(uint32_t sm0, int kg, int row, const float v[8], bool relu) {
    uint32_t hi[4], lo[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const float a = relu ? fmaxf(v[2 * j], 0.f) : v[2 * j], b = relu ? fmaxf(v[2 * j + 1], 0.f) : v[2 * j + 1];
        hi[j] = ptx::pack_f16(a, b, false);
        const float2 h = __half22float2(*reinterpret_cast<const __half2*>(&hi[j]));
        lo[j] = ptx::pack_f16(a - h.x, b - h.y, false);
    }
    const uint32_t off = kg * kKgBytes + row * 16;
    ptx::st_shared_v4(sm0 + kAHi + off, hi[0], hi[1], hi[2], hi[3]);
    ptx::st_shared_v4(sm0 + kALo + off, lo[0], lo[1], lo[2], lo[3]);
}

// channel c of the reference's embedding of a 3-vector with F octaves: [x | sin(2^f x) (3) cos(2^f x) (3)]_f
template <int F>
__device__ __forceinline__ float embed_channel(const float x[3], int c) {
    if (c < 3) return x[c];
    if (c >= 3 + 6 * F) return 0.f;
    const int f = (c - 3) / 6, r = (c - 3) - 6 * f;
    const float a = x[r < 3 ? r : r - 3] * static_cast<float>(1 << f);
    return r < 3 ? sinf(a) : cosf(a);
}

// cycle counters of block 0 when debug bit 4 is set: issuer [0] total, [1] wait w_full, [2] wait a_ready |
// epilogue warp 0: [8] total, [9] wait acc_ready | producer [10] wait w_empty
__device__ long long g_dbg_x3[12];

// mbarriers of this kernel (8 B each, inside the FwBars block): w_full[4] | w_empty[4] | a_ready[4] (one per 64-column
// quarter of the layer output == 2 K-chunks of the next layer) | acc_ready
struct X3Bars {
    uint32_t base;
    __device__ __forceinline__ uint32_t w_full(int s) const { return base + 8 * s; }
    __device__ __forceinline__ uint32_t w_empty(int s) const { return base + 8 * (kFwStages + s); }
    __device__ __forceinline__ uint32_t a_ready(int j) const { return base + 8 * (2 * kFwStages + j); }
    __device__ __forceinline__ uint32_t acc_ready() const { return base + 8 * (2 * kFwStages + 4); }
};
static_assert(kFwSmDens - kFwSmBar >= 8 * (2 * kFwStages + 5) + 8, "barrier block too small for the x3 kernel");

// Pipelining across layers: the accumulator alternates between the two 256-column halves of TMEM (layer l -> half
// l & 1), and the epilogue walks the layer output in four 64-column quarters, every warp taking 16 columns of each, and
// signals a_ready[j] after quarter j.  Quarter j is K-chunks 2j, 2j+1 of the next layer, so the next layer's products
// start one quarter into this layer's epilogue and run under the other three.
// kDens: density only (raw_out = sigma [n_points]): the trunk T0..T7 and the density dot product, as in mlp_tc.cu.
template <bool kDens>
__global__ void __launch_bounds__(kThreads, 1)
mlp_fwd_x3_kernel(MlpInput in, const uint8_t* __restrict__ image, Image im, float* __restrict__ raw_out, int flags) {
    extern __shared__ __align__(1024) uint8_t smem[];
    const uint32_t sm0 = ptx::smem_u32(smem);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const X3Bars bars{sm0 + kFwSmBar};
    float* s_misc = reinterpret_cast<float*>(smem + kFwSmMisc);      // [0,128) T9 bias | [128,384) w_density | [384] b_density
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(smem + kFwSmBar + 8 * (2 * kFwStages + 5));
    const int C = im.ins_num + 5;

    if (threadIdx.x == 0) {
        for (int s = 0; s < kFwStages; ++s) { ptx::mbar_init(bars.w_full(s), 1); ptx::mbar_init(bars.w_empty(s), 1); }
        for (int j = 0; j < 4; ++j) ptx::mbar_init(bars.a_ready(j), 4 * kTile);
        ptx::mbar_init(bars.acc_ready(), 1);
        ptx::fence_barrier_init();
    }
    if (warp == kMmaWarp) ptx::tmem_alloc(ptx::smem_u32(const_cast<uint32_t*>(tmem_slot)), 512);
    {
        const float* gb = reinterpret_cast<const float*>(image + im.bias_off);
        for (int i = threadIdx.x; i < 128 + 256 + 1; i += kThreads) s_misc[i] = gb[kBiasT9 + i];
        uint32_t* ones = reinterpret_cast<uint32_t*>(smem + kFwSmOnes);   // rows of [1,1,1,0,0,0,0,0] fp16
        for (int i = threadIdx.x; i < 32; i += kThreads) ones[i] = (i & 3) == 0 ? 0x3C003C00u : ((i & 3) == 1 ? 0x00003C00u : 0u);
    }
    ptx::fence_proxy_async_smem();
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem = *tmem_slot;

    const int64_t n_tiles = (in.n_points + kTile - 1) / kTile;

    if (warp == kProducerWarp) {
        if (lane == 0) {
            int stage = 0; uint32_t phase = 0;
            long long dbg_w = 0;
            for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
                for (int l = 0; l < (kDens ? 8 : kLayers); ++l) {
                    const uint32_t bytes = kFwChunkK * im.n[l] * 2;
                    const int parts = 2 * (layer_k(l) / kFwChunkK);
                    const uint8_t* src = image + im.chunk_off[l];
                    for (int c = 0; c < parts + (l < 9); ++c) {
                        const long long w0 = clock64();
                        ptx::mbar_wait(bars.w_empty(stage), phase ^ 1);
                        dbg_w += clock64() - w0;
                        const uint32_t nb = c < parts ? bytes : kFwBiasBlkBytes;
                        ptx::mbar_arrive_expect_tx(bars.w_full(stage), nb);
                        ptx::bulk_g2s(sm0 + kFwSmRing + stage * kFwStageBytes,
                                      c < parts ? src + static_cast<int64_t>(c) * bytes
                                                : image + im.fwb_off + static_cast<int64_t>(l) * kFwBiasBlkBytes,
                                      nb, bars.w_full(stage));
                        if (++stage == kFwStages) { stage = 0; phase ^= 1; }
                    }
                }
            }
            if ((flags & 16) && blockIdx.x == 0) g_dbg_x3[10] = dbg_w;
            for (int i = 0; i < kFwStages; ++i) {          // drain: the last releases must land before the CTA exits
                ptx::mbar_wait(bars.w_empty(stage), phase ^ 1);
                if (++stage == kFwStages) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp == kMmaWarp) {
        // ===== the issuer: hi stage -> A_hi W_hi, A_lo W_hi; lo stage -> A_hi W_lo; bias stage -> ones x bias =====
        if (lane == 0) {
            int stage = 0; uint32_t phase = 0, a_par = 0;
            constexpr uint32_t kKg16 = kKgBytes >> 4;
            const uint64_t desc_hi = static_cast<uint64_t>(0x4008u) << 32;
            const uint32_t ah0 = ((sm0 + kAHi) >> 4) | (kKg16 << 16), al0 = ((sm0 + kALo) >> 4) | (kKg16 << 16);
            const uint64_t ones_d = (static_cast<uint64_t>(0x4000u) << 32) | ((sm0 + kFwSmOnes) >> 4);
            const uint32_t ring_lo = (sm0 + kFwSmRing) >> 4;
            long long dbg_f = 0, dbg_a = 0;
            const long long dbg_start = clock64();
            for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
                for (int l = 0; l < (kDens ? 8 : kLayers); ++l) {
                    const uint32_t N = im.n[l];
                    const uint32_t idesc = ptx::idesc_f16(kTile, N);
                    const int chunks = layer_k(l) / kFwChunkK;
                    const uint32_t d_tmem = tmem + (l & 1) * 256;
                    uint32_t a_off = layer_akg0(l) * kKg16;
                    int next_q = 0;                                                  // a_ready quarters not yet waited for
#pragma unroll 1
                    for (int c = 0; c < chunks; ++c, a_off += 4 * kKg16) {
                        // quarter j feeds chunks 2j, 2j+1; the last chunk of a short layer (T0) collects the rest
                        const int need = (c == chunks - 1) ? 3 : (c >> 1);
                        if (next_q <= need) {
                            const long long w1 = clock64();
                            for (; next_q <= need && next_q < 4; ++next_q) ptx::mbar_wait(bars.a_ready(next_q), a_par);
                            dbg_a += clock64() - w1;
                        }
                        // hi stage
                        long long w0 = clock64();
                        ptx::mbar_wait(bars.w_full(stage), phase);
                        dbg_f += clock64() - w0;
                        ptx::tc_fence_after();
                        uint32_t b_lo = (ring_lo + stage * (kFwStageBytes >> 4)) | (N << 16);
                        ptx::umma_bf16(d_tmem, desc_hi | (ah0 + a_off), desc_hi | b_lo, idesc, c != 0);
                        ptx::umma_bf16(d_tmem, desc_hi | (ah0 + a_off + 2 * kKg16), desc_hi | (b_lo + 2 * N), idesc, 1u);
                        ptx::umma_bf16(d_tmem, desc_hi | (al0 + a_off), desc_hi | b_lo, idesc, 1u);
                        ptx::umma_bf16(d_tmem, desc_hi | (al0 + a_off + 2 * kKg16), desc_hi | (b_lo + 2 * N), idesc, 1u);
                        ptx::umma_commit(bars.w_empty(stage));
                        if (++stage == kFwStages) { stage = 0; phase ^= 1; }
                        // lo stage
                        w0 = clock64();
                        ptx::mbar_wait(bars.w_full(stage), phase);
                        dbg_f += clock64() - w0;
                        ptx::tc_fence_after();
                        b_lo = (ring_lo + stage * (kFwStageBytes >> 4)) | (N << 16);
                        ptx::umma_bf16(d_tmem, desc_hi | (ah0 + a_off), desc_hi | b_lo, idesc, 1u);
                        ptx::umma_bf16(d_tmem, desc_hi | (ah0 + a_off + 2 * kKg16), desc_hi | (b_lo + 2 * N), idesc, 1u);
                        if (l == 9 && c == chunks - 1) ptx::umma_commit(bars.acc_ready());
                        ptx::umma_commit(bars.w_empty(stage));
                        if (++stage == kFwStages) { stage = 0; phase ^= 1; }
                    }
                    a_par ^= 1;
                    if (l < 9) {                                                     // bias stage
                        const long long w0 = clock64();
                        ptx::mbar_wait(bars.w_full(stage), phase);
                        dbg_f += clock64() - w0;
                        ptx::tc_fence_after();
                        const uint32_t b_lo = (ring_lo + stage * (kFwStageBytes >> 4)) | (N << 16);
                        ptx::umma_bf16(d_tmem, ones_d, desc_hi | b_lo, idesc, 1u);
                        ptx::umma_commit(bars.acc_ready());
                        ptx::umma_commit(bars.w_empty(stage));
                        if (++stage == kFwStages) { stage = 0; phase ^= 1; }
                    }
                }
            }
            if ((flags & 16) && blockIdx.x == 0) { g_dbg_x3[0] = clock64() - dbg_start; g_dbg_x3[1] = dbg_f; g_dbg_x3[2] = dbg_a; }
        }
    } else if (warp < kEpiWarps) {
        // ===== epilogue / embedding: 16 warps; thread == (point == TMEM lane, 16-column slot wq of every quarter) =====
        const int q = warp & 3, wq = warp >> 2, row = q * 32 + lane;
        const uint32_t t_lane = tmem + (static_cast<uint32_t>(q * 32) << 16);
        float* stage_out = reinterpret_cast<float*>(smem + kAHi);
        float* dens_part = reinterpret_cast<float*>(smem + kAHi + kDensKg * kKgBytes);
        uint32_t acc_par = 0;
        long long dbg_acc = 0;
        const long long dbg_start = clock64();
        for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
            const int64_t tile0 = tile * kTile;
            const int64_t p = tile0 + row;
            const bool valid = p < in.n_points;
            float dir[3] = {0.f, 0.f, 0.f}, pos[3] = {0.f, 0.f, 0.f};
            if (in.x == nullptr && valid) {
                const int64_t ray = p / in.S;
                const float zz = __ldg(in.z + p);
                float d[3];
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    d[c] = __ldg(in.rays_d + ray * 3 + c);
                    pos[c] = __fadd_rn(__ldg(in.rays_o + ray * 3 + c), __fmul_rn(d[c], zz));
                }
                // viewdirs = rays_d / ||rays_d|| as torch does it (tester.py / render.py): sqrt of the sum, then a division
                const float nrm = sqrtf(__fadd_rn(__fadd_rn(__fmul_rn(d[0], d[0]), __fmul_rn(d[1], d[1])), __fmul_rn(d[2], d[2])));
#pragma unroll
                for (int c = 0; c < 3; ++c) dir[c] = d[c] / nrm;
            }
            // pos-embedding: 8 k-groups of EXT, two per warp slot
#pragma unroll 1
            for (int g = 2 * wq; g < 2 * wq + 2; ++g) {
                float v[8];
#pragma unroll
                for (int e = 0; e < 8; ++e) {
                    const int c = g * 8 + e;
                    if (in.x == nullptr) v[e] = embed_channel<kPosFreqs>(pos, c);
                    else v[e] = (valid && c < 63) ? __ldg(in.x + p * 90 + c) : 0.f;
                }
                store_split(sm0, kExtKg + g, row, v, false);
            }
            ptx::fence_proxy_async_smem();
            ptx::tc_fence_before();
#pragma unroll
            for (int j = 0; j < 4; ++j) ptx::mbar_arrive(bars.a_ready(j));

            float dens = 0.f;
            for (int l = 0; l < kLayers - 1; ++l) {
                const long long w0 = clock64();
                ptx::mbar_wait(bars.acc_ready(), acc_par);
                dbg_acc += clock64() - w0;
                acc_par ^= 1;
                ptx::tc_fence_after();
                const uint32_t t_base = t_lane + (l & 1) * 256;
#pragma unroll 1
                for (int j = 0; j < 4; ++j) {
                    const int col0 = 64 * j + 16 * wq;
                    uint32_t r[16];
                    ptx::tmem_ld16(t_base + col0, r);
                    ptx::tmem_ld_wait();
#pragma unroll
                    for (int g = 0; g < 2; ++g) {
                        float v[8];
#pragma unroll
                        for (int e = 0; e < 8; ++e) v[e] = __uint_as_float(r[g * 8 + e]);
                        if (l == 7) {
#pragma unroll
                            for (int e = 0; e < 8; ++e) dens = fmaf(fmaxf(v[e], 0.f), s_misc[128 + col0 + g * 8 + e], dens);
                        }
                        if (!(kDens && l == 7)) store_split(sm0, (col0 >> 3) + g, row, v, true);
                    }
                    if constexpr (kDens) {
                        if (l == 7) {
                            if (j < 3) continue;                       // nothing is handed on: no operand, no a_ready
                            dens_part[wq * kTile + row] = dens;
                            ptx::named_bar_sync(1, 4 * kTile);
                            if (wq == 0 && valid)
                                raw_out[p] = ((dens_part[row] + dens_part[kTile + row]) + (dens_part[2 * kTile + row] + dens_part[3 * kTile + row])) + s_misc[384];
                            break;
                        }
                    }
                    if (l == 7 && j == 3) {
                        // dir-embedding replaces the pos-embedding in EXT (T5 is long done with it): one k-group per slot
                        float v[8];
#pragma unroll
                        for (int e = 0; e < 8; ++e) {
                            const int c = wq * 8 + e;
                            if (in.x == nullptr) v[e] = embed_channel<kDirFreqs>(dir, c);
                            else v[e] = (valid && c < 27) ? __ldg(in.x + p * 90 + 63 + c) : 0.f;
                        }
                        store_split(sm0, kExtKg + wq, row, v, false);
                        dens_part[wq * kTile + row] = dens;
                    }
                    ptx::fence_proxy_async_smem();
                    ptx::tc_fence_before();
                    ptx::mbar_arrive(bars.a_ready(j));
                }
                if constexpr (kDens) {
                    if (l == 7) break;                                 // T8 / T9 are not evaluated
                }
            }
            if constexpr (kDens) continue;
            // ---- T9: heads -> fp32 raw row [rgb3 | sigma | ins K+1], staged in smem, bulk-stored (warp slot 0) ----
            const long long w9 = clock64();
            ptx::mbar_wait(bars.acc_ready(), acc_par);
            dbg_acc += clock64() - w9;
            acc_par ^= 1;
            ptx::tc_fence_after();
            if (wq == 0) dens = (dens_part[row] + dens_part[kTile + row]) + (dens_part[2 * kTile + row] + dens_part[3 * kTile + row]);
            // the other slots go on to the next tile's embedding, which overwrites the partials: not before they are read
            ptx::named_bar_sync(1, 4 * kTile);
            if (wq == 0) {
                const uint32_t t_base = t_lane + 256;                        // T9 is layer 9: the odd half
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
                ptx::named_bar_sync(2, kTile);
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
                // T0's epilogue of the next tile (the next writer of this region) cannot start before all 512 threads,
                // incl. the bulk-store issuer after its read-wait, arrived on a_ready again
                ptx::named_bar_sync(2, kTile);
            }
        }
        if ((flags & 16) && blockIdx.x == 0 && threadIdx.x == 0) { g_dbg_x3[8] = clock64() - dbg_start; g_dbg_x3[9] = dbg_acc; }
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == kMmaWarp) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc(tmem, 512);
    }
}

}  // namespace x3
}  // namespace tc

int persistent_ctas(int sm_count);
int tc_debug_flags();

int64_t x3_image_bytes(int ins_num) { return tc::x3::make_image_x3(ins_num).total_bytes; }

int pack_x3(const float* const* params, int ins_num, void* packed, cudaStream_t st) {
    const tc::Image im = tc::x3::make_image_x3(ins_num);
    tc::Params pr;
    for (int i = 0; i < kNumParams; ++i) pr.p[i] = params[i];
    uint8_t* image = static_cast<uint8_t*>(packed);
    float* scratch = reinterpret_cast<float*>(image + im.scratch_off);
    tc::compose_kernel<<<dim3(16, 8, 3), 256, 0, st>>>(pr, scratch);
    const int64_t elems = im.bias_off / 2;
    tc::x3::pack_chunks_kernel<<<static_cast<unsigned>((elems + 255) / 256), 256, 0, st>>>(pr, im, image);
    tc::pack_bias_kernel<<<(tc::kBiasFloats + 255) / 256, 256, 0, st>>>(pr, im, image);
    tc::x3::pack_bias3_kernel<<<9, 256, 0, st>>>(pr, im, image);
    DM_LAUNCH_CHECK("pack_x3");
    return DMNERF_OK;
}

int launch_mlp_x3(const MlpInput& in, const void* packed, int ins_num, float* raw_out, void* stash, cudaStream_t st, bool dens) {
    DM_REQUIRE(!stash, DMNERF_EUNSUPPORTED,
               "mlp_fwd(tc32): the split-fp16 chain is inference-only (no activation stash; train in DMNERF_MODE_BF16 or FP32)");
    static int sm_count = 0;
    if (sm_count == 0) {
        int dev = 0;
        DM_CUDA(cudaGetDevice(&dev));
        DM_CUDA(cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev));
    }
    DM_REQUIRE((reinterpret_cast<uintptr_t>(packed) & 15) == 0 && (reinterpret_cast<uintptr_t>(raw_out) & 15) == 0,
               DMNERF_EINVAL, "mlp_fwd(tc32): packed image and raw_out must be 16-byte aligned");
    const int64_t tiles = (in.n_points + tc::kTile - 1) / tc::kTile;
    const int sms = persistent_ctas(sm_count);
    const unsigned grid = static_cast<unsigned>(tiles < sms ? tiles : sms);
    if (dens) {
        DM_CUDA(cudaFuncSetAttribute(tc::x3::mlp_fwd_x3_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, tc::kFwSmTotal));
        tc::x3::mlp_fwd_x3_kernel<true><<<grid, tc::x3::kThreads, tc::kFwSmTotal, st>>>(in, static_cast<const uint8_t*>(packed),
                                                                                      tc::x3::make_image_x3(ins_num), raw_out, tc_debug_flags());
    } else {
        DM_CUDA(cudaFuncSetAttribute(tc::x3::mlp_fwd_x3_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, tc::kFwSmTotal));
        tc::x3::mlp_fwd_x3_kernel<false><<<grid, tc::x3::kThreads, tc::kFwSmTotal, st>>>(in, static_cast<const uint8_t*>(packed),
                                                                                       tc::x3::make_image_x3(ins_num), raw_out, tc_debug_flags());
    }

    DM_LAUNCH_CHECK("mlp_fwd_x3_kernel");
    return DMNERF_OK;
}

}  // namespace dmnerf

using namespace dmnerf;

// cycle counters of block 0 of the last DMNERF_MODE_TC32 forward that ran with debug bit 4 set (layout at g_dbg_x3)
extern "C" int dmnerf_debug_counters_tc32(long long* host_out12) {
    DM_REQUIRE(host_out12, DMNERF_EINVAL, "debug_counters_tc32: null pointer");
    DM_CUDA(cudaDeviceSynchronize());
    DM_CUDA(cudaMemcpyFromSymbol(host_out12, tc::x3::g_dbg_x3, 12 * sizeof(long long)));
    return DMNERF_OK;
}
