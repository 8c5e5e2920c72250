// mlp_tc2.cu -- a1+a2+a9 in DMNERF_MODE_BF16, pipelined: the forward of mlp_tc.cu re-scheduled so that the tensor
// pipe and the epilogue warps overlap across layers (design and hazards: mlp_tc2.cuh).  Same math, same operand
// layouts, same stash; only the order of work and the bias handling differ.
//
// Warp roles (576 threads): warps 0-7 epilogue of tile X, 8-15 of tile Y (thread = point = TMEM lane; warp & 3 =
// lane quarter, (warp >> 2) & 1 = which 64 columns of the 128-column half), warp 16 weight producer, warp 17 MMA
// issuer + TMEM owner.  Barriers per tile t and half j: acc_done[t][j] (tcgen05.commit -> epilogue),
// h_ready[t][j] (epilogue -> MMA issuer: H[j] of this layer is in shared memory and accumulator half j is drained;
// one arrival per warp after its threads' proxy fence).
#include "../mlp_tc_common.cuh"

namespace dmnerf {

int tc_debug_flags();

namespace tc2 {

using namespace tc;

// ---- shared memory map ------------------------------------------------------------------------------------
constexpr int kSm2A = 0;                                        // 2 tiles x 40 k-groups x 2048 B
constexpr int kSm2Ones = 2 * kABytes;                           // 163,840: 2 k-groups (K=16) x 128 rows x 16 B
constexpr int kSm2Ring = kSm2Ones + 2 * kKgBytes;               // 167,936
constexpr int kSm2Sched = kSm2Ring + kStagesV2 * kChunkBytes;   // 217,088
constexpr int kSm2Recs = kSm2Sched + kMaxChunks * 12;           // 219,776: uint4 per chunk for the MMA issuers
constexpr int kSm2Misc = kSm2Recs + kMaxChunks * 16;            // T9 bias (128) | w_density (256) | b_density
constexpr int kSm2Dens = kSm2Misc + (128 + 256 + 4) * 4;        // density partials of the upper halves: [2][128] floats
constexpr int kSm2Bar = kSm2Dens + 2 * kTile * 4;
constexpr int kSm2Total = kSm2Bar + 8 * (2 * kStagesV2 + 8) + 16;
static_assert(sizeof(Chunk) == 12, "schedule entry size");
static_assert(kSm2Total <= 227 * 1024, "shared memory budget");

// ---- pack: v2 chunk stream + schedule table ----------------------------------------------------------------
// One block per chunk.  Element (g, n, e) of a chunk = bf16 of W[out = 128 j + n][k], k from the chunk's A k-group;
// the bias step holds bias_hi in K row 0 and bias_lo = bf16(bias - bias_hi) in row 1.
__global__ void pack_v2_kernel(Params pr, Image im, uint8_t* __restrict__ image) {
    const float* scratch = reinterpret_cast<const float*>(image + im.scratch_off);
    Chunk* sched = reinterpret_cast<Chunk*>(image + im.v2_sched_off);
    __shared__ Chunk ck;
    if (threadIdx.x == 0) {
        if (blockIdx.x == 0) build_schedule(im.npad, sched);
        Chunk tmp[kMaxChunks];
        build_schedule(im.npad, tmp);
        ck = tmp[blockIdx.x];
    }
    __syncthreads();
    const int l = ck.layer, nrows = ck.nrows, kk = ck.ksteps * 16;
    __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(image + im.v2_off + ck.off);
    for (int i = threadIdx.x; i < kk * nrows; i += blockDim.x) {
        const int g = i / (nrows * 8), n = (i >> 3) % nrows, e = i & 7;
        const int kin = g * 8 + e;
        const int row = l == 9 ? n : 128 * ck.j + n;
        float v;
        if (ck.flags & kFOnes) {
            const float b = l < 8 ? pr.p[2 * l + 1][row] : scratch[65536 + row];
            const float hi = __bfloat162float(__float2bfloat16_rn(b));
            v = kin == 0 ? hi : (kin == 1 ? b - hi : 0.f);
        } else {
            int k;
            if (ck.a_kg >= kExtKg) k = (l == 0 ? 0 : 256) + (ck.a_kg - kExtKg) * 8 + kin;
            else k = ck.a_kg * 8 + kin;
            v = logical_weight(pr, scratch, l, row, k, im.ins_num);
        }
        dst[i] = __float2bfloat16_rn(v);
    }
}

int pack_v2(const float* const* params, int ins_num, void* packed, cudaStream_t st) {
    const Image im = make_image(ins_num);
    Params pr;
    for (int i = 0; i < kNumParams; ++i) pr.p[i] = params[i];
    pack_v2_kernel<<<im.v2_chunks, 256, 0, st>>>(pr, im, static_cast<uint8_t*>(packed));
    DM_LAUNCH_CHECK("pack_v2_kernel");
    return DMNERF_OK;
}

// cycle counters of block 0 (bring-up aid, read through dmnerf_debug_counters): 0 MMA thread total, 1 waiting w_full,
// 2 waiting h_ready, 3 epilogue thread 0 total, 4 waiting acc_done, 5 producer waiting w_empty
__device__ long long g_dbg[8];
#define DBG_T0() const long long dbg_t0 = clock64()
#define DBG_ADD(slot) dbg[slot] += clock64() - dbg_t0

constexpr int kThreadsV2 = 19 * 32;         // 16 epilogue warps, producer (16), MMA issuers (17: tile X, 18: tile Y)

struct Bars2 {
    uint32_t base;
    __device__ __forceinline__ uint32_t w_full(int s) const { return base + 8 * s; }
    __device__ __forceinline__ uint32_t w_empty(int s) const { return base + 8 * (kStagesV2 + s); }
    __device__ __forceinline__ uint32_t h_ready(int t, int j) const { return base + 8 * (2 * kStagesV2 + 2 * t + j); }
    __device__ __forceinline__ uint32_t acc_done(int t, int j) const { return base + 8 * (2 * kStagesV2 + 4 + 2 * t + j); }
};

// per-warp arrival on h_ready: every thread has fenced its smem writes towards the async proxy, then one lane arrives
__device__ __forceinline__ void warp_arrive(uint32_t bar, int lane) {
    ptx::fence_proxy_async_smem();
    ptx::tc_fence_before();
    __syncwarp();
    if (lane == 0) ptx::mbar_arrive(bar);
}

template <bool kStash>
__global__ void __launch_bounds__(kThreadsV2, 1)
mlp_fwd_bf16_v2_kernel(MlpInput in, const uint8_t* __restrict__ image, Image im, float* __restrict__ raw_out,
                       uint8_t* __restrict__ stash) {
    extern __shared__ __align__(1024) uint8_t smem[];
    const uint32_t sm0 = ptx::smem_u32(smem);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const Bars2 bars{sm0 + kSm2Bar};
    const Chunk* sched = reinterpret_cast<const Chunk*>(smem + kSm2Sched);
    float* s_misc = reinterpret_cast<float*>(smem + kSm2Misc);          // [0,128) T9 bias | [128,384) w_density | [384] b_density
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(smem + kSm2Bar + 8 * (2 * kStagesV2 + 8));
    const int C = im.ins_num + 5;
    const int n_chunks = im.v2_chunks;

    if (threadIdx.x == 0) {
        for (int s = 0; s < kStagesV2; ++s) { ptx::mbar_init(bars.w_full(s), 1); ptx::mbar_init(bars.w_empty(s), 2); }
        for (int t = 0; t < 2; ++t)
            for (int j = 0; j < 2; ++j) { ptx::mbar_init(bars.h_ready(t, j), 8); ptx::mbar_init(bars.acc_done(t, j), 1); }
        ptx::fence_barrier_init();
    }
    if (warp == kMmaWarp) ptx::tmem_alloc(ptx::smem_u32(const_cast<uint32_t*>(tmem_slot)), 512);
    {
        const uint32_t* gs = reinterpret_cast<const uint32_t*>(image + im.v2_sched_off);
        uint32_t* ss = reinterpret_cast<uint32_t*>(smem + kSm2Sched);
        for (int i = threadIdx.x; i < n_chunks * 3; i += kThreadsV2) ss[i] = gs[i];
        // lean 16-byte records for the MMA issuers: x = A offset (16-byte units, tile relative), y = flags | ksteps << 8 |
        // j << 12 | N << 16, z = instruction descriptor
        const Chunk* gc = reinterpret_cast<const Chunk*>(image + im.v2_sched_off);
        uint4* recs = reinterpret_cast<uint4*>(smem + kSm2Recs);
        for (int i = threadIdx.x; i < n_chunks; i += kThreadsV2) {
            const Chunk ck = gc[i];
            recs[i] = make_uint4(static_cast<uint32_t>(ck.a_kg) * (kKgBytes >> 4),
                                 ck.flags | (static_cast<uint32_t>(ck.ksteps) << 8) | (static_cast<uint32_t>(ck.j) << 12) |
                                     (static_cast<uint32_t>(ck.nrows) << 16),
                                 ptx::idesc_bf16(kTile, ck.nrows), 0u);
        }
        const float* gb = reinterpret_cast<const float*>(image + im.bias_off);
        for (int i = threadIdx.x; i < 128 + 256 + 1; i += kThreadsV2) s_misc[i] = gb[kBiasT9 + i];
        // the "ones" operand of the bias steps: [2 k-groups][128 rows][8 bf16], features 0 and 1 = 1.0
        uint32_t* ones = reinterpret_cast<uint32_t*>(smem + kSm2Ones);
        for (int i = threadIdx.x; i < 2 * kKgBytes / 4; i += kThreadsV2) ones[i] = (i < kKgBytes / 4 && (i & 3) == 0) ? 0x3F803F80u : 0u;
    }
    ptx::fence_proxy_async_smem();
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem = *tmem_slot;
    const int64_t n_pairs = (in.n_points + kPair - 1) / kPair;

    if (warp == kProducerWarp) {
        // ===== weight producer: the chunk stream in schedule order, once per tile pair =====
        if (lane == 0) {
            int stage = 0; uint32_t phase = 0;
            long long dbg[8] = {0, 0, 0, 0, 0, 0, 0, 0};
            const uint8_t* src = image + im.v2_off;
            for (int64_t pair = blockIdx.x; pair < n_pairs; pair += gridDim.x) {
                for (int c = 0; c < n_chunks; ++c) {
                    const Chunk ck = sched[c];
                    const uint32_t bytes = static_cast<uint32_t>(ck.ksteps) * 16 * ck.nrows * 2;
                    { DBG_T0(); ptx::mbar_wait(bars.w_empty(stage), phase ^ 1); DBG_ADD(5); }
                    ptx::mbar_arrive_expect_tx(bars.w_full(stage), bytes);
                    ptx::bulk_g2s(sm0 + kSm2Ring + stage * kChunkBytes, src + ck.off, bytes, bars.w_full(stage));
                    if (++stage == kStagesV2) { stage = 0; phase ^= 1; }
                }
            }
            for (int i = 0; i < kStagesV2; ++i) {
                ptx::mbar_wait(bars.w_empty(stage), phase ^ 1);
                if (++stage == kStagesV2) { stage = 0; phase ^= 1; }
            }
            if (blockIdx.x == 0) g_dbg[5] = dbg[5];
        }
    } else if (warp >= kMmaWarp) {
        // ===== MMA issuers: warp 17 issues (and commits) tile X's MMAs, warp 18 tile Y's.  A naive single-thread issue
        // loop is bound by that thread's dependent-instruction latency (~180 cycles per MMA measured), not by the
        // tensor pipe (64 cycles per M=128,N=128,K=16), hence: one issuer per tile on different schedulers, 64-wide K
        // chunks (4 MMAs per barrier round trip), one 16-byte schedule record per chunk, and descriptors formed by
        // adding 16-byte offsets to constant bases. =====
        if (lane == 0) {
            const int t = warp - kMmaWarp;
            uint32_t stage = 0, phase = 0, hp0 = 0, hp1 = 0;
            const uint4* recs = reinterpret_cast<const uint4*>(smem + kSm2Recs);
            const uint32_t a_base_lo = ((sm0 + kSm2A + t * kABytes) >> 4) | (static_cast<uint32_t>(kKgBytes >> 4) << 16);
            const uint32_t ones_lo = ((sm0 + kSm2Ones) >> 4) | (static_cast<uint32_t>(kKgBytes >> 4) << 16);
            const uint32_t ring_lo = (sm0 + kSm2Ring) >> 4;
            const uint32_t d_base = tmem + t * 256;
            const uint64_t desc_hi = static_cast<uint64_t>(0x4008u) << 32;       // SBO = 128 B, descriptor version 1
            const uint32_t h0 = bars.h_ready(t, 0), h1 = bars.h_ready(t, 1), ad0 = bars.acc_done(t, 0), ad1 = bars.acc_done(t, 1);
            const long long dbg_start = clock64();
            long long dbg_wait = 0, dbg_hwait = 0;
            for (int64_t pair = blockIdx.x; pair < n_pairs; pair += gridDim.x) {
                uint4 r = recs[0];
#pragma unroll 1
                for (int c = 0; c < n_chunks; ++c) {
                    const uint4 rn = recs[c + 1 < n_chunks ? c + 1 : 0];          // next record: off the critical path
                    if (!ptx::mbar_try_wait(bars.w_full(stage), phase)) {
                        const long long w0 = clock64();
                        ptx::mbar_wait(bars.w_full(stage), phase);
                        dbg_wait += clock64() - w0;
                    }
                    if (r.y & (kFWait0 | kFWait1)) {
                        const long long w0 = clock64();
                        if (r.y & kFWait0) { ptx::mbar_wait(h0, hp0); hp0 ^= 1; }
                        if (r.y & kFWait1) { ptx::mbar_wait(h1, hp1); hp1 ^= 1; }
                        dbg_hwait += clock64() - w0;
                    }
                    ptx::tc_fence_after();
                    const uint32_t nn = r.y >> 16, ksteps = (r.y >> 8) & 15;
                    const uint32_t a_lo = (r.y & kFOnes) ? ones_lo : a_base_lo + r.x;
                    const uint32_t b_lo = ring_lo + stage * (kChunkBytes >> 4) + (nn << 16);
                    const uint32_t d = d_base + ((r.y >> 12) & 1) * 128;
                    ptx::umma_bf16(d, desc_hi | a_lo, desc_hi | b_lo, r.z, (r.y & kFFirst) ? 0u : 1u);
                    if (ksteps > 1) {
                        ptx::umma_bf16(d, desc_hi | (a_lo + 2 * (kKgBytes >> 4)), desc_hi | (b_lo + nn * 2), r.z, 1u);
                        if (ksteps > 2) {
                            ptx::umma_bf16(d, desc_hi | (a_lo + 4 * (kKgBytes >> 4)), desc_hi | (b_lo + nn * 4), r.z, 1u);
                            ptx::umma_bf16(d, desc_hi | (a_lo + 6 * (kKgBytes >> 4)), desc_hi | (b_lo + nn * 6), r.z, 1u);
                        }
                    }
                    if (r.y & kFCommit0) ptx::umma_commit(ad0);
                    if (r.y & kFCommit1) ptx::umma_commit(ad1);
                    ptx::umma_commit(bars.w_empty(stage));
                    if (++stage == kStagesV2) { stage = 0; phase ^= 1; }
                    r = rn;
                }
            }
            if (blockIdx.x == 0 && t == 0) { g_dbg[0] = clock64() - dbg_start; g_dbg[1] = dbg_wait; g_dbg[2] = dbg_hwait; }
        }
    } else {
        // ===== epilogue / embedding warps =====
        const int t = warp >> 3, q = warp & 3, hh = (warp >> 2) & 1, row = q * 32 + lane;
        const uint32_t a_base = sm0 + kSm2A + t * kABytes;
        const uint32_t t_base = tmem + t * 256 + (static_cast<uint32_t>(q * 32) << 16);
        float* stage_out = reinterpret_cast<float*>(smem + kSm2A + t * kABytes);
        float* dens_part = reinterpret_cast<float*>(smem + kSm2Dens) + t * kTile;
        const float* s_wd = s_misc + 128;
        uint32_t acc_par[2] = {0, 0};
        long long dbg[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        const long long dbg_start = clock64();
        for (int64_t pair = blockIdx.x; pair < n_pairs; pair += gridDim.x) {
            const int64_t tile0 = pair * kPair + t * kTile;
            const int64_t p = tile0 + row;
            const bool valid = p < in.n_points;
            uint8_t* gst = kStash ? stash + ((pair * 2 + t) * 2 + (row >> 6)) * kStActBlkBytes + (row & 63) * 16 : nullptr;
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
                if (hh == 0) embed_to_ext<kPosFreqs, 0, 4>(a_base, row, pos, gst, kStPos);
                else         embed_to_ext<kPosFreqs, 4, 8>(a_base, row, pos, gst, kStPos);
            } else {
                row_to_ext(a_base, row, valid ? in.x + p * 90 : nullptr, 63, hh * 4, hh * 4 + 4, gst, kStPos);
            }
            // pos-embedding in place (and the previous pair's head epilogue finished): T0's cells may go
            ptx::fence_proxy_async_smem();
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) { ptx::mbar_arrive(bars.h_ready(t, 0)); ptx::mbar_arrive(bars.h_ready(t, 1)); }

            float dens = 0.f;
            for (int l = 0; l < kLayers - 1; ++l) {
                const int gkg0 = l == 8 ? kStHid : kStH0 + 32 * l;
#pragma unroll 1
                for (int j = 0; j < 2; ++j) {
                    { DBG_T0(); ptx::mbar_wait(bars.acc_done(t, j), acc_par[j]); DBG_ADD(4); }
                    acc_par[j] ^= 1;
                    ptx::tc_fence_after();
                    const int col0 = 128 * j + 64 * hh;
                    uint32_t r0[32], r1[32];
                    ptx::tmem_ld32(t_base + col0, r0);
                    ptx::tmem_ld32(t_base + col0 + 32, r1);
                    ptx::tmem_ld_wait();
#pragma unroll
                    for (int g = 0; g < 8; ++g) {
                        float v[8];
#pragma unroll
                        for (int e = 0; e < 8; ++e) v[e] = __uint_as_float(g < 4 ? r0[g * 8 + e] : r1[(g - 4) * 8 + e]);
                        if (l == 7) {
#pragma unroll
                            for (int e = 0; e < 8; ++e) dens = fmaf(fmaxf(v[e], 0.f), s_wd[col0 + g * 8 + e], dens);
                        }
                        store_kg(a_base, (col0 >> 3) + g, row, v, true, gst, gkg0 + (col0 >> 3) + g);
                    }
                    if (l == 7 && j == 0) {
                        // dir-embedding replaces the pos-embedding in EXT (T5 is long done with it); T8's cells that read it
                        // are issued after h_ready[t][1] of this layer, i.e. after every warp has passed this point
                        if (in.x == nullptr) {
                            if (hh == 0) embed_to_ext<kDirFreqs, 0, 2>(a_base, row, dir, gst, kStDir);
                            else         embed_to_ext<kDirFreqs, 2, 4>(a_base, row, dir, gst, kStDir);
                        } else {
                            row_to_ext(a_base, row, valid ? in.x + p * 90 + 63 : nullptr, 27, hh * 2, hh * 2 + 2, gst, kStDir);
                        }
                    }
                    if (l == 7 && j == 1 && hh == 1) dens_part[row] = dens;
                    warp_arrive(bars.h_ready(t, j), lane);
                }
            }
            // ---- T9: heads -> fp32 raw row [rgb3 | sigma | ins K+1], staged in smem, bulk-stored (hh == 0 warps) ----
            ptx::mbar_wait(bars.acc_done(t, 0), acc_par[0]);
            acc_par[0] ^= 1;
            ptx::tc_fence_after();
            if (hh == 0) {
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
                            if (col == 3) v = dens + s_misc[128 + 256];
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
                // H is next written by T0's epilogue of the next pair; its cells wait for h_ready, which needs the
                // arrival of these warps (after the bulk store's read-wait)
                ptx::named_bar_sync(1 + t, kTile);
            }
        }
        if (blockIdx.x == 0 && threadIdx.x == 0) { g_dbg[3] = clock64() - dbg_start; g_dbg[4] = dbg[4]; }
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == kMmaWarp) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc(tmem, 512);
    }
}


}  // namespace tc2

int launch_mlp_bf16_v2(const MlpInput& in, const void* packed, int ins_num, float* raw_out, void* stash, cudaStream_t st) {
    static int sm_count = 0;
    if (sm_count == 0) {
        int dev = 0;
        DM_CUDA(cudaGetDevice(&dev));
        DM_CUDA(cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev));
    }
    DM_REQUIRE((reinterpret_cast<uintptr_t>(packed) & 15) == 0 && (reinterpret_cast<uintptr_t>(raw_out) & 15) == 0,
               DMNERF_EINVAL, "mlp_fwd(bf16): packed image and raw_out must be 16-byte aligned");
    const int64_t pairs = (in.n_points + tc::kPair - 1) / tc::kPair;
    const unsigned grid = static_cast<unsigned>(pairs < sm_count ? pairs : sm_count);
    const tc::Image im = tc::make_image(ins_num);
    if (stash) {
        DM_REQUIRE((reinterpret_cast<uintptr_t>(stash) & 15) == 0, DMNERF_EINVAL, "mlp_fwd(bf16): stash must be 16-byte aligned");
        DM_CUDA(cudaFuncSetAttribute(tc2::mlp_fwd_bf16_v2_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, tc2::kSm2Total));
        tc2::mlp_fwd_bf16_v2_kernel<true><<<grid, tc2::kThreadsV2, tc2::kSm2Total, st>>>(
            in, static_cast<const uint8_t*>(packed), im, raw_out, static_cast<uint8_t*>(stash));
    } else {
        DM_CUDA(cudaFuncSetAttribute(tc2::mlp_fwd_bf16_v2_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, tc2::kSm2Total));
        tc2::mlp_fwd_bf16_v2_kernel<false><<<grid, tc2::kThreadsV2, tc2::kSm2Total, st>>>(
            in, static_cast<const uint8_t*>(packed), im, raw_out, nullptr);
    }
    DM_LAUNCH_CHECK("mlp_fwd_bf16_v2_kernel");
    return DMNERF_OK;
}

namespace tc2 {
int debug_counters_v2(long long* host_out8) {
    DM_CUDA(cudaMemcpyFromSymbol(host_out8, g_dbg, 8 * sizeof(long long)));
    return DMNERF_OK;
}
}  // namespace tc2
}  // namespace dmnerf
