// mlp_tc3.cu -- a1+a2+a9 in DMNERF_MODE_BF16 on CTA PAIRS (tcgen05 cta_group::2): the forward of mlp_tc.cu with
//   * clusters of two CTAs (two SMs of one TPC); every MMA is M=256: the two CTAs' 128-point tiles against ONE weight
//     chunk whose 256 output rows are split between the CTAs' shared memories.  Per SM this halves the weight bytes
//     streamed from L2 and the B-operand bytes read from shared memory (A 4 KB + B 4 KB instead of 4 + 8 KB per MMA),
//     which is what leaves shared-memory bandwidth for an epilogue running CONCURRENTLY with MMAs (the N=128
//     single-CTA pipelining of mlp_tc2.cu was bound by exactly that bandwidth);
//   * ping-pong between each CTA's two tiles X and Y: the issuer walks X.L0, Y.L0, X.L1, Y.L1, ...; while the pair
//     MMA of tile X's layer l+1 cannot start (its epilogue is running), tile Y's layer-l MMA keeps the tensor pipe
//     busy, and vice versa.  Tiles are independent pipelines; each re-streams the layer's chunks (per SM the L2 traffic
//     equals the lockstep kernel's because every CTA loads only half of each chunk);
//   * biases inside the GEMM (a K=16 step against a constant ones operand, bias as a bf16 hi+lo pair), so the epilogue
//     is LDTM -> relu+pack -> STS.128.
// Roles per CTA (18 warps): warps 0-7 epilogue of tile X, 8-15 of tile Y, warp 16 weight producer (this CTA's chunk
// halves), warp 17 = MMA issuer in the leader CTA (rank 0) / "chunk landed" relay in the peer CTA.
// Barriers (same offsets in both CTAs): w_full[s] local TMA bytes, plus -- in the leader -- 4 bytes stored asynchronously
// by the peer's relay when ITS half has landed; w_empty[s] and acc_ready[t] by multicast tcgen05.commit; a_ready[t]
// (leader only) one pending arrival (the issuer's own expect_tx, re-armed after every wait) + 4 bytes per epilogue warp
// of tile t from BOTH CTAs (st.async ... mbarrier::complete_tx; a remote mbarrier.arrive blocks the lane ~800 clk).
#include "../mlp_tc_common.cuh"

namespace dmnerf {
namespace tc3 {

using namespace tc;

constexpr int kSm3A = 0;                                        // 2 tiles x 40 k-groups x 2048 B
constexpr int kSm3Ones = 2 * kABytes;                           // 163,840: one 128-byte core matrix (descriptor LBO = SBO = 0)
constexpr int kSm3Ring = kSm3Ones + 128;
constexpr int kSm3Recs = kSm3Ring + kStagesV3 * kSlotBytes;     // uint4 per chunk (issuer), then Chunk table
constexpr int kSm3Sched = kSm3Recs + kMaxChunks * 16;
constexpr int kSm3Misc = kSm3Sched + kMaxChunks * 12 + 64;      // layer_start[11] | T9 bias (128) | w_density (256) | b_density
constexpr int kSm3Dens = kSm3Misc + (16 + 128 + 256 + 4) * 4;
constexpr int kSm3Sig = kSm3Dens + 2 * kTile * 4;               // landing slots of the async-store signals: [2 CTAs][16 warps] + ring
constexpr int kSm3Gen = kSm3Sig + (2 * kEpiWarps + kStagesV3 + 1) * 4;   // per ring stage: how many of its uses have been consumed
constexpr int kSm3Bar = kSm3Gen + (kStagesV3 + 1) * 4;
constexpr int kSm3Total = kSm3Bar + 8 * (3 * kStagesV3 + 4) + 16;
constexpr int kSigBytes3 = kEpiWarps * 4;                       // 8 warps per tile x 2 CTAs x 4 bytes complete one a_ready phase
static_assert(sizeof(Chunk) == 12, "schedule entry size");
static_assert(kSm3Total <= 227 * 1024, "shared memory budget");

// ---- pack ----------------------------------------------------------------------------------------------------
// One block per (chunk, CTA rank).  Element (g, n, e) = bf16 of W[out = rows * rank + n][k].
__global__ void pack_v3_kernel(Params pr, Image im, uint8_t* __restrict__ image) {
    const float* scratch = reinterpret_cast<const float*>(image + im.scratch_off);
    __shared__ Chunk ck;
    const int rank = blockIdx.y;
    if (threadIdx.x == 0) {
        Chunk tmp[kMaxChunks];
        int ls[16];
        build_schedule(im.npad2, tmp, ls);
        ck = tmp[blockIdx.x];
        if (blockIdx.x == 0 && rank == 0) {
            Chunk* sched = reinterpret_cast<Chunk*>(image + im.v3_sched_off);
            int* gls = reinterpret_cast<int*>(image + im.v3_sched_off + kMaxChunks * sizeof(Chunk));
            for (int i = 0; i < im.v3_chunks; ++i) sched[i] = tmp[i];
            for (int i = 0; i < 11; ++i) gls[i] = ls[i];
        }
    }
    __syncthreads();
    const int l = ck.layer, rows = ck.rows, kk = ck.ksteps * 16;
    __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(image + im.v3_off[rank] + ck.off);
    for (int i = threadIdx.x; i < kk * rows; i += blockDim.x) {
        const int g = i / (rows * 8), n = (i >> 3) % rows, e = i & 7;
        const int kin = g * 8 + e;
        const int row = rows * rank + n;
        float v;
        if (ck.flags & kOnes) {
            const float b = l < 8 ? pr.p[2 * l + 1][row] : scratch[65536 + row];
            const float hi = __bfloat162float(__float2bfloat16_rn(b));
            v = kin == 0 ? hi : (kin == 1 ? b - hi : 0.f);
        } else {
            int k;
            if (ck.a_kg >= kExtKg) k = (l == 0 ? 0 : 256) + (ck.a_kg - kExtKg) * 8 + kin;
            else k = ck.a_kg * 8 + kin;
            v = (l == 9 && row >= im.npad) ? 0.f : logical_weight(pr, scratch, l, row, k, im.ins_num);
        }
        dst[i] = __float2bfloat16_rn(v);
    }
}

int pack_v3(const float* const* params, int ins_num, void* packed, cudaStream_t st) {
    const Image im = make_image(ins_num);
    Params pr;
    for (int i = 0; i < kNumParams; ++i) pr.p[i] = params[i];
    pack_v3_kernel<<<dim3(im.v3_chunks, 2), 256, 0, st>>>(pr, im, static_cast<uint8_t*>(packed));
    DM_LAUNCH_CHECK("pack_v3_kernel");
    return DMNERF_OK;
}

// cycle counters of block 0 (bring-up aid, dmnerf_debug_counters with debug bit 3 set): issuer 0 total, 1 wait w_full,
// 2 wait peer_full, 3 wait a_ready | epilogue warp 0: 4 total, 5 wait acc_ready | producer: 6 wait w_empty
__device__ long long g_dbg3[8];

// One issuer warp per tile (true) or one for both (false).  Two issuers need the ring-order guard below (an mbarrier
// parity wait cannot tell "this use of the stage" from "the use before the previous one"; without the guard the
// second issuer runs through a stage that has not been loaded yet and the launch faults).  Measured: 7.0 ms with one
// issuer, 7.7 ms with two (the guard + spills cost more than the parallel issue gains), so one it is.
constexpr bool kTwoIssuers = false;
// Round-2 experiment: LOCKSTEP instead of ping-pong -- every chunk is streamed once per layer and feeds both tiles of both
// CTAs (8 pair-MMAs per ring step) before it is released, as in mlp_tc.cu but with the B operand split across the pair
// (half the weight bytes per SM, a ring three K=64 chunks deep).
constexpr bool kLockstep = false;     // measured: 6.85 ms vs 6.06 ms (default kernel), see DESIGN.md 4.1b

struct Bars3 {
    uint32_t base;
    __device__ __forceinline__ uint32_t w_full(int s) const { return base + 8 * s; }
    __device__ __forceinline__ uint32_t w_empty(int s) const { return base + 8 * (kStagesV3 + s); }
    __device__ __forceinline__ uint32_t peer_full(int s) const { return base + 8 * (2 * kStagesV3 + s); }
    __device__ __forceinline__ uint32_t a_ready(int t) const { return base + 8 * (3 * kStagesV3 + t); }
    __device__ __forceinline__ uint32_t acc_ready(int t) const { return base + 8 * (3 * kStagesV3 + 2 + t); }
};

// "this warp's share of tile t's operand is in shared memory": a 4-byte asynchronous store onto the tx-count of the
// leader's a_ready[t] (armed by the issuer with expect_tx), not a remote mbarrier.arrive -- that one blocks the
// issuing lane for ~800 clk (measured in mlp_tc4).
__device__ __forceinline__ void warp_signal_leader(uint32_t sig_remote, uint32_t bar_remote, int lane) {
    ptx::fence_proxy_async_smem();
    ptx::tc_fence_before();
    __syncwarp();
    if (lane == 0) ptx::st_async_signal(sig_remote, bar_remote, 1u);
}

template <bool kStash>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kFwThreads, 1)
mlp_fwd_bf16_v3_kernel(MlpInput in, const uint8_t* __restrict__ image, Image im, float* __restrict__ raw_out,
                       uint8_t* __restrict__ stash) {
    extern __shared__ __align__(1024) uint8_t smem[];
    const uint32_t sm0 = ptx::smem_u32(smem);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = ptx::cluster_ctarank();
    const int64_t cluster_id = blockIdx.x >> 1, n_clusters = gridDim.x >> 1;
    const Bars3 bars{sm0 + kSm3Bar};
    const Chunk* sched = reinterpret_cast<const Chunk*>(smem + kSm3Sched);
    const uint4* recs = reinterpret_cast<const uint4*>(smem + kSm3Recs);
    const int* layer_start = reinterpret_cast<const int*>(smem + kSm3Misc);
    float* s_misc = reinterpret_cast<float*>(smem + kSm3Misc) + 16;     // [0,128) T9 bias | [128,384) w_density | [384] b_density
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(smem + kSm3Bar + 8 * (3 * kStagesV3 + 4));
    const int C = im.ins_num + 5;
    const int n_chunks = im.v3_chunks;

    if (threadIdx.x == 0) {
        for (int s = 0; s < kStagesV3; ++s) { ptx::mbar_init(bars.w_full(s), 1); ptx::mbar_init(bars.w_empty(s), 1); }
        for (int t = 0; t < 2; ++t) { ptx::mbar_init(bars.a_ready(t), 1); ptx::mbar_init(bars.acc_ready(t), 1); }
        ptx::fence_barrier_init();
        for (int t = 0; t < 2; ++t) ptx::mbar_arrive_expect_tx(bars.a_ready(t), kSigBytes3);      // arm phase 0 (used in the leader)
        for (int i = 0; i < kStagesV3; ++i) reinterpret_cast<volatile uint32_t*>(smem + kSm3Gen)[i] = 0;
    }
    if (warp == kMmaWarp) ptx::tmem_alloc2(ptx::smem_u32(const_cast<uint32_t*>(tmem_slot)), 512);
    {
        const Chunk* gc = reinterpret_cast<const Chunk*>(image + im.v3_sched_off);
        const uint32_t* gs = reinterpret_cast<const uint32_t*>(gc);
        uint32_t* ss = reinterpret_cast<uint32_t*>(smem + kSm3Sched);
        for (int i = threadIdx.x; i < n_chunks * 3; i += kFwThreads) ss[i] = gs[i];
        uint4* rr = reinterpret_cast<uint4*>(smem + kSm3Recs);
        for (int i = threadIdx.x; i < n_chunks; i += kFwThreads) {
            const Chunk ck = gc[i];
            rr[i] = make_uint4(static_cast<uint32_t>(ck.a_kg) * (kKgBytes >> 4),
                               ck.flags | (static_cast<uint32_t>(ck.ksteps) << 8) | (static_cast<uint32_t>(ck.rows) << 16),
                               ptx::idesc_bf16(2 * kTile, 2 * ck.rows), 0u);
        }
        const int* gls = reinterpret_cast<const int*>(image + im.v3_sched_off + kMaxChunks * sizeof(Chunk));
        int* sls = reinterpret_cast<int*>(smem + kSm3Misc);
        for (int i = threadIdx.x; i < 11; i += kFwThreads) sls[i] = gls[i];
        const float* gb = reinterpret_cast<const float*>(image + im.bias_off);
        for (int i = threadIdx.x; i < 128 + 256 + 1; i += kFwThreads) s_misc[i] = gb[kBiasT9 + i];
        uint32_t* ones = reinterpret_cast<uint32_t*>(smem + kSm3Ones);
        for (int i = threadIdx.x; i < 32; i += kFwThreads) ones[i] = (i & 3) == 0 ? 0x3F803F80u : 0u;
    }
    ptx::fence_proxy_async_smem();
    ptx::tc_fence_before();
    ptx::cluster_sync();
    ptx::tc_fence_after();
    const uint32_t tmem = *tmem_slot;
    const int64_t n_quads = (in.n_points + 2 * kPair - 1) / (2 * kPair);

    if (warp == kProducerWarp) {
        // ===== weight producer: this CTA's halves of the chunks, layer by layer, once per tile (X then Y) =====
        if (lane == 0) {
            int stage = 0; uint32_t phase = 0;
            long long dbg_w = 0;
            const uint8_t* src = image + im.v3_off[rank];
            for (int64_t quad = cluster_id; quad < n_quads; quad += n_clusters) {
                for (int l = 0; l < kLayers; ++l) {
                    for (int t = 0; t < (kLockstep ? 1 : 2); ++t) {
                        for (int c = layer_start[l]; c < layer_start[l + 1]; ++c) {
                            const Chunk ck = sched[c];
                            const uint32_t bytes = static_cast<uint32_t>(ck.ksteps) * 16 * ck.rows * 2;
                            const long long w0 = clock64();
                            ptx::mbar_wait(bars.w_empty(stage), phase ^ 1);
                            dbg_w += clock64() - w0;
                            // the leader's barrier also counts the peer's 4-byte "my half has landed" store
                            ptx::mbar_arrive_expect_tx(bars.w_full(stage), bytes + (rank == 0 ? 4u : 0u));
                            ptx::bulk_g2s(sm0 + kSm3Ring + stage * kSlotBytes, src + ck.off, bytes, bars.w_full(stage));
                            if (++stage == kStagesV3) { stage = 0; phase ^= 1; }
                        }
                    }
                }
            }
            if (blockIdx.x == 0) g_dbg3[6] = dbg_w;
            for (int i = 0; i < kStagesV3; ++i) {
                ptx::mbar_wait(bars.w_empty(stage), phase ^ 1);
                if (++stage == kStagesV3) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp >= kMmaWarp) {
        if (lane == 0 && rank == 0) {
            // ===== MMA issuers (leader CTA): warp 17 drives tile X, warp 18 tile Y.  The ring carries, per layer, the
            // chunks of X's pass and then the chunks of Y's pass; each issuer consumes only its own slots. =====
            if (!kTwoIssuers && warp != kMmaWarp) goto issuers_done;
            {
            const int t_first = kTwoIssuers ? warp - kMmaWarp : 0, t_last = kTwoIssuers ? t_first : 1;
            uint32_t a_par2[2] = {0, 0};
            long long dbg_f = 0, dbg_a = 0;
            const long long dbg_start = clock64();
            const uint32_t lbo_a = static_cast<uint32_t>(kKgBytes >> 4) << 16;
            const uint64_t ones_d = (static_cast<uint64_t>(0x4000u) << 32) | ((sm0 + kSm3Ones) >> 4);
            const uint32_t ring_lo = (sm0 + kSm3Ring) >> 4;
            const uint64_t desc_hi = static_cast<uint64_t>(0x4008u) << 32;       // SBO = 128 B, descriptor version 1
            volatile uint32_t* gen = reinterpret_cast<volatile uint32_t*>(smem + kSm3Gen);
            uint32_t g = 0;                                                       // ring slots handed out so far
            for (int64_t quad = cluster_id; quad < n_quads; quad += n_clusters) {
                for (int l = 0; l < kLayers; ++l) {
                    const int c0 = layer_start[l], c1 = layer_start[l + 1];
                    const uint32_t g_layer = g;
                    if (kLockstep) {
                        g += c1 - c0;
                        uint32_t stage = g_layer % kStagesV3, phase = (g_layer / kStagesV3) & 1u;
#pragma unroll 1
                        for (int c = c0; c < c1; ++c) {
                            const uint4 r = recs[c];
                            const long long w0 = clock64();
                            ptx::mbar_wait(bars.w_full(stage), phase);
                            const long long w1 = clock64();
                            if (r.y & kFirst) {
                                for (int t = 0; t < 2; ++t) {
                                    ptx::mbar_wait(bars.a_ready(t), a_par2[t]);
                                    a_par2[t] ^= 1;
                                    ptx::mbar_arrive_expect_tx(bars.a_ready(t), kSigBytes3);
                                }
                            }
                            dbg_f += w1 - w0; dbg_a += clock64() - w1;
                            ptx::tc_fence_after();
                            const uint32_t nh = r.y >> 16, ksteps = (r.y >> 8) & 15;
                            const uint32_t b_lo = ring_lo + stage * (kSlotBytes >> 4) + (nh << 16);
                            for (int t = 0; t < 2; ++t) {
                                const uint32_t d = tmem + t * 256;
                                if (r.y & kOnes) {
                                    ptx::umma2_bf16(d, ones_d, desc_hi | b_lo, r.z, 1u);
                                } else {
                                    const uint32_t a_lo = (((sm0 + kSm3A + t * kABytes) >> 4) | lbo_a) + r.x;
                                    ptx::umma2_bf16(d, desc_hi | a_lo, desc_hi | b_lo, r.z, (r.y & kFirst) ? 0u : 1u);
                                    ptx::umma2_bf16(d, desc_hi | (a_lo + 2 * (kKgBytes >> 4)), desc_hi | (b_lo + nh * 2), r.z, 1u);
                                    if (ksteps > 2) {
                                        ptx::umma2_bf16(d, desc_hi | (a_lo + 4 * (kKgBytes >> 4)), desc_hi | (b_lo + nh * 4), r.z, 1u);
                                        ptx::umma2_bf16(d, desc_hi | (a_lo + 6 * (kKgBytes >> 4)), desc_hi | (b_lo + nh * 6), r.z, 1u);
                                    }
                                }
                                if (r.y & kLast) ptx::umma_commit2(bars.acc_ready(t), 3);
                            }
                            ptx::umma_commit2(bars.w_empty(stage), 3);
                            if (++stage == kStagesV3) { stage = 0; phase ^= 1; }
                        }
                        continue;
                    }
                    g += 2 * (c1 - c0);
                  for (int t = t_first; t <= t_last; ++t) {
                    const uint32_t a_base_lo = ((sm0 + kSm3A + t * kABytes) >> 4) | lbo_a;
                    const uint32_t d = tmem + t * 256;
                    uint32_t& a_par = a_par2[t];
                    const uint32_t gi = g_layer + t * (c1 - c0);
                    uint32_t stage = gi % kStagesV3, use = gi / kStagesV3, phase = use & 1u;
#pragma unroll 1
                    for (int c = c0; c < c1; ++c) {
                        const uint4 r = recs[c];
                        const long long w0 = clock64();
                        if (kTwoIssuers) {
                            // a parity wait can only tell "this use" from "the previous use": the other issuer may still
                            // be waiting for the previous use of this stage, so queue up behind it first
                            uint32_t spins = 0;
                            while (gen[stage] != use) {
                                if (++spins > (1u << 24)) { printf("dmnerf: ring order timeout (stage %u use %u)\n", stage, use); __trap(); }
                            }
                        }
                        ptx::mbar_wait(bars.w_full(stage), phase);
                        if (kTwoIssuers) gen[stage] = use + 1;
                        const long long w1 = clock64();
                        if (r.y & kFirst) {
                            ptx::mbar_wait(bars.a_ready(t), a_par);
                            a_par ^= 1;
                            ptx::mbar_arrive_expect_tx(bars.a_ready(t), kSigBytes3);    // arm the next phase
                        }
                        dbg_f += w1 - w0; dbg_a += clock64() - w1;
                        ptx::tc_fence_after();
                        const uint32_t nh = r.y >> 16, ksteps = (r.y >> 8) & 15;
                        const uint32_t b_lo = ring_lo + stage * (kSlotBytes >> 4) + (nh << 16);
                        if (r.y & kOnes) {
                            ptx::umma2_bf16(d, ones_d, desc_hi | b_lo, r.z, 1u);
                        } else {
                            const uint32_t a_lo = a_base_lo + r.x;
                            ptx::umma2_bf16(d, desc_hi | a_lo, desc_hi | b_lo, r.z, (r.y & kFirst) ? 0u : 1u);
                            ptx::umma2_bf16(d, desc_hi | (a_lo + 2 * (kKgBytes >> 4)), desc_hi | (b_lo + nh * 2), r.z, 1u);
                            if (ksteps > 2) {
                                ptx::umma2_bf16(d, desc_hi | (a_lo + 4 * (kKgBytes >> 4)), desc_hi | (b_lo + nh * 4), r.z, 1u);
                                ptx::umma2_bf16(d, desc_hi | (a_lo + 6 * (kKgBytes >> 4)), desc_hi | (b_lo + nh * 6), r.z, 1u);
                            }
                        }
                        if (r.y & kLast) ptx::umma_commit2(bars.acc_ready(t), 3);
                        ptx::umma_commit2(bars.w_empty(stage), 3);
                        if (++stage == kStagesV3) { stage = 0; phase ^= 1; ++use; }
                    }
                  }
                }
            }
            if (blockIdx.x == 0 && t_first == 0) { g_dbg3[0] = clock64() - dbg_start; g_dbg3[1] = dbg_f; g_dbg3[2] = 0; g_dbg3[3] = dbg_a; }
            }
            issuers_done:;
        } else if (lane == 0 && warp == kMmaWarp) {
            // ===== relay (peer CTA): "my half of this chunk has landed" -> 4 bytes on the leader's w_full tx-count =====
            uint32_t stage = 0, phase = 0;
            const uint32_t remote_bar0 = ptx::mapa(bars.w_full(0), 0);
            const uint32_t remote_sig0 = ptx::mapa(sm0 + kSm3Sig + 2 * kEpiWarps * 4, 0);
            for (int64_t quad = cluster_id; quad < n_quads; quad += n_clusters) {
                for (int l = 0; l < kLayers; ++l) {
                    const int cnt = (kLockstep ? 1 : 2) * (layer_start[l + 1] - layer_start[l]);
                    for (int c = 0; c < cnt; ++c) {
                        ptx::mbar_wait(bars.w_full(stage), phase);
                        ptx::st_async_signal(remote_sig0 + 4 * stage, remote_bar0 + 8 * stage, 1u);
                        if (++stage == kStagesV3) { stage = 0; phase ^= 1; }
                    }
                }
            }
        }
    } else {
        // ===== epilogue / embedding warps (both CTAs): thread == (point == TMEM lane, column half) =====
        const int t = warp >> 3, q = warp & 3, half = (warp >> 2) & 1, row = q * 32 + lane;
        const uint32_t a_base = sm0 + kSm3A + t * kABytes;
        const uint32_t t_base = tmem + t * 256 + (static_cast<uint32_t>(q * 32) << 16);
        float* stage_out = reinterpret_cast<float*>(smem + kSm3A + t * kABytes);
        float* dens_part = reinterpret_cast<float*>(smem + kSm3Dens) + t * kTile;
        const float* s_wd = s_misc + 128;
        const uint32_t sig_remote = ptx::mapa(sm0 + kSm3Sig + (rank * kEpiWarps + warp) * 4, 0);
        const uint32_t bar_remote = ptx::mapa(bars.a_ready(t), 0);
        uint32_t acc_par = 0;
        long long dbg_acc = 0;
        const long long dbg_start = clock64();
        for (int64_t quad = cluster_id; quad < n_quads; quad += n_clusters) {
            const int64_t tileg = quad * 4 + rank * 2 + t;
            const int64_t tile0 = tileg * kTile;
            const int64_t p = tile0 + row;
            const bool valid = p < in.n_points;
            uint8_t* gst = kStash ? stash + (tileg * 2 + (row >> 6)) * kStActBlkBytes + (row & 63) * 16 : nullptr;
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
                if (half == 0) embed_to_ext<kPosFreqs, 0, 4>(a_base, row, pos, gst, kStPos);
                else           embed_to_ext<kPosFreqs, 4, 8>(a_base, row, pos, gst, kStPos);
            } else {
                row_to_ext(a_base, row, valid ? in.x + p * 90 : nullptr, 63, half * 4, half * 4 + 4, gst, kStPos);
            }
            warp_signal_leader(sig_remote, bar_remote, lane);

            float dens = 0.f;
            for (int l = 0; l < kLayers - 1; ++l) {
                const int gkg0 = l == 8 ? kStHid : kStH0 + 32 * l;
                const long long w0 = clock64();
                ptx::mbar_wait(bars.acc_ready(t), acc_par);
                dbg_acc += clock64() - w0;
                acc_par ^= 1;
                ptx::tc_fence_after();
#pragma unroll 1
                for (int bb = 0; bb < 2; ++bb) {
                    const int col0 = half * 128 + bb * 64;
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
                }
                if (l == 7) {
                    // dir-embedding replaces the pos-embedding in EXT (this tile's T5 pass is long done with it)
                    if (in.x == nullptr) {
                        if (half == 0) embed_to_ext<kDirFreqs, 0, 2>(a_base, row, dir, gst, kStDir);
                        else           embed_to_ext<kDirFreqs, 2, 4>(a_base, row, dir, gst, kStDir);
                    } else {
                        row_to_ext(a_base, row, valid ? in.x + p * 90 + 63 : nullptr, 27, half * 2, half * 2 + 2, gst, kStDir);
                    }
                    if (half == 1) dens_part[row] = dens;
                }
                warp_signal_leader(sig_remote, bar_remote, lane);
            }
            // ---- T9: heads -> fp32 raw row [rgb3 | sigma | ins K+1], staged in smem, bulk-stored (half 0 warps) ----
            const long long w9 = clock64();
            ptx::mbar_wait(bars.acc_ready(t), acc_par);
            dbg_acc += clock64() - w9;
            acc_par ^= 1;
            ptx::tc_fence_after();
            if (half == 0) {
                dens += dens_part[row];
                for (int cb = 0; cb * 16 < im.npad2; ++cb) {
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
                ptx::named_bar_sync(1 + t, kTile);
            }
        }
        if (blockIdx.x == 0 && threadIdx.x == 0) { g_dbg3[4] = clock64() - dbg_start; g_dbg3[5] = dbg_acc; }
    }
    __syncwarp();
    ptx::tc_fence_before();
    ptx::cluster_sync();
    if (warp == kMmaWarp) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc2(tmem, 512);
    }
}


}  // namespace tc3

int launch_mlp_bf16_v3(const MlpInput& in, const void* packed, int ins_num, float* raw_out, void* stash, cudaStream_t st) {
    static int sm_count = 0;
    if (sm_count == 0) {
        int dev = 0;
        DM_CUDA(cudaGetDevice(&dev));
        DM_CUDA(cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev));
    }
    DM_REQUIRE((reinterpret_cast<uintptr_t>(packed) & 15) == 0 && (reinterpret_cast<uintptr_t>(raw_out) & 15) == 0,
               DMNERF_EINVAL, "mlp_fwd(bf16): packed image and raw_out must be 16-byte aligned");
    const int64_t quads = (in.n_points + 2 * tc::kPair - 1) / (2 * tc::kPair);
    const int64_t max_clusters = sm_count / 2;
    const unsigned grid = 2 * static_cast<unsigned>(quads < max_clusters ? quads : max_clusters);
    const tc::Image im = tc::make_image(ins_num);
    if (stash) {
        DM_REQUIRE((reinterpret_cast<uintptr_t>(stash) & 15) == 0, DMNERF_EINVAL, "mlp_fwd(bf16): stash must be 16-byte aligned");
        DM_CUDA(cudaFuncSetAttribute(tc3::mlp_fwd_bf16_v3_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, tc3::kSm3Total));
        tc3::mlp_fwd_bf16_v3_kernel<true><<<grid, tc::kFwThreads, tc3::kSm3Total, st>>>(
            in, static_cast<const uint8_t*>(packed), im, raw_out, static_cast<uint8_t*>(stash));
    } else {
        DM_CUDA(cudaFuncSetAttribute(tc3::mlp_fwd_bf16_v3_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, tc3::kSm3Total));
        tc3::mlp_fwd_bf16_v3_kernel<false><<<grid, tc::kFwThreads, tc3::kSm3Total, st>>>(
            in, static_cast<const uint8_t*>(packed), im, raw_out, nullptr);
    }
    DM_LAUNCH_CHECK("mlp_fwd_bf16_v3_kernel");
    return DMNERF_OK;
}

namespace tc3 {
int debug_counters_v3(long long* host_out8) {
    DM_CUDA(cudaMemcpyFromSymbol(host_out8, g_dbg3, 8 * sizeof(long long)));
    return DMNERF_OK;
}

}  // namespace tc3
}  // namespace dmnerf
