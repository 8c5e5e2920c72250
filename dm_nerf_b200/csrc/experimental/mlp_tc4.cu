// mlp_tc4.cu -- a1+a2+a9 in DMNERF_MODE_BF16, CTA pairs with ONE resident tile per CTA and double-buffered
// accumulators: the variant in which a layer's epilogue and the next layer's MMAs overlap.
//
// Why this shape (the measurements are in DESIGN.md 4.1): with two resident tiles per SM (mlp_tc.cu) all 512 TMEM
// columns hold the two N=256 accumulators and 160 of the 227 KB of shared memory hold the two activation tiles, so
// a layer's MMAs cannot start before the previous epilogue has drained the accumulator (no second buffer) and a tile
// cannot run ahead of the other (no ring depth).  Here a cluster of two CTAs works on two 128-point tiles, one per
// CTA, as ONE M=256 tcgen05.mma.cta_group::2 per K step:
//   * each CTA holds its tile's activations (80 KB) and HALF of every weight chunk (its 128 of the 256 output rows),
//     so per SM the streamed weight bytes per tile are the same as in mlp_tc.cu, and an 8-stage x 16 KB ring fits;
//   * TMEM holds TWO accumulator buffers (layer l uses buffer l & 1), so layer l+1's MMAs may write while layer l's
//     epilogue still reads;
//   * the activations are updated in place, 64 columns (= one K=64 weight chunk) at a time: every epilogue warp
//     converts 16 columns of column block j = 0..3 in turn and arrives on a_ready[j]; the issuer starts chunk j of the
//     next layer as soon as block j is complete in BOTH CTAs.  The first block is ready after a quarter of the
//     epilogue, and from then on the tensor pipe is the only critical path.
// Weight streams, chunk schedule and the GEMM-side biases are those of mlp_tc3 (pack_v3, mlp_tc3.cuh).
// Roles per CTA (18 warps): 0-15 epilogue (lane quarter q = w & 3, column slice cg = w >> 2), 16 weight producer,
// 17 MMA issuer (leader CTA) / "chunk landed" relay (peer CTA).
#include "../mlp_tc_common.cuh"

namespace dmnerf {
namespace tc4 {

using namespace tc;
using tc3::Chunk;
using tc3::kFirst;
using tc3::kLast;
using tc3::kOnes;
using tc3::kMaxChunks;
using tc3::kSlotBytes;

// A ring STEP is up to three consecutive-in-use chunks of one layer (<= 36 KB per CTA, 8-9 MMAs ~ 1.1 k clk): one ring
// step costs the issuing thread ~300 clk of control code plus 88 clk per barrier it waits on (tools/issue_cost.py), so
// the K=64 chunk (4 MMAs = 512 clk) of mlp_tc3 is too small a unit for ONE issuer serving the pair.
constexpr int kStages4 = 3;
constexpr int kSlot4 = 2 * kSlotBytes + 4096;                   // 36,864: two K=64 half-chunks + the bias half-chunk
constexpr int kMaxSteps = 32;
constexpr int kSm4A = 0;                                        // 40 k-groups x 2048 B
constexpr int kSm4Ones = kABytes;                               // 81,920: one 128-byte core matrix (LBO = SBO = 0)
constexpr int kSm4Ring = kSm4Ones + 128;
constexpr int kSm4Recs = kSm4Ring + kStages4 * kSlot4;          // uint4 per chunk (issuer), then the Chunk table
constexpr int kSm4Sched = kSm4Recs + kMaxChunks * 16;
constexpr int kSm4Steps = kSm4Sched + kMaxChunks * 12 + 64;     // uint4 per step: n|dep|first|last, chunk ids, bytes; [kMaxSteps] = count
constexpr int kSm4Misc = kSm4Steps + (kMaxSteps + 1) * 16;      // layer_start[11] | T9 bias (128) | w_density (256) | b_density
constexpr int kSm4Dens = kSm4Misc + (16 + 128 + 256 + 4) * 4;   // density partials [4 column slices][128 rows]
constexpr int kSm4Sig = kSm4Dens + 4 * kTile * 4;               // landing slots of the async-store signals: [2 CTAs][16 warps] + ring
constexpr int kSm4Bar = kSm4Sig + (2 * kEpiWarps + kStages4 + 1) * 4;
constexpr int kNumBars4 = 2 * kStages4 + 3 + 2;
constexpr int kSigBytes = 2 * kEpiWarps * 4;                    // tx bytes that complete one a_ready phase
constexpr int kSm4Total = kSm4Bar + 8 * kNumBars4 + 16;
static_assert(kSm4Total <= 227 * 1024, "shared memory budget");
static_assert(kSm4Ring % 128 == 0 && kSm4Bar % 8 == 0 && kSm4Steps % 16 == 0, "alignment");

struct Bars4 {
    uint32_t base;
    __device__ __forceinline__ uint32_t w_full(int s) const { return base + 8 * s; }
    __device__ __forceinline__ uint32_t w_empty(int s) const { return base + 8 * (kStages4 + s); }
    __device__ __forceinline__ uint32_t a_ready(int b) const { return base + 8 * (2 * kStages4 + b); }      // 0: H cols 0-127, 1: 128-255, 2: EXT
    __device__ __forceinline__ uint32_t acc_ready(int b) const { return base + 8 * (2 * kStages4 + 3 + b); }
};

// cycle counters of block 0 (debug bit 5 | dmnerf_debug_counters): issuer 0 total, 1 wait w_full, 3 wait a_ready |
// epilogue warp 0: 4 total, 5 wait acc_ready, 2 signalling, 7 LDTM | producer: 6 wait w_empty
__device__ long long g_dbg4[8];

// "this warp's part of block b is in shared memory": the leader's a_ready[b] has ONE pending arrival (the issuer's own
// arrive.expect_tx of kSigBytes, re-armed after every wait) and completes when all 32 epilogue warps of the pair have
// landed their 4-byte asynchronous store.
__device__ __forceinline__ void warp_signal_leader(uint32_t sig_remote, uint32_t bar_remote, int lane) {
    ptx::fence_proxy_async_smem();
    ptx::tc_fence_before();
    __syncwarp();
    if (lane == 0) ptx::st_async_signal(sig_remote, bar_remote, 1u);
}

template <bool kStash>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
mlp_fwd_bf16_v4_kernel(MlpInput in, const uint8_t* __restrict__ image, Image im, float* __restrict__ raw_out,
                       uint8_t* __restrict__ stash) {
    extern __shared__ __align__(1024) uint8_t smem[];
    const uint32_t sm0 = ptx::smem_u32(smem);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = ptx::cluster_ctarank();
    const int64_t cluster_id = blockIdx.x >> 1, n_clusters = gridDim.x >> 1;
    const Bars4 bars{sm0 + kSm4Bar};
    const Chunk* sched = reinterpret_cast<const Chunk*>(smem + kSm4Sched);
    const uint4* recs = reinterpret_cast<const uint4*>(smem + kSm4Recs);
    const int* layer_start = reinterpret_cast<const int*>(smem + kSm4Misc);
    float* s_misc = reinterpret_cast<float*>(smem + kSm4Misc) + 16;     // [0,128) T9 bias | [128,384) w_density | [384] b_density
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(smem + kSm4Bar + 8 * kNumBars4);
    const int C = im.ins_num + 5;
    const int n_chunks = im.v3_chunks;

    if (threadIdx.x == 0) {
        for (int s = 0; s < kStages4; ++s) { ptx::mbar_init(bars.w_full(s), 1); ptx::mbar_init(bars.w_empty(s), 1); }
        for (int b = 0; b < 3; ++b) ptx::mbar_init(bars.a_ready(b), 1);
        for (int b = 0; b < 2; ++b) ptx::mbar_init(bars.acc_ready(b), 1);
        ptx::fence_barrier_init();
        for (int b = 0; b < 3; ++b) ptx::mbar_arrive_expect_tx(bars.a_ready(b), kSigBytes);     // arm phase 0 (used in the leader)
        // the ring steps of a tile: chunks grouped by layer so that each step waits for at most one a_ready barrier
        const Chunk* gc = reinterpret_cast<const Chunk*>(image + im.v3_sched_off);
        const int* gls = reinterpret_cast<const int*>(image + im.v3_sched_off + kMaxChunks * sizeof(Chunk));
        uint4* steps = reinterpret_cast<uint4*>(smem + kSm4Steps);
        int ns = 0;
        for (int l = 0; l < kLayers; ++l) {
            int ext = -1, bias = -1, h[4] = {-1, -1, -1, -1}, nh = 0;
            for (int c = gls[l]; c < gls[l + 1]; ++c) {
                const Chunk ck = gc[c];
                if (ck.flags & kOnes) bias = c;
                else if (ck.a_kg >= kExtKg) ext = c;
                else if (nh < 4) h[nh++] = c;
            }
            const int first_step = ns;
            auto add = [&](uint32_t dep, int c0, int c1, int c2) {
                uint32_t n = 0, bytes = 0;
                const int cs[3] = {c0, c1, c2};
                for (int i = 0; i < 3; ++i)
                    if (cs[i] >= 0) { ++n; bytes += static_cast<uint32_t>(gc[cs[i]].ksteps) * 16 * gc[cs[i]].rows * 2; }
                steps[ns++] = make_uint4(n | (dep << 4) | (static_cast<uint32_t>(l) << 12),
                                         static_cast<uint32_t>(c0 & 0xFFFF) | (static_cast<uint32_t>(c1 & 0xFFFF) << 16),
                                         static_cast<uint32_t>(c2 & 0xFFFF), bytes);
            };
            if (l == 0) {
                add(2, ext, bias, -1);
            } else {
                add(0, h[0], h[1], -1);
                if (ext >= 0) { add(1, h[2], h[3], -1); add(l == 8 ? 2 : 7, ext, bias, -1); }      // T5 (pos-emb, long ready), T8 (dir-emb)
                else add(1, h[2], h[3], bias);                                                      // bias = -1 for the heads
            }
            steps[first_step].x |= 1u << 8;
            steps[ns - 1].x |= 1u << 9;
        }
        steps[kMaxSteps].x = ns;
    }
    if (warp == kMmaWarp) ptx::tmem_alloc2(ptx::smem_u32(const_cast<uint32_t*>(tmem_slot)), 512);
    {
        const Chunk* gc = reinterpret_cast<const Chunk*>(image + im.v3_sched_off);
        const uint32_t* gs = reinterpret_cast<const uint32_t*>(gc);
        uint32_t* ss = reinterpret_cast<uint32_t*>(smem + kSm4Sched);
        for (int i = threadIdx.x; i < n_chunks * 3; i += kThreads) ss[i] = gs[i];
        uint4* rr = reinterpret_cast<uint4*>(smem + kSm4Recs);
        for (int i = threadIdx.x; i < n_chunks; i += kThreads) {
            const Chunk ck = gc[i];
            // x: A offset (16-byte units) | y: flags, k-steps, B rows per CTA | z: instruction descriptor | w: bytes per CTA
            rr[i] = make_uint4(static_cast<uint32_t>(ck.a_kg) * (kKgBytes >> 4),
                               ck.flags | (static_cast<uint32_t>(ck.ksteps) << 8) | (static_cast<uint32_t>(ck.rows) << 16),
                               ptx::idesc_bf16(2 * kTile, 2 * ck.rows), static_cast<uint32_t>(ck.ksteps) * 16 * ck.rows * 2);
        }
        const int* gls = reinterpret_cast<const int*>(image + im.v3_sched_off + kMaxChunks * sizeof(Chunk));
        int* sls = reinterpret_cast<int*>(smem + kSm4Misc);
        for (int i = threadIdx.x; i < 11; i += kThreads) sls[i] = gls[i];
        const float* gb = reinterpret_cast<const float*>(image + im.bias_off);
        for (int i = threadIdx.x; i < 128 + 256 + 1; i += kThreads) s_misc[i] = gb[kBiasT9 + i];
        uint32_t* ones = reinterpret_cast<uint32_t*>(smem + kSm4Ones);
        for (int i = threadIdx.x; i < 32; i += kThreads) ones[i] = (i & 3) == 0 ? 0x3F803F80u : 0u;
    }
    ptx::fence_proxy_async_smem();
    ptx::tc_fence_before();
    ptx::cluster_sync();
    ptx::tc_fence_after();
    const uint32_t tmem = *tmem_slot;
    const int64_t n_duos = (in.n_points + kPair - 1) / kPair;          // two tiles (one per CTA) per cluster trip

    const uint4* steps = reinterpret_cast<const uint4*>(smem + kSm4Steps);
    const int n_steps = static_cast<int>(steps[kMaxSteps].x);
    if (warp == kProducerWarp) {
        // ===== weight producer: this CTA's halves of the chunks of every ring step, once per tile =====
        if (lane == 0) {
            int stage = 0; uint32_t phase = 0;
            long long dbg_w = 0;
            const uint8_t* src = image + im.v3_off[rank];
            for (int64_t duo = cluster_id; duo < n_duos; duo += n_clusters) {
                for (int st = 0; st < n_steps; ++st) {
                    const uint4 sp = steps[st];
                    const long long w0 = clock64();
                    ptx::mbar_wait(bars.w_empty(stage), phase ^ 1);
                    dbg_w += clock64() - w0;
                    // the leader's barrier also counts the peer's 4-byte "my half has landed" store
                    ptx::mbar_arrive_expect_tx(bars.w_full(stage), sp.w + (rank == 0 ? 4u : 0u));
                    uint32_t dst = sm0 + kSm4Ring + stage * kSlot4;
                    const uint32_t ids[3] = {sp.y & 0xFFFFu, sp.y >> 16, sp.z & 0xFFFFu};
                    for (uint32_t i = 0; i < (sp.x & 15u); ++i) {
                        const Chunk ck = sched[ids[i]];
                        const uint32_t bytes = static_cast<uint32_t>(ck.ksteps) * 16 * ck.rows * 2;
                        ptx::bulk_g2s(dst, src + ck.off, bytes, bars.w_full(stage));
                        dst += bytes;
                    }
                    if (++stage == kStages4) { stage = 0; phase ^= 1; }
                }
            }
            if (blockIdx.x == 0) g_dbg4[6] = dbg_w;
            for (int i = 0; i < kStages4; ++i) {           // drain the last multicast releases before the CTA may exit
                ptx::mbar_wait(bars.w_empty(stage), phase ^ 1);
                if (++stage == kStages4) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp == kMmaWarp) {
        if (lane == 0 && rank == 0) {
            // ===== MMA issuer (leader CTA): M=256 pair MMAs, accumulator buffer = layer & 1 =====
            uint32_t stage = 0, phase = 0, a_par = 0;        // bit b of a_par: parity to wait for on a_ready[b]
            long long dbg_f = 0, dbg_a = 0;
            const long long dbg_start = clock64();
            const uint32_t lbo_a = static_cast<uint32_t>(kKgBytes >> 4) << 16;
            const uint32_t a_base_lo = ((sm0 + kSm4A) >> 4) | lbo_a;
            const uint64_t ones_d = (static_cast<uint64_t>(0x4000u) << 32) | ((sm0 + kSm4Ones) >> 4);
            const uint32_t ring_lo = (sm0 + kSm4Ring) >> 4;
            const uint64_t desc_hi = static_cast<uint64_t>(0x4008u) << 32;       // SBO = 128 B, descriptor version 1
            for (int64_t duo = cluster_id; duo < n_duos; duo += n_clusters) {
#pragma unroll 1
                for (int st = 0; st < n_steps; ++st) {
                    const uint4 sp = steps[st];
                    const uint32_t dep = (sp.x >> 4) & 15u, l = sp.x >> 12;
                    const uint32_t d = tmem + (l & 1) * 256;
                    const long long w0 = clock64();
                    ptx::mbar_wait(bars.w_full(stage), phase);
                    const long long w1 = clock64();
                    if (dep < 3) {
                        ptx::mbar_wait(bars.a_ready(dep), (a_par >> dep) & 1);
                        a_par ^= 1u << dep;
                        ptx::mbar_arrive_expect_tx(bars.a_ready(dep), kSigBytes);               // arm the next phase
                    }
                    dbg_f += w1 - w0; dbg_a += clock64() - w1;
                    ptx::tc_fence_after();
                    uint32_t b_off = ring_lo + stage * (kSlot4 >> 4);
                    uint32_t acc = (sp.x >> 8) & 1u ? 0u : 1u;                                  // first MMA of a layer overwrites
                    const uint32_t ids[3] = {sp.y & 0xFFFFu, sp.y >> 16, sp.z & 0xFFFFu};
#pragma unroll
                    for (uint32_t i = 0; i < 3; ++i) {
                        if (i < (sp.x & 15u)) {
                            const uint4 r = recs[ids[i]];
                            const uint32_t nh = r.y >> 16, ksteps = (r.y >> 8) & 15;
                            const uint32_t b_lo = b_off + (nh << 16);
                            if (r.y & kOnes) {
                                ptx::umma2_bf16(d, ones_d, desc_hi | b_lo, r.z, 1u);
                            } else {
                                const uint32_t a_lo = a_base_lo + r.x;
                                ptx::umma2_bf16(d, desc_hi | a_lo, desc_hi | b_lo, r.z, acc);
                                ptx::umma2_bf16(d, desc_hi | (a_lo + 2 * (kKgBytes >> 4)), desc_hi | (b_lo + nh * 2), r.z, 1u);
                                if (ksteps > 2) {
                                    ptx::umma2_bf16(d, desc_hi | (a_lo + 4 * (kKgBytes >> 4)), desc_hi | (b_lo + nh * 4), r.z, 1u);
                                    ptx::umma2_bf16(d, desc_hi | (a_lo + 6 * (kKgBytes >> 4)), desc_hi | (b_lo + nh * 6), r.z, 1u);
                                }
                            }
                            acc = 1u;
                            b_off += r.w >> 4;
                        }
                    }
                    if ((sp.x >> 9) & 1u) ptx::umma_commit2(bars.acc_ready(l & 1), 3);
                    ptx::umma_commit2(bars.w_empty(stage), 3);
                    if (++stage == kStages4) { stage = 0; phase ^= 1; }
                }
            }
            if (blockIdx.x == 0) { g_dbg4[0] = clock64() - dbg_start; g_dbg4[1] = dbg_f; g_dbg4[3] = dbg_a; }
        } else if (lane == 0) {
            // ===== relay (peer CTA): "my half of this step has landed" -> 4 bytes on the leader's w_full tx-count =====
            uint32_t stage = 0, phase = 0;
            const uint32_t remote_bar0 = ptx::mapa(bars.w_full(0), 0);
            const uint32_t remote_sig0 = ptx::mapa(sm0 + kSm4Sig + 2 * kEpiWarps * 4, 0);
            for (int64_t duo = cluster_id; duo < n_duos; duo += n_clusters) {
                for (int st = 0; st < n_steps; ++st) {
                    ptx::mbar_wait(bars.w_full(stage), phase);
                    ptx::st_async_signal(remote_sig0 + 4 * stage, remote_bar0 + 8 * stage, 1u);
                    if (++stage == kStages4) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else {
        // ===== epilogue / embedding warps (both CTAs): thread == (point == TMEM lane, 16-column slice cg of each block) =====
        const int q = warp & 3, cg = warp >> 2, row = q * 32 + lane;
        const uint32_t a_base = sm0 + kSm4A;
        const uint32_t t_lane = tmem + (static_cast<uint32_t>(q * 32) << 16);
        float* stage_out = reinterpret_cast<float*>(smem + kSm4A);
        float* dens_part = reinterpret_cast<float*>(smem + kSm4Dens);
        const float* s_wd = s_misc + 128;
        const uint32_t sig_remote = ptx::mapa(sm0 + kSm4Sig + (rank * kEpiWarps + warp) * 4, 0);
        const uint32_t bar_remote0 = ptx::mapa(bars.a_ready(0), 0);
        uint32_t acc_par = 0;                                // bit b: parity to wait for on acc_ready[b]
        long long dbg_acc = 0;
        const long long dbg_start = clock64();
        for (int64_t duo = cluster_id; duo < n_duos; duo += n_clusters) {
            const int64_t tileg = duo * 2 + rank;
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
                if (cg == 0)      embed_to_ext<kPosFreqs, 0, 2>(a_base, row, pos, gst, kStPos);
                else if (cg == 1) embed_to_ext<kPosFreqs, 2, 4>(a_base, row, pos, gst, kStPos);
                else if (cg == 2) embed_to_ext<kPosFreqs, 4, 6>(a_base, row, pos, gst, kStPos);
                else              embed_to_ext<kPosFreqs, 6, 8>(a_base, row, pos, gst, kStPos);
            } else {
                row_to_ext(a_base, row, valid ? in.x + p * 90 : nullptr, 63, cg * 2, cg * 2 + 2, gst, kStPos);
            }
            warp_signal_leader(sig_remote, bar_remote0 + 8 * 2, lane);

            float dens = 0.f;
            for (int l = 0; l < kLayers - 1; ++l) {
                const int buf = l & 1;
                const int gkg0 = l == 8 ? kStHid : kStH0 + 32 * l;
                const long long w0 = clock64();
                ptx::mbar_wait(bars.acc_ready(buf), (acc_par >> buf) & 1);
                dbg_acc += clock64() - w0;
                acc_par ^= 1u << buf;
                ptx::tc_fence_after();
#pragma unroll 1
                for (int hf = 0; hf < 2; ++hf) {
                    // this warp's 16-column slices of column blocks 2hf and 2hf+1: two loads in flight, one wait
                    const int colA = 128 * hf + 16 * cg, colB = colA + 64;
                    uint32_t ra[16], rb[16];
                    ptx::tmem_ld16(t_lane + buf * 256 + colA, ra);
                    ptx::tmem_ld16(t_lane + buf * 256 + colB, rb);
                    ptx::tmem_ld_wait();
#pragma unroll
                    for (int g = 0; g < 4; ++g) {
                        const int col = (g < 2 ? colA : colB) + (g & 1) * 8;
                        float v[8];
#pragma unroll
                        for (int e = 0; e < 8; ++e) v[e] = __uint_as_float(g < 2 ? ra[(g & 1) * 8 + e] : rb[(g & 1) * 8 + e]);
                        if (l == 7) {
#pragma unroll
                            for (int e = 0; e < 8; ++e) dens = fmaf(fmaxf(v[e], 0.f), s_wd[col + e], dens);
                        }
                        store_kg(a_base, col >> 3, row, v, true, gst, gkg0 + (col >> 3));
                    }
                    warp_signal_leader(sig_remote, bar_remote0 + 8 * hf, lane);     // columns 128hf .. 128hf+127 are in place
                }
                if (l == 7) {
                    // dir-embedding replaces the pos-embedding in EXT (T5's MMAs are long done with it)
                    if (in.x == nullptr) {
                        if (cg == 0)      embed_to_ext<kDirFreqs, 0, 1>(a_base, row, dir, gst, kStDir);
                        else if (cg == 1) embed_to_ext<kDirFreqs, 1, 2>(a_base, row, dir, gst, kStDir);
                        else if (cg == 2) embed_to_ext<kDirFreqs, 2, 3>(a_base, row, dir, gst, kStDir);
                        else              embed_to_ext<kDirFreqs, 3, 4>(a_base, row, dir, gst, kStDir);
                    } else {
                        row_to_ext(a_base, row, valid ? in.x + p * 90 + 63 : nullptr, 27, cg, cg + 1, gst, kStDir);
                    }
                    dens_part[cg * kTile + row] = dens;
                    warp_signal_leader(sig_remote, bar_remote0 + 8 * 2, lane);
                }
            }
            // ---- T9: heads -> fp32 raw row [rgb3 | sigma | ins K+1], staged in smem, bulk-stored (slice-0 warps) ----
            const long long w9 = clock64();
            ptx::mbar_wait(bars.acc_ready(1), (acc_par >> 1) & 1);
            dbg_acc += clock64() - w9;
            acc_par ^= 2u;
            ptx::tc_fence_after();
            if (cg == 0) {
                dens = dens_part[row] + dens_part[kTile + row] + dens_part[2 * kTile + row] + dens_part[3 * kTile + row];
                for (int cb = 0; cb * 16 < im.npad2; ++cb) {
                    uint32_t r[16];
                    ptx::tmem_ld16(t_lane + 256 + cb * 16, r);
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
                ptx::named_bar_sync(1, kTile);
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
                // the next write into this region is T0's epilogue of the next tile, which cannot start before these
                // warps (incl. the bulk-store issuer, after its read-wait) arrived on a_ready[4] again
                ptx::named_bar_sync(1, kTile);
            }
        }
        if (blockIdx.x == 0 && threadIdx.x == 0) {
            g_dbg4[4] = clock64() - dbg_start; g_dbg4[5] = dbg_acc;

        }
    }
    __syncwarp();
    ptx::tc_fence_before();
    ptx::cluster_sync();
    if (warp == kMmaWarp) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc2(tmem, 512);
    }
}

}  // namespace tc4

int launch_mlp_bf16_v4(const MlpInput& in, const void* packed, int ins_num, float* raw_out, void* stash, cudaStream_t st) {
    static int sm_count = 0;
    if (sm_count == 0) {
        int dev = 0;
        DM_CUDA(cudaGetDevice(&dev));
        DM_CUDA(cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev));
    }
    DM_REQUIRE((reinterpret_cast<uintptr_t>(packed) & 15) == 0 && (reinterpret_cast<uintptr_t>(raw_out) & 15) == 0,
               DMNERF_EINVAL, "mlp_fwd(bf16): packed image and raw_out must be 16-byte aligned");
    const int64_t duos = (in.n_points + tc::kPair - 1) / tc::kPair;
    const int64_t max_clusters = sm_count / 2;
    const unsigned grid = 2 * static_cast<unsigned>(duos < max_clusters ? duos : max_clusters);
    const tc::Image im = tc::make_image(ins_num);
    if (stash) {
        DM_REQUIRE((reinterpret_cast<uintptr_t>(stash) & 15) == 0, DMNERF_EINVAL, "mlp_fwd(bf16): stash must be 16-byte aligned");
        DM_CUDA(cudaFuncSetAttribute(tc4::mlp_fwd_bf16_v4_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, tc4::kSm4Total));
        tc4::mlp_fwd_bf16_v4_kernel<true><<<grid, tc::kThreads, tc4::kSm4Total, st>>>(
            in, static_cast<const uint8_t*>(packed), im, raw_out, static_cast<uint8_t*>(stash));
    } else {
        DM_CUDA(cudaFuncSetAttribute(tc4::mlp_fwd_bf16_v4_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, tc4::kSm4Total));
        tc4::mlp_fwd_bf16_v4_kernel<false><<<grid, tc::kThreads, tc4::kSm4Total, st>>>(
            in, static_cast<const uint8_t*>(packed), im, raw_out, nullptr);
    }
    DM_LAUNCH_CHECK("mlp_fwd_bf16_v4_kernel");
    return DMNERF_OK;
}

namespace tc4 {
int debug_counters_v4(long long* host_out8) {
    DM_CUDA(cudaMemcpyFromSymbol(host_out8, g_dbg4, 8 * sizeof(long long)));
    return DMNERF_OK;
}
}  // namespace tc4

}  // namespace dmnerf
