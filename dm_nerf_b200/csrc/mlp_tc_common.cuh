// mlp_tc_common.cuh -- constants, packed-image map and small device helpers shared by the tensor-core MLP
// kernels: forward (mlp_tc.cu) and training backward (mlp_tc_bwd.cu).
#pragma once
#include "mlp_layout.cuh"
#include "tc_ptx.cuh"
#include <cuda_bf16.h>
#ifdef DMNERF_EXPERIMENTAL            // parked forward variants (csrc/experimental/): not part of the product library
#include "experimental/mlp_tc2.cuh"
#include "experimental/mlp_tc3.cuh"
#endif

namespace dmnerf {
namespace tc {

constexpr int kTile = 128;
constexpr int kPair = 2 * kTile;
constexpr int kLayers = 10;
constexpr int kChunkK = 32;                       // K columns per streamed weight chunk
constexpr int kStages = 3;
constexpr int kStageBytes = kChunkK * 256 * 2;    // 16 KB
constexpr int kKgBytes = kTile * 16;              // one k-group of the A operand: 128 rows x 16 B
constexpr int kABytes = 40 * kKgBytes;            // 32 k-groups H + 8 k-groups EXT = 81,920 B per tile
constexpr int kExtKg = 32;                        // first k-group of the EXT region
constexpr int kBiasFloats = 2704;                 // 9*256 | T9 bias (128) | w_density (256) | b_density (+pad)
constexpr int kBiasT9 = 9 * 256;
constexpr int kBiasWd = kBiasT9 + 128;
constexpr int kBiasBd = kBiasWd + 256;
constexpr int kEpiWarps = 16;                     // 8 per tile: 4 lane quarters x 2 column halves
constexpr int kProducerWarp = 16, kMmaWarp = 17;
constexpr int kThreads = 18 * 32;

// shared memory map (bytes)
constexpr int kSmA = 0;
constexpr int kSmRing = 2 * kABytes;                           // 163,840
constexpr int kSmBias = kSmRing + kStages * kStageBytes;       // 212,992
constexpr int kSmBar = kSmBias + kBiasFloats * 4;              // 223,808
constexpr int kSmDens = kSmBar + 128;                          // density partials of column-half 1: [2][128] floats
constexpr int kSmTotal = kSmDens + 2 * kTile * 4;

// forward kernel (mlp_tc.cu): every layer ends with one more K=16 step against a constant "ones" operand whose B
// rows hold the bias as a bf16 hi+lo pair (Image::fwb_off), so the epilogue is LDTM -> relu+pack -> STS only.  A ring
// step costs the issuing thread ~300 clk of control code whatever it issues (tools/issue_cost.py: an mbarrier
// try_wait that succeeds at once is 88 clk alone), so each tile has its own issuer warp (2 MMAs = 256 clk per step).
constexpr int kFwChunkK = 32;
constexpr int kFwStages = 4;
constexpr int kFwStageBytes = kFwChunkK * 256 * 2;              // 16 KB
constexpr int kFwBiasBlkBytes = 16 * 256 * 2;                   // [2 k-groups][256][8] bf16 per layer T0..T8
constexpr int kFwThreads = 19 * 32;                             // 16 epilogue warps | producer | one MMA issuer per tile
constexpr int kFwSmOnes = 2 * kABytes;                          // 163,840: ONE 8x8 core matrix (128 B), elements 0,1 of each row = 1;
                                                                //   the ones-operand descriptor has LBO = SBO = 0 (broadcast)
constexpr int kFwSmRing = kFwSmOnes + 128;                      // 163,968
constexpr int kFwSmMisc = kFwSmRing + kFwStages * kFwStageBytes;  // 229,504: T9 bias (128) | w_density (256) | b_density
constexpr int kFwSmBar = kFwSmMisc + 400 * 4;
constexpr int kFwSmDens = kFwSmBar + 8 * (2 * kFwStages + 4) + 16;
constexpr int kFwSmTotal = kFwSmDens + 2 * kTile * 4;
static_assert(kFwSmTotal <= 227 * 1024, "forward shared memory budget");

__host__ __device__ inline int layer_k(int l) { return l == 0 ? 64 : (l == 5 ? 320 : (l == 8 ? 288 : 256)); }
__host__ __device__ inline int layer_akg0(int l) { return l == 0 ? kExtKg : 0; }

// dgrad chain (training): layer d = 0..8 computes dX = G * W for forward layers T9, T8(rgb rows only: the
// instance branch reads h.detach()), T7, ..., T1; K = forward outputs feeding a gradient, N = 256 forward inputs.
constexpr int kDgLayers = 9;
// fp32 copies kept for the chain rule through the two composed pairs (unpack_grads): floats
constexpr int kF32Rgbf = 0, kF32Insf = 65536, kF32Rgb0 = 131072, kF32Ins0 = 131072 + 32768,
              kF32BRgbf = 131072 + 65536, kF32BInsf = kF32BRgbf + 256, kF32Floats = kF32BInsf + 256;

#ifdef DMNERF_EXPERIMENTAL
static_assert(tc2::kExtKg == kExtKg, "mlp_tc2.cuh out of sync");
#endif

struct Image {
    int64_t chunk_off[kLayers];   // byte offset of the layer's first chunk
    int n[kLayers];               // MMA N of the layer
    int64_t bias_off;             // fp32 block (kBiasFloats)
    int64_t scratch_off;          // fp32 scratch: Wc_rgb[128][256], Wc_ins[128][256], bc[256]
    int64_t dg_off[kDgLayers];    // dgrad chunk streams ([k-group][256 rows][8 bf16] per 32-wide K chunk)
    int dg_k[kDgLayers];
    int64_t f32_off;              // kF32Floats fp32 copies (rgb/ins feature layers) for unpack_grads
    int64_t v2_off;               // chunk stream of the pipelined forward (mlp_tc2.cuh), in issue order
    int64_t v2_sched_off;         // tc2::Chunk[v2_chunks]
    int v2_chunks;
    int64_t fwb_off;              // 9 bias blocks of the forward (kFwBiasBlkBytes each)
    int64_t v3_off[2];            // per-CTA-rank chunk streams of the CTA-pair forward (mlp_tc3.cuh)
    int64_t v3_sched_off;         // tc3::Chunk[v3_chunks] then int layer_start[11]
    int v3_chunks, npad2;         // npad2 = C rounded up to 32 (N of the pair's head MMA)
    int64_t total_bytes;
    int npad, ins_num, kp9;       // kp9 = C rounded up to 32: K of the first dgrad layer
};

__host__ __device__ inline Image make_image(int ins_num) {
    Image im;
    im.ins_num = ins_num;
    im.npad = ((ins_num + 5) + 15) & ~15;
    int64_t off = 0;
    for (int l = 0; l < kLayers; ++l) {
        im.n[l] = (l == 9) ? im.npad : 256;
        im.chunk_off[l] = off;
        off += static_cast<int64_t>(layer_k(l)) * im.n[l] * 2;
    }
    im.bias_off = off;       off += kBiasFloats * 4;
    im.scratch_off = off;    off += (2 * 128 * 256 + 256) * 4;
    im.kp9 = ((ins_num + 5) + 31) & ~31;
    for (int d = 0; d < kDgLayers; ++d) {
        im.dg_k[d] = d == 0 ? im.kp9 : (d == 1 ? 128 : 256);
        im.dg_off[d] = off;
        off += static_cast<int64_t>(im.dg_k[d]) * 256 * 2;
    }
    im.f32_off = off;        off += static_cast<int64_t>(kF32Floats) * 4;
    im.fwb_off = off;        off += 9 * kFwBiasBlkBytes;
    im.npad2 = ((ins_num + 5) + 31) & ~31;
    im.v2_off = im.v2_sched_off = im.v3_off[0] = im.v3_off[1] = im.v3_sched_off = 0;
    im.v2_chunks = im.v3_chunks = 0;
#ifdef DMNERF_EXPERIMENTAL
    uint32_t v2_bytes = 0;
    im.v2_chunks = tc2::build_schedule(im.npad, nullptr, &v2_bytes);
    im.v2_off = off;         off += v2_bytes;
    im.v2_sched_off = off;   off += static_cast<int64_t>(tc2::kMaxChunks) * sizeof(tc2::Chunk);
    uint32_t v3_bytes = 0;
    im.v3_chunks = tc3::build_schedule(im.npad2, nullptr, nullptr, &v3_bytes);
    im.v3_off[0] = off;      off += v3_bytes;
    im.v3_off[1] = off;      off += v3_bytes;
    im.v3_sched_off = off;   off += static_cast<int64_t>(tc3::kMaxChunks) * sizeof(tc3::Chunk) + 16 * sizeof(int);
#endif
    im.total_bytes = off;
    return im;
}

// ---- pack helpers ----
struct Params { const float* p[kNumParams]; };


__device__ inline float logical_weight(const Params& pr, const float* scratch, int l, int n, int k, int ins_num) {
    if (l == 0) return k < 63 ? pr.p[0][n * 63 + k] : 0.f;
    if (l == 5) return k < 319 ? pr.p[10][n * 319 + k] : 0.f;
    if (l <= 7) return pr.p[2 * l][n * 256 + k];
    if (l == 8) {
        if (n < 128) {
            if (k < 256) return scratch[n * 256 + k];
            return k < 283 ? pr.p[P_RGB0_W][n * 283 + k] : 0.f;
        }
        return k < 256 ? scratch[32768 + (n - 128) * 256 + k] : 0.f;
    }
    // l == 9: block-diagonal heads.  rows 0..2 rgb (k<128), row 3 density (filled in the epilogue), rows 4.. ins
    if (n < 3) return k < 128 ? pr.p[P_RGBL_W][n * 128 + k] : 0.f;
    if (n >= 4 && n < 5 + ins_num) return k >= 128 ? pr.p[P_INSL_W][(n - 4) * 128 + (k - 128)] : 0.f;
    return 0.f;
}


// ---- device helpers ---------------------------------------------------------------------------------------
// mbarrier addresses (8 B each) inside the barrier block
struct FwBars {
    uint32_t base;
    __device__ __forceinline__ uint32_t w_full(int s) const { return base + 8 * s; }
    __device__ __forceinline__ uint32_t w_empty(int s) const { return base + 8 * (kFwStages + s); }
    __device__ __forceinline__ uint32_t a_ready(int t) const { return base + 8 * (2 * kFwStages + t); }
    __device__ __forceinline__ uint32_t acc_ready(int t) const { return base + 8 * (2 * kFwStages + 2 + t); }
};

struct Bars {
    uint32_t base;
    __device__ __forceinline__ uint32_t w_full(int s) const { return base + 8 * s; }
    __device__ __forceinline__ uint32_t w_empty(int s) const { return base + 8 * (kStages + s); }
    __device__ __forceinline__ uint32_t a_ready(int t) const { return base + 8 * (2 * kStages + t); }
    __device__ __forceinline__ uint32_t acc_ready(int t) const { return base + 8 * (2 * kStages + 2 + t); }
};

__device__ __forceinline__ uint64_t op_desc(uint32_t addr, uint32_t lbo, uint32_t sbo, int swap) {
    return swap ? ptx::smem_desc(addr, sbo, lbo) : ptx::smem_desc(addr, lbo, sbo);
}

// 8 consecutive features of one row -> one 16-byte store into k-group `kg` of the A operand; when `gst` is
// given (training; already offset to this thread's row inside its 64-point stash block), the same 16 bytes
// also go to slab `gkg` of the HBM stash (a warp writes 512 contiguous bytes).
template <bool F16 = false>
__device__ __forceinline__ void store_kg(uint32_t a_base, int kg, int row, const float v[8], bool relu,
                                         uint8_t* gst = nullptr, int gkg = 0) {
    const uint32_t w0 = ptx::pack_h2<F16>(v[0], v[1], relu), w1 = ptx::pack_h2<F16>(v[2], v[3], relu);
    const uint32_t w2 = ptx::pack_h2<F16>(v[4], v[5], relu), w3 = ptx::pack_h2<F16>(v[6], v[7], relu);
    ptx::st_shared_v4(a_base + kg * kKgBytes + row * 16, w0, w1, w2, w3);
    if (gst) *reinterpret_cast<uint4*>(gst + gkg * kStKgBytes) = make_uint4(w0, w1, w2, w3);
}

// same, and folds the 8 "is positive" flags of the ReLU outputs into the mask word m (mlp_layout.cuh: word j of
// k-group g -> bits g*4+j and 16+g*4+j).  After cvt.rn.relu a bf16 half is in [0, 0x7FFF], so adding 0x7FFF sets its
// top bit exactly when it is non-zero and never carries into the other half.
__device__ __forceinline__ uint32_t store_kg_mask(uint32_t a_base, int kg, int row, const float v[8], uint8_t* gst, int gkg,
                                                  int g, uint32_t m) {
    const uint32_t w[4] = {ptx::pack_bf16(v[0], v[1], true), ptx::pack_bf16(v[2], v[3], true),
                           ptx::pack_bf16(v[4], v[5], true), ptx::pack_bf16(v[6], v[7], true)};
    ptx::st_shared_v4(a_base + kg * kKgBytes + row * 16, w[0], w[1], w[2], w[3]);
    __stcs(reinterpret_cast<uint4*>(gst + gkg * kStKgBytes), make_uint4(w[0], w[1], w[2], w[3]));
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int k = g * 4 + j;
        m |= ((w[j] + 0x7FFF7FFFu) >> (15 - k)) & (0x00010001u << k);
    }
    return m;
}

// Embedding of a 3-vector into EXT k-groups [KG0, KG1): channel c = [x | f: sin(3) cos(3)], 8 channels per
// k-group.  The first needed octave uses full-range sincosf on x*2^f (exact scaling); higher octaves follow by
// angle doubling (error << bf16 resolution).  Channels outside the range are dead code after unrolling.
template <int F, int KG0, int KG1, bool F16 = false>
__device__ __forceinline__ void embed_to_ext(uint32_t a_base, int row, const float x[3], uint8_t* gst = nullptr,
                                             int gkg0 = 0) {
    constexpr int CH = 3 + 6 * F;
    constexpr int C0 = KG0 * 8, C1 = KG1 * 8;
    constexpr int F0 = C0 < 3 ? 0 : (C0 - 3) / 6;           // first octave that touches [C0, C1)
    float ch[(KG1 - KG0) * 8];
#pragma unroll
    for (int i = 0; i < (KG1 - KG0) * 8; ++i) ch[i] = 0.f;
    float s[3], c[3];
#pragma unroll
    for (int d = 0; d < 3; ++d) {
        if (C0 == 0) ch[d] = x[d];
        sincosf(x[d] * static_cast<float>(1 << F0), &s[d], &c[d]);
    }
#pragma unroll
    for (int f = F0; f < F; ++f) {
#pragma unroll
        for (int d = 0; d < 3; ++d) {
            const int cs = 3 + 6 * f + d, cc = cs + 3;
            if (cs >= C0 && cs < C1 && cs < CH) ch[cs - C0] = s[d];
            if (cc >= C0 && cc < C1 && cc < CH) ch[cc - C0] = c[d];
            const float s2 = 2.f * s[d] * c[d];
            const float c2 = (c[d] - s[d]) * (c[d] + s[d]);
            s[d] = s2; c[d] = c2;
        }
    }
#pragma unroll
    for (int g = KG0; g < KG1; ++g) store_kg<F16>(a_base, kExtKg + g, row, &ch[(g - KG0) * 8], false, gst, gkg0 + g);
}

template <bool F16 = false>
__device__ __forceinline__ void row_to_ext(uint32_t a_base, int row, const float* __restrict__ src, int count, int kg0,
                                           int kg1, uint8_t* gst = nullptr, int gkg0 = 0) {
    for (int g = kg0; g < kg1; ++g) {
        float v[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) v[e] = (src && g * 8 + e < count) ? __ldg(src + g * 8 + e) : 0.f;
        store_kg<F16>(a_base, kExtKg + g, row, v, false, gst, gkg0 + g);
    }
}


}  // namespace tc
}  // namespace dmnerf
