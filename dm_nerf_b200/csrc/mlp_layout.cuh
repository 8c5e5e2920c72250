// mlp_layout.cuh -- packed weight images and activation-stash layout shared by the MLP kernels.
//
// The DM_NeRF network (networks/dm_nerf.py:58-106) as the kernels see it, "big" layers first:
//   id  module                     K (padded)        N    act   input
//    0  mlps.0                      63 (64)          256  relu  pos-embedding
//   1-4 mlps.1..4                   256              256  relu  h
//    5  mlps.5                      256+63 (320)     256  relu  [h | pos-embedding]   (skip, :86-87)
//   6,7 mlps.6,7                    256              256  relu  h
//    8  ins_feature_linear          256              256  -     h7 (detached, :95)
//    9  ins_feature_linears.0       256              128  relu  ins_feat
//   10  rgb_feature_linear          256              256  -     h7
//   11  rgb_feature_linears.0       256+27 (288)     128  relu  [rgb_feat | dir-embedding]
// and three small heads evaluated as dot products: density (256->1), rgb (128->3), ins (128->K+1).
#pragma once
#include "common.cuh"

namespace dmnerf {

constexpr int kBigLayers = 12;

struct BigLayer {
    int w_idx;     // index of the weight tensor in the 30-entry state_dict order (bias = w_idx+1)
    int k_src;     // columns of the torch weight
    int k_pad;     // padded reduction length (multiple of 16)
    int n;         // outputs
    int k_grad;    // leading input columns that need a gradient (dgrad width); 0 = none
};

__host__ __device__ inline BigLayer big_layer(int id) {
    switch (id) {
        case 0:  return {0, 63, 64, 256, 0};
        case 5:  return {10, 319, 320, 256, 256};
        case 8:  return {P_INSF_W, 256, 256, 256, 0};      // h.detach(): no gradient into the trunk
        case 9:  return {P_INS0_W, 256, 256, 128, 256};
        case 10: return {P_RGBF_W, 256, 256, 256, 256};
        case 11: return {P_RGB0_W, 283, 288, 128, 256};
        default: return {2 * id, 256, 256, 256, 256};      // mlps.1-4, 6, 7
    }
}

// ---- fp32 packed image (DMNERF_MODE_FP32) ------------------------------------------------------------
// per big layer:  Wt [k_pad][n]  (transposed, zero padded)   | bias [n] | Wg [n][k_grad] (row-major copy
// trimmed to the columns that receive a gradient; used by the backward's dgrad)
// then heads:     w_density[256], b_density[1] (padded to 4), W_rgb[3][128], b_rgb[4], W_ins[K+1][128], b_ins[K+1 pad4]
struct Fp32Image {
    int64_t wt[kBigLayers];
    int64_t bias[kBigLayers];
    int64_t wg[kBigLayers];
    int64_t w_dens, b_dens, w_rgb, b_rgb, w_ins, b_ins;
    int64_t total_floats;
    int ins_num;
};

__host__ __device__ inline int64_t round4(int64_t x) { return (x + 3) & ~int64_t(3); }

__host__ __device__ inline Fp32Image fp32_image(int ins_num) {
    Fp32Image im;
    int64_t off = 0;
    for (int l = 0; l < kBigLayers; ++l) {
        const BigLayer b = big_layer(l);
        im.wt[l] = off;   off += static_cast<int64_t>(b.k_pad) * b.n;
        im.bias[l] = off; off += b.n;
        im.wg[l] = off;   off += static_cast<int64_t>(b.n) * b.k_grad;
    }
    im.w_dens = off; off += 256;
    im.b_dens = off; off += 4;
    im.w_rgb = off;  off += 3 * 128;
    im.b_rgb = off;  off += 4;
    im.w_ins = off;  off += static_cast<int64_t>(ins_num + 1) * 128;
    im.b_ins = off;  off += round4(ins_num + 1);
    im.total_floats = off;
    im.ins_num = ins_num;
    return im;
}

// ---- activation stash (training mode), per tile of 64 points: rows of 64 floats --------------------
constexpr int kTileM = 64;
constexpr int kActPos = 0;                    // 64 rows  pos-embedding (row 63 = 0)
constexpr int kActDir = 64;                   // 32 rows  dir-embedding (rows 27.. = 0)
constexpr int kActH0 = 96;                    // 8 x 256  post-ReLU trunk activations h0..h7
constexpr int kActInsF = kActH0 + 8 * 256;    // 256      ins_feature_linear output
constexpr int kActInsH = kActInsF + 256;      // 128      relu(ins_feature_linears.0)
constexpr int kActRgbF = kActInsH + 128;      // 256      rgb_feature_linear output
constexpr int kActRgbH = kActRgbF + 256;      // 128      relu(rgb_feature_linears.0)
constexpr int kActRows = kActRgbH + 128;      // 2912

// ---- bf16 stash of the tensor-core TRAINING path (DMNERF_MODE_BF16), per block of 64 points ----------
// The unit is one k-group slab [64 rows][8 bf16] = 1024 B; a region of F features is F/8 consecutive slabs,
// i.e. exactly a no-swizzle UMMA operand image: MN-major
// with the 64 points as the reduction dim for wgrad -- so wgrad moves whole regions with ONE 1-D bulk copy.
//   activations (written by the forward):  POS 8 | DIR 4 | H0..H7 8 x 32 | HID 32           = 300 slabs
//   gradients   (written by dgrad)      :  G0..G7 8 x 32 | G8 32 | G9 16 (C padded to 128)  = 304 slabs
// POS channel 63 (the pad) is unused; rows past n_points hold finite garbage in the activations and exact
// zeros in the gradients, so they add nothing to any weight gradient.
constexpr int kStBlk = 64;
constexpr int kStKgBytes = kStBlk * 16;
constexpr int kStPos = 0, kStDir = 8, kStH0 = 12, kStHid = 12 + 8 * 32, kStActKg = 300;
constexpr int kStG0 = 0, kStG8 = 8 * 32, kStG9 = 9 * 32, kStGradKg = 304;
constexpr int64_t kStActBlkBytes = static_cast<int64_t>(kStActKg) * kStKgBytes;    // 307,200
constexpr int64_t kStGradBlkBytes = static_cast<int64_t>(kStGradKg) * kStKgBytes;  // 311,296
// blocks are allocated for whole tile QUADS (512 points: two tiles in each CTA of a pair) of the persistent kernels
__host__ __device__ inline int64_t tc_stash_blocks(int64_t n_points) { return ((n_points + 511) / 512) * 8; }
// ReLU sign masks (1 bit per activation), appended to the activation stash after its last block: the dgrad chain only
// needs to know WHICH of h0..h7 / hid were positive, so it reads 288 B per point here instead of 4.6 KB of bf16
// activations.  Per 64-point block: [9 layers a: h0..h7, hid][2 column halves][4 words c][64 rows] uint32; word c of
// (a, half) covers features half*128 + c*32 .. +31 of layer a for one point: bit k (k = 0..15) is the feature
// c*32 + (k>>2)*8 + 2*(k&3), bit 16+k the feature after it (the order the epilogues' packed bf16 pairs come in).
constexpr int kStMaskLayers = 9;
constexpr int64_t kStMaskBlkBytes = static_cast<int64_t>(kStMaskLayers) * 2 * 4 * kStBlk * 4;   // 18,432
__host__ __device__ inline int64_t tc_mask_offset(int64_t n_points) { return tc_stash_blocks(n_points) * kStActBlkBytes; }
__host__ __device__ inline int64_t tc_mask_word(int64_t blk, int a, int half, int c, int row64) {
    return ((((blk * kStMaskLayers + a) * 2 + half) * 4 + c) * kStBlk + row64) * 4;
}

// Input of one MLP launch: either rays (points o + d*z and both embeddings computed in-kernel) or
// already-embedded rows x[m,90].
struct MlpInput {
    const float* rays_o;   // [n_rays,3]       (rays mode)
    const float* rays_d;   // [n_rays,3]
    const float* z;        // [n_rays,S]
    const float* x;        // [m,90]           (embedded mode; rays_* == nullptr)
    int64_t n_points;
    int S;
};

}  // namespace dmnerf
