// mlp_tc2.cuh -- chunk schedule of the PIPELINED tensor-core forward (mlp_tc2.cu), shared by its pack kernels and
// the kernel itself.
//
// Why: in the lockstep kernel (mlp_tc.cu) the tensor pipe idles while a layer's epilogue runs and vice versa
// (ncu: tensor pipe 44.6 % active, ~1.7k of ~8.2k cycles per layer are pure MMA<->epilogue handoff latency).
// Here every layer's N=256 outputs are split into two halves j (TMEM columns 128j..128j+127) and its K=256 hidden
// inputs into two halves g, giving four MMA "cells" (j,g) issued in the order (0,0) (0,1) (1,0) (1,1).  The
// activations are updated IN PLACE in shared memory, so the epilogue E_j of output half j may start once row j
// (its accumulator) and column j (every cell that still reads the old H[j]) are complete: E_0 after cell (1,0),
// E_1 after (1,1).  The next layer's cells (.,g) only wait for E_g, so E_0 overlaps cell (1,1) and E_1 overlaps the
// next layer's (0,0): the tensor pipe stays busy across layers.  M=128,N=128,K=16 issues in 72 cycles (measured;
// 128 for N=256), a 12 % tax that buys the overlap.  Biases ride in the GEMM: one extra K=16 step per cell row
// against a constant "ones" operand, its B rows holding bias_hi / bias_lo (bf16 pair ~ 16 mantissa bits), so the
// epilogue is LDTM -> relu+pack -> STS.128 only.
#pragma once
#include <stdint.h>
#include <cuda_runtime.h>

namespace dmnerf {
namespace tc2 {

constexpr int kExtKg = 32;                 // == tc::kExtKg (static_assert in mlp_tc_common.cuh)

constexpr int kChunkBytes = 16384;         // ring slot: K=64 x N=128 bf16 (T9: K=64 x npad <= 128)
constexpr int kStagesV2 = 3;
constexpr int kMaxChunks = 112;

// schedule entry flags
constexpr uint32_t kFFirst = 1;            // first MMA of this accumulator half in this layer (accumulate = 0)
constexpr uint32_t kFWait0 = 2;            // wait h_ready[t][0] before this chunk
constexpr uint32_t kFWait1 = 4;            // wait h_ready[t][1]
constexpr uint32_t kFCommit0 = 8;          // after this chunk: commit acc_done[t][0]
constexpr uint32_t kFCommit1 = 16;         // after this chunk: commit acc_done[t][1]
constexpr uint32_t kFOnes = 32;            // A operand = the shared "ones" block (bias step, K=16)

struct Chunk {
    uint16_t a_kg;        // first k-group of the A operand inside the tile's smem image (ignored for kFOnes)
    uint8_t ksteps;       // K / 16 (4; 2 for the dir-embedding columns; 1 for the bias step)
    uint8_t j;            // accumulator half
    uint8_t flags;
    uint8_t layer;
    uint16_t nrows;       // MMA N (128, or npad for the head layer)
    uint32_t off;         // byte offset in the v2 chunk stream
};

// A-operand k-groups inside a tile's smem image: H 0..31 | EXT 32..39
// Layers: T0 (A = pos-embedding, EXT) | T1..T4 | T5 (+ pos-embedding) | T6, T7 | T8 (+ dir-embedding, EXT) | T9 heads.
__host__ __device__ inline int build_schedule(int npad, Chunk* out, uint32_t* stream_bytes = nullptr) {
    int n = 0;
    uint32_t off = 0;
    auto add = [&](int layer, int j, int a_kg, int ksteps, uint32_t flags, int nrows) {
        if (out) out[n] = Chunk{static_cast<uint16_t>(a_kg), static_cast<uint8_t>(ksteps), static_cast<uint8_t>(j),
                                static_cast<uint8_t>(flags), static_cast<uint8_t>(layer), static_cast<uint16_t>(nrows), off};
        off += static_cast<uint32_t>(ksteps) * 16 * nrows * 2;
        ++n;
    };
    for (int l = 0; l < 9; ++l) {
        for (int j = 0; j < 2; ++j) {
            uint32_t first = kFFirst;
            if (l == 0) {                                   // K = 64 from EXT; E_j may start right after cell j
                add(l, j, kExtKg, 4, first | (j == 0 ? kFWait0 : kFWait1), 128);
                add(l, j, 0, 1, kFOnes | (j == 0 ? kFCommit0 : kFCommit1), 128);
                continue;
            }
            // ---- input half g = 0 (+ the pos-embedding columns of the skip layer, available early)
            uint32_t w = j == 0 ? kFWait0 : 0;              // (1,0) comes after (0,1), which already waited for E_1
            if (l == 5) { add(l, j, kExtKg, 4, first | w, 128); first = 0; w = 0; }
            for (int c = 0; c < 2; ++c) { add(l, j, c * 8, 4, first | w, 128); first = 0; w = 0; }
            add(l, j, 0, 1, kFOnes | (j == 1 ? kFCommit0 : 0), 128);     // after cell (1,0): row 0 and column 0 complete
            // ---- input half g = 1 (+ the dir-embedding columns of T8, written during E_0 of T7)
            w = j == 0 ? kFWait1 : 0;
            for (int c = 0; c < 2; ++c) {
                const bool last = c == 1 && l != 8;
                add(l, j, 16 + c * 8, 4, w | (last && j == 1 ? kFCommit1 : 0), 128);
                w = 0;
            }
            if (l == 8) add(l, j, kExtKg, 2, j == 1 ? kFCommit1 : 0, 128);
        }
    }
    // T9: heads, N = npad, one accumulator half
    for (int c = 0; c < 4; ++c)
        add(9, 0, c * 8, 4, (c == 0 ? kFFirst | kFWait0 : 0) | (c == 2 ? kFWait1 : 0) | (c == 3 ? kFCommit0 : 0), npad);
    if (stream_bytes) *stream_bytes = off;
    return n;
}

}  // namespace tc2
}  // namespace dmnerf
