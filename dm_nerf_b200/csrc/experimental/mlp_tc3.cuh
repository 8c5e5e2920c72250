// mlp_tc3.cuh -- chunk schedule of the CTA-pair forward (mlp_tc3.cu): per layer, the K chunks (64 wide) of one
// tile's MMA pass plus one K=16 bias step.  Each CTA of the pair streams only ITS half of every weight chunk (its 128
// of the 256 output rows; npad2/2 for the head layer), so a chunk is [K/8 k-groups][rows][8 bf16] <= 16 KB per CTA.
#pragma once
#include <stdint.h>
#include <cuda_runtime.h>

namespace dmnerf {
namespace tc3 {

constexpr int kExtKg = 32;
constexpr int kSlotBytes = 16384;
constexpr int kStagesV3 = 3;
constexpr int kMaxChunks = 64;

constexpr uint32_t kFirst = 1;     // first chunk of a (tile, layer) pass: wait a_ready, accumulate = 0
constexpr uint32_t kLast = 2;      // last chunk: commit acc_ready
constexpr uint32_t kOnes = 4;      // A operand = the shared "ones" block (bias step)

struct Chunk {
    uint16_t a_kg;
    uint8_t ksteps, flags, layer, pad;
    uint16_t rows;        // B rows held by ONE CTA (N / 2)
    uint32_t off;         // byte offset inside one CTA's stream
};

// layer_start[l] .. layer_start[l+1]: the chunks of layer l (l = 0..9); returns the chunk count.
__host__ __device__ inline int build_schedule(int npad2, Chunk* out, int* layer_start, uint32_t* stream_bytes = nullptr) {
    int n = 0;
    uint32_t off = 0;
    auto add = [&](int layer, int a_kg, int ksteps, uint32_t flags, int rows) {
        if (out) out[n] = Chunk{static_cast<uint16_t>(a_kg), static_cast<uint8_t>(ksteps), static_cast<uint8_t>(flags),
                                static_cast<uint8_t>(layer), 0, static_cast<uint16_t>(rows), off};
        off += static_cast<uint32_t>(ksteps) * 16 * rows * 2;
        ++n;
    };
    for (int l = 0; l < 9; ++l) {
        if (layer_start) layer_start[l] = n;
        uint32_t first = kFirst;
        if (l == 0 || l == 5) { add(l, kExtKg, 4, first, 128); first = 0; }          // pos-embedding columns
        if (l != 0) for (int c = 0; c < 4; ++c) { add(l, c * 8, 4, first, 128); first = 0; }
        if (l == 8) add(l, kExtKg, 2, 0, 128);                                        // dir-embedding columns
        add(l, 0, 1, kOnes | kLast, 128);                                             // bias
    }
    if (layer_start) layer_start[9] = n;
    for (int c = 0; c < 4; ++c) add(9, c * 8, 4, (c == 0 ? kFirst : 0) | (c == 3 ? kLast : 0), npad2 / 2);
    if (layer_start) layer_start[10] = n;
    if (stream_bytes) *stream_bytes = off;
    return n;
}

}  // namespace tc3
}  // namespace dmnerf
