// chunk_pipe.cuh -- per-warp bulk-copy pipeline for the warp-per-ray kernels (compositing, penalizer).
//
// A ray's raw row is S*C contiguous floats.  Each warp walks its rays in 32-sample chunks (32*C floats =
// 128*C bytes, a multiple of 16) and keeps STAGES chunks in flight with 1-D bulk async copies (TMA engine,
// UBLKCP) into its private shared-memory ring, one mbarrier per stage.  Producer and consumer are the same
// warp, so releasing a stage is just the __syncwarp() that ends the chunk's reads; lane 0 then re-arms it.
// This replaces scalar __ldg staging (one chunk in flight per warp) and is what lifts these kernels from ~30 %
// to the HBM roofline: bytes in flight per SM = warps * STAGES * 128*C.
// Requirements (checked by the launchers): base pointer 16-byte aligned and S*C % 4 == 0.
#pragma once
#include "common.cuh"
#include "tc_ptx.cuh"

namespace dmnerf {

template <int STAGES>
struct ChunkPipe {
    const float* base;        // raw [n, S, C]
    uint32_t tile0;           // smem address of this warp's ring (STAGES * chunk_bytes)
    uint32_t bar0;            // smem address of this warp's STAGES mbarriers
    float* tile_ptr;          // generic pointer to the ring
    int64_t ray0, ray_stride, total;
    int S, C, CH, nchunk, steps, chunk_floats, lane;   // CH = samples per chunk (32, or 16 for wide rows)
    int64_t issued, consumed;
    bool there_and_back;      // per ray: chunks 0..n-1, then n-1..0 (recompute pass + reverse gradient pass)
    bool stores;              // the consumer bulk-stores tiles back out: wait for those reads before refilling
    bool reverse;             // per ray: chunks n-1..0 only (single reverse gradient pass)

    __device__ __forceinline__ void init(const float* raw, float* ring, uint64_t* bars, int64_t first_ray, int64_t stride,
                                         int64_t n, int S_, int C_, int lane_, bool there_and_back_ = false,
                                         bool stores_ = false, bool reverse_ = false, int ch = 32) {
        base = raw; tile_ptr = ring; tile0 = ptx::smem_u32(ring); bar0 = ptx::smem_u32(bars);
        ray0 = first_ray; ray_stride = stride; S = S_; C = C_; lane = lane_;
        there_and_back = there_and_back_; stores = stores_; reverse = reverse_;
        CH = ch;
        nchunk = (S + CH - 1) / CH;
        steps = there_and_back ? 2 * nchunk : nchunk;
        chunk_floats = CH * C;
        const int64_t my_rays = first_ray < n ? (n - first_ray + stride - 1) / stride : 0;
        total = my_rays * steps;
        issued = consumed = 0;
        if (lane == 0) {
            for (int s = 0; s < STAGES; ++s) ptx::mbar_init(bar0 + 8 * s, 1);
            ptx::fence_barrier_init();
        }
        __syncwarp();
        for (int s = 0; s < STAGES - 1; ++s) issue();
    }
    __device__ __forceinline__ int chunk_of(int64_t k) const {
        const int c = static_cast<int>(k % steps);
        if (reverse) return nchunk - 1 - c;
        return c < nchunk ? c : 2 * nchunk - 1 - c;
    }
    __device__ __forceinline__ void issue() {
        if (issued < total && lane == 0) {
            if (stores) ptx::bulk_wait_read0();
            const int64_t ray = ray0 + (issued / steps) * ray_stride;
            const int c0 = chunk_of(issued) * CH;
            const int cnt = min(CH, S - c0);
            const uint32_t bytes = static_cast<uint32_t>(cnt) * C * 4;
            const int st = static_cast<int>(issued % STAGES);
            ptx::mbar_arrive_expect_tx(bar0 + 8 * st, bytes);
            ptx::bulk_g2s(tile0 + st * chunk_floats * 4, base + (ray * S + c0) * static_cast<int64_t>(C), bytes, bar0 + 8 * st);
        }
        ++issued;
    }
    // Blocks until the next chunk has landed; returns its tile.  ray / c0 describe it; `back` = second (reverse) pass.
    // The caller must have ended its reads of the previous tile with __syncwarp() (that stage is refilled here).
    __device__ __forceinline__ float* acquire(int64_t& ray, int& c0, bool* back = nullptr) {
        issue();                                            // refill the stage released by the previous chunk
        const int st = static_cast<int>(consumed % STAGES);
        ptx::mbar_wait(bar0 + 8 * st, static_cast<uint32_t>((consumed / STAGES) & 1));
        ray = ray0 + (consumed / steps) * ray_stride;
        c0 = chunk_of(consumed) * CH;
        if (back) *back = static_cast<int>(consumed % steps) >= nchunk;
        ++consumed;
        return tile_ptr + st * chunk_floats;
    }
    // The chunk the NEXT acquire() will return (valid while !done()): lets the consumer start dependent global loads
    // (e.g. the z values of that chunk) one chunk ahead.
    __device__ __forceinline__ void peek(int64_t& ray, int& c0) const {
        ray = ray0 + (consumed / steps) * ray_stride;
        c0 = chunk_of(consumed) * CH;
    }
    // Bulk-store the (modified) tile just acquired to dst; every lane must have finished writing it.
    __device__ __forceinline__ void store_tile(float* dst, const float* tile, uint32_t bytes) {
        ptx::fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) {
            ptx::bulk_s2g(dst, ptx::smem_u32(tile), bytes);
            ptx::bulk_commit();
        }
    }
    __device__ __forceinline__ void finish() {
        if (stores && lane == 0) ptx::bulk_wait0();
    }
    __device__ __forceinline__ bool done() const { return consumed >= total; }
};

}  // namespace dmnerf
