// composite.cu -- a5/a8 volume compositing, one warp per ray.
//
// Forward (render_train, networks/render.py:6-28; manipulator_render, networks/manipulator.py:86-105):
//   delta_s = (z[s+1]-z[s]) * |d|   (last: 1e10*|d|)          alpha_s = 1 - exp(-relu(sigma_s)*delta_s)
//   T_s = prod_{r<s} (1 - alpha_r + 1e-10)   (warp-shuffle exclusive product scan + running carry)
//   w_s = alpha_s*T_s     rgb = sum w*sigmoid(raw_rgb)    depth = sum w*z    ins = sigmoid(sum w*raw_ins)
// Data movement: each 32-sample chunk of raw (32*C contiguous floats) is staged once into shared
// memory with fully coalesced loads; lanes then act as samples (scan) and as channels (logit sums).
// HBM traffic = the algorithmic bytes (raw read once); roofline: HBM.
#include "chunk_pipe.cuh"

namespace dmnerf {
int tc_debug_flags();

constexpr int kCompWarps = 8;

__device__ __forceinline__ void stage_chunk(float* tile, const float* __restrict__ src, int cnt, int C, int CP,
                                            float invC, int lane) {
    const int total = cnt * C;
    for (int i = lane; i < total; i += 32) {
        int row = __float2int_rz((static_cast<float>(i) + 0.5f) * invC);
        int col = i - row * C;
        tile[row * CP + col] = __ldg(src + i);
    }
}

__device__ __forceinline__ void unstage_chunk(const float* tile, float* __restrict__ dst, int cnt, int C, int CP,
                                              float invC, int lane) {
    const int total = cnt * C;
    for (int i = lane; i < total; i += 32) {
        int row = __float2int_rz((static_cast<float>(i) + 0.5f) * invC);
        int col = i - row * C;
        dst[i] = tile[row * CP + col];
    }
}

__global__ void __launch_bounds__(kCompWarps * 32)
composite_fwd_kernel(const float* __restrict__ raw, const float* __restrict__ z, const float* __restrict__ rays_d,
                     int64_t n, int S, int C, int flags, float* __restrict__ rgb, float* __restrict__ depth,
                     float* __restrict__ weights, float* __restrict__ ins, float* __restrict__ trans_out) {
    extern __shared__ float smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t ray = static_cast<int64_t>(blockIdx.x) * kCompWarps + warp;
    if (ray >= n) return;
    const int CP = C | 1;
    float* tile = smem + warp * (32 * CP + 32);
    float* wbuf = tile + 32 * CP;
    const float invC = 1.0f / static_cast<float>(C);
    const int L = C - 4;  // instance logits incl. the "empty" class

    const float dx = rays_d[ray * 3 + 0], dy = rays_d[ray * 3 + 1], dz = rays_d[ray * 3 + 2];
    const float norm = sqrtf(dx * dx + dy * dy + dz * dz);
    const float* zrow = z + ray * S;
    const float* rrow = raw + ray * S * static_cast<int64_t>(C);

    float carry = 1.0f, a_r = 0.f, a_g = 0.f, a_b = 0.f, a_d = 0.f;
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    for (int c0 = 0; c0 < S; c0 += 32) {
        const int cnt = min(32, S - c0);
        stage_chunk(tile, rrow + static_cast<int64_t>(c0) * C, cnt, C, CP, invC, lane);
        __syncwarp();
        const int s = c0 + lane;
        const bool valid = lane < cnt;
        float zc = valid ? zrow[s] : 0.f;
        float dist = 0.f, sigma = 0.f;
        if (valid) {
            dist = ((s + 1 < S) ? (zrow[s + 1] - zc) : 1e10f) * norm;
            sigma = tile[lane * CP + 3];
        }
        float alpha, trans, w;
        scan_chunk(sigma, dist, valid, lane, carry, alpha, trans, w);
        if (valid) {
            if (weights) weights[ray * S + s] = w;
            if (trans_out) trans_out[ray * S + s] = trans;
            a_r += w * sigmoidf_ref(tile[lane * CP + 0]);
            a_g += w * sigmoidf_ref(tile[lane * CP + 1]);
            a_b += w * sigmoidf_ref(tile[lane * CP + 2]);
            a_d += w * zc;
        }
        wbuf[lane] = w;
        __syncwarp();
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int j = lane + 32 * q;
            if (j < L) {
                float a = acc[q];
                const float* col = tile + 4 + j;
                for (int r = 0; r < cnt; ++r) a += wbuf[r] * col[r * CP];
                acc[q] = a;
            }
        }
        __syncwarp();
    }
    a_r = warp_sum(a_r); a_g = warp_sum(a_g); a_b = warp_sum(a_b); a_d = warp_sum(a_d);
    if (lane == 0) {
        rgb[ray * 3 + 0] = a_r; rgb[ray * 3 + 1] = a_g; rgb[ray * 3 + 2] = a_b;
        depth[ray] = a_d;
    }
    const int Lout = L - ((flags & DMNERF_COMP_DROP_EMPTY) ? 1 : 0);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const int j = lane + 32 * q;
        if (j < Lout) ins[ray * Lout + j] = (flags & DMNERF_COMP_RAW_INS) ? acc[q] : sigmoidf_ref(acc[q]);
    }
}

// Forward, bulk-copy staged (the default whenever raw is 16-byte aligned and S*C % 4 == 0): persistent warps,
// each walking rays gw, gw+W, ... with kCompStages chunks in flight (chunk_pipe.cuh).  Same arithmetic and
// summation order per ray as composite_fwd_kernel.  Logit sums: lanes are channels for full 32-channel passes;
// a remainder of <= 4 channels (K=32: the single "empty" class) is summed sample-parallel instead of spending a
// whole pass on one lane.
constexpr int kCompStages = 2;        // ring depth per warp.  Measured (tools/sweep_composite.py): resident warps matter more
                                      // than depth -- 2 stages x 20 warps/SM beats 3 x 12 by 15-30 %, 4 stages is worse still
constexpr int kCompBulkWarps = 4;

template <int ST>
__global__ void __launch_bounds__(kCompBulkWarps * 32)
composite_fwd_bulk_kernel(const float* __restrict__ raw, const float* __restrict__ z, const float* __restrict__ rays_d,
                          int64_t n, int S, int C, int flags, float* __restrict__ rgb, float* __restrict__ depth,
                          float* __restrict__ weights, float* __restrict__ ins, float* __restrict__ trans_out, int CH) {
    extern __shared__ __align__(128) float smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int per_warp = ST * CH * C + 32;
    float* ring = smem + warp * per_warp;
    float* wbuf = ring + ST * CH * C;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kCompBulkWarps * per_warp) + warp * ST;
    ChunkPipe<ST> pipe;
    pipe.init(raw, ring, bars, static_cast<int64_t>(blockIdx.x) * kCompBulkWarps + warp,
              static_cast<int64_t>(gridDim.x) * kCompBulkWarps, n, S, C, lane, false, false, false, CH);
    const int L = C - 4;
    const int Lout = L - ((flags & DMNERF_COMP_DROP_EMPTY) ? 1 : 0);
    const int rem = (L & 31) <= 4 ? (L & 31) : 0;          // channels summed sample-parallel
    const int Lp = L - rem;                                // channels summed channel-parallel (lanes = channels)

    float carry = 1.0f, a_r = 0.f, a_g = 0.f, a_b = 0.f, a_d = 0.f, norm = 0.f;
    float acc[4] = {0.f, 0.f, 0.f, 0.f}, racc[4] = {0.f, 0.f, 0.f, 0.f};
    // z (and the ray direction) do not come through the bulk pipe: they are loaded one chunk ahead so that their
    // latency hides behind the current chunk's arithmetic instead of sitting between acquire() and the scan
    float z_pre = 0.f, zn_pre = 0.f, norm_pre = 0.f;
    auto prefetch = [&]() {
        int64_t r2; int c2;
        pipe.peek(r2, c2);
        const int s2 = c2 + lane;
        if (s2 < S) {
            z_pre = __ldg(z + r2 * S + s2);
            zn_pre = (s2 + 1 < S) ? __ldg(z + r2 * S + s2 + 1) : 0.f;
        }
        if (c2 == 0) {
            const float dx = __ldg(rays_d + r2 * 3 + 0), dy = __ldg(rays_d + r2 * 3 + 1), dz = __ldg(rays_d + r2 * 3 + 2);
            norm_pre = sqrtf(dx * dx + dy * dy + dz * dz);
        }
    };
    if (!pipe.done()) prefetch();
    while (!pipe.done()) {
        int64_t ray; int c0;
        const float* tile = pipe.acquire(ray, c0);
        const float z_cur = z_pre, zn_cur = zn_pre;
        if (c0 == 0) {
            norm = norm_pre;
            carry = 1.0f; a_r = a_g = a_b = a_d = 0.f;
#pragma unroll
            for (int q = 0; q < 4; ++q) acc[q] = racc[q] = 0.f;
        }
        if (!pipe.done()) prefetch();
        const int cnt = min(CH, S - c0);
        const int s = c0 + lane;
        const bool valid = lane < cnt;
        float zc = 0.f, dist = 0.f, sigma = 0.f;
        if (valid) {
            zc = z_cur;
            dist = ((s + 1 < S) ? (zn_cur - zc) : 1e10f) * norm;
            sigma = tile[lane * C + 3];
        }
        float alpha, trans, w;
        scan_chunk(sigma, dist, valid, lane, carry, alpha, trans, w);
        if (valid) {
            if (weights) weights[ray * S + s] = w;
            if (trans_out) trans_out[ray * S + s] = trans;
            a_r += w * sigmoidf_ref(tile[lane * C + 0]);
            a_g += w * sigmoidf_ref(tile[lane * C + 1]);
            a_b += w * sigmoidf_ref(tile[lane * C + 2]);
            a_d += w * zc;
#pragma unroll
            for (int q = 0; q < 4; ++q)
                if (q < rem) racc[q] += w * tile[lane * C + 4 + Lp + q];
        }
        wbuf[lane] = w;
        __syncwarp();
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int j = lane + 32 * q;
            if (j < Lp) {
                float a = acc[q];
                const float* col = tile + 4 + j;
#pragma unroll 8
                for (int r = 0; r < cnt; ++r) a += wbuf[r] * col[r * C];
                acc[q] = a;
            }
        }
        if (c0 + CH >= S) {                                  // last chunk of this ray: write the outputs
            a_r = warp_sum(a_r); a_g = warp_sum(a_g); a_b = warp_sum(a_b); a_d = warp_sum(a_d);
            if (lane == 0) {
                rgb[ray * 3 + 0] = a_r; rgb[ray * 3 + 1] = a_g; rgb[ray * 3 + 2] = a_b;
                depth[ray] = a_d;
            }
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int j = lane + 32 * q;
                if (j < Lp && j < Lout) ins[ray * Lout + j] = (flags & DMNERF_COMP_RAW_INS) ? acc[q] : sigmoidf_ref(acc[q]);
            }
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                if (q < rem) {
                    const float v = warp_sum(racc[q]);
                    if (lane == 0 && Lp + q < Lout)
                        ins[ray * Lout + Lp + q] = (flags & DMNERF_COMP_RAW_INS) ? v : sigmoidf_ref(v);
                }
            }
        }
        __syncwarp();                                        // all reads of this tile done -> stage may be refilled
    }
}

// Backward.  Phase A recomputes alpha/T/w (kept in shared memory per ray) and the logit sums,
// phase C walks the chunks in reverse with a running suffix sum  R_s = sum_{r>s} dw_r*w_r:
//   dalpha_s = dw_s*T_s - R_s/(1-alpha_s+1e-10)        dsigma_s = [sigma>0]*dalpha_s*delta_s*(1-alpha_s)
__global__ void __launch_bounds__(kCompWarps * 32)
composite_bwd_kernel(const float* __restrict__ raw, const float* __restrict__ z, const float* __restrict__ rays_d,
                     int64_t n, int S, int C, int flags, int ins_detach, const float* __restrict__ d_rgb,
                     const float* __restrict__ d_depth, const float* __restrict__ d_weights,
                     const float* __restrict__ d_ins, float* __restrict__ d_raw) {
    extern __shared__ float smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t ray = static_cast<int64_t>(blockIdx.x) * kCompWarps + warp;
    if (ray >= n) return;
    const int CP = C | 1;
    float* tile = smem + warp * (32 * CP + 32 + 3 * S);
    float* wbuf = tile + 32 * CP;
    float* s_alpha = wbuf + 32;
    float* s_trans = s_alpha + S;
    float* s_w = s_trans + S;
    const float invC = 1.0f / static_cast<float>(C);
    const int L = C - 4;
    const int Lout = L - ((flags & DMNERF_COMP_DROP_EMPTY) ? 1 : 0);

    const float dx = rays_d[ray * 3 + 0], dy = rays_d[ray * 3 + 1], dz = rays_d[ray * 3 + 2];
    const float norm = sqrtf(dx * dx + dy * dy + dz * dz);
    const float* zrow = z + ray * S;
    const float* rrow = raw + ray * S * static_cast<int64_t>(C);
    float* grow = d_raw + ray * S * static_cast<int64_t>(C);

    // ---- phase A
    float carry = 1.0f;
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    for (int c0 = 0; c0 < S; c0 += 32) {
        const int cnt = min(32, S - c0);
        const bool need_logits = (d_ins != nullptr);
        if (need_logits) stage_chunk(tile, rrow + static_cast<int64_t>(c0) * C, cnt, C, CP, invC, lane);
        __syncwarp();
        const int s = c0 + lane;
        const bool valid = lane < cnt;
        float dist = 0.f, sigma = 0.f;
        if (valid) {
            float zc = zrow[s];
            dist = ((s + 1 < S) ? (zrow[s + 1] - zc) : 1e10f) * norm;
            sigma = need_logits ? tile[lane * CP + 3] : rrow[static_cast<int64_t>(s) * C + 3];
        }
        float alpha, trans, w;
        scan_chunk(sigma, dist, valid, lane, carry, alpha, trans, w);
        if (valid) { s_alpha[s] = alpha; s_trans[s] = trans; s_w[s] = w; }
        wbuf[lane] = w;
        __syncwarp();
        if (need_logits) {
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int j = lane + 32 * q;
                if (j < L) {
                    float a = acc[q];
                    const float* col = tile + 4 + j;
                    for (int r = 0; r < cnt; ++r) a += wbuf[r] * col[r * CP];
                    acc[q] = a;
                }
            }
        }
        __syncwarp();
    }
    // ---- phase B: gradient w.r.t. the accumulated logits
    float dA[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const int j = lane + 32 * q;
        float g = 0.f;
        if (d_ins != nullptr && j < Lout) {
            g = d_ins[ray * Lout + j];
            if (!(flags & DMNERF_COMP_RAW_INS)) {
                float p = sigmoidf_ref(acc[q]);
                g *= p * (1.0f - p);
            }
        }
        dA[q] = g;
    }
    const float g_r = d_rgb ? d_rgb[ray * 3 + 0] : 0.f;
    const float g_g = d_rgb ? d_rgb[ray * 3 + 1] : 0.f;
    const float g_b = d_rgb ? d_rgb[ray * 3 + 2] : 0.f;
    const float g_d = d_depth ? d_depth[ray] : 0.f;

    // ---- phase C
    float suffix = 0.f;
    const int last0 = ((S - 1) / 32) * 32;
    for (int c0 = last0; c0 >= 0; c0 -= 32) {
        const int cnt = min(32, S - c0);
        stage_chunk(tile, rrow + static_cast<int64_t>(c0) * C, cnt, C, CP, invC, lane);
        __syncwarp();
        const int s = c0 + lane;
        const bool valid = lane < cnt;
        float dw = 0.f, w = 0.f, alpha = 0.f, trans = 0.f, zc = 0.f;
        float c_r = 0.f, c_g = 0.f, c_b = 0.f;
        if (valid) {
            w = s_w[s]; alpha = s_alpha[s]; trans = s_trans[s]; zc = zrow[s];
            c_r = sigmoidf_ref(tile[lane * CP + 0]);
            c_g = sigmoidf_ref(tile[lane * CP + 1]);
            c_b = sigmoidf_ref(tile[lane * CP + 2]);
            dw = g_r * c_r + g_g * c_g + g_b * c_b + g_d * zc;
            if (d_weights) dw += d_weights[ray * S + s];
        }
        if (!ins_detach && d_ins != nullptr) {
            for (int r = 0; r < cnt; ++r) {
                float part = 0.f;
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int j = lane + 32 * q;
                    if (j < L) part += dA[q] * tile[r * CP + 4 + j];
                }
                part = warp_sum(part);
                if (lane == r) dw += part;
            }
        }
        // suffix sums of g = dw*w over later samples
        float g = valid ? dw * w : 0.f;
        float incl = g;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            float v = __shfl_down_sync(0xffffffffu, incl, o);
            if (lane + o < 32) incl += v;
        }
        const float after = suffix + (incl - g);
        suffix += __shfl_sync(0xffffffffu, incl, 0);
        if (valid) {
            const float t = (1.0f - alpha) + 1e-10f;
            const float dalpha = dw * trans - after / t;
            const float sigma = tile[lane * CP + 3];
            const float dist = ((s + 1 < S) ? (zrow[s + 1] - zc) : 1e10f) * norm;
            const float dsigma = (sigma > 0.f) ? dalpha * dist * (1.0f - alpha) : 0.f;
            tile[lane * CP + 0] = g_r * w * c_r * (1.0f - c_r);
            tile[lane * CP + 1] = g_g * w * c_g * (1.0f - c_g);
            tile[lane * CP + 2] = g_b * w * c_b * (1.0f - c_b);
            tile[lane * CP + 3] = dsigma;
        }
        wbuf[lane] = w;
        __syncwarp();
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int j = lane + 32 * q;
            if (j < L) {
                float* col = tile + 4 + j;
                for (int r = 0; r < cnt; ++r) col[r * CP] = dA[q] * wbuf[r];
            }
        }
        __syncwarp();
        unstage_chunk(tile, grow + static_cast<int64_t>(c0) * C, cnt, C, CP, invC, lane);
        __syncwarp();
    }
}

// Backward, bulk-copy staged: same phases as composite_bwd_kernel, but a ray's chunks arrive through the per-warp
// bulk-copy ring twice (forward for the recompute pass, then in reverse for the gradient pass) and every finished
// d_raw chunk leaves with one bulk store from the tile it was built in.
template <int ST>
__global__ void __launch_bounds__(kCompBulkWarps * 32)
composite_bwd_bulk_kernel(const float* __restrict__ raw, const float* __restrict__ z, const float* __restrict__ rays_d,
                          int64_t n, int S, int C, int flags, int ins_detach, const float* __restrict__ d_rgb,
                          const float* __restrict__ d_depth, const float* __restrict__ d_weights,
                          const float* __restrict__ d_ins, float* __restrict__ d_raw) {
    extern __shared__ __align__(128) float smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int per_warp = ST * 32 * C + 32 + 5 * S;
    float* ring = smem + warp * per_warp;
    float* wbuf = ring + ST * 32 * C;
    float* s_alpha = wbuf + 32;
    float* s_trans = s_alpha + S;
    float* s_w = s_trans + S;
    float* s_z = s_w + S;
    float* s_dist = s_z + S;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kCompBulkWarps * per_warp + (kCompBulkWarps * per_warp & 1)) + warp * ST;
    ChunkPipe<ST> pipe;
    pipe.init(raw, ring, bars, static_cast<int64_t>(blockIdx.x) * kCompBulkWarps + warp,
              static_cast<int64_t>(gridDim.x) * kCompBulkWarps, n, S, C, lane, true, true);
    const int L = C - 4;
    const int Lout = L - ((flags & DMNERF_COMP_DROP_EMPTY) ? 1 : 0);
    const bool need_logits = d_ins != nullptr;

    float carry = 1.0f, norm = 0.f, suffix = 0.f, g_r = 0.f, g_g = 0.f, g_b = 0.f, g_d = 0.f;
    float acc[4] = {0.f, 0.f, 0.f, 0.f}, dA[4] = {0.f, 0.f, 0.f, 0.f};
    // z, the ray direction and the per-ray output gradients do not come through the bulk pipe: z is loaded one
    // chunk ahead (phase A only; phase C re-reads it from shared memory), the gradients at the start of the ray
    float z_pre = 0.f, zn_pre = 0.f, norm_pre = 0.f;
    auto prefetch = [&]() {
        int64_t r2; int c2;
        pipe.peek(r2, c2);
        const int s2 = c2 + lane;
        if (s2 < S) {
            z_pre = __ldg(z + r2 * S + s2);
            zn_pre = (s2 + 1 < S) ? __ldg(z + r2 * S + s2 + 1) : 0.f;
        }
        if (c2 == 0) {
            const float dx = __ldg(rays_d + r2 * 3 + 0), dy = __ldg(rays_d + r2 * 3 + 1), dz = __ldg(rays_d + r2 * 3 + 2);
            norm_pre = sqrtf(dx * dx + dy * dy + dz * dz);
        }
    };
    if (!pipe.done()) prefetch();
    while (!pipe.done()) {
        int64_t ray; int c0; bool back;
        float* tile = pipe.acquire(ray, c0, &back);
        const int cnt = min(32, S - c0);
        const int s = c0 + lane;
        const bool valid = lane < cnt;
        if (!back) {
            // ---- phase A: recompute alpha / T / w and the logit sums
            const float z_cur = z_pre, zn_cur = zn_pre;
            if (c0 == 0) {
                norm = norm_pre;
                carry = 1.0f;
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    acc[q] = 0.f;
                    const int j = lane + 32 * q;
                    dA[q] = (need_logits && j < Lout) ? __ldg(d_ins + ray * Lout + j) : 0.f;
                }
                g_r = d_rgb ? __ldg(d_rgb + ray * 3 + 0) : 0.f;
                g_g = d_rgb ? __ldg(d_rgb + ray * 3 + 1) : 0.f;
                g_b = d_rgb ? __ldg(d_rgb + ray * 3 + 2) : 0.f;
                g_d = d_depth ? __ldg(d_depth + ray) : 0.f;
            }
            if (c0 + 32 < S) prefetch();                 // next chunk is still phase A of this ray
            float dist = 0.f, sigma = 0.f;
            if (valid) {
                dist = ((s + 1 < S) ? (zn_cur - z_cur) : 1e10f) * norm;
                sigma = tile[lane * C + 3];
                s_z[s] = z_cur; s_dist[s] = dist;
            }
            float alpha, trans, w;
            scan_chunk(sigma, dist, valid, lane, carry, alpha, trans, w);
            if (valid) { s_alpha[s] = alpha; s_trans[s] = trans; s_w[s] = w; }
            wbuf[lane] = w;
            __syncwarp();
            if (need_logits) {
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int j = lane + 32 * q;
                    if (j < L) {
                        float a = acc[q];
                        const float* col = tile + 4 + j;
#pragma unroll 8
                        for (int r = 0; r < cnt; ++r) a += wbuf[r] * col[r * C];
                        acc[q] = a;
                    }
                }
            }
            if (c0 + 32 >= S) {
                // ---- phase B: gradient w.r.t. the accumulated logits, per-ray output gradients
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int j = lane + 32 * q;
                    if (need_logits && j < Lout && !(flags & DMNERF_COMP_RAW_INS)) {
                        const float p = sigmoidf_ref(acc[q]);
                        dA[q] *= p * (1.0f - p);
                    }
                }
                suffix = 0.f;
            }
            __syncwarp();
            continue;
        }
        // ---- phase C (reverse): running suffix sum R_s = sum_{r>s} dw_r w_r
        float dw = 0.f, w = 0.f, alpha = 0.f, trans = 0.f, zc = 0.f, c_r = 0.f, c_g = 0.f, c_b = 0.f;
        if (valid) {
            w = s_w[s]; alpha = s_alpha[s]; trans = s_trans[s]; zc = s_z[s];
            c_r = sigmoidf_ref(tile[lane * C + 0]);
            c_g = sigmoidf_ref(tile[lane * C + 1]);
            c_b = sigmoidf_ref(tile[lane * C + 2]);
            dw = g_r * c_r + g_g * c_g + g_b * c_b + g_d * zc;
            if (d_weights) dw += d_weights[ray * S + s];
        }
        if (!ins_detach && need_logits) {
            for (int r = 0; r < cnt; ++r) {
                float part = 0.f;
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int j = lane + 32 * q;
                    if (j < L) part += dA[q] * tile[r * C + 4 + j];
                }
                part = warp_sum(part);
                if (lane == r) dw += part;
            }
        }
        const float g = valid ? dw * w : 0.f;
        float incl = g;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const float v = __shfl_down_sync(0xffffffffu, incl, o);
            if (lane + o < 32) incl += v;
        }
        const float after = suffix + (incl - g);
        suffix += __shfl_sync(0xffffffffu, incl, 0);
        __syncwarp();
        if (valid) {
            const float t = (1.0f - alpha) + 1e-10f;
            const float dalpha = dw * trans - after / t;
            const float sigma = tile[lane * C + 3];
            const float dist = s_dist[s];
            const float dsigma = (sigma > 0.f) ? dalpha * dist * (1.0f - alpha) : 0.f;
            tile[lane * C + 0] = g_r * w * c_r * (1.0f - c_r);
            tile[lane * C + 1] = g_g * w * c_g * (1.0f - c_g);
            tile[lane * C + 2] = g_b * w * c_b * (1.0f - c_b);
            tile[lane * C + 3] = dsigma;
        }
        wbuf[lane] = w;
        __syncwarp();
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int j = lane + 32 * q;
            if (j < L) {
                float* col = tile + 4 + j;
                const float a = dA[q];
#pragma unroll 8
                for (int r = 0; r < cnt; ++r) col[r * C] = a * wbuf[r];
            }
        }
        pipe.store_tile(d_raw + (ray * S + c0) * static_cast<int64_t>(C), tile, static_cast<uint32_t>(cnt) * C * 4);
        if (c0 == 0 && !pipe.done()) prefetch();         // next chunk opens the next ray's phase A
        __syncwarp();
    }
    pipe.finish();
}

// Backward in ONE reverse pass, for callers that kept the forward's per-sample weights and transmittances and its
// composited instance map (the training path does): what the recompute pass of composite_bwd_bulk_kernel rebuilds --
// T_s, w_s, and sigmoid'(sum_s w_s logit_s) = p (1 - p) with p the forward's own output -- is read back instead
// (8 B per sample), so every raw chunk crosses L2 -> shared memory once instead of twice and the warp never walks
// a ray forwards.  alpha is recomputed from sigma and the sample spacing with the forward's formula.
template <int ST>
__global__ void __launch_bounds__(kCompBulkWarps * 32)
composite_bwd_rev_kernel(const float* __restrict__ raw, const float* __restrict__ z, const float* __restrict__ rays_d,
                         int64_t n, int S, int C, int flags, int ins_detach, const float* __restrict__ d_rgb,
                         const float* __restrict__ d_depth, const float* __restrict__ d_weights,
                         const float* __restrict__ d_ins, float* __restrict__ d_raw, const float* __restrict__ fwd_w,
                         const float* __restrict__ fwd_t, const float* __restrict__ fwd_ins, int CH) {
    extern __shared__ __align__(128) float smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int per_warp = ST * CH * C + 32;
    float* ring = smem + warp * per_warp;
    float* wbuf = ring + ST * CH * C;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kCompBulkWarps * per_warp + (kCompBulkWarps * per_warp & 1)) + warp * ST;
    ChunkPipe<ST> pipe;
    pipe.init(raw, ring, bars, static_cast<int64_t>(blockIdx.x) * kCompBulkWarps + warp,
              static_cast<int64_t>(gridDim.x) * kCompBulkWarps, n, S, C, lane, false, true, true, CH);
    const int L = C - 4;
    const int Lout = L - ((flags & DMNERF_COMP_DROP_EMPTY) ? 1 : 0);
    const bool need_logits = d_ins != nullptr;
    const int last_c0 = ((S - 1) / CH) * CH;

    float norm = 0.f, suffix = 0.f, g_r = 0.f, g_g = 0.f, g_b = 0.f, g_d = 0.f;
    float dA[4] = {0.f, 0.f, 0.f, 0.f};
    // per-sample scalars of the NEXT chunk (z, z of the following sample, saved w and T) and, when that chunk opens a
    // ray, the ray's norm: loaded one chunk ahead, outside the bulk pipe
    float z_pre = 0.f, zn_pre = 0.f, w_pre = 0.f, t_pre = 0.f, norm_pre = 0.f;
    auto prefetch = [&]() {
        int64_t r2; int c2;
        pipe.peek(r2, c2);
        const int s2 = c2 + lane;
        if (s2 < S) {
            z_pre = __ldg(z + r2 * S + s2);
            zn_pre = (s2 + 1 < S) ? __ldg(z + r2 * S + s2 + 1) : 0.f;
            w_pre = __ldg(fwd_w + r2 * S + s2);
            t_pre = __ldg(fwd_t + r2 * S + s2);
        }
        if (c2 == last_c0) {
            const float dx = __ldg(rays_d + r2 * 3 + 0), dy = __ldg(rays_d + r2 * 3 + 1), dz = __ldg(rays_d + r2 * 3 + 2);
            norm_pre = sqrtf(dx * dx + dy * dy + dz * dz);
        }
    };
    if (!pipe.done()) prefetch();
    while (!pipe.done()) {
        int64_t ray; int c0;
        float* tile = pipe.acquire(ray, c0);
        const int cnt = min(CH, S - c0);
        const int s = c0 + lane;
        const bool valid = lane < cnt;
        const float zc = z_pre, zn_cur = zn_pre, w = valid ? w_pre : 0.f, trans = t_pre;
        if (c0 == last_c0) {                               // a new ray starts at its LAST chunk
            norm = norm_pre;
            suffix = 0.f;
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int j = lane + 32 * q;
                float a = (need_logits && j < Lout) ? __ldg(d_ins + ray * Lout + j) : 0.f;
                if (need_logits && j < Lout && !(flags & DMNERF_COMP_RAW_INS)) {
                    const float p = __ldg(fwd_ins + ray * Lout + j);
                    a *= p * (1.0f - p);
                }
                dA[q] = a;
            }
            g_r = d_rgb ? __ldg(d_rgb + ray * 3 + 0) : 0.f;
            g_g = d_rgb ? __ldg(d_rgb + ray * 3 + 1) : 0.f;
            g_b = d_rgb ? __ldg(d_rgb + ray * 3 + 2) : 0.f;
            g_d = d_depth ? __ldg(d_depth + ray) : 0.f;
        }
        if (!pipe.done()) prefetch();
        float dw = 0.f, alpha = 0.f, dist = 0.f, sigma = 0.f, c_r = 0.f, c_g = 0.f, c_b = 0.f;
        if (valid) {
            dist = ((s + 1 < S) ? (zn_cur - zc) : 1e10f) * norm;
            sigma = tile[lane * C + 3];
            alpha = 1.0f - expf(-fmaxf(sigma, 0.0f) * dist);
            c_r = sigmoidf_ref(tile[lane * C + 0]);
            c_g = sigmoidf_ref(tile[lane * C + 1]);
            c_b = sigmoidf_ref(tile[lane * C + 2]);
            dw = g_r * c_r + g_g * c_g + g_b * c_b + g_d * zc;
            if (d_weights) dw += d_weights[ray * S + s];
        }
        if (!ins_detach && need_logits) {
            for (int r = 0; r < cnt; ++r) {
                float part = 0.f;
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int j = lane + 32 * q;
                    if (j < L) part += dA[q] * tile[r * C + 4 + j];
                }
                part = warp_sum(part);
                if (lane == r) dw += part;
            }
        }
        const float g = valid ? dw * w : 0.f;
        float incl = g;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const float v = __shfl_down_sync(0xffffffffu, incl, o);
            if (lane + o < 32) incl += v;
        }
        const float after = suffix + (incl - g);
        suffix += __shfl_sync(0xffffffffu, incl, 0);
        __syncwarp();
        if (valid) {
            const float t = (1.0f - alpha) + 1e-10f;
            const float dalpha = dw * trans - after / t;
            const float dsigma = (sigma > 0.f) ? dalpha * dist * (1.0f - alpha) : 0.f;
            tile[lane * C + 0] = g_r * w * c_r * (1.0f - c_r);
            tile[lane * C + 1] = g_g * w * c_g * (1.0f - c_g);
            tile[lane * C + 2] = g_b * w * c_b * (1.0f - c_b);
            tile[lane * C + 3] = dsigma;
        }
        wbuf[lane] = w;
        __syncwarp();
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int j = lane + 32 * q;
            if (j < L) {
                float* col = tile + 4 + j;
                const float a = dA[q];
#pragma unroll 8
                for (int r = 0; r < cnt; ++r) col[r * C] = a * wbuf[r];
            }
        }
        pipe.store_tile(d_raw + (ray * S + c0) * static_cast<int64_t>(C), tile, static_cast<uint32_t>(cnt) * C * 4);
        __syncwarp();
    }
    pipe.finish();
}

}  // namespace dmnerf

using namespace dmnerf;

// weights [n,s] of render_train (render.py:9-14) from the density alone: the same dists / alpha / transmittance scan as the
// compositing kernels, chunked like them (CH samples per scan step), so the result is bit-identical to their `weights`.
namespace dmnerf {
__global__ void __launch_bounds__(128)
weights_from_sigma_kernel(const float* __restrict__ sigma, const float* __restrict__ z, const float* __restrict__ rays_d,
                          int64_t n, int S, int CH, float* __restrict__ weights) {
    const int lane = threadIdx.x & 31;
    const int64_t ray = static_cast<int64_t>(blockIdx.x) * 4 + (threadIdx.x >> 5);
    if (ray >= n) return;
    const float dx = __ldg(rays_d + ray * 3 + 0), dy = __ldg(rays_d + ray * 3 + 1), dz = __ldg(rays_d + ray * 3 + 2);
    const float norm = sqrtf(dx * dx + dy * dy + dz * dz);
    const float* zrow = z + ray * S;
    float carry = 1.0f;
    for (int c0 = 0; c0 < S; c0 += CH) {
        const int cnt = min(CH, S - c0);
        const int s = c0 + lane;
        const bool valid = lane < cnt;
        float dist = 0.f, sg = 0.f;
        if (valid) {
            const float zc = __ldg(zrow + s);
            dist = ((s + 1 < S) ? (__ldg(zrow + s + 1) - zc) : 1e10f) * norm;
            sg = __ldg(sigma + ray * S + s);
        }
        float alpha, trans, w;
        scan_chunk(sg, dist, valid, lane, carry, alpha, trans, w);
        if (valid) weights[ray * S + s] = w;
    }
}
}  // namespace dmnerf

extern "C" int dmnerf_weights_from_sigma(const float* sigma, const float* z, const float* rays_d, int64_t n, int s, int c,
                                         float* weights, void* stream) {
    if (n == 0) return DMNERF_OK;
    DM_REQUIRE(sigma && z && rays_d && weights, DMNERF_EINVAL, "weights_from_sigma: null pointer");
    DM_REQUIRE(n >= 0 && s >= 1 && s <= 1024 && c >= 6 && c <= 128, DMNERF_EUNSUPPORTED,
               "weights_from_sigma: n_samples %d / channels %d unsupported", s, c);
    // the scan step dmnerf_composite_fwd would use for raw [n,s,c] (16 samples for wide rows on its bulk path)
    const int ch = (c > 64 && (static_cast<int64_t>(s) * c) % 4 == 0) ? 16 : 32;
    weights_from_sigma_kernel<<<static_cast<unsigned>((n + 3) / 4), 128, 0, as_stream(stream)>>>(sigma, z, rays_d, n, s, ch, weights);
    DM_LAUNCH_CHECK("weights_from_sigma_kernel");
    return DMNERF_OK;
}

extern "C" int dmnerf_composite_fwd(const float* raw, const float* z, const float* rays_d, int64_t n, int s, int c,
                                    int flags, float* rgb, float* depth, float* weights, float* ins, float* trans,
                                    void* stream) {
    if (n == 0) return DMNERF_OK;
    DM_REQUIRE(raw && z && rays_d && rgb && depth && ins, DMNERF_EINVAL, "composite_fwd: null pointer");
    DM_REQUIRE(n >= 0 && s >= 1 && s <= 1024, DMNERF_EUNSUPPORTED, "composite_fwd: n_samples %d outside [1,1024]", s);
    DM_REQUIRE(c >= 6 && c <= 128, DMNERF_EUNSUPPORTED, "composite_fwd: channels %d outside [6,128]", c);
    if (n == 0) return DMNERF_OK;
    if ((reinterpret_cast<uintptr_t>(raw) & 15) == 0 && (static_cast<int64_t>(s) * c) % 4 == 0) {
        static int sm_count = 0;
        if (sm_count == 0) {
            int dev = 0;
            DM_CUDA(cudaGetDevice(&dev));
            DM_CUDA(cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev));
        }
        // debug bits 8-9 (1..3 -> 2..4 stages) and 10-12 (blocks per SM) override the launch shape for tuning sweeps
        // wide rows (K > 59): 16-sample chunks halve the ring, which is what bounds the resident warps per SM
        const int ch = c > 64 ? 16 : 32;
        auto smem_for = [&](int st) { return static_cast<size_t>(kCompBulkWarps) * ((st * ch * c + 32) * sizeof(float) + st * 8); };
        const int dbg = tc_debug_flags();
        int st = kCompStages;
        if ((dbg >> 8) & 3) st = ((dbg >> 8) & 3) + 1;
        const size_t smem = smem_for(st);
        int per_sm = static_cast<int>((220 * 1024) / (smem + 1024));
        per_sm = per_sm < 1 ? 1 : (per_sm > 6 ? 6 : per_sm);
        if ((dbg >> 10) & 7) per_sm = (dbg >> 10) & 7;
        const int64_t want = (n + kCompBulkWarps - 1) / kCompBulkWarps;
        const int64_t cap = static_cast<int64_t>(sm_count) * per_sm;
        const unsigned grid = static_cast<unsigned>(want < cap ? want : cap);
        auto launch = [&](auto kern) {
            cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            kern<<<grid, kCompBulkWarps * 32, smem, as_stream(stream)>>>(raw, z, rays_d, n, s, c, flags, rgb, depth, weights, ins, trans, ch);
        };
        if (st == 2) launch(composite_fwd_bulk_kernel<2>);
        else if (st == 3) launch(composite_fwd_bulk_kernel<3>);
        else launch(composite_fwd_bulk_kernel<4>);
        DM_LAUNCH_CHECK("composite_fwd_bulk_kernel");
        return DMNERF_OK;
    }
    const int CP = c | 1;
    const size_t smem = static_cast<size_t>(kCompWarps) * (32 * CP + 32) * sizeof(float);
    DM_CUDA(cudaFuncSetAttribute(composite_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const unsigned grid = static_cast<unsigned>((n + kCompWarps - 1) / kCompWarps);
    composite_fwd_kernel<<<grid, kCompWarps * 32, smem, as_stream(stream)>>>(raw, z, rays_d, n, s, c, flags, rgb,
                                                                            depth, weights, ins, trans);
    DM_LAUNCH_CHECK("composite_fwd_kernel");
    return DMNERF_OK;
}

extern "C" int dmnerf_composite_bwd(const float* raw, const float* z, const float* rays_d, int64_t n, int s, int c,
                                    int flags, int ins_detach, const float* d_rgb, const float* d_depth,
                                    const float* d_weights, const float* d_ins, float* d_raw, const float* fwd_weights,
                                    const float* fwd_trans, const float* fwd_ins, void* stream) {
    if (n == 0) return DMNERF_OK;
    DM_REQUIRE(raw && z && rays_d && d_raw, DMNERF_EINVAL, "composite_bwd: null pointer");
    DM_REQUIRE(n >= 0 && s >= 1 && s <= 1024, DMNERF_EUNSUPPORTED, "composite_bwd: n_samples %d outside [1,1024]", s);
    DM_REQUIRE(c >= 6 && c <= 128, DMNERF_EUNSUPPORTED, "composite_bwd: channels %d outside [6,128]", c);
    if (n == 0) return DMNERF_OK;
    if ((reinterpret_cast<uintptr_t>(raw) & 15) == 0 && (reinterpret_cast<uintptr_t>(d_raw) & 15) == 0 &&
        (static_cast<int64_t>(s) * c) % 4 == 0 && s % 4 == 0) {
        static int sm_count = 0;
        if (sm_count == 0) {
            int dev = 0;
            DM_CUDA(cudaGetDevice(&dev));
            DM_CUDA(cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev));
        }
        const bool rev = fwd_weights && fwd_trans && (fwd_ins || !d_ins || (flags & DMNERF_COMP_RAW_INS));
        const int ch = (rev && c > 64) ? 16 : 32;
        auto smem_for = [&](int st) { return static_cast<size_t>(kCompBulkWarps) * ((st * ch * c + 32 + (rev ? 0 : 5 * s)) * sizeof(float) + st * 8) + 8; };
        const int dbg = tc_debug_flags();
        int st = kCompStages;
        if ((dbg >> 8) & 3) st = ((dbg >> 8) & 3) + 1;
        const size_t smem = smem_for(st);
        if (smem <= 220 * 1024) {
            int per_sm = static_cast<int>((220 * 1024) / (smem + 1024));
            per_sm = per_sm < 1 ? 1 : (per_sm > 6 ? 6 : per_sm);
            if ((dbg >> 10) & 7) per_sm = (dbg >> 10) & 7;
            const int64_t want = (n + kCompBulkWarps - 1) / kCompBulkWarps;
            const int64_t cap = static_cast<int64_t>(sm_count) * per_sm;
            const unsigned grid = static_cast<unsigned>(want < cap ? want : cap);
            auto launch = [&](auto kern) {
                cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
                kern<<<grid, kCompBulkWarps * 32, smem, as_stream(stream)>>>(raw, z, rays_d, n, s, c, flags, ins_detach, d_rgb,
                                                                             d_depth, d_weights, d_ins, d_raw);
            };
            auto launch_rev = [&](auto kern) {
                cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
                kern<<<grid, kCompBulkWarps * 32, smem, as_stream(stream)>>>(raw, z, rays_d, n, s, c, flags, ins_detach, d_rgb,
                                                                             d_depth, d_weights, d_ins, d_raw, fwd_weights,
                                                                             fwd_trans, fwd_ins, ch);
            };
            if (rev) {
                if (st == 2) launch_rev(composite_bwd_rev_kernel<2>);
                else if (st == 3) launch_rev(composite_bwd_rev_kernel<3>);
                else launch_rev(composite_bwd_rev_kernel<4>);
            } else if (st == 2) launch(composite_bwd_bulk_kernel<2>);
            else if (st == 3) launch(composite_bwd_bulk_kernel<3>);
            else launch(composite_bwd_bulk_kernel<4>);
            DM_LAUNCH_CHECK("composite_bwd_bulk_kernel");
            return DMNERF_OK;
        }
    }
    const int CP = c | 1;
    const size_t smem = static_cast<size_t>(kCompWarps) * (32 * CP + 32 + 3 * s) * sizeof(float);
    DM_REQUIRE(smem <= 227 * 1024, DMNERF_EUNSUPPORTED, "composite_bwd: shared memory %zu too large", smem);
    DM_CUDA(cudaFuncSetAttribute(composite_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const unsigned grid = static_cast<unsigned>((n + kCompWarps - 1) / kCompWarps);
    composite_bwd_kernel<<<grid, kCompWarps * 32, smem, as_stream(stream)>>>(
        raw, z, rays_d, n, s, c, flags, ins_detach, d_rgb, d_depth, d_weights, d_ins, d_raw);
    DM_LAUNCH_CHECK("composite_bwd_kernel");
    return DMNERF_OK;
}
