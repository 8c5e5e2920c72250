// hungarian.cu -- SURVEY 8(f) rank 2: the linear-sum assignment of the instance loss (networks/evaluator.py:41-53,
// which calls scipy.optimize.linear_sum_assignment on a [valid, ins_num] score matrix every training step) on the
// device, so that the training step has no host round trip.
//
// scipy (pinned by the reference's environment; algorithm = Crouse, "On implementing 2D rectangular assignment
// algorithms", 2016: shortest augmenting paths with dual variables, in double precision) is restated here with one
// WARP per score matrix: lane j owns columns j, j+32, j+64, j+96 (ins_num <= 123), so the "scan all unvisited columns,
// relax, take the minimum" step of the path search is a handful of per-lane operations plus one warp arg-min.
// The optimum is unique unless two path costs tie exactly; scipy's tie order depends on its scan order of the
// remaining-columns array, which is not reproduced (ties then give another assignment of the SAME total cost).
//
// Output (the reference's `col_ind` after evaluator.py:47-52): col[r] for r < n_valid = column matched to row r, then
// the unmatched columns in ascending order.
#include "common.cuh"

namespace dmnerf {

constexpr int kHungMaxN = 128;
constexpr int kHungCols = kHungMaxN / 32;       // columns per lane

// order-preserving map double -> uint64 (all finite values and +inf)
__device__ __forceinline__ unsigned long long dkey(double x) {
    const unsigned long long b = static_cast<unsigned long long>(__double_as_longlong(x));
    return (b >> 63) ? ~b : (b | 0x8000000000000000ULL);
}

// launches whose score matrix had a row without any finite entry (NaN / inf: a diverged step) since the last query
__device__ int g_hung_infeasible = 0;

__global__ void __launch_bounds__(32)
hungarian_kernel(const float* __restrict__ scores, int64_t mat_stride, int ld, const int* __restrict__ n_valid_ptr, int nc,
                 int* __restrict__ col_out) {
    extern __shared__ float s_cost[];            // [nr][nc] staged once: the path search reads one row per step
    // The path search keeps everything a lane needs about ITS columns (shortest path cost, dual v, owner row, visited
    // bit) in registers: a step is two shared-memory loads (u[i], the cost row), three dependent double adds per
    // column and the warp arg-min.  Shared memory holds what is indexed by data: u, the matching, the path, and a
    // copy of the path costs published once per augmentation for the dual update.
    __shared__ double s_u[kHungMaxN], s_spc[kHungMaxN];
    __shared__ int s_col4row[kHungMaxN], s_row4col[kHungMaxN], s_path[kHungMaxN], s_sr_list[kHungMaxN];
    const int lane = threadIdx.x;
    const float* cost = scores + static_cast<int64_t>(blockIdx.x) * mat_stride;
    int* out = col_out + static_cast<int64_t>(blockIdx.x) * nc;
    int nr = *n_valid_ptr;
    nr = nr < 0 ? 0 : (nr > nc ? nc : nr);

    for (int j = lane; j < kHungMaxN; j += 32) { s_row4col[j] = -1; s_u[j] = 0.0; s_col4row[j] = -1; }
    for (int i = 0; i < nr; ++i)
        for (int j = lane; j < nc; j += 32) s_cost[i * nc + j] = __ldg(cost + static_cast<int64_t>(i) * ld + j);
    __syncwarp();

    const double kInf = __longlong_as_double(0x7ff0000000000000LL);
    double v[kHungCols], spc[kHungCols];
    int r4c[kHungCols];
#pragma unroll
    for (int k = 0; k < kHungCols; ++k) { v[k] = 0.0; r4c[k] = -1; }
    for (int cur = 0; cur < nr; ++cur) {
#pragma unroll
        for (int k = 0; k < kHungCols; ++k) spc[k] = kInf;
        unsigned vis = 0;                                    // bit k: column lane + 32k is on the alternating tree
        int n_sr = 0, sink = -1, i = cur;
        double min_val = 0.0;
        while (sink < 0) {
            if (lane == 0) s_sr_list[n_sr] = i;              // rows on the current alternating tree
            ++n_sr;
            const double ui = s_u[i];
            const float* crow = s_cost + i * nc;
            float cv[kHungCols];
#pragma unroll
            for (int k = 0; k < kHungCols; ++k) cv[k] = lane + 32 * k < nc ? crow[lane + 32 * k] : 0.f;
            // relax every unvisited column from row i; lane-local minimum (an unassigned column wins a tie)
            double best = kInf; int best_j = -1, best_r4c = -1;
#pragma unroll
            for (int k = 0; k < kHungCols; ++k) {
                const int j = lane + 32 * k;
                if (j < nc && !((vis >> k) & 1u)) {
                    const double r = min_val + static_cast<double>(cv[k]) - ui - v[k];
                    if (r < spc[k]) { spc[k] = r; s_path[j] = i; }
                    if (spc[k] < best || (spc[k] == best && r4c[k] < 0 && best_r4c >= 0)) { best = spc[k]; best_j = j; best_r4c = r4c[k]; }
                }
            }
            // warp arg-min by three hardware reductions: high word, low word of the order-preserving key, then
            // (assigned?, column, owner row) among the ties -- every lane ends up with the column AND its owner
            const unsigned long long key = best_j >= 0 ? dkey(best) : ~0ULL;
            const unsigned hi = static_cast<unsigned>(key >> 32), lo = static_cast<unsigned>(key);
            const unsigned mhi = __reduce_min_sync(0xffffffffu, hi);
            const unsigned mlo = __reduce_min_sync(0xffffffffu, hi == mhi ? lo : 0xffffffffu);
            const bool mine = best_j >= 0 && hi == mhi && lo == mlo;
            const unsigned cand = ((best_r4c < 0 ? 0u : 1u) << 16) | (static_cast<unsigned>(best_j) << 8) | static_cast<unsigned>(best_r4c + 1);
            const unsigned sel = __reduce_min_sync(0xffffffffu, mine ? cand : 0xffffffffu);
            const unsigned long long mkey = (static_cast<unsigned long long>(mhi) << 32) | mlo;
            if (sel == 0xffffffffu || mkey == dkey(kInf)) { sink = -2; break; }     // infeasible (cannot happen for finite costs)
            const int j = static_cast<int>((sel >> 8) & 255u);
            const int owner = static_cast<int>(sel & 255u) - 1;
            min_val = __longlong_as_double(static_cast<long long>((mkey >> 63) ? (mkey & 0x7fffffffffffffffULL) : ~mkey));
            if (lane == (j & 31)) vis |= 1u << (j >> 5);
            if (owner < 0) sink = j; else i = owner;
        }
        if (sink < 0) {
            // infeasible: scipy raises ValueError here.  A kernel cannot, so the result is DEFINED -- the identity
            // permutation, every slot of `out` written, nothing out of range -- and counted for
            // dmnerf_hungarian_infeasible(), which the host checks outside captured graphs.
            for (int j = lane; j < nc; j += 32) out[j] = j;
            if (lane == 0) atomicAdd(&g_hung_infeasible, 1);
            return;
        }
        // dual update (uses the matching BEFORE augmentation); the path costs are published for the row duals
#pragma unroll
        for (int k = 0; k < kHungCols; ++k) s_spc[lane + 32 * k] = spc[k];
        __syncwarp();
        for (int t = 1 + lane; t < n_sr; t += 32) {                        // rows of the tree other than cur
            const int r = s_sr_list[t];
            s_u[r] += min_val - s_spc[s_col4row[r]];
        }
        if (lane == 0) s_u[cur] += min_val;
#pragma unroll
        for (int k = 0; k < kHungCols; ++k)
            if ((vis >> k) & 1u) v[k] -= min_val - spc[k];
        __syncwarp();
        // augment along the path sink -> cur
        if (lane == 0) {
            int j = sink;
            while (true) {
                const int pi = s_path[j];
                s_row4col[j] = pi;
                const int prev = s_col4row[pi];
                s_col4row[pi] = j;
                j = prev;
                if (pi == cur) break;
            }
        }
        __syncwarp();
#pragma unroll
        for (int k = 0; k < kHungCols; ++k) r4c[k] = s_row4col[lane + 32 * k];
    }
    // output: matched columns of rows 0..nr-1, then the unmatched columns ascending
    for (int r = lane; r < nr; r += 32) out[r] = s_col4row[r];
    int base = nr;
#pragma unroll
    for (int k = 0; k < kHungCols; ++k) {
        const int j = lane + 32 * k;
        const bool un = j < nc && s_row4col[j] < 0;
        const unsigned m = __ballot_sync(0xffffffffu, un);
        if (un) out[base + __popc(m & ((1u << lane) - 1u))] = j;
        base += __popc(m);
    }
}

}  // namespace dmnerf

using namespace dmnerf;

// ---- label-wise sums for the assignment costs --------------------------------------------------------------------
// The reference's cost matrices (evaluator.py:55-63) reduce a [K_gt, K_pred, N] broadcast over the rays; with a ONE-HOT
// ground truth that is, per predicted channel, the sum of a per-ray quantity over the rays of each ground-truth slot:
//   out[g][j] = sum over rays n with slot[n] == g of x[n][j]
// (N*K additions instead of the N*K*K multiply-adds of a dense [K,N]x[N,K] product -- this runs on the ONE SM the
// persistent MLP kernels leave free, where the dense product took 0.46 ms at K = 93).  One block per 32 channels; its
// eight warps take an eighth of the rays each into their own shared-memory accumulators, which are then added in a fixed
// order: the result does not depend on scheduling.  Column sums of x are sum_g out[g][j] (every ray has one slot).
namespace dmnerf {
constexpr int kSegWarps = 8;
__global__ void __launch_bounds__(kSegWarps * 32)
segment_sum_kernel(const float* __restrict__ x, const int64_t* __restrict__ slot, int64_t n, int k, int n_slots,
                   float* __restrict__ out) {
    extern __shared__ float s_acc[];             // [kSegWarps][n_slots][32]
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int j = blockIdx.x * 32 + lane;
    float* acc = s_acc + static_cast<size_t>(warp) * n_slots * 32;
    for (int g = 0; g < n_slots; ++g) acc[g * 32 + lane] = 0.f;
    const int64_t per = (n + kSegWarps - 1) / kSegWarps;
    const int64_t lo = warp * per, hi = lo + per < n ? lo + per : n;
    constexpr int U = 16;                                          // loads in flight per lane
    for (int64_t r = lo; r < hi; r += U) {
        float v[U];
        int g[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const bool ok = r + u < hi;
            g[u] = ok ? static_cast<int>(__ldg(slot + r + u)) : -1;
            v[u] = (ok && j < k) ? __ldg(x + (r + u) * k + j) : 0.f;
        }
#pragma unroll
        for (int u = 0; u < U; ++u)
            if (g[u] >= 0 && g[u] < n_slots) acc[g[u] * 32 + lane] += v[u];    // in ray order: a defined summation order
    }
    __syncthreads();
    if (j < k) {
        for (int g = warp; g < n_slots; g += kSegWarps) {
            float t = s_acc[g * 32 + lane];
            for (int w = 1; w < kSegWarps; ++w) t += s_acc[(static_cast<size_t>(w) * n_slots + g) * 32 + lane];
            out[static_cast<int64_t>(g) * k + j] = t;
        }
    }
}

// The three label-wise sums the assignment costs need of ONE prediction map p [n,k] in one pass:
//   out[0][g][j] = sum p,  out[1][g][j] = sum log(p + 1e-8),  out[2][g][j] = sum log(1 - p + 1e-8)   (evaluator.py:55-66)
// over the rays of slot g, which the caller lists: perm [n] = ray indices ordered by slot (stable), offs [n_slots + 1] =
// where each slot's rays start.  One warp per slot walks its rays in that order, lanes = channels: every (slot, channel)
// sum is an independent chain in a defined order, n_slots warps run in parallel, nothing is staged.  Same logf as ATen's
// log kernel.
constexpr int kSeg3Warps = 4;
__global__ void __launch_bounds__(kSeg3Warps * 32)
segment_sum3_kernel(const float* __restrict__ p, const int64_t* __restrict__ perm, const int64_t* __restrict__ offs, int k,
                    int n_slots, float* __restrict__ out) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int g = blockIdx.x * kSeg3Warps + warp;
    if (g >= n_slots) return;
    const int64_t lo = __ldg(offs + g), hi = __ldg(offs + g + 1);
    const int64_t plane = static_cast<int64_t>(n_slots) * k;
    for (int j = lane; j < k; j += 32) {
        float a0 = 0.f, a1 = 0.f, a2 = 0.f;
        constexpr int U = 8;
        for (int64_t r = lo; r < hi; r += U) {
            float v[U];
#pragma unroll
            for (int u = 0; u < U; ++u) v[u] = r + u < hi ? __ldg(p + __ldg(perm + r + u) * k + j) : -1.f;
#pragma unroll
            for (int u = 0; u < U; ++u) {
                if (v[u] >= 0.f) {                 // probabilities: the padding marker cannot occur in the data
                    a0 += v[u];
                    a1 += logf(v[u] + 1e-8f);
                    a2 += logf(1.0f - v[u] + 1e-8f);
                }
            }
        }
        out[static_cast<int64_t>(g) * k + j] = a0;
        out[plane + static_cast<int64_t>(g) * k + j] = a1;
        out[2 * plane + static_cast<int64_t>(g) * k + j] = a2;
    }
}

}  // namespace dmnerf
using namespace dmnerf;

extern "C" int dmnerf_segment_sum3(const float* p, const int64_t* perm, const int64_t* offs, int64_t n, int k, int n_slots,
                                   float* out, void* stream) {
    DM_REQUIRE(p && perm && offs && out, DMNERF_EINVAL, "segment_sum3: null pointer");
    DM_REQUIRE(n >= 0 && k >= 1 && n_slots >= 1, DMNERF_EUNSUPPORTED, "segment_sum3: k=%d n_slots=%d", k, n_slots);
    segment_sum3_kernel<<<(n_slots + kSeg3Warps - 1) / kSeg3Warps, kSeg3Warps * 32, 0, as_stream(stream)>>>(p, perm, offs, k,
                                                                                                         n_slots, out);
    DM_LAUNCH_CHECK("segment_sum3_kernel");
    return DMNERF_OK;
}

// ---- the instance loss around the assignment, as three launches instead of ~130 small ATen kernels ------------------
// (networks/evaluator.py:19-37, 55-74 with a one-hot ground truth).  s3 = [tp | sum log p | sum log(1-p)] per (slot, channel)
// from segment_sum3_kernel; cnt[g] = rays of slot g.
namespace dmnerf {

// scores[g][j] = ce + siou (evaluator.py:64-66, 68-73), also kept separately for the loss; colsum[0][j] = sum_n p[n][j],
// colsum[1][j] = sum_n log(1 - p[n][j] + 1e-8) (both as sums over the slots: every ray has exactly one).
__global__ void __launch_bounds__(128)
assign_scores_kernel(const float* __restrict__ s3, const float* __restrict__ cnt, float n, int k, float* __restrict__ scores,
                     float* __restrict__ ce, float* __restrict__ siou, float* __restrict__ colsum) {
    const int g = blockIdx.x;
    const int64_t plane = static_cast<int64_t>(k) * k;
    for (int j = threadIdx.x; j < k; j += blockDim.x) {
        float P = 0.f, Q = 0.f;
        for (int r = 0; r < k; ++r) { P += s3[r * k + j]; Q += s3[2 * plane + r * k + j]; }
        if (g == 0) { colsum[j] = P; colsum[k + j] = Q; }
        const float tp = s3[g * k + j];
        const float c = -((s3[plane + g * k + j] - s3[2 * plane + g * k + j]) + Q) / n;
        const float fp = P - tp, fn = cnt[g] - tp;
        const float si = 1.0f - tp / (tp + fp + fn + 1e-6f);
        ce[g * k + j] = c;
        siou[g * k + j] = si;
        scores[g * k + j] = c + si;
    }
}

// losses = {total, valid_ce, invalid_ce, valid_siou} (evaluator.py:27-37) for the assignment col (the first n_valid entries:
// column matched to row r; the rest: the unmatched columns); rowof[j] = the row column j is matched to, or -1.
__global__ void __launch_bounds__(128)
assign_loss_kernel(const float* __restrict__ ce, const float* __restrict__ siou, const float* __restrict__ colsum,
                   const int* __restrict__ col, const int* __restrict__ n_valid_ptr, float n, int k, float* __restrict__ losses,
                   int* __restrict__ rowof) {
    const int nv = *n_valid_ptr;
    for (int j = threadIdx.x; j < k; j += blockDim.x) rowof[j] = -1;
    __syncthreads();
    for (int r = threadIdx.x; r < nv && r < k; r += blockDim.x) rowof[col[r]] = r;
    if (threadIdx.x == 0) {
        float a = 0.f, b = 0.f, c = 0.f;
        for (int r = 0; r < nv && r < k; ++r) { a += ce[r * k + col[r]]; b += siou[r * k + col[r]]; }
        for (int r = nv; r < k; ++r) c += colsum[col[r]];
        const float vce = a / static_cast<float>(nv), vsi = b / static_cast<float>(nv);
        const float ice = k > nv ? c / (n * static_cast<float>(k - nv)) : 0.f;
        losses[0] = vce + ice + vsi; losses[1] = vce; losses[2] = ice; losses[3] = vsi;
    }
}

// d(total loss) / d p[n][j]
__global__ void __launch_bounds__(256)
assign_grad_kernel(const float* __restrict__ p, const int64_t* __restrict__ slot, const float* __restrict__ s3,
                   const float* __restrict__ colsum, const float* __restrict__ cnt, const int* __restrict__ rowof,
                   const int* __restrict__ n_valid_ptr, int64_t n, int k, float* __restrict__ grad) {
    const int64_t idx = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (idx >= n * k) return;
    const int64_t ray = idx / k;
    const int j = static_cast<int>(idx - ray * k);
    const int nv = *n_valid_ptr;
    const int r = rowof[j];
    const float fn_ = static_cast<float>(n);
    float g;
    if (r < 0) {
        g = k > nv ? 1.0f / (fn_ * static_cast<float>(k - nv)) : 0.f;
    } else {
        const float pv = p[idx];
        const bool m = static_cast<int>(slot[ray]) == r;
        const float pe = pv + 1e-8f, qe = 1.0f - pv + 1e-8f;
        const float d_ce = -((m ? 1.0f / pe + 1.0f / qe : 0.f) - 1.0f / qe) / fn_;
        const float tp = s3[r * k + j];
        const float D = tp + (colsum[j] - tp) + (cnt[r] - tp) + 1e-6f;
        const float d_si = -((m ? D : 0.f) - tp * (m ? 0.f : 1.0f)) / (D * D);
        g = (d_ce + d_si) / static_cast<float>(nv);
    }
    grad[idx] = g;
}

}  // namespace dmnerf

extern "C" int dmnerf_assign_scores(const float* s3, const float* cnt, int64_t n, int k, float* scores, float* ce, float* siou,
                                    float* colsum, void* stream) {
    DM_REQUIRE(s3 && cnt && scores && ce && siou && colsum, DMNERF_EINVAL, "assign_scores: null pointer");
    DM_REQUIRE(k >= 1 && k <= kHungMaxN && n >= 1, DMNERF_EUNSUPPORTED, "assign_scores: k=%d n=%lld", k, (long long)n);
    assign_scores_kernel<<<k, 128, 0, as_stream(stream)>>>(s3, cnt, static_cast<float>(n), k, scores, ce, siou, colsum);
    DM_LAUNCH_CHECK("assign_scores_kernel");
    return DMNERF_OK;
}

extern "C" int dmnerf_assign_loss_grad(const float* p, const int64_t* slot, const float* s3, const float* ce, const float* siou,
                                       const float* colsum, const float* cnt, const int* col, const int* n_valid, int64_t n,
                                       int k, float* losses, int* rowof, float* grad, void* stream) {
    DM_REQUIRE(p && slot && s3 && ce && siou && colsum && cnt && col && n_valid && losses && rowof, DMNERF_EINVAL,
               "assign_loss_grad: null pointer");
    DM_REQUIRE(k >= 1 && k <= kHungMaxN && n >= 1, DMNERF_EUNSUPPORTED, "assign_loss_grad: k=%d n=%lld", k, (long long)n);
    assign_loss_kernel<<<1, 128, 0, as_stream(stream)>>>(ce, siou, colsum, col, n_valid, static_cast<float>(n), k, losses, rowof);
    if (grad) {
        const int64_t total = n * k;
        assign_grad_kernel<<<static_cast<unsigned>((total + 255) / 256), 256, 0, as_stream(stream)>>>(p, slot, s3, colsum, cnt,
                                                                                                  rowof, n_valid, n, k, grad);
    }
    DM_LAUNCH_CHECK("assign_loss_grad");
    return DMNERF_OK;
}

extern "C" int dmnerf_segment_sum(const float* x, const int64_t* slot, int64_t n, int k, int n_slots, float* out, void* stream) {
    DM_REQUIRE(x && slot && out, DMNERF_EINVAL, "segment_sum: null pointer");
    DM_REQUIRE(n >= 0 && k >= 1 && n_slots >= 1 && n_slots <= 128, DMNERF_EUNSUPPORTED,
               "segment_sum: k=%d n_slots=%d (need k >= 1, 1 <= n_slots <= 128)", k, n_slots);
    const size_t smem = static_cast<size_t>(kSegWarps) * n_slots * 32 * sizeof(float);
    DM_CUDA(cudaFuncSetAttribute(segment_sum_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    segment_sum_kernel<<<(k + 31) / 32, kSegWarps * 32, smem, as_stream(stream)>>>(x, slot, n, k, n_slots, out);
    DM_LAUNCH_CHECK("segment_sum_kernel");
    return DMNERF_OK;
}

extern "C" int dmnerf_hungarian_infeasible(int* count_out, void* stream) {
    DM_REQUIRE(count_out, DMNERF_EINVAL, "hungarian_infeasible: null pointer");
    DM_CUDA(cudaStreamSynchronize(as_stream(stream)));
    DM_CUDA(cudaMemcpyFromSymbol(count_out, g_hung_infeasible, sizeof(int)));
    const int zero = 0;
    DM_CUDA(cudaMemcpyToSymbol(g_hung_infeasible, &zero, sizeof(int)));
    return DMNERF_OK;
}

extern "C" int dmnerf_hungarian(const float* scores, int n_mats, int64_t mat_stride, int ld, const int* n_valid, int n_cols,
                                int* col_out, void* stream) {
    DM_REQUIRE(scores && n_valid && col_out, DMNERF_EINVAL, "hungarian: null pointer");
    DM_REQUIRE(n_mats >= 1 && n_cols >= 1 && n_cols <= kHungMaxN && ld >= n_cols, DMNERF_EUNSUPPORTED,
               "hungarian: n_mats=%d n_cols=%d ld=%d (need 1 <= n_cols <= 128 <= ... ld >= n_cols)", n_mats, n_cols, ld);
    const size_t smem = static_cast<size_t>(n_cols) * n_cols * sizeof(float);
    DM_CUDA(cudaFuncSetAttribute(hungarian_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    hungarian_kernel<<<n_mats, 32, smem, as_stream(stream)>>>(scores, mat_stride, ld, n_valid, n_cols, col_out);
    DM_LAUNCH_CHECK("hungarian_kernel");
    return DMNERF_OK;
}
