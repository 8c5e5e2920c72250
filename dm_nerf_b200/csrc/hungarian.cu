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
