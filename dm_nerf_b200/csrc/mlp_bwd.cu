// mlp_bwd.cu -- a12: parameter gradients of the DM_NeRF MLP (autograd of networks/dm_nerf.py:80-106)
// in fp32 on CUDA cores, one tile of 64 points per CTA iteration, walking the layers in reverse with
// every intermediate gradient kept in shared memory.  Inputs: d(raw) and the activation stash the
// fp32 forward wrote (mlp_layout.cuh).  Per layer:
//   wgrad  dW[n][k] += sum_p dY[n][p] * X[k][p]   (register tile 32n x 2k per thread, X streamed in
//                                                  64-row slabs through a cp.async double buffer,
//                                                  result merged with coalesced fp32 REDs)
//   bgrad  db[n]    += sum_p dY[n][p]
//   dgrad  dX[k][p]  = sum_n dY[n][p] * W[n][k]   (same dense_layer routine as the forward, fed with
//                                                  the row-major "Wg" copy), then the ReLU mask of X.
// Detach points honoured: ins_feature_linear's input gradient is dropped (dm_nerf.py:95); no
// gradient flows to the embeddings (rays / z are data, render.py:68).
#include "mlp_layout.cuh"

namespace dmnerf {

constexpr int kBwdThreads = 256;
constexpr int kChunkKb = 16;
constexpr int kXsStride = 68;                          // padded row stride of an X slab (floats)
constexpr int kSlabFloats = 64 * kXsStride;

__device__ __forceinline__ void cpa16(void* smem_dst, const void* gmem_src) {
    const unsigned s = static_cast<unsigned>(__cvta_generic_to_shared(smem_dst));
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gmem_src));
}
__device__ __forceinline__ void cpa_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N>
__device__ __forceinline__ void cpa_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N)); }

// out[n][m] = sum_k in[k][m] * w[k][n]  (n < 256, m < 64), k over `kred` rows (multiple of 16).
__device__ __forceinline__ void dgrad_layer(const float* __restrict__ in, int kred, const float* __restrict__ w,
                                            float* __restrict__ out, float* __restrict__ wst) {
    const int t = threadIdx.x, tx = t & 15, ty = t >> 4;
    float acc[4][16];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 16; ++j) acc[i][j] = 0.f;
    const int nchunks = kred / kChunkKb;
    auto load_chunk = [&](int kc, int stage) {
        const float* src = w + static_cast<int64_t>(kc) * kChunkKb * 256;
        float* dst = wst + stage * (kChunkKb * 256);
        for (int i = t; i < kChunkKb * 64; i += kBwdThreads) cpa16(dst + i * 4, src + i * 4);
        cpa_commit();
    };
    load_chunk(0, 0);
    for (int kc = 0; kc < nchunks; ++kc) {
        if (kc + 1 < nchunks) { load_chunk(kc + 1, (kc + 1) & 1); cpa_wait<1>(); }
        else                  { cpa_wait<0>(); }
        __syncthreads();
        const float* arow = in + kc * kChunkKb * kTileM;
        const float* wrow = wst + (kc & 1) * (kChunkKb * 256);
#pragma unroll
        for (int kk = 0; kk < kChunkKb; ++kk) {
            const float4 a = *reinterpret_cast<const float4*>(arow + kk * kTileM + ty * 4);
            const float av[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float4 wv4 = *reinterpret_cast<const float4*>(wrow + kk * 256 + tx * 4 + 64 * j);
                const float wv[4] = {wv4.x, wv4.y, wv4.z, wv4.w};
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int c = 0; c < 4; ++c) acc[i][j * 4 + c] = fmaf(av[i], wv[c], acc[i][j * 4 + c]);
            }
        }
        __syncthreads();
    }
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const int n = tx * 4 + 64 * j + c;
            *reinterpret_cast<float4*>(out + n * kTileM + ty * 4) =
                make_float4(acc[0][j * 4 + c], acc[1][j * 4 + c], acc[2][j * 4 + c], acc[3][j * 4 + c]);
        }
    __syncthreads();
}

// g[k][p] *= (x[k][p] > 0) for 256 rows; x is the stashed post-ReLU activation (global, [256][64]).
__device__ __forceinline__ void relu_mask_256(float* __restrict__ g, const float* __restrict__ x) {
    for (int i = threadIdx.x; i < 256 * kTileM / 4; i += kBwdThreads) {
        const float4 xv = __ldg(reinterpret_cast<const float4*>(x) + i);
        float4 gv = reinterpret_cast<float4*>(g)[i];
        gv.x = xv.x > 0.f ? gv.x : 0.f; gv.y = xv.y > 0.f ? gv.y : 0.f;
        gv.z = xv.z > 0.f ? gv.z : 0.f; gv.w = xv.w > 0.f ? gv.w : 0.f;
        reinterpret_cast<float4*>(g)[i] = gv;
    }
    __syncthreads();
}

// db[n] += sum_p dy[n][p]; rotated column order keeps the 32 rows of a warp on distinct banks.
__device__ __forceinline__ void bias_grad(const float* __restrict__ dy, int rows, float* __restrict__ db) {
    const int t = threadIdx.x;
    if (t < rows) {
        float a = 0.f;
        for (int i = 0; i < kTileM; ++i) a += dy[t * kTileM + ((i + t) & 63)];
        atomicAdd(db + t, a);
    }
}

__device__ __forceinline__ void load_slab(float* __restrict__ xs, const float* __restrict__ src, int rows) {
    for (int i = threadIdx.x; i < rows * 16; i += kBwdThreads) {
        const int r = i >> 4, piece = i & 15;
        cpa16(xs + r * kXsStride + piece * 4, src + r * kTileM + piece * 4);
    }
    cpa_commit();
}

// dW[n*ld + col0 + k] += sum_p dy[n][p] * X[k][p]  for n < 8*NJ, k < x_rows (only k < col_limit is written).
// X rows come from the stash (global, row stride 64).  Thread (warp w, lane): n = w + 8*j, k = lane + 32*i.
template <int NJ>
__device__ __forceinline__ void wgrad_block(const float* __restrict__ dy, const float* __restrict__ xsrc, int x_rows,
                                            float* __restrict__ dW, int ld, int col0, int col_limit,
                                            float* __restrict__ xs) {
    const int t = threadIdx.x, w = t >> 5, lane = t & 31;
    const int nslabs = (x_rows + 63) / 64;
    load_slab(xs, xsrc, min(64, x_rows));
    for (int s = 0; s < nslabs; ++s) {
        const int rows = min(64, x_rows - s * 64);
        if (s + 1 < nslabs) {
            load_slab(xs + ((s + 1) & 1) * kSlabFloats, xsrc + static_cast<int64_t>(s + 1) * 64 * kTileM,
                      min(64, x_rows - (s + 1) * 64));
            cpa_wait<1>();
        } else {
            cpa_wait<0>();
        }
        __syncthreads();
        const float* xb = xs + (s & 1) * kSlabFloats;
        const bool two = rows > 32;
        float acc0[NJ], acc1[NJ];
#pragma unroll
        for (int j = 0; j < NJ; ++j) { acc0[j] = 0.f; acc1[j] = 0.f; }
        for (int p = 0; p < kTileM; p += 4) {
            const float4 x0 = *reinterpret_cast<const float4*>(xb + lane * kXsStride + p);
            float4 x1 = make_float4(0.f, 0.f, 0.f, 0.f);
            if (two) x1 = *reinterpret_cast<const float4*>(xb + (lane + 32) * kXsStride + p);
#pragma unroll
            for (int j = 0; j < NJ; ++j) {
                const float4 g = *reinterpret_cast<const float4*>(dy + (w + 8 * j) * kTileM + p);
                acc0[j] = fmaf(g.x, x0.x, acc0[j]); acc0[j] = fmaf(g.y, x0.y, acc0[j]);
                acc0[j] = fmaf(g.z, x0.z, acc0[j]); acc0[j] = fmaf(g.w, x0.w, acc0[j]);
                acc1[j] = fmaf(g.x, x1.x, acc1[j]); acc1[j] = fmaf(g.y, x1.y, acc1[j]);
                acc1[j] = fmaf(g.z, x1.z, acc1[j]); acc1[j] = fmaf(g.w, x1.w, acc1[j]);
            }
        }
        const int k0 = s * 64 + lane, k1 = k0 + 32;
#pragma unroll
        for (int j = 0; j < NJ; ++j) {
            float* rowp = dW + static_cast<int64_t>(w + 8 * j) * ld + col0;
            if (lane < rows && k0 < col_limit) atomicAdd(rowp + k0, acc0[j]);
            if (two && lane + 32 < rows && k1 < col_limit) atomicAdd(rowp + k1, acc1[j]);
        }
        __syncthreads();
    }
}

struct GradPtrs { float* g[kNumParams]; };

}  // namespace dmnerf

namespace dmnerf {

// Shared memory (floats): G[256][64] | G2[256][64] | XS[2][64][68] | WST[2][16][256] | DO[C][64]
__host__ __device__ inline int bwd_smem_floats(int C) { return 2 * 256 * kTileM + 2 * kSlabFloats + 2 * kChunkKb * 256 + C * kTileM; }

__global__ void __launch_bounds__(kBwdThreads, 1)
mlp_bwd_fp32_kernel(const float* __restrict__ packed, Fp32Image im, const float* __restrict__ acts,
                    const float* __restrict__ d_raw, int64_t n_points, GradPtrs gp) {
    extern __shared__ __align__(16) float sm[];
    float* G = sm;
    float* G2 = G + 256 * kTileM;
    float* XS = G2 + 256 * kTileM;
    float* WST = XS + 2 * kSlabFloats;
    float* DO = WST + 2 * kChunkKb * 256;
    const int t = threadIdx.x;
    const int C = im.ins_num + 5, L = im.ins_num + 1;

    for (int64_t tile = blockIdx.x; tile * kTileM < n_points; tile += gridDim.x) {
        const int64_t p0 = tile * kTileM;
        const float* A = acts + tile * static_cast<int64_t>(kActRows) * kTileM;
        // ---- d(raw) tile, transposed to [C][64]; rows past the end contribute nothing --------------------
        {
            const int64_t limit = (n_points - p0) * C;
            const float* src = d_raw + p0 * C;
            for (int i = t; i < kTileM * C; i += kBwdThreads) {
                const int pt = i / C, c = i - pt * C;
                DO[c * kTileM + pt] = (i < limit) ? __ldg(src + i) : 0.f;
            }
        }
        __syncthreads();

        // ======== instance branch (its gradient stops at ins_feature_linear: h.detach()) ========
        // ins_linear: dW[c][k] += sum_p DO[4+c][p]*insH[k][p];  dY9[k][p] = sum_c DO[4+c][p]*W[c][k] * [insH>0]
        {
            const float* insH = A + kActInsH * kTileM;
            for (int i = t; i < 128 * kTileM / 4; i += kBwdThreads)      // stage insH in XS (dense [128][64])
                reinterpret_cast<float4*>(XS)[i] = __ldg(reinterpret_cast<const float4*>(insH) + i);
            __syncthreads();
            for (int o = t; o < L * 128; o += kBwdThreads) {
                const int c = o >> 7, k = o & 127;
                float a = 0.f;
                for (int i = 0; i < kTileM; ++i) { const int pp = (i + t) & 63; a = fmaf(DO[(4 + c) * kTileM + pp], XS[k * kTileM + pp], a); }
                atomicAdd(gp.g[P_INSL_W] + o, a);
            }
            if (t < L) {
                float a = 0.f;
                for (int i = 0; i < kTileM; ++i) a += DO[(4 + t) * kTileM + i];
                atomicAdd(gp.g[P_INSL_B] + t, a);
            }
            const float* w_ins = packed + im.w_ins;
            for (int o = t; o < 128 * kTileM; o += kBwdThreads) {
                const int k = o >> 6, pt = o & 63;
                float a = 0.f;
                for (int c = 0; c < L; ++c) a = fmaf(DO[(4 + c) * kTileM + pt], __ldg(w_ins + c * 128 + k), a);
                G2[o] = XS[o] > 0.f ? a : 0.f;
            }
            __syncthreads();
        }
        // layer 9 (ins_feature_linears.0): X = insF
        wgrad_block<16>(G2, A + kActInsF * kTileM, 256, gp.g[P_INS0_W], 256, 0, 256, XS);
        bias_grad(G2, 128, gp.g[P_INS0_B]);
        dgrad_layer(G2, 128, packed + im.wg[9], G, WST);                 // G = d(insF), no activation
        // layer 8 (ins_feature_linear): X = h7; no input gradient (detach)
        wgrad_block<32>(G, A + (kActH0 + 7 * 256) * kTileM, 256, gp.g[P_INSF_W], 256, 0, 256, XS);
        bias_grad(G, 256, gp.g[P_INSF_B]);
        __syncthreads();

        // ======== colour branch ========
        {
            const float* rgbH = A + kActRgbH * kTileM;
            for (int i = t; i < 128 * kTileM / 4; i += kBwdThreads)
                reinterpret_cast<float4*>(XS)[i] = __ldg(reinterpret_cast<const float4*>(rgbH) + i);
            __syncthreads();
            for (int o = t; o < 3 * 128; o += kBwdThreads) {
                const int c = o >> 7, k = o & 127;
                float a = 0.f;
                for (int i = 0; i < kTileM; ++i) { const int pp = (i + t) & 63; a = fmaf(DO[c * kTileM + pp], XS[k * kTileM + pp], a); }
                atomicAdd(gp.g[P_RGBL_W] + o, a);
            }
            if (t < 3) {
                float a = 0.f;
                for (int i = 0; i < kTileM; ++i) a += DO[t * kTileM + i];
                atomicAdd(gp.g[P_RGBL_B] + t, a);
            }
            const float* w_rgb = packed + im.w_rgb;
            for (int o = t; o < 128 * kTileM; o += kBwdThreads) {
                const int k = o >> 6, pt = o & 63;
                float a = 0.f;
#pragma unroll
                for (int c = 0; c < 3; ++c) a = fmaf(DO[c * kTileM + pt], __ldg(w_rgb + c * 128 + k), a);
                G2[o] = XS[o] > 0.f ? a : 0.f;
            }
            __syncthreads();
        }
        // layer 11 (rgb_feature_linears.0): X = [rgbF | dir], torch weight is [128][283]
        wgrad_block<16>(G2, A + kActRgbF * kTileM, 256, gp.g[P_RGB0_W], 283, 0, 256, XS);
        wgrad_block<16>(G2, A + kActDir * kTileM, 32, gp.g[P_RGB0_W], 283, 256, 27, XS);
        bias_grad(G2, 128, gp.g[P_RGB0_B]);
        dgrad_layer(G2, 128, packed + im.wg[11], G, WST);                // G = d(rgbF)
        // layer 10 (rgb_feature_linear): X = h7
        wgrad_block<32>(G, A + (kActH0 + 7 * 256) * kTileM, 256, gp.g[P_RGBF_W], 256, 0, 256, XS);
        bias_grad(G, 256, gp.g[P_RGBF_B]);
        dgrad_layer(G, 256, packed + im.wg[10], G2, WST);                // G2 = d(h7) from the colour branch
        // density head: d(h7)[k][p] += DO[3][p]*w_d[k];  dW_d[k] += sum_p DO[3][p]*h7[k][p]
        {
            const float* h7 = A + (kActH0 + 7 * 256) * kTileM;
            const float* wd = packed + im.w_dens;
            for (int o = t; o < 256 * kTileM; o += kBwdThreads) {
                const int k = o >> 6, pt = o & 63;
                G2[o] = fmaf(DO[3 * kTileM + pt], __ldg(wd + k), G2[o]);
            }
            {
                float a = 0.f;                                           // thread t <-> row k = t
                for (int i = 0; i < kTileM; ++i) a = fmaf(DO[3 * kTileM + i], __ldg(h7 + t * kTileM + i), a);
                atomicAdd(gp.g[P_DENS_W] + t, a);
            }
            if (t == 0) {
                float a = 0.f;
                for (int i = 0; i < kTileM; ++i) a += DO[3 * kTileM + i];
                atomicAdd(gp.g[P_DENS_B], a);
            }
            __syncthreads();
            relu_mask_256(G2, h7);                                       // dY7
        }

        // ======== trunk, layers 7..0 ========
        float* dy = G2;
        float* dx = G;
        for (int l = 7; l >= 0; --l) {
            float* dWl = gp.g[2 * l];
            if (l == 0) {
                wgrad_block<32>(dy, A + kActPos * kTileM, 64, dWl, 63, 0, 63, XS);
            } else if (l == 5) {
                wgrad_block<32>(dy, A + (kActH0 + 4 * 256) * kTileM, 256, dWl, 319, 0, 256, XS);
                wgrad_block<32>(dy, A + kActPos * kTileM, 64, dWl, 319, 256, 63, XS);
            } else {
                wgrad_block<32>(dy, A + (kActH0 + (l - 1) * 256) * kTileM, 256, dWl, 256, 0, 256, XS);
            }
            bias_grad(dy, 256, gp.g[2 * l + 1]);
            if (l > 0) {
                dgrad_layer(dy, 256, packed + im.wg[l], dx, WST);
                relu_mask_256(dx, A + (kActH0 + (l - 1) * 256) * kTileM);
                float* tmp = dy; dy = dx; dx = tmp;
            }
        }
        __syncthreads();
    }
}

int launch_mlp_bwd_fp32(const float* packed, const float* acts, const float* d_raw, int64_t n_points, int ins_num,
                        float* const* grads, cudaStream_t st) {
    static int sm_count = 0;
    if (sm_count == 0) {
        int dev = 0;
        DM_CUDA(cudaGetDevice(&dev));
        DM_CUDA(cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev));
    }
    GradPtrs gp;
    for (int i = 0; i < kNumParams; ++i) gp.g[i] = grads[i];
    const size_t smem = static_cast<size_t>(bwd_smem_floats(ins_num + 5)) * sizeof(float);
    DM_REQUIRE(smem <= 232448, DMNERF_EUNSUPPORTED, "mlp_bwd: %zu bytes of shared memory needed", smem);
    DM_CUDA(cudaFuncSetAttribute(mlp_bwd_fp32_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int64_t tiles = (n_points + kTileM - 1) / kTileM;
    const unsigned grid = static_cast<unsigned>(tiles < sm_count ? tiles : sm_count);
    mlp_bwd_fp32_kernel<<<grid, kBwdThreads, smem, st>>>(packed, fp32_image(ins_num), acts, d_raw, n_points, gp);
    DM_LAUNCH_CHECK("mlp_bwd_fp32_kernel");
    return DMNERF_OK;
}

}  // namespace dmnerf
