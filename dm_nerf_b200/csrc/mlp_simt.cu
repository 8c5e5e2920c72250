// mlp_simt.cu -- a1+a2+a9 in DMNERF_MODE_FP32: positional encoding + the whole DM_NeRF MLP
// (networks/dm_nerf.py:8-55, 80-106) for a tile of 64 points per CTA, fp32 FFMA on CUDA cores.
// This is the <=1e-4 parity mode; activations never leave shared memory between layers, weights
// stream L2 -> smem through a cp.async double buffer.  Roofline: fp32 FFMA (not tensor); the bf16
// tcgen05 path lives in mlp_tc.cu.
//
// Shared memory map (floats): bufA[256][64] bufB[256][64] ext[64][64] dext[32][64] bufC[128][64]
// wst[2][16][256]  = 221,184 B.  Activations are stored feature-major ([k][point]) so the inner
// loop reads one float4 of 4 points (broadcast) and four float4 of weights per k.
#include "mlp_layout.cuh"

namespace dmnerf {

constexpr int kSimtThreads = 256;
constexpr int kChunkK = 16;
constexpr int kSmemFloats = 256 * 64 * 2 + 64 * 64 + 32 * 64 + 128 * 64 + 2 * kChunkK * 256;

// ---- packing ----------------------------------------------------------------------------------------
// dst[k*n + j] = (k < k_src) ? src[j*k_src + k] : 0         (transpose + zero pad)
__global__ void pack_transpose_kernel(const float* __restrict__ src, int n, int k_src, int k_pad,
                                      float* __restrict__ dst) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n * k_pad) return;
    const int k = idx / n, j = idx - k * n;
    dst[idx] = (k < k_src) ? src[static_cast<int64_t>(j) * k_src + k] : 0.0f;
}
// dst[j*k_keep + k] = src[j*k_src + k]  for k < k_keep           (row-major trim)
__global__ void pack_trim_kernel(const float* __restrict__ src, int n, int k_src, int k_keep,
                                 float* __restrict__ dst) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n * k_keep) return;
    const int j = idx / k_keep, k = idx - j * k_keep;
    dst[idx] = src[static_cast<int64_t>(j) * k_src + k];
}
__global__ void pack_copy_kernel(const float* __restrict__ src, int count, int padded, float* __restrict__ dst) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= padded) return;
    dst[idx] = idx < count ? src[idx] : 0.0f;
}

int pack_fp32(const float* const* params, int ins_num, float* packed, cudaStream_t st) {
    const Fp32Image im = fp32_image(ins_num);
    auto blocks = [](int64_t nelem) { return static_cast<unsigned>((nelem + 255) / 256); };
    for (int l = 0; l < kBigLayers; ++l) {
        const BigLayer b = big_layer(l);
        pack_transpose_kernel<<<blocks(b.n * b.k_pad), 256, 0, st>>>(params[b.w_idx], b.n, b.k_src, b.k_pad,
                                                                    packed + im.wt[l]);
        pack_copy_kernel<<<blocks(b.n), 256, 0, st>>>(params[b.w_idx + 1], b.n, b.n, packed + im.bias[l]);
        if (b.k_grad > 0)
            pack_trim_kernel<<<blocks(b.n * b.k_grad), 256, 0, st>>>(params[b.w_idx], b.n, b.k_src, b.k_grad,
                                                                    packed + im.wg[l]);
    }
    const int L = ins_num + 1;
    pack_copy_kernel<<<1, 256, 0, st>>>(params[P_DENS_W], 256, 256, packed + im.w_dens);
    pack_copy_kernel<<<1, 256, 0, st>>>(params[P_DENS_B], 1, 4, packed + im.b_dens);
    pack_copy_kernel<<<blocks(384), 256, 0, st>>>(params[P_RGBL_W], 384, 384, packed + im.w_rgb);
    pack_copy_kernel<<<1, 256, 0, st>>>(params[P_RGBL_B], 3, 4, packed + im.b_rgb);
    pack_copy_kernel<<<blocks(L * 128), 256, 0, st>>>(params[P_INSL_W], L * 128, L * 128, packed + im.w_ins);
    pack_copy_kernel<<<1, 256, 0, st>>>(params[P_INSL_B], L, static_cast<int>(round4(L)), packed + im.b_ins);
    DM_LAUNCH_CHECK("pack_fp32");
    return DMNERF_OK;
}

// ---- device building blocks -------------------------------------------------------------------------
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gmem_src) {
    const unsigned s = static_cast<unsigned>(__cvta_generic_to_shared(smem_dst));
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gmem_src));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N)); }

// One dense layer on the tile:  out[n][m] = act( sum_k in[k][m] * Wt[k][n] + bias[n] ),  m in [0,64).
// The reduction runs over in0 (k0 rows) then in1 (k1 rows); k0, k1 multiples of 16.  NJ = n/64.
// Thread (tx = t&15, ty = t>>4) owns points 4*ty..4*ty+3 and outputs 4*tx + 64*j + {0..3}.
template <int NJ>
__device__ __forceinline__ void dense_layer(const float* __restrict__ in0, int k0, const float* __restrict__ in1,
                                            int k1, const float* __restrict__ wt, const float* __restrict__ bias,
                                            float* __restrict__ out, bool relu, float* __restrict__ wst) {
    constexpr int N = NJ * 64;
    const int t = threadIdx.x, tx = t & 15, ty = t >> 4;
    float acc[4][NJ * 4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < NJ * 4; ++j) acc[i][j] = 0.f;

    const int nchunks = (k0 + k1) / kChunkK;
    constexpr int kPieces = kChunkK * N / 4;  // 16-byte pieces per chunk
    auto load_chunk = [&](int kc, int stage) {
        const float* src = wt + static_cast<int64_t>(kc) * kChunkK * N;
        float* dst = wst + stage * (kChunkK * 256);
        for (int i = t; i < kPieces; i += kSimtThreads) cp_async16(dst + i * 4, src + i * 4);
        cp_async_commit();
    };
    load_chunk(0, 0);
    for (int kc = 0; kc < nchunks; ++kc) {
        if (kc + 1 < nchunks) { load_chunk(kc + 1, (kc + 1) & 1); cp_async_wait<1>(); }
        else                  { cp_async_wait<0>(); }
        __syncthreads();
        const int kbase = kc * kChunkK;
        const float* arow = (kbase < k0) ? (in0 + kbase * kTileM) : (in1 + (kbase - k0) * kTileM);
        const float* wrow = wst + (kc & 1) * (kChunkK * 256);
#pragma unroll
        for (int kk = 0; kk < kChunkK; ++kk) {
            const float4 a = *reinterpret_cast<const float4*>(arow + kk * kTileM + ty * 4);
            const float av[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
            for (int j = 0; j < NJ; ++j) {
                const float4 w = *reinterpret_cast<const float4*>(wrow + kk * N + tx * 4 + 64 * j);
                const float wv[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int c = 0; c < 4; ++c) acc[i][j * 4 + c] = fmaf(av[i], wv[c], acc[i][j * 4 + c]);
            }
        }
        __syncthreads();
    }
#pragma unroll
    for (int j = 0; j < NJ; ++j)
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const int n = tx * 4 + 64 * j + c;
            const float b = __ldg(bias + n);
            float4 v;
            v.x = acc[0][j * 4 + c] + b; v.y = acc[1][j * 4 + c] + b;
            v.z = acc[2][j * 4 + c] + b; v.w = acc[3][j * 4 + c] + b;
            if (relu) { v.x = fmaxf(v.x, 0.f); v.y = fmaxf(v.y, 0.f); v.z = fmaxf(v.z, 0.f); v.w = fmaxf(v.w, 0.f); }
            *reinterpret_cast<float4*>(out + n * kTileM + ty * 4) = v;
        }
    __syncthreads();
}

// copy rows x 64 floats from shared memory to the activation stash (coalesced float4)
__device__ __forceinline__ void stash_rows(float* __restrict__ dst, const float* __restrict__ src, int rows) {
    const int total4 = rows * kTileM / 4;
    for (int i = threadIdx.x; i < total4; i += kSimtThreads)
        reinterpret_cast<float4*>(dst)[i] = reinterpret_cast<const float4*>(src)[i];
}

}  // namespace dmnerf

namespace dmnerf {

// Embedding of one scalar coordinate group: channel layout [x(3) | f0: sin(3) cos(3) | f1: ... ]
// (networks/dm_nerf.py:17-32).  x*2^f is exact in fp32; sincosf is the full-range libm path.
__device__ __forceinline__ void embed_freq(float* __restrict__ dst, int pt, const float v[3], int f) {
    const float scale = static_cast<float>(1 << f);
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        float s, co;
        sincosf(v[c] * scale, &s, &co);
        dst[(3 + 6 * f + c) * kTileM + pt] = s;
        dst[(3 + 6 * f + 3 + c) * kTileM + pt] = co;
    }
}


__global__ void __launch_bounds__(kSimtThreads, 1)
mlp_fwd_fp32_kernel(MlpInput in, const float* __restrict__ packed, Fp32Image im, float* __restrict__ raw_out,
                    float* __restrict__ acts) {
    extern __shared__ __align__(16) float sm[];
    float* bufA = sm;
    float* bufB = bufA + 256 * kTileM;
    float* ext = bufB + 256 * kTileM;
    float* dext = ext + 64 * kTileM;
    float* bufC = dext + 32 * kTileM;
    float* wst = bufC + 128 * kTileM;
    const int t = threadIdx.x;
    const int C = im.ins_num + 5;
    const int L = im.ins_num + 1;

    for (int64_t tile = blockIdx.x; tile * kTileM < in.n_points; tile += gridDim.x) {
        const int64_t p0 = tile * kTileM;
        float* stash = acts ? acts + tile * static_cast<int64_t>(kActRows) * kTileM : nullptr;
        // ---- input stage -------------------------------------------------------------------------
        if (in.x == nullptr) {
            const int pt = t & 63, part = t >> 6;
            const int64_t p = p0 + pt;
            float pos[3] = {0.f, 0.f, 0.f}, dir[3] = {0.f, 0.f, 0.f};
            if (p < in.n_points) {
                const int64_t ray = p / in.S;
                const float zz = in.z[p];
                float d[3];
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    d[c] = in.rays_d[ray * 3 + c];
                    pos[c] = __fadd_rn(in.rays_o[ray * 3 + c], __fmul_rn(d[c], zz));   // o + d*z, un-fused
                }
                const float nrm = sqrtf(__fadd_rn(__fadd_rn(__fmul_rn(d[0], d[0]), __fmul_rn(d[1], d[1])),
                                                  __fmul_rn(d[2], d[2])));
#pragma unroll
                for (int c = 0; c < 3; ++c) dir[c] = __fdiv_rn(d[c], nrm);
            }
            for (int f = part; f < kPosFreqs; f += 4) embed_freq(ext, pt, pos, f);
            embed_freq(dext, pt, dir, part);           // 4 direction bands, one per part
            if (part == 0) {
#pragma unroll
                for (int c = 0; c < 3; ++c) { ext[c * kTileM + pt] = pos[c]; dext[c * kTileM + pt] = dir[c]; }
                ext[63 * kTileM + pt] = 0.f;
            } else if (part == 1) {
#pragma unroll
                for (int c = 27; c < 32; ++c) dext[c * kTileM + pt] = 0.f;
            }
        } else {
            const float* src = in.x + p0 * 90;
            const int64_t limit = (in.n_points - p0) * 90;
            for (int i = t; i < kTileM * 90; i += kSimtThreads) {
                const int pt = i / 90, ch = i - pt * 90;
                const float v = (i < limit) ? __ldg(src + i) : 0.f;
                if (ch < 63) ext[ch * kTileM + pt] = v; else dext[(ch - 63) * kTileM + pt] = v;
            }
            if (t < 64) ext[63 * kTileM + t] = 0.f;
            for (int i = t; i < 5 * kTileM; i += kSimtThreads) dext[27 * kTileM + i] = 0.f;
        }
        __syncthreads();
        if (stash) { stash_rows(stash + kActPos * kTileM, ext, 64); stash_rows(stash + kActDir * kTileM, dext, 32); }

        // ---- trunk ----------------------------------------------------------------------------------
        float* cur = bufA;
        float* nxt = bufB;
        for (int l = 0; l < 8; ++l) {
            const float* wt = packed + im.wt[l];
            const float* bs = packed + im.bias[l];
            if (l == 0)      dense_layer<4>(ext, 64, nullptr, 0, wt, bs, nxt, true, wst);
            else if (l == 5) dense_layer<4>(cur, 256, ext, 64, wt, bs, nxt, true, wst);
            else             dense_layer<4>(cur, 256, nullptr, 0, wt, bs, nxt, true, wst);
            float* tmp = cur; cur = nxt; nxt = tmp;
            if (stash) stash_rows(stash + (kActH0 + 256 * l) * kTileM, cur, 256);
        }
        // cur = h7, nxt = scratch
        // ---- instance branch (reads h7; the detach only matters for the backward) -------------------
        dense_layer<4>(cur, 256, nullptr, 0, packed + im.wt[8], packed + im.bias[8], nxt, false, wst);
        if (stash) stash_rows(stash + kActInsF * kTileM, nxt, 256);
        dense_layer<2>(nxt, 256, nullptr, 0, packed + im.wt[9], packed + im.bias[9], bufC, true, wst);
        if (stash) stash_rows(stash + kActInsH * kTileM, bufC, 128);
        // ---- density head straight off h7 -> out tile (staged in wst: [64][C]) ------------------------
        float* otile = wst;
        if (t < kTileM) {
            const float* wd = packed + im.w_dens;
            float a = 0.f;
            for (int k = 0; k < 256; ++k) a = fmaf(cur[k * kTileM + t], __ldg(wd + k), a);
            otile[t * C + 3] = a + __ldg(packed + im.b_dens);
        }
        // instance logits from bufC
        for (int o = t; o < kTileM * L; o += kSimtThreads) {
            const int c = o >> 6, pt = o & 63;
            const float* w = packed + im.w_ins + c * 128;
            float a = 0.f;
            for (int k = 0; k < 128; ++k) a = fmaf(bufC[k * kTileM + pt], __ldg(w + k), a);
            otile[pt * C + 4 + c] = a + __ldg(packed + im.b_ins + c);
        }
        __syncthreads();
        // NOTE: otile aliases the weight staging buffers; the colour branch below needs them, so the
        // partial tile is parked in bufC's tail?  bufC is still live (ins_hid) only until here -> copy.
        for (int i = t; i < kTileM * C; i += kSimtThreads) bufC[i] = otile[i];
        __syncthreads();
        // ---- colour branch -------------------------------------------------------------------------------
        dense_layer<4>(cur, 256, nullptr, 0, packed + im.wt[10], packed + im.bias[10], nxt, false, wst);
        if (stash) stash_rows(stash + kActRgbF * kTileM, nxt, 256);
        dense_layer<2>(nxt, 256, dext, 32, packed + im.wt[11], packed + im.bias[11], cur, true, wst);
        if (stash) stash_rows(stash + kActRgbH * kTileM, cur, 128);
        for (int o = t; o < kTileM * 3; o += kSimtThreads) {
            const int c = o >> 6, pt = o & 63;
            const float* w = packed + im.w_rgb + c * 128;
            float a = 0.f;
            for (int k = 0; k < 128; ++k) a = fmaf(cur[k * kTileM + pt], __ldg(w + k), a);
            bufC[pt * C + c] = a + __ldg(packed + im.b_rgb + c);
        }
        __syncthreads();
        // ---- coalesced store of the [64][C] tile -----------------------------------------------------------
        {
            const int64_t limit = (in.n_points - p0) * C;
            float* dst = raw_out + p0 * C;
            for (int i = t; i < kTileM * C; i += kSimtThreads)
                if (i < limit) dst[i] = bufC[i];
        }
        __syncthreads();
    }
}

int launch_mlp_fp32(const MlpInput& in, const float* packed, int ins_num, float* raw_out, float* acts,
                    cudaStream_t st) {
    static int sm_count = 0;
    if (sm_count == 0) {
        int dev = 0;
        DM_CUDA(cudaGetDevice(&dev));
        DM_CUDA(cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev));
    }
    const size_t smem = kSmemFloats * sizeof(float);
    DM_CUDA(cudaFuncSetAttribute(mlp_fwd_fp32_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int64_t tiles = (in.n_points + kTileM - 1) / kTileM;
    const unsigned grid = static_cast<unsigned>(tiles < sm_count ? tiles : sm_count);
    mlp_fwd_fp32_kernel<<<grid, kSimtThreads, smem, st>>>(in, packed, fp32_image(ins_num), raw_out, acts);
    DM_LAUNCH_CHECK("mlp_fwd_fp32_kernel");
    return DMNERF_OK;
}

}  // namespace dmnerf
