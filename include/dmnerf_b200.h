/* dmnerf_b200.h -- C ABI of libdmnerf_b200.so (hand-written sm_100a CUDA for DM-NeRF's
 * per-ray hot path).  This is the drop-in boundary: plain pointers and sizes only, no torch
 * types.  The reference (ChenLu-china/DM-NeRF) is pure Python with no FFI layer, so each entry
 * point cites the reference *function* it stands in for (paths relative to the reference root);
 * dm_nerf_b200/networks/*.py is the ctypes host side that mirrors the reference's call surface.
 *
 * Conventions
 *  - every pointer is a DEVICE pointer to fp32 row-major contiguous data unless stated;
 *  - the caller owns every buffer; nothing is allocated inside;
 *  - calls enqueue on `stream` (a cudaStream_t passed as void*) and never synchronise;
 *  - return 0 on success, a DMNERF_E* code otherwise; text via dmnerf_last_error();
 *  - supported network shape (the only one config.py:23-26,56-59,153 ever builds):
 *    D=8, W=256, skips=[4], L_pos=10 (63 ch), L_dir=4 (27 ch), 1 <= ins_num <= 123,
 *    n_samples <= 1024.  Anything else is DMNERF_EUNSUPPORTED (there is no fallback).
 */
#ifndef DMNERF_B200_H
#define DMNERF_B200_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define DMNERF_OK            0
#define DMNERF_EINVAL        1   /* null pointer / bad size / misaligned */
#define DMNERF_EUNSUPPORTED  2   /* shape outside the supported contract */
#define DMNERF_ECUDA         3   /* a CUDA runtime call failed */

#define DMNERF_MODE_FP32     0   /* CUDA-core fp32 FFMA chain: <=1e-4 parity mode            */
#define DMNERF_MODE_BF16     1   /* tcgen05 bf16 tensor-core chain (fp32 accumulate in TMEM) */
#define DMNERF_MODE_F16      2   /* same chain with fp16 operands (11-bit significands, 8x finer than bf16; activations
                                    saturate at 65504): INFERENCE ONLY -- no acts stash, no backward              */

#define DMNERF_MODE_TC32     3   /* fp32-accurate tensor-core chain: every operand an fp16 hi+lo pair, three tcgen05
                                    products per layer (hi*hi + hi*lo + lo*hi), fp32 accumulate and fp32 embeddings;
                                    meets the fp32 mode's 1e-4 bound at ~1/3 of the bf16 chain's speed.  INFERENCE ONLY */

#define DMNERF_NUM_PARAMS   30   /* tensors in DM_NeRF.state_dict(), networks/dm_nerf.py:65-78 */

/* composite flags */
#define DMNERF_COMP_DROP_EMPTY 1 /* render_train: ins_map[..., :-1] (render.py:26)            */
#define DMNERF_COMP_RAW_INS    2 /* do not sigmoid the accumulated instance logits            */

int  dmnerf_version(void);
int  dmnerf_last_error(char* buf, int cap);        /* copies the last message, returns its length */
int  dmnerf_built_arch(void);                      /* 100 => sm_100a image present               */
int  dmnerf_has_experimental(void);                /* 1 when built with -DDMNERF_EXPERIMENTAL (parked forward variants) */

/* ---- weights -------------------------------------------------------------------------------
 * params: HOST array of DMNERF_NUM_PARAMS device pointers in state_dict order:
 *   mlps.0..7 {weight,bias}, rgb_feature_linear, ins_feature_linear, rgb_feature_linears.0,
 *   ins_feature_linears.0, density_linear, ins_linear, rgb_linear  (networks/dm_nerf.py:65-78).
 * The packed image is a derived cache (K-major tiles, padded, fp32 or bf16 by mode); refresh it
 * after every optimizer step. */
int64_t dmnerf_packed_bytes(int ins_num, int mode);
int  dmnerf_weights_pack(const float* const* params, int ins_num, int mode, void* packed, void* stream);

/* ---- a1+a2+a9: embedding + DM_NeRF MLP ------------------------------------------------------
 * Replaces Embedder.embed + DM_NeRF.forward (networks/dm_nerf.py:37-38, 80-106) as called from
 * dm_nerf (networks/render.py:49-61, 71-83) and manipulator_nerf (networks/manipulator.py:108-134).
 * rays_o, rays_d [n_rays,3]; z [n_rays,n_samples]; raw_out [n_rays,n_samples,5+ins_num].
 * Points o + d*z, view dirs d/|d| and both embeddings are computed in-kernel.
 * acts (may be NULL): training-mode activation stash of dmnerf_acts_bytes(n_points, ins_num, mode) bytes for
 * the backward (fp32 rows in DMNERF_MODE_FP32; bf16 UMMA operand images per 64-point block in DMNERF_MODE_BF16,
 * 1024-byte aligned). */
int  dmnerf_mlp_fwd(const float* rays_o, const float* rays_d, const float* z, int64_t n_rays, int n_samples,
                    const void* packed, int ins_num, int mode, float* raw_out, void* acts, void* stream);
/* Same network on already-embedded rows x[m,90] (DM_NeRF.forward(x) called directly,
 * tools/mesh_generator.py:45). */
int  dmnerf_mlp_fwd_embedded(const float* x, int64_t m, const void* packed, int ins_num, int mode,
                             float* raw_out, void* acts, void* stream);
/* Density only (inference): sigma_out [n_rays, n_samples] = raw[..., 3] of dmnerf_mlp_fwd, bit for bit, from the trunk
 * mlps.0..7 + density_linear alone (networks/dm_nerf.py:82-92); the colour / instance layers are not evaluated.  What a
 * coarse RENDER pass needs: its rgb / instance outputs are never read (networks/tester.py:73-76 keeps *_fine), only its
 * weights steer the fine samples (render.py:63-70).  Tensor-core modes (BF16 / F16 / TC32) only. */
int  dmnerf_mlp_density(const float* rays_o, const float* rays_d, const float* z, int64_t n_rays, int n_samples,
                        const void* packed, int ins_num, int mode, float* sigma_out, void* stream);
/* weights [n,s] of render_train (render.py:9-14: dists, alpha, transmittance, w = alpha T) from sigma [n,s] alone -- the same
 * bits as the `weights` output of dmnerf_composite_fwd on a raw [n,s,c] with that density (c = 5 + ins_num picks its scan step). */
int  dmnerf_weights_from_sigma(const float* sigma, const float* z, const float* rays_d, int64_t n, int s, int c,
                               float* weights, void* stream);
int64_t dmnerf_acts_bytes(int64_t n_points, int ins_num, int mode);
/* a12: parameter gradients of the MLP given d(raw) [n_points, 5+ins_num] (the autograd of
 * networks/dm_nerf.py:80-106 that total_loss.backward() runs, train_dmsr.py:63).  packed: the image of
 * the SAME weights and the SAME mode the forward used; acts: the stash that forward wrote.
 * grads: HOST array of DMNERF_NUM_PARAMS device pointers (same order/shapes as params), ACCUMULATED
 * into (caller zeroes).  Honors h.detach() on the instance branch (networks/dm_nerf.py:95): the
 * instance head's gradient stops at ins_feature_linear.  workspace: dmnerf_bwd_workspace_bytes() bytes
 * (1024-byte aligned; NULL allowed when that is 0).  DMNERF_MODE_BF16 runs the dgrad chain and the
 * split-K wgrad GEMMs on tcgen05 (csrc/mlp_tc_bwd.cu); DMNERF_MODE_FP32 is the CUDA-core parity path. */
int  dmnerf_mlp_bwd(const void* packed, const void* acts, const float* d_raw, int64_t n_points,
                    int ins_num, int mode, float* const* grads, void* workspace, void* stream);
int64_t dmnerf_bwd_workspace_bytes(int64_t n_points, int ins_num, int mode);
/* NO counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight loops, harnesses that forbid it).  dmnerf_mlp_bwd in DMNERF_MODE_BF16 with its
 * workspace taken from a file-scope `static __device__` array inside the library instead of from the caller: the call
 * allocates NOTHING (no cudaMalloc / cudaFree / allocating library routine); inputs and outputs are still the caller's.
 * Limit: n_points <= dmnerf_noalloc_max_points() (2048, e.g. 32 rays x 64 samples); larger sizes return
 * DMNERF_EUNSUPPORTED with a message.  Cost: every GPU that loads the library holds the array (~13 MB) until exit.
 * Every call on a device shares the one array and finds the previous call's contents in it (what must start at zero is
 * cleared on the caller's stream): calls that OVERLAP on the GPU -- other streams, other host threads -- race; serialise
 * them or use dmnerf_mlp_bwd with a workspace per stream.  Same arguments, alignment rules and results as
 * dmnerf_mlp_bwd(..., DMNERF_MODE_BF16, ...). */
int  dmnerf_mlp_bwd_bf16_noalloc(const void* packed, const void* acts, const float* d_raw, int64_t n_points,
                                 int ins_num, float* const* grads, void* stream);
int64_t dmnerf_noalloc_max_points(void);

/* ---- a5/a8: volume compositing ---------------------------------------------------------------
 * Replaces render_train (networks/render.py:6-28) [flags=DROP_EMPTY] and manipulator_render
 * (networks/manipulator.py:86-105) [flags=0].  raw [n,s,c]; z [n,s]; rays_d [n,3];
 * rgb [n,3]; depth [n]; weights [n,s] (may be NULL); ins [n, c-4-(DROP_EMPTY?1:0)]; trans [n,s] (may be NULL):
 * the per-sample transmittances T_s, kept by training callers for the one-pass backward. */
int  dmnerf_composite_fwd(const float* raw, const float* z, const float* rays_d, int64_t n, int s, int c,
                          int flags, float* rgb, float* depth, float* weights, float* ins, float* trans, void* stream);
/* Backward of the above.  d_ins touches only raw[...,4:] when ins_detach!=0 (render.py:22-24).
 * Any of d_rgb/d_depth/d_weights/d_ins may be NULL (= zero).  d_raw [n,s,c] is overwritten.
 * fwd_weights / fwd_trans / fwd_ins: the forward's weights, trans and ins outputs for the SAME inputs, or NULL.  With
 * all three the gradient is taken in one reverse pass over raw (nothing is recomputed forwards); otherwise the kernel
 * first rebuilds them.  Same result either way. */
int  dmnerf_composite_bwd(const float* raw, const float* z, const float* rays_d, int64_t n, int s, int c,
                          int flags, int ins_detach, const float* d_rgb, const float* d_depth,
                          const float* d_weights, const float* d_ins, float* d_raw, const float* fwd_weights,
                          const float* fwd_trans, const float* fwd_ins, void* stream);

/* ---- a4: hierarchical sampler ------------------------------------------------------------------
 * Replaces sample_pdf (networks/helpers.py:123-155).  bins [n,n_bins]; weights [n,n_bins-1];
 * u: uniforms, row r at u + r*u_stride (u_stride = n_samples for a dense [n,n_samples] tensor, 0 to
 * broadcast one row, e.g. torch.linspace(0,1,n_samples) for det); u == NULL -> in-kernel linspace(0,1).
 * z_out [n,n_samples]; inds_out (int64 [n,n_samples], the searchsorted(right=True) result) may be NULL.
 * The CDF is summed in a fixed sequential fp32 order (see csrc/sampler.cu). */
int  dmnerf_sample_pdf(const float* bins, const float* weights, const float* u, int64_t u_stride, int64_t n,
                       int n_bins, int n_samples, float* z_out, int64_t* inds_out, void* stream);
/* Strict variant: the caller supplies cdf [n,n_bins] (e.g. computed by torch); searchsorted, gather
 * and lerp only -- indices and samples are bit-exact with ATen given identical cdf and u. */
int  dmnerf_sample_cdf(const float* bins, const float* cdf, const float* u, int64_t u_stride, int64_t n,
                       int n_bins, int n_samples, float* z_out, int64_t* inds_out, void* stream);
/* Fused coarse tail of dm_nerf (networks/render.py:66-70): mids of z_coarse, sample_pdf on
 * weights[:,1:-1] (weights is the full [n,sc] compositing output), sorted merge of z_coarse and the
 * new samples.  z_fine [n, sc+n_imp]; z_samples [n,n_imp] may be NULL. */
int  dmnerf_fine_z(const float* z_coarse, const float* weights, const float* u, int64_t u_stride, int64_t n,
                   int sc, int n_imp, float* z_fine, float* z_samples, void* stream);
/* The whole coarse tail of a render pass in one launch (SURVEY 8b `dmnerf_coarse_tail`; networks/render.py:9-14 + 63-70):
 * compositing weights from the coarse DENSITIES sigma [n,sc] (= raw_coarse[...,3], e.g. dmnerf_mlp_density), then
 * dmnerf_fine_z on them.  z_coarse: row r at z_coarse + r*z_stride (z_stride = sc for a dense [n,sc] tensor, 0 to share one
 * row); rays_d [n,3]; c = ins_num + 5 (selects the scan step of the compositing kernel for raw [n,sc,c], so the weights
 * carry its bits); u / u_stride as in dmnerf_sample_pdf.  z_fine [n, sc+n_imp]; weights_out [n,sc] may be NULL.
 * Same z_fine, bit for bit, as dmnerf_weights_from_sigma followed by dmnerf_fine_z. */
int  dmnerf_coarse_tail(const float* sigma, const float* z_coarse, int64_t z_stride, const float* rays_d, const float* u,
                        int64_t u_stride, int64_t n, int sc, int c, int n_imp, float* z_fine, float* weights_out,
                        void* stream);
/* Sorted merge of `parts` ascending-unsorted rows: out[n, sum(len)] = sort(cat(...)) along the
 * last dim (networks/manipulator.py:190,195).  srcs/lens are HOST arrays (<= 8 parts). */
int  dmnerf_sort_cat(const float* const* srcs, const int* lens, int parts, int64_t n, float* out, void* stream);
/* Stratified jitter of the coarse depths (networks/render.py:40-47). */
int  dmnerf_jitter_z(const float* z, const float* t_rand, int64_t n, int s, float* z_out, void* stream);

/* ---- a7: emptiness penalizer --------------------------------------------------------------------
 * Replaces ins_penalizer/emptiness_penalizer (networks/penalizer.py:5-62).  partials: 4 floats
 * (device) = {sum_before, count_before, sum_middle, count_middle}; zeroed by the call.
 * loss (device scalar) = sum_before/((K+1)*max(count_before,1e-8)) + sum_middle/max(count_middle,1e-8).
 * flags: DMNERF_PEN_ACCUMULATE (backward only) = d_raw already holds another gradient of the same raw (the compositing
 * backward's) and the penalizer's is ADDED to the logits it weighs, nothing else is touched; 0 otherwise. */
#define DMNERF_PEN_ACCUMULATE 2
int  dmnerf_penalizer_fwd(const float* raw, const float* z, const float* depth, const float* rays_d,
                          int64_t n, int s, int c, float tolerance, float deta_w, int flags, float* partials,
                          float* loss, void* stream);
/* d_raw[n,s,c] = d_loss * dLoss/draw (channels 0..3 are written as 0). d_loss: device scalar. */
int  dmnerf_penalizer_bwd(const float* raw, const float* z, const float* depth, const float* rays_d,
                          int64_t n, int s, int c, float tolerance, float deta_w, int flags, const float* partials,
                          const float* d_loss, float* d_raw, void* stream);

/* ---- a10: exchanger -------------------------------------------------------------------------------
 * Replaces exchanger (networks/manipulator.py:18-83) for ONE move label; call once per label in
 * order (the reference's in-place order dependence is preserved through ori_label).
 * ori_raw [n,s,c] is edited in place; tar_raw [n,s,c]; ori_acc/tar_acc [n,c-4] (already sigmoided
 * accumulated maps); ori_label int64 [n,s] in/out (first_label!=0: initialise it from ori_raw);
 * tar_label int64 [n,s] out.  A caller that does not keep the labels (manipulator, manipulator.py:177,201, drops them)
 * may pass tar_label = NULL, and ori_label = NULL when first_label != 0: the 16 bytes per sample are then not written. */
int  dmnerf_exchange(float* ori_raw, const float* tar_raw, const float* ori_acc, const float* tar_acc,
                     int move_label, int first_label, int64_t n, int s, int c, int64_t* ori_label,
                     int64_t* tar_label, void* stream);

/* ---- adjacent (8f-1): pinhole ray generation --------------------------------------------------------
 * Replaces get_rays_k (networks/helpers.py:50-61) for pixels [pix0, pix0+n) of an HxW frame
 * (row-major, pixel = j*W+i).  K: 9 floats, c2w: 12 floats (3x4 row-major), both HOST. */
int  dmnerf_rays_k(int H, int W, const float* K_host, const float* c2w_host, int64_t pix0,
                   const int64_t* pix_idx /* device, may be NULL: then pixels pix0..pix0+n-1 */, int64_t n,
                   float* rays_o, float* rays_d, void* stream);
/* Same as dmnerf_rays_k with K [3,3] and c2w [>=3,4] (row-major fp32) in DEVICE memory: no camera value is baked into
 * the launch, so a captured training step replays for any view (train_dmsr.py:30-36 draws a new image per step). */
int  dmnerf_rays_k_dev(int H, int W, const float* K_dev, const float* c2w_dev, int64_t pix0,
                       const int64_t* pix_idx, int64_t n, float* rays_o, float* rays_d, void* stream);
/* ---- adjacent (8f-1): whole-range deterministic render, paced from C (SURVEY 8b `dmnerf_render_full`) ---------------
 * Replaces the chunk loop of render_test (networks/tester.py:62-76; perturb = 0, det sampling) for pixels
 * [pix0, pix0+count) of an HxW frame: per pass of <= chunk rays (count is cut into equal passes of whole 128-point tiles)
 * dmnerf_rays_k -> coarse pass -> fine depths -> fine query -> compositing -> label argmax, every launch enqueued by this
 * one call and the compositing kernel writing straight into rgb [count,3], depth [count], ins [count,ins_num];
 * labels uint8 [count] may be NULL.  Tensor-core modes evaluate the density-only coarse trunk (dmnerf_mlp_density +
 * dmnerf_coarse_tail); DMNERF_MODE_FP32 the full coarse query + compositing + dmnerf_fine_z.  Results equal the
 * step-by-step entry points bit for bit.  K_host: 9 floats, c2w_host: 12 floats (3x4), both HOST; z_row: DEVICE, the
 * n_samples coarse depths shared by every ray (z_val_sample, networks/helpers.py:114-119); u_row: DEVICE, the n_importance
 * uniforms shared by every ray (helpers.py:137), NULL = linspace(0,1) evaluated in the kernel; packed_*: dmnerf_weights_pack
 * images of the two models for `mode`; workspace: DEVICE scratch, 256-byte aligned, at least
 * dmnerf_render_workspace_bytes(count, chunk, ...) bytes (-1: unsupported arguments).  Nothing is allocated. */
int64_t dmnerf_render_workspace_bytes(int64_t count, int64_t chunk, int n_samples, int n_importance, int ins_num, int mode);
int  dmnerf_render_full(int H, int W, const float* K_host, const float* c2w_host, int64_t pix0, int64_t count,
                        const float* z_row, const float* u_row, int n_samples, int n_importance,
                        const void* packed_coarse, const void* packed_fine, int ins_num, int mode, int64_t chunk, void* workspace,
                        int64_t workspace_bytes, float* rgb, float* depth, float* ins, uint8_t* labels, void* stream);
/* Standalone Embedder.embed (networks/dm_nerf.py:37-38): x [n,3] -> out [n, 3+6*n_freqs]. */
int  dmnerf_embed(const float* x, int64_t n, int n_freqs, float* out, void* stream);
/* Per-ray label image: plain argmax of a composited instance map whose entries are probabilities already (first
 * maximum wins) = torch.argmax(ins, -1) at networks/evaluator.py:132, networks/manipulator.py:321,478; uint8 out. */
int  dmnerf_label_argmax(const float* ins, int64_t n, int k, uint8_t* labels, void* stream);

/* Linear-sum assignment on the device (replaces scipy.optimize.linear_sum_assignment at networks/evaluator.py:45-46
 * plus the unmatched-column bookkeeping of :47-52).  scores: n_mats matrices, matrix m at scores + m*mat_stride,
 * row-major with leading dimension ld; rows 0 .. *n_valid-1 (n_valid: DEVICE int, <= n_cols) are assigned to distinct
 * columns of 0 .. n_cols-1 (n_cols <= 128) minimising the summed score.  col_out [n_mats, n_cols] int32: the column
 * of each assigned row, then the unmatched columns in ascending order.  No host synchronisation. */
int  dmnerf_hungarian(const float* scores, int n_mats, int64_t mat_stride, int ld, const int* n_valid, int n_cols,
                      int* col_out, void* stream);
/* A score matrix with a row that has no finite entry (NaN / inf, e.g. a diverged step) has no assignment; scipy raises
 * ValueError (evaluator.py:45).  The kernel then writes the identity permutation (every slot defined and in range) and
 * counts the launch.  This call SYNCHRONISES the stream, returns the count since the last query and resets it. */
int  dmnerf_hungarian_infeasible(int* count_out, void* stream);
/* Label-wise sums for the assignment's cost matrices (networks/evaluator.py:55-63: the [K_gt, K_pred, N] broadcast over
 * a one-hot ground truth): out[g][j] = sum over rays n with slot[n] == g of x[n][j].  x [n,k]; slot int64 [n] (values
 * outside [0,n_slots) are skipped); out [n_slots,k] is overwritten.  Fixed summation order (reproducible). */
int  dmnerf_segment_sum(const float* x, const int64_t* slot, int64_t n, int k, int n_slots, float* out, void* stream);
/* The three sums the cost matrices need of one prediction map p [n,k] (probabilities) in one pass: out [3,n_slots,k] =
 * label-wise sums of p, log(p + 1e-8), log(1 - p + 1e-8) (evaluator.py:55-66).  perm int64 [n]: ray indices ordered by
 * slot; offs int64 [n_slots+1]: first position of each slot in perm (offs[n_slots] = n).  One warp per slot, fixed order. */
int  dmnerf_segment_sum3(const float* p, const int64_t* perm, const int64_t* offs, int64_t n, int k, int n_slots,
                         float* out, void* stream);
/* The instance loss around the assignment (networks/evaluator.py:19-37, 55-74) without its ~130 small tensor ops:
 * dmnerf_assign_scores: s3 [3,k,k] (dmnerf_segment_sum3), cnt [k] rays per slot -> scores = ce + siou [k,k] (the
 * Hungarian's input), ce, siou [k,k], colsum [2,k] (column sums of p and of log(1-p+1e-8)).
 * dmnerf_assign_loss_grad: given the assignment col int32 [k] (dmnerf_hungarian's output) and n_valid (device int32):
 * losses[4] = {total, valid_ce, invalid_ce, valid_siou}; rowof int32 [k] scratch; grad [n,k] = d total / d p (NULL: loss only). */
int  dmnerf_assign_scores(const float* s3, const float* cnt, int64_t n, int k, float* scores, float* ce, float* siou,
                          float* colsum, void* stream);
int  dmnerf_assign_loss_grad(const float* p, const int64_t* slot, const float* s3, const float* ce, const float* siou,
                             const float* colsum, const float* cnt, const int* col, const int* n_valid, int64_t n, int k,
                             float* losses, int* rowof, float* grad, void* stream);

/* ---- self test of the tcgen05 building blocks (used by tests/) ---------------------------------------
 * D[128,n] = A[128,k] * B[n,k]^T in bf16 with fp32 accumulate; a,b fp32 in, d fp32 out. */
int  dmnerf_selftest_umma(const float* a, const float* b, int n, int k, float* d, void* stream);
/* D[128,n] = A[64,128]^T * B[64,n]: both operands MN-major, the 64 rows being the reduction dim -- the operand
 * form of the wgrad GEMMs.  a,b fp32 in (row-major [64][128], [64][n]), d fp32 out [128][n]. */
int  dmnerf_selftest_umma_mn(const float* a, const float* b, int n, float* d, void* stream);
/* CTA-pair form (tcgen05 cta_group::2, cluster of two CTAs): D[256,n] = A[256,k] * B[n,k]^T, each CTA holding its 128
 * rows of A and its half of B's rows. */
int  dmnerf_selftest_umma2(const float* a, const float* b, int n, int k, float* d, void* stream);
/* Micro-benchmark: `warps` (4..16, multiple of 4) warps of one CTA read all 512 TMEM columns `iters` times with
 * tcgen05.ld 32x32b.x32 (batched != 0: two loads per wait).  out (device, 3 x int64): cycles, bytes read. */
int  dmnerf_selftest_tmem_read_bw(int warps, int iters, int batched, long long* out, void* stream);
/* Micro-benchmark: cycles for `count` back-to-back tcgen05.mma (M=128, N=n, K=16, bf16) issued by one thread. */
int  dmnerf_selftest_umma_rate(int n, int count, long long* out, void* stream);
/* Bring-up aid: cycle counters block 0 of the last pipelined-forward launch wrote (8 x int64, HOST pointer;
 * synchronises the device): MMA issuer total / waiting for weights / waiting for activations, epilogue thread total /
 * waiting for accumulators, producer waiting for a free ring slot. */
int  dmnerf_debug_counters(long long* host_out8);
int  dmnerf_debug_counters_tc32(long long* host_out12);   /* same for the DMNERF_MODE_TC32 forward (12 counters) */
int  dmnerf_selftest_issue_cost(int mode, int count, long long* out2_dev, void* stream);
/* Bring-up switches for the tensor path (bit0: swap LBO/SBO in the forward's descriptors; bit1: in wgrad's).
 * flags < 0 only queries.  Returns the previous value. */
int  dmnerf_debug_flags(int flags);
/* The optimizer step of the training loop (train_dmsr.py:62-72: torch.optim.Adam(betas=(0.9,0.999)).step() + lr decay)
 * as one kernel over a flat fp32 parameter buffer: p, g, m, v are DEVICE buffers of n floats (n % 4 == 0, 16-byte
 * aligned; g is read-only).  step_dev: device int32 holding the 1-based step count t; lr_dev: device float with the
 * learning rate, or NULL to use `lr`.  g is scaled by grad_scale first (1/world_size after a summing all-reduce).
 * Same arithmetic as torch.optim.Adam with weight_decay = 0, amsgrad = False. */
int  dmnerf_adam_step(float* p, const float* g, float* m, float* v, int64_t n, const float* lr_dev, float lr,
                      const int* step_dev, float beta1, float beta2, float eps, float grad_scale, void* stream);
/* The pixel draw of the training loops (networks/helpers.py:99-111: np.random.choice(H*W, N, replace=False)): n DISTINCT
 * indices in [0,total) = the first n values of a keyed pseudo-random permutation (4-round Feistel, cycle-walked).  The
 * key mixes `seed` with the device int32 at step_dev, so a captured CUDA graph draws new pixels on every replay.
 * idx_out: device int64 [n].  Not the reference's numpy stream (helpers.get_select_full(numpy_rng=True) reproduces that). */
int  dmnerf_draw_pixels(int64_t* idx_out, int64_t n, int64_t total, const int* step_dev, uint64_t seed, void* stream);
/* SM partitioning for overlapped training (train_dmsr.py:24-72 as one step; networks/evaluator.py:45-46 is the
 * assignment that runs beside the ray kernels): the persistent one-CTA-per-SM MLP kernels (forward, dgrad, wgrad) take a
 * whole SM's shared memory, so a small kernel on another stream (dmnerf_hungarian) can only run concurrently if some SMs
 * are left free.  n >= 0 sets how many SMs those kernels leave unused from now on (process-wide; grids are fixed at launch,
 * so a captured CUDA graph keeps the value it was captured with); n < 0 only queries.  Returns the previous value. */
int  dmnerf_reserve_sms(int n);

#ifdef __cplusplus
}
#endif
#endif
