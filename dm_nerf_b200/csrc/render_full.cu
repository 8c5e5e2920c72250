// render_full.cu -- the chunk loop of render_test (networks/tester.py:62-76) paced from C: rays of a pixel range are
// generated, queried, sampled, queried again, composited and labelled pass by pass with every launch enqueued from this
// one call (no interpreter between launches, no staging copies: the compositing kernel writes straight into the
// caller's frame buffers).  Deterministic render (perturb = 0, det sampling), the configuration test_dmsr.py uses.
//
// Per pass of <= chunk rays (each step is the library's own entry point, so the results are those of the step-by-step
// path bit for bit):
//   rays_k -> [tensor modes] density-only coarse trunk -> coarse_tail (weights from sigma + sample_pdf + merge, one launch)
//          -> [fp32 mode]    full coarse query -> compositing (weights) -> fine_z
//          -> fine query -> compositing (DROP_EMPTY) into rgb / depth / ins -> label argmax
// What a render caller keeps is the fine pass (tester.py:73-76); the coarse pass only places the fine samples
// (render.py:63-70).
#include "common.cuh"

namespace dmnerf {

__global__ void expand_row_kernel(const float* __restrict__ row, int s, int64_t total, float* __restrict__ out) {
    const int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (i < total) out[i] = __ldg(row + static_cast<int>(i % s));
}

struct RenderWs {
    int64_t ro, rd, zc, sigma, zf, raw, total;   // byte offsets, 256-byte aligned
};

static inline int64_t up256(int64_t v) { return (v + 255) / 256 * 256; }

static RenderWs render_ws(int64_t chunk, int sc, int n_imp, int ins_num, int mode) {
    const int64_t c = ins_num + 5, s = sc + n_imp;
    RenderWs w{};
    int64_t o = 0;
    w.ro = o;    o += up256(chunk * 3 * 4);
    w.rd = o;    o += up256(chunk * 3 * 4);
    w.zc = o;    o += up256(chunk * sc * 4);
    w.sigma = o; o += up256(chunk * sc * 4);           // densities (tensor modes) or compositing weights (fp32 mode)
    w.zf = o;    o += up256(chunk * s * 4);
    w.raw = o;   o += up256(chunk * s * c * 4);         // raw_fine; the fp32 mode parks raw_coarse here first
    (void)mode;
    w.total = o;
    return w;
}

// passes of equal size (whole 128-point tiles) instead of full chunks plus a short tail: the persistent MLP kernels deal
// 256-point tile pairs round-robin to the SMs, so a short tail costs whole rounds for a fraction of the work
static int64_t balanced_chunk(int64_t count, int64_t chunk) {
    const int64_t passes = (count + chunk - 1) / chunk;
    int64_t per = (count + passes - 1) / passes;
    per = (per + 127) / 128 * 128;
    return per < count ? per : count;
}

}  // namespace dmnerf

using namespace dmnerf;

extern "C" int64_t dmnerf_render_workspace_bytes(int64_t count, int64_t chunk, int n_samples, int n_importance,
                                                 int ins_num, int mode) {
    if (count < 0 || chunk < 1 || n_samples < 4 || n_importance < 2 || ins_num < 1 || ins_num > kMaxIns) return -1;
    if (count == 0) return 0;
    return render_ws(balanced_chunk(count, chunk), n_samples, n_importance, ins_num, mode).total;
}

extern "C" int dmnerf_render_full(int H, int W, const float* K_host, const float* c2w_host, int64_t pix0, int64_t count,
                                  const float* z_row, const float* u_row, int n_samples, int n_importance,
                                  const void* packed_coarse,
                                  const void* packed_fine, int ins_num, int mode, int64_t chunk, void* workspace,
                                  int64_t workspace_bytes, float* rgb, float* depth, float* ins, uint8_t* labels,
                                  void* stream) {
    if (count == 0) return DMNERF_OK;
    DM_REQUIRE(K_host && c2w_host && z_row && packed_coarse && packed_fine && workspace && rgb && depth && ins,
               DMNERF_EINVAL, "render_full: null pointer");
    DM_REQUIRE(H > 0 && W > 0 && pix0 >= 0 && count > 0 && pix0 + count <= static_cast<int64_t>(H) * W, DMNERF_EINVAL,
               "render_full: pixels [%lld, %lld) outside the %dx%d frame", static_cast<long long>(pix0),
               static_cast<long long>(pix0 + count), H, W);
    DM_REQUIRE(chunk >= 1, DMNERF_EINVAL, "render_full: chunk %lld", static_cast<long long>(chunk));
    DM_REQUIRE(ins_num >= 1 && ins_num <= kMaxIns, DMNERF_EUNSUPPORTED, "render_full: ins_num %d outside [1,%d]", ins_num,
               kMaxIns);
    DM_REQUIRE(mode == DMNERF_MODE_FP32 || mode == DMNERF_MODE_BF16 || mode == DMNERF_MODE_F16 || mode == DMNERF_MODE_TC32,
               DMNERF_EUNSUPPORTED, "render_full: unknown mode %d", mode);
    DM_REQUIRE(n_samples >= 4 && n_samples <= 512 && n_importance >= 2 && n_samples + n_importance <= 1024,
               DMNERF_EUNSUPPORTED, "render_full: n_samples %d / n_importance %d unsupported", n_samples, n_importance);
    const int64_t per = balanced_chunk(count, chunk);
    const RenderWs ws = render_ws(per, n_samples, n_importance, ins_num, mode);
    DM_REQUIRE(workspace_bytes >= ws.total, DMNERF_EINVAL, "render_full: workspace of %lld bytes, %lld needed "
               "(dmnerf_render_workspace_bytes)", static_cast<long long>(workspace_bytes), static_cast<long long>(ws.total));
    DM_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 255) == 0, DMNERF_EINVAL, "render_full: workspace must be "
               "256-byte aligned");
    char* base = static_cast<char*>(workspace);
    float* ro = reinterpret_cast<float*>(base + ws.ro);
    float* rd = reinterpret_cast<float*>(base + ws.rd);
    float* zc = reinterpret_cast<float*>(base + ws.zc);
    float* sg = reinterpret_cast<float*>(base + ws.sigma);
    float* zf = reinterpret_cast<float*>(base + ws.zf);
    float* raw = reinterpret_cast<float*>(base + ws.raw);
    const int c = ins_num + 5, s = n_samples + n_importance;
    cudaStream_t st = as_stream(stream);
    {
        const int64_t total = per * n_samples;
        expand_row_kernel<<<static_cast<unsigned>((total + 255) / 256), 256, 0, st>>>(z_row, n_samples, total, zc);
        DM_LAUNCH_CHECK("expand_row_kernel");
    }
    for (int64_t a = 0; a < count; a += per) {
        const int64_t nb = (count - a < per) ? (count - a) : per;
        float* o_rgb = rgb + a * 3;
        float* o_depth = depth + a;
        float* o_ins = ins + a * ins_num;
        if (int rc = dmnerf_rays_k(H, W, K_host, c2w_host, pix0 + a, nullptr, nb, ro, rd, stream)) return rc;
        if (mode == DMNERF_MODE_FP32) {
            if (int rc = dmnerf_mlp_fwd(ro, rd, zc, nb, n_samples, packed_coarse, ins_num, mode, raw, nullptr, stream))
                return rc;
            // the coarse maps land in the output slots and are overwritten by the fine pass below
            if (int rc = dmnerf_composite_fwd(raw, zc, rd, nb, n_samples, c, DMNERF_COMP_DROP_EMPTY, o_rgb, o_depth, sg,
                                              o_ins, nullptr, stream))
                return rc;
            if (int rc = dmnerf_fine_z(zc, sg, u_row, 0, nb, n_samples, n_importance, zf, nullptr, stream)) return rc;
        } else {
            if (int rc = dmnerf_mlp_density(ro, rd, zc, nb, n_samples, packed_coarse, ins_num, mode, sg, stream)) return rc;
            if (int rc = dmnerf_coarse_tail(sg, zc, n_samples, rd, u_row, 0, nb, n_samples, c, n_importance, zf, nullptr,
                                            stream))
                return rc;
        }
        if (int rc = dmnerf_mlp_fwd(ro, rd, zf, nb, s, packed_fine, ins_num, mode, raw, nullptr, stream)) return rc;
        if (int rc = dmnerf_composite_fwd(raw, zf, rd, nb, s, c, DMNERF_COMP_DROP_EMPTY, o_rgb, o_depth, nullptr, o_ins,
                                          nullptr, stream))
            return rc;
        if (labels)
            if (int rc = dmnerf_label_argmax(o_ins, nb, ins_num, labels + a, stream)) return rc;
    }
    return DMNERF_OK;
}
