/* dmnerf_oracle_c.c -- plain-C restatement of the index-exact stages, TEST INFRASTRUCTURE ONLY.
 *
 * torch's CPU sum/cumsum association is implementation-defined (cumsum accumulates in double,
 * sum is a vectorised cascade), so "the reference's CDF bits" differ between torch builds and
 * between CPU and CUDA.  The CUDA sampler therefore fixes one documented order -- running DOUBLE
 * accumulators rounded to fp32 per element, i.e. what ATen's CPU cumsum does and the correctly
 * rounded value its cascade sum approximates -- and THIS file restates the reference algorithm
 * (networks/helpers.py:123-155) in that order, so the kernel's indices can be checked bit-for-bit.
 * Compiled with -ffp-contract=off (no FMA).
 * Only tests/ and bench.py's cpu_baseline leg may load it. */
#include <math.h>
#include <stdint.h>

/* helpers.py:123-155; bins [n,nb], weights [n,nb-1], u [n,ns] (u_stride 0 = shared row). */
void oracle_sample_pdf(const float* bins, const float* weights, const float* u, int64_t u_stride, int64_t n,
                       int nb, int ns, float* z_out, int64_t* inds_out) {
    /* rays are independent and the function keeps no state outside its stack: bench.py's stage-level CPU baseline calls
     * it from one host thread per core on disjoint ray ranges (ctypes releases the GIL; libgomp is not in this image) */
    for (int64_t r = 0; r < n; ++r) {
        float cdf[1024], w[1024];
        const float* b = bins + r * nb;
        double tot = 0.0;
        for (int j = 0; j < nb - 1; ++j) { w[j] = weights[r * (nb - 1) + j] + 1e-5f; tot += (double)w[j]; }   /* :125 */
        const float total = (float)tot;
        double c = 0.0;
        cdf[0] = 0.f;                                                                                     /* :128 */
        for (int j = 0; j < nb - 1; ++j) {                                                                /* :126-127 */
            const float pdf = w[j] / total;
            c += (double)pdf;
            cdf[j + 1] = (float)c;
        }
        for (int i = 0; i < ns; ++i) {
            const float uu = u[r * u_stride + i];
            int lo = 0, hi = nb;                                     /* searchsorted(right=True), :139 */
            while (lo < hi) { int mid = (lo + hi) >> 1; if (cdf[mid] <= uu) lo = mid + 1; else hi = mid; }
            const int below = lo - 1 > 0 ? lo - 1 : 0;               /* :140 */
            const int above = lo < nb - 1 ? lo : nb - 1;             /* :141 */
            float denom = cdf[above] - cdf[below];                   /* :150 */
            if (denom < 1e-5f) denom = 1.0f;                         /* :151 */
            const float t = (uu - cdf[below]) / denom;               /* :152 */
            z_out[r * ns + i] = b[below] + t * (b[above] - b[below]);/* :153 */
            if (inds_out) inds_out[r * ns + i] = lo;
        }
    }
}

/* render.py:6-28 with sequential (left-to-right) products and sums; flags bit0 = drop empty class. */
void oracle_composite(const float* raw, const float* z, const float* rays_d, int64_t n, int s, int c, int flags,
                      float* rgb, float* depth, float* weights, float* ins) {
    const int L = c - 4, Lout = L - (flags & 1);
    for (int64_t r = 0; r < n; ++r) {
        const float* d = rays_d + r * 3;
        const float norm = sqrtf(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
        float T = 1.f, acc[3] = {0, 0, 0}, dep = 0.f, a_ins[128];
        for (int j = 0; j < L; ++j) a_ins[j] = 0.f;
        for (int k = 0; k < s; ++k) {
            const float* row = raw + (r * s + k) * c;
            const float dist = (k + 1 < s ? z[r * s + k + 1] - z[r * s + k] : 1e10f) * norm;
            const float sg = row[3] > 0.f ? row[3] : 0.f;
            const float alpha = 1.f - expf(-sg * dist);
            const float w = alpha * T;
            T = T * (1.f - alpha + 1e-10f);
            if (weights) weights[r * s + k] = w;
            for (int j = 0; j < 3; ++j) acc[j] += w * (1.f / (1.f + expf(-row[j])));
            dep += w * z[r * s + k];
            for (int j = 0; j < L; ++j) a_ins[j] += w * row[4 + j];
        }
        for (int j = 0; j < 3; ++j) rgb[r * 3 + j] = acc[j];
        depth[r] = dep;
        for (int j = 0; j < Lout; ++j) ins[r * Lout + j] = 1.f / (1.f + expf(-a_ins[j]));
    }
}
