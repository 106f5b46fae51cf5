// Shared by metrics.cu (cm_metrics_finalize) and sod_metrics.cu (the fused region-table kernel): per-image ratios of
// utils/metrics.py:9-56 from the confusion counts + the batch sums, one block of 256 threads, fixed summation order.
#pragma once
#include "common.cuh"

namespace cm {

__device__ __forceinline__ void metrics_finalize_block(const unsigned long long* __restrict__ counts, const double* __restrict__ abs_err, int batch,
                                                       long long pixels, double* __restrict__ per_image, double* __restrict__ partial,
                                                       double (*red)[14] /* shared [8][14] */) {
    double s[14];
#pragma unroll
    for (int k = 0; k < 14; ++k) s[k] = 0.0;
    const float epsf = 1e-8f;
    const double n = (double)pixels;
    for (int i = threadIdx.x; i < batch; i += 256) {
        // The reference forms the five ratios in float32 (0/1 masks are astype(np.float32) and
        // float32 + 1e-8 stays float32, utils/metrics.py:11-17); MAE and accuracy are float64.
        const float tp = (float)__ldcg(counts + 4 * i), sp = (float)__ldcg(counts + 4 * i + 1), sg = (float)__ldcg(counts + 4 * i + 2);
        const double eq = (double)__ldcg(counts + 4 * i + 3);
        const float fp = __fsub_rn(sp, tp), fn = __fsub_rn(sg, tp);
        const float iou = __fdiv_rn(tp, __fadd_rn(__fsub_rn(__fadd_rn(sp, sg), tp), epsf));            // :14-17
        const float dice = __fdiv_rn(__fmul_rn(2.0f, tp), __fadd_rn(__fadd_rn(sp, sg), epsf));          // :26
        const float prec = __fdiv_rn(tp, __fadd_rn(__fadd_rn(tp, fp), epsf));                           // :39
        const float rec = __fdiv_rn(tp, __fadd_rn(__fadd_rn(tp, fn), epsf));                            // :40
        const float f1 = __fdiv_rn(__fmul_rn(2.0f, __fmul_rn(prec, rec)), __fadd_rn(__fadd_rn(prec, rec), epsf));   // :41
        double m[7];
        m[0] = iou; m[1] = dice; m[2] = prec; m[3] = rec; m[4] = f1;
        m[5] = __ldcg(abs_err + i) / n;                      // :47
        m[6] = eq / n;                              // :54
#pragma unroll
        for (int k = 0; k < 7; ++k) {
            if (per_image) per_image[7 * (size_t)i + k] = m[k];
            s[k] += m[k];
            s[7 + k] += m[k] * m[k];
        }
    }
#pragma unroll
    for (int k = 0; k < 14; ++k) s[k] = warp_sum(s[k]);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0)
        for (int k = 0; k < 14; ++k) red[warp][k] = s[k];
    __syncthreads();
    if (threadIdx.x < 14 && partial) {
        double t = 0.0;
        for (int wv = 0; wv < 8; ++wv) t += red[wv][threadIdx.x];
        partial[1 + threadIdx.x] = t;
    }
    if (threadIdx.x == 0 && partial) partial[0] = (double)batch;
}

}  // namespace cm
