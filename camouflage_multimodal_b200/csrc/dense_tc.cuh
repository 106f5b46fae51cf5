// Dense layer on the 5th-generation tensor cores (sm_100a): fp16 operands staged by TMA (128B swizzle) into a
// 4-stage shared-memory ring, tcgen05.mma accumulating fp32 in TMEM, epilogue read back with tcgen05.ld.
//   out[m, n] = LN?( act( alpha * (sum_k A[m,k] W[n,k] + bias[n]) + residual[m,n] ) )
// A [M,K] and W [N,K] are K-major fp16 (W is an nn.Linear weight).  M is a capacity; *m_dev the live rows.
#pragma once
#include <cuda_fp16.h>

#include "common.cuh"

namespace cm {

struct Dense {
    const op_t* A = nullptr;   // [M, lda] fp16, lda % 8 == 0
    int lda = 0;
    const op_t* W = nullptr;   // [N, K] fp16 (row stride K, K % 8 == 0)
    int M = 0, N = 0, K = 0;
    const int* m_dev = nullptr;
    const float* bias = nullptr;        // [N]
    float alpha = 1.f;
    int relu = 0;
    const op_t* residual = nullptr;   // [M or res_mod, ldr] fp16, added after alpha
    const float* residual_f32 = nullptr;       // same role in fp32 (at most one of the two)
    int ldr = 0;
    int res_mod = 0;
    const float* ln_gamma = nullptr;    // LayerNorm over the N outputs (needs N <= 256: one tile per row)
    const float* ln_beta = nullptr;
    op_t* out_h = nullptr;
    int ldo_h = 0;
    // split_out: out_h rows are 3N wide, [hi | lo | hi] with hi = fp16(v), lo = fp16(v - hi): fed to the next
    // layer against [W_hi | W_hi | W_lo] (launch_split_weight) the three products recover ~16 mantissa bits.
    int split_out = 0;
    float* out_f32 = nullptr;
    int ldo_f32 = 0;
    // scores mode (sm_gs = 16 or 32, N == 256): output columns [0,128) are 128/sm_gs head groups of sm_gs slots with
    // sm_valid live keys each -> per-group softmax, written fp16 to out_h [M,128]; columns [128,256) -> raw fp32
    // to out_f32 [M,128]; attn_mean [M, sm_valid] (optional) = mean over the head groups of the probabilities.
    int sm_gs = 0, sm_valid = 0;
    float* attn_mean = nullptr;
    // split-K (training weight gradients: few output tiles, a reduction over every row of the batch): grid.z = k_splits,
    // slice z covers the k-blocks [z * ceil(num_kb / k_splits), ...) and writes its raw fp32 partial tile (no bias / alpha /
    // residual) to out_f32 + z * split_stride; the caller sums the slices in a fixed order.
    int k_splits = 1;
    size_t split_stride = 0;
    // operands are bf16 instead of fp16 (same 2-byte K-major layout; kind::f16 takes either): the training GEMMs use it --
    // gradients span far more than fp16's exponent range.  fp32 outputs only.
    int bf16_operands = 0;
};

int launch_dense(const Dense& d, cudaStream_t st);

// fp32 -> fp16 row copy with zero padding of columns [cols, ld_out): out[r, c] = in[r, c]
int launch_cast_f16(const float* in, int ld_in, int rows, int cols, const int* rows_dev, op_t* out, int ld_out,
                     cudaStream_t st);

// fp32 weight [rows, cols] (optional per-row scale) -> fp16 [rows, 3*kpad] = [hi | hi | lo] thirds, zero padded
int launch_split_weight(const float* in, int rows, int cols, int kpad, const float* row_scale, op_t* out, cudaStream_t st);

}  // namespace cm
