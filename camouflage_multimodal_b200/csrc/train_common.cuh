// Building blocks shared by the training-mode forward/backward passes (gnn_train.cu, fusion_train.cu): a strided
// fp32 GEMM (NT / NN / TN through strides, optional split-K with a deterministic second-stage reduction), column
// reductions with fp64 partials, and small elementwise helpers.  fp32 CUDA-core arithmetic throughout: gradients are
// compared against torch autograd at 1e-4, and the training step is not the bench's hot loop.
#pragma once
#include "common.cuh"

namespace cm {

// C[M,N] = op(A)[M,K] * op(B)[K,N] (+ bias[n]) (relu) (* keep[m,n] ? keep_scale : 0)
//   A(m,k) = A[m*sam + k*sak], B(k,n) = B[k*sbk + n*sbn]; *_k_contig says which stride is 1 (coalescing only).
// splits > 1: grid.z slices K; slice z writes its partial tile to part + z*M*N (row stride N) and the epilogue
// arguments are applied by splitk_reduce (fixed summation order => deterministic).
struct Gemm {
    const float* A = nullptr;
    long sam = 0, sak = 0;
    bool a_k_contig = true;
    const float* B = nullptr;
    long sbk = 0, sbn = 0;
    bool b_k_contig = true;
    float* C = nullptr;
    long ldc = 0;
    int M = 0, N = 0, K = 0;
    const float* bias = nullptr;
    int relu = 0;
    const uint8_t* keep = nullptr;   // [M, N] dropout keep mask or NULL
    float keep_scale = 1.f;
    float alpha = 1.f;
    int accumulate = 0;              // C += result
    int splits = 1;
    float* part = nullptr;           // [splits, M, N] scratch when splits > 1
};
int launch_gemm(const Gemm& g, cudaStream_t st, const char* name);

// Y = X W^T (+bias)(relu)(dropout): X [M,K] row-major (ldx), W [N,K] (nn.Linear weight)
inline Gemm gemm_nt(const float* X, int ldx, const float* W, float* Y, int ldy, int M, int N, int K, const float* bias, int relu) {
    Gemm g;
    g.A = X; g.sam = ldx; g.sak = 1; g.a_k_contig = true;
    g.B = W; g.sbk = 1; g.sbn = K; g.b_k_contig = true;
    g.C = Y; g.ldc = ldy; g.M = M; g.N = N; g.K = K; g.bias = bias; g.relu = relu;
    return g;
}
// dX = dY W: dY [M,N] (ldy), W [N,K] -> dX [M,K]
inline Gemm gemm_nn(const float* dY, int ldy, const float* W, float* dX, int ldx, int M, int N, int K) {
    Gemm g;
    g.A = dY; g.sam = ldy; g.sak = 1; g.a_k_contig = true;
    g.B = W; g.sbk = K; g.sbn = 1; g.b_k_contig = false;
    g.C = dX; g.ldc = ldx; g.M = M; g.N = K; g.K = N;
    return g;
}
// dW = dY^T X: dY [R,N] (ldy), X [R,K] (ldx) -> dW [N,K]; reduction over the R rows, split-K
inline Gemm gemm_tn(const float* dY, int ldy, const float* X, int ldx, float* dW, int R, int N, int K, float* part, int splits) {
    Gemm g;
    g.A = dY; g.sam = 1; g.sak = ldy; g.a_k_contig = false;
    g.B = X; g.sbk = ldx; g.sbn = 1; g.b_k_contig = false;
    g.C = dW; g.ldc = K; g.M = N; g.N = K; g.K = R; g.part = part; g.splits = splits;
    return g;
}

// out[c] = sum_r a[r*lda + c] (fp64 partials); part: [col_parts(rows)][cols] doubles
int col_parts(int rows);
int launch_colsum(const float* a, int lda, int rows, int cols, double* part, float* out, int accumulate, cudaStream_t st);

// elementwise helpers
int launch_add(const float* a, const float* b, float* out, long count, cudaStream_t st);                 // out = a + b (b may be NULL)
int launch_relu_drop_bwd(const float* act, float scale, long count, float* d, cudaStream_t st);           // d = act > 0 ? d*scale : 0

// LayerNorm over `dim` columns (eps 1e-5, biased variance), one warp per row.  stats [rows,2] = mean, rstd.
int launch_ln_fwd(const float* x, const float* gamma, const float* beta, int rows, int dim, float* y, float* stats, cudaStream_t st);
// dx from dy; writes prod = dy * xhat (for dgamma = colsum(prod), dbeta = colsum(dy))
int launch_ln_bwd(const float* x, const float* stats, const float* gamma, const float* dy, int rows, int dim, float* dx, float* prod,
                  cudaStream_t st);

}  // namespace cm
