// Internal dense-layer interface shared by the GNN and fusion stages:
//   C[m, n] = act( alpha * (sum_k A[m,k] * W[n,k] + bias[n]) + residual[m,n] )
// W is a torch nn.Linear weight ([out_features, in_features], row-major).  m_dev (optional) points at the
// device-side row count so ragged batches never need a host round trip; rows >= *m_dev are not touched.
#pragma once
#include "common.cuh"

namespace cm {

struct Linear {
    const float* A = nullptr;
    int lda = 0;
    const float* W = nullptr;
    const float* bias = nullptr;       // [N] or NULL
    const float* residual = nullptr;   // [M, ldr] or NULL
    int ldr = 0;
    int res_mod = 0;                   // > 0: residual row = row % res_mod (rows broadcast over graphs)
    float* C = nullptr;
    int ldc = 0;
    int M = 0, N = 0, K = 0;           // M is the capacity (grid size); *m_dev the live row count
    const int* m_dev = nullptr;
    float alpha = 1.f;
    int relu = 0;
};

int launch_linear(const Linear& a, cudaStream_t st);

// out[r, :] = LayerNorm(in[r, :]) * gamma + beta over `dim` columns (eps 1e-5, biased variance).
int launch_layernorm(const float* in, float* out, const float* gamma, const float* beta, int rows, int dim,
                     const int* rows_dev, cudaStream_t st);

// out[g, :] = mean over rows seg_ptr[g] .. seg_ptr[g+1]-1 of in (ld = dim); empty segment -> 0.
// With seg_ptr == NULL the segments are `fixed_len` rows each.
int launch_segment_mean(const float* in, int dim, const int* seg_ptr, int fixed_len, int n_seg, float* out, int ldo,
                        cudaStream_t st);

// Second layer of a bundle of small MLP heads: head t reads hid[r, t*hdim .. (t+1)*hdim) and writes cdim[t]
// outputs with weight w2[t] [cdim, hdim]; optional sigmoid; prob_out = softmax(head 0)[:, 1].
struct Heads2 {
    int n_heads = 0, hdim = 0;
    const float* w2[4] = {nullptr, nullptr, nullptr, nullptr};
    const float* b2[4] = {nullptr, nullptr, nullptr, nullptr};
    int cdim[4] = {0, 0, 0, 0};
    float* out[4] = {nullptr, nullptr, nullptr, nullptr};
    int sigmoid[4] = {0, 0, 0, 0};
    float* prob_out = nullptr;
};
int launch_heads2(const float* hid, int ld, int rows, const int* rows_dev, const Heads2& h, cudaStream_t st);

}  // namespace cm
