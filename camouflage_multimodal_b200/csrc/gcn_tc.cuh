// Fused GCNConv layer for hidden == 128 on the 5th-generation tensor cores (sm_100a):
//   out = relu( (A_hat x) W'^T + b' )        A_hat = D^-1/2 (A + I) D^-1/2 over the CSR rows, W'/b' BatchNorm-folded
// The neighbour sum is formed in fp32 straight into the swizzled shared-memory A operand (fp16 hi + lo split, the
// same three-product scheme as dense_tc's split layers), so the aggregated tensor never goes to HBM.  With
// `chain` the CTA carries on through fc_shared (-> emb) and, with `do_heads`, the three head MLPs (-> logits).
#pragma once
#include "common.cuh"

namespace cm {

struct GcnLayer {
    const float* x = nullptr;          // [M, 128] fp32 layer input
    const int* row_ptr = nullptr;      // CSR over target nodes
    const int* col = nullptr;
    const float* w = nullptr;          // edge weights (NULL = 1)
    const float* self_w = nullptr;     // self-loop weights (NULL = 1)
    const float* dis = nullptr;        // deg^-1/2
    const int* n_dev = nullptr;        // live rows (NULL = M)
    int M = 0;
    const op_t* w_split = nullptr;     // [128, 384] fp16 = [W_hi | W_hi | W_lo]
    const float* bias = nullptr;       // [128]
    float* out = nullptr;              // [M, 128] fp32 (plain layer)
    // GAT layer (gat = 1): x is [M, gat_c] fp32 input features, a_src / a_dst [M, 4] the per-head attention logits,
    // w_split [128, 3 x 64] = [W_hi | W_hi | W_lo] over the 4 x gat_c aggregated columns; dis_out receives deg^-1/2
    int gat = 0, gat_c = 0;
    const float* a_src = nullptr;
    const float* a_dst = nullptr;
    float* dis_out = nullptr;
    int chain = 0, do_heads = 0;
    const op_t* fc_split = nullptr;    // [128, 384]
    const float* fc_bias = nullptr;    // [128] or NULL
    float* emb = nullptr;              // [M, 128] fp32 or NULL
    const op_t* h1_split = nullptr;    // [192, 384]: mask | instance | edge hidden layers
    const float* h1_bias = nullptr;    // [192]
    const float* w2[3] = {nullptr, nullptr, nullptr};   // [2,64], [2,64], [1,64]
    const float* b2[3] = {nullptr, nullptr, nullptr};
    float *out_mask = nullptr, *out_inst = nullptr, *out_edge = nullptr, *prob_out = nullptr;
};

int launch_gcn_layer_tc(const GcnLayer& p, cudaStream_t st);

}  // namespace cm
