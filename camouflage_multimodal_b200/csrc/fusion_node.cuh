// Fused region-node side of the KG-folded cross-attention fusion (fusion_node.cu): hidden_dim 256, rg_dim 128,
// 128 folded key slots.
#pragma once
#include <cuda.h>

#include "common.cuh"

namespace cm {

struct FusionNodeArgs {
    CUtensorMap maps[7];   // x, Wr, [U1;U2], Wv2, VO^T, W1, W2 (filled by launch_fusion_node)
    int M;                 // row capacity of x (grid size); *m_dev the live rows
    const int* m_dev;
    const float *b_rgproj, *ccat, *b_v2, *b_out1, *ln_g, *ln_b, *b_ffn1, *b_ffn2;
    const float* x32 = nullptr;        // region-node kernel: fp32 input rows [M,128]; converted to the fp16 A operand in the
                                       // kernel (no separate cast pass); NULL = fp16 rows through the x tensor map
    const float* res_rows = nullptr;   // KG-row chain: fp32 residual rows [res_mod, 256] (projected KG rows)
    int res_mod = 1;
    int sm_gs, sm_valid;   // slots per head (16 / 32) and live KG rows
    float* s2;             // [M,128] raw kg->rg scores
    op_t* v2;              // [M,256] V of the kg->rg attention
    float* y;              // [M,256] rg stream after the FFN residual
    float* attn_mean;      // [M, sm_valid] head-averaged rg->kg probabilities or NULL
    // MODE 2 (part_acc != NULL): the kg->rg partials and the column sums of y per (tile + image) slot instead of s2 / v2 / y
    float *part_m = nullptr, *part_l = nullptr;   // [slots,128] tile maximum of the raw scores, sum of exp(s - max) over the image's rows
    float* part_acc = nullptr;                    // [slots,16,256] sum of exp(s - max) V2
    float* part_y = nullptr;                      // [slots,256] column sums of y
    const int* node_ptr = nullptr;                // [n_graphs + 1]
    int n_graphs = 0;
    long long* dbg = nullptr;   // phase time stamps of CTA 0 (tools/fusion_phase_profile.py): [0..7] MMA issue, [8..15] MMA waited-for, [16..23] epilogue start, [24..31] epilogue done, [32] kernel start
};

int launch_fusion_node(FusionNodeArgs& a, const void* x, const void* w_rgproj, const void* wcat, const void* wv2, const void* vot,
                       const void* ffn1, const void* ffn2, cudaStream_t st);

int launch_fusion_kgrows(FusionNodeArgs& a, const void* x, const void* wo, const void* ffn1, const void* ffn2, cudaStream_t st);

}  // namespace cm
