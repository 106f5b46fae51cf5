// Tensor-core path of the cross-attention fusion for a knowledge graph shared by the whole batch (fusion_tc.cu).
#pragma once
#include "common.cuh"

namespace cm {

struct FusionFolded {   // byte offsets into the folded blob
    int gs, n_kg;
    size_t w_rgproj, wcat, wv2, vot, ffn1, ffn2, wo2_h, kffn1_h, kffn2_h, ccat, kp, k1v1, q2, total;
};

bool fusion_fold_supported(const cm_fusion_params* p, int n_kg);
FusionFolded fusion_folded_layout(const cm_fusion_params* p, int n_kg);
int fusion_fold(const cm_fusion_params* p, const float* kg, int n_kg, void* folded, cudaStream_t st);
size_t fusion_tc_workspace_bytes(const cm_fusion_params* p, int n_graphs, int node_cap, int n_kg);
int fusion_forward_tc(const cm_fusion_params* p, const void* folded, const float* rg, const int32_t* node_ptr, int n_graphs,
                      int node_cap, int n_kg, void* workspace, float* fused, float* mask_out, float* instance_out, float* edge_out,
                      float* score_out, float* attn_rg2kg, float* attn_kg2rg, cudaStream_t st);

}  // namespace cm
