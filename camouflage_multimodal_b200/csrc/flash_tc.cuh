// Large-KG tensor-core path of the cross-attention fusion (flash_tc.cu): fused QK^T.softmax.V on tcgen05 / TMEM / TMA.
#pragma once
#include "common.cuh"

namespace cm {

constexpr int FUSION_FLASH_MIN_KG = 128;   // below this the folded / per-row kernels win (one key block would be mostly padding)

int launch_flash(const op_t* Q, int Mq, const op_t* K, int Mk, const op_t* Vt, int ldv, int D, const int* seg_ptr, int n_seg, float scale,
                 op_t* out, int ldo, cudaStream_t st);

struct FlashTail {   // graph-level buffers inside the flash workspace for the fusion MLP + heads that follow
    float *both, *f1, *fused, *hid;
};
bool fusion_flash_supported(const cm_fusion_params* p, int n_kg);
size_t fusion_flash_workspace_bytes(const cm_fusion_params* p, int n_graphs, int node_cap, int n_kg);
// Shared KG rows [n_kg, kg_dim]; fills tail->both [G, 2*hidden] (mean-pooled region stream | mean-pooled KG stream).
int fusion_forward_flash(const cm_fusion_params* p, const float* rg, const int32_t* node_ptr, int n_graphs, int node_cap, const float* kg,
                         int n_kg, void* workspace, FlashTail* tail, cudaStream_t st);

}  // namespace cm
