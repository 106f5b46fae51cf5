// KG side + graph-level layers of the folded cross-attention fusion (fusion_kg.cu).
#pragma once
#include "common.cuh"

namespace cm {

constexpr int KG_MAX_CHUNKS = 16;   // node chunks per image (flash-style partials, merged in fixed order)

struct KgSideArgs {
    const float* s2;      // [M,128] raw kg->rg scores
    const op_t* v2;       // [M,D]
    const float* y;       // [M,D] rg stream
    float *part_m, *part_l;   // [G, KG_MAX_CHUNKS, 128]
    float* part_acc;      // [G, KG_MAX_CHUNKS, gs, D]
    float* part_y;        // [G, KG_MAX_CHUNKS, D]
    float* attn_kg2rg;    // optional output
    int node_cap;
};

struct GraphTailArgs {
    int D, n_kg, num_classes, want_heads;
    const float* z;       // [G * n_kg, D] KG stream after the FFN residual
    const float* both;    // [G, 2D]: columns [0, D) = node mean of the rg stream
    const float *fl1_w, *fl1_b, *fl2_w, *fl2_b;   // fusion_layer.0 / .3 as stored ([out, in])
    const float *hw1[4], *hb1[4], *hw2[4], *hb2[4];
    float *fused, *mask_out, *inst_out, *edge_out, *score_out;
};

// the fused tail (kg_tail_kernel): hidden 256, K <= 16 rows
struct KgTailArgs {
    const float *part_m, *part_l, *part_acc, *part_y;
    int tile_partials;                    // 0: KG_MAX_CHUNKS chunk partials per image (kg2rg_partial); 1: one partial per 128-row tile of
                                          // the node kernel that holds rows of the image, at slot tile + image (fusion_node.cu MODE 2)
    const float* v2_bias;                 // tile partials: the bias of V2 [256], added to the normalised context (NULL: already inside V2)
    const int* node_ptr;
    int n_kg, gs, heads;
    const op_t *wo2, *w1, *w2;            // fp16 [256,256], [512,256], [256,512] (nn.Linear layout)
    const float *kp, *bo2, *ln_g, *ln_b, *b1, *b2;
    float* stat;                          // [G,128,2] merged softmax statistics (for the attention map) or NULL
    GraphTailArgs gt;                     // fusion MLP + heads (z / both unused)
};
int launch_kg_tail(const KgTailArgs& a, const KgSideArgs& k, int n_graphs, cudaStream_t st);

int launch_kg2rg_partial(const KgSideArgs& k, int D, int heads, int gs, const int* node_ptr, int n_graphs, cudaStream_t st);
// ctx2 [G * n_kg, D] fp16, both [G, 2D] (first half), stat [G,128,2] (needed only for the attention map)
int launch_kg_merge(const KgSideArgs& k, int D, int heads, int gs, int n_kg, const int* node_ptr, int n_graphs, op_t* ctx2, float* both, float* stat,
                    cudaStream_t st);
int launch_graph_tail(const GraphTailArgs& a, int n_graphs, cudaStream_t st);

}  // namespace cm
