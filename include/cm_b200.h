/*
 * cm_b200.h -- C ABI of the B200-native (sm_100a) region-graph hot path.
 *
 * The reference (rajan-dubey8/camouflage-multimodal) is pure Python and has no FFI of its own
 * (SURVEY.md F2); these entry points are what a binding for the path would call.  Each one names the
 * reference code it replaces (paths relative to the reference tree).  Conventions:
 *   - every pointer is a DEVICE pointer on the current CUDA device unless its name ends in _host;
 *   - outputs are caller-allocated at the stated capacity; nothing is allocated or freed here;
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream); calls are
 *     asynchronous with respect to the host and do not synchronise;
 *   - return value 0 = launched; negative = argument / launch error, text in cm_last_error();
 *   - data-dependent overflows (label >= label_cap, more nodes/edges than the output capacity) are
 *     reported through the device-side `status` word (bit mask CM_STATUS_*), never by writing out of
 *     bounds;
 *   - ragged batches: image b owns nodes node_ptr[b] .. node_ptr[b+1]-1 (global node ids) and directed
 *     edges edge_ptr[b] .. edge_ptr[b+1]-1.
 * There is no CPU fallback: every entry point fails with CM_ERR_DEVICE on a device that is not
 * compute capability 10.x.
 */
#ifndef CM_B200_H
#define CM_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CM_OK 0
#define CM_ERR_ARG (-1)
#define CM_ERR_DEVICE (-2)
#define CM_ERR_LAUNCH (-3)
#define CM_ERR_WORKSPACE (-4)

#define CM_STATUS_LABEL_OVERFLOW 1u /* a label >= label_cap was seen (ignored like a negative label) */
#define CM_STATUS_NODE_OVERFLOW 2u  /* more nodes than node_cap */
#define CM_STATUS_EDGE_OVERFLOW 4u  /* more directed edges than edge_cap */

#define CM_NODE_FEATURES 15 /* extract_rg_embeddings.py:195-206 */

const char* cm_last_error(void);
int cm_abi_version(void);
/* 0 when the current device is sm_100 (B200); CM_ERR_DEVICE otherwise. */
int cm_device_check(void);

/* Optional per-kernel timing for bench.py: when enabled, every kernel launch made by this library is
 * bracketed by CUDA events on its launching stream.  cm_timing_report synchronises those events, writes one
 * line "name launches total_ms" per kernel name into buf, clears the records and returns the bytes written. */
void cm_timing_enable(int on);
int cm_timing_report(char* buf, int cap);

/* ------------------------------------------------------------------------------------------------
 * A1 + A2: region adjacency + label compaction.
 * Replaces skimage.graph.rag_mean_color(...).edges() (models/region_graph/extract_rg_embeddings.py:216,
 * 220) and the region_id_map bookkeeping + edge remap (:148-149, 155-161, 210-211, 221-225).
 *
 * labels      [B,H,W] int32, any values; labels < 0 or >= label_cap are never nodes.
 * id_of_label [B,label_cap] int32 out: node id LOCAL to the image (0..N_b-1, ascending label order) or -1.
 * node_ptr    [B+1] int32 out, edge_ptr [B+1] int32 out (exclusive prefix sums over images).
 * node_label  [node_cap] int32 out: label of every global node.
 * node_bbox   [node_cap,4] int32 out: x0, y0, x1, y1 (inclusive) of the region.
 * batch_vec   [node_cap] int64 out or NULL: image index of every node (PyG `batch`).
 * csr_row_ptr [node_cap+1] int32 out: CSR over GLOBAL node ids; row i lists the neighbours of i ascending.
 * csr_col     [edge_cap] int32 out: global node id of the neighbour.
 * edge_index  [2,edge_cap] int64 out or NULL (row stride = edge_cap): directed edges (i -> j), both
 *             directions of every RAG edge, sorted lexicographically by (row 0, row 1) -- the canonical
 *             order on which parity with the reference's networkx order is defined.  Node ids are LOCAL
 *             to the image when local_ids != 0 (one Data per image), GLOBAL otherwise (Batch.from_data_list,
 *             train.py:231).
 * workspace   cm_rag_workspace_bytes(B, label_cap) bytes of scratch (zeroed by the call).
 * ------------------------------------------------------------------------------------------------ */
size_t cm_rag_workspace_bytes(int batch, int label_cap);
int cm_rag_build(const int32_t* labels, int batch, int h, int w, int label_cap, void* workspace,
                 size_t workspace_bytes, int32_t* id_of_label, int32_t* node_ptr, int32_t* edge_ptr,
                 int32_t* node_label, int32_t* node_bbox, int64_t* batch_vec, int32_t* csr_row_ptr,
                 int32_t* csr_col, int64_t* edge_index, int local_ids, int node_cap, int edge_cap,
                 uint32_t* status, void* stream);

/* A3: the 15 region features (extract_rg_embeddings.py:154-211).
 * image [B,H,W,3] float32 in [0,1]; canny [B,H,W] uint8 (non-zero = edge pixel, the output of
 * skimage.feature.canny(gray, sigma=2), :152); x [node_cap,15] float32 out.
 * norm_side is the reference's hard-coded 256.0 (:172-174).  label_cap = the bound passed to cm_rag_build
 * (labels >= label_cap are never nodes); 0 selects the bounding-box kernel that needs no per-label table. */
int cm_node_features(const float* image, const int32_t* labels, const uint8_t* canny, int batch, int h,
                     int w, int label_cap, const int32_t* node_ptr, const int32_t* node_label,
                     const int32_t* node_bbox, float norm_side, float* x, int node_cap, void* stream);

/* A3 on 8-bit images: image_u8 [B,H,W,3] uint8 as a decoder yields it (PIL convert('RGB') + resize,
 * extract_rg_embeddings.py:260-261); the image value is byte / 255.0, which is what the reference's
 * ToTensor -> Normalize -> de-normalise -> clip round trip (:263-276) hands create_region_graph up to float32
 * rounding.  All moments are integers, so the features are exact up to the final fp64 divisions.
 * node_bbox (from cm_rag_build) selects the region-parallel kernel (one warp per region over its bounding box);
 * node_bbox == NULL selects the pixel-parallel tile kernel, which requires label_cap > 0 and
 * (h / 4 or h / 8 rows) * w <= 66051 pixels per CTA (CM_ERR_ARG otherwise). */
int cm_node_features_u8(const uint8_t* image_u8, const int32_t* labels, const uint8_t* canny, int batch, int h,
                        int w, int label_cap, const int32_t* node_ptr, const int32_t* node_label,
                        const int32_t* node_bbox, float norm_side, float* x, int node_cap, void* stream);

/* Label maps arrive from the host as int16 when they fit (SLIC with n_segments=500 yields < 600 labels; the
 * reference holds them as int64, extract_rg_embeddings.py:144): widen n labels to the int32 the kernels read.
 * Both buffers 16-byte aligned. */
int cm_labels_widen_i16(const int16_t* labels16, int64_t n, int32_t* labels32, void* stream);

/* A4: Gaussian-kernel edge weights (extract_rg_embeddings.py:227-235), one per directed CSR entry,
 * float32 arithmetic.  n_nodes_dev points at the device-side node count (node_ptr + B). */
int cm_edge_weights(const float* x, const int32_t* csr_row_ptr, const int32_t* csr_col,
                    const int32_t* n_nodes_dev, int node_cap, float* edge_w, void* stream);

/* CSR keyed by TARGET node from an arbitrary PyG-style edge_index [2,E] int64 (source row 0, target row 1)
 * with optional per-edge weights -- what RegionGraphGNN.forward receives (extract_rg_embeddings.py:54-56)
 * when the graph did not come from cm_rag_build.  Self loops are removed from the CSR and returned in
 * self_w (last one per node wins, default 1: PyG add_remaining_self_loops).  Row entries keep the order of
 * the input edges.  workspace: cm_csr_workspace_bytes(n_nodes, n_edges). */
size_t cm_csr_workspace_bytes(int n_nodes, int n_edges);
int cm_csr_from_edge_index(const int64_t* edge_index, int64_t row_stride, const float* edge_w, int n_nodes,
                           int n_edges, void* workspace, size_t workspace_bytes, int32_t* csr_row_ptr,
                           int32_t* csr_col, float* csr_w, float* self_w, uint32_t* status, void* stream);

/* Dense-layer building block of A5-A11 on the tcgen05 tensor cores (fp16 operands, fp32 accumulation in TMEM):
 *   out[m,n] = LN?( relu?( alpha * (sum_k a[m,k] w[n,k] + bias[n]) + residual[m,n] ) )
 * a [m,lda] fp16 and w [n,k] fp16 (an nn.Linear weight), k % 8 == 0, lda % 8 == 0, 16-byte aligned; m is a
 * capacity and *m_dev (optional) the live row count; bias / ln_gamma / ln_beta fp32 [n] (LayerNorm eps 1e-5,
 * n <= 256); residual fp16 [m,ldr]; out_f16 and/or out_f32 [m,n]. */
int cm_linear_f16(const void* a_f16, int lda, const void* w_f16, int m, int n, int k, const int32_t* m_dev,
                   const float* bias, float alpha, int relu, const void* residual_f16, int ldr,
                   const float* ln_gamma, const float* ln_beta, void* out_f16, float* out_f32, void* stream);

/* ------------------------------------------------------------------------------------------------
 * A5-A7: RegionGraphGNN forward (extract_rg_embeddings.py:27-135), eval mode.
 * The raw parameters are the reference state-dict tensors (PyG names); cm_gnn_fold turns them
 * into the folded device blob the forward kernels read (BatchNorm folded into the preceding conv, the GAT
 * linear folded into its attention vectors) -- call it once per weight set.
 * ------------------------------------------------------------------------------------------------ */
typedef struct cm_gnn_params {
    int in_channels, hidden, heads, num_classes; /* 15, 128, 4, 2 */
    const float* conv1_lin_weight;               /* [heads*hidden, in] */
    const float* conv1_att_src;                  /* [heads*hidden] */
    const float* conv1_att_dst;                  /* [heads*hidden] */
    const float* conv1_bias;                     /* [hidden] */
    const float* conv_lin_weight[3];             /* conv2..4 lin.weight [hidden, hidden] */
    const float* conv_bias[3];                   /* [hidden] */
    const float* bn_weight[4];                   /* bn1..4 */
    const float* bn_bias[4];
    const float* bn_mean[4];
    const float* bn_var[4];
    float bn_eps;
    const float* fc_shared_weight; /* [hidden, hidden] */
    const float* fc_shared_bias;
    const float* head1_weight[3]; /* fc_{mask,instance,edge}_1.weight [hidden/2, hidden] */
    const float* head1_bias[3];
    const float* head2_weight[3]; /* [c, hidden/2], c = num_classes, num_classes, 1 */
    const float* head2_bias[3];
} cm_gnn_params;

size_t cm_gnn_folded_bytes(const cm_gnn_params* p);
int cm_gnn_fold(const cm_gnn_params* p, void* folded, size_t folded_bytes, void* stream);

/* Scratch for one forward over at most node_cap nodes. */
size_t cm_gnn_workspace_bytes(const cm_gnn_params* p, int node_cap);
/* x [node_cap,15]; CSR by target (csr_w NULL = unweighted GCN, self_w NULL = all ones);
 * n_nodes_dev NULL => exactly node_cap nodes.  Outputs (any may be NULL): node_emb [node_cap,hidden],
 * mask_out [node_cap,num_classes], instance_out [node_cap,num_classes], edge_out [node_cap,1],
 * mask_prob [node_cap] = softmax(mask_out,1)[:,1] (models/region_graph/test.py:238),
 * graph_emb [n_graphs,hidden] = global_mean_pool over node_ptr segments (:129-133). */
int cm_gnn_forward(const cm_gnn_params* p, const void* folded, const float* x, const int32_t* csr_row_ptr,
                   const int32_t* csr_col, const float* csr_w, const float* self_w,
                   const int32_t* n_nodes_dev, int node_cap, const int32_t* node_ptr, int n_graphs,
                   void* workspace, size_t workspace_bytes, float* node_emb, float* mask_out,
                   float* instance_out, float* edge_out, float* mask_prob, float* graph_emb, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Training-mode RegionGraphGNN: forward under model.train() and the backward pass
 * (extract_rg_embeddings.py:54-90; the step is models/region_graph/train.py:345-357).
 * BatchNorm uses the statistics of ALL nodes in the batch (Batch.from_data_list, train.py:231) and returns them
 * so the caller can update running_mean / running_var; dropout (F.dropout p=0.3 after bn1..bn3, p=0.2 after
 * fc_shared and the three head hidden layers) is applied through caller-supplied keep masks -- PyTorch's Philox
 * stream cannot be reproduced, so the masks are data (NULL = dropout off, which is what gradient parity checks use).
 * The CSR must be structurally symmetric (every region adjacency is); a missing reverse edge sets
 * CM_STATUS_EDGE_OVERFLOW in *status.  n_nodes / n_edges are host-side exact counts.
 * The workspace carries the saved activations from cm_gnn_train_forward to cm_gnn_backward.
 * ------------------------------------------------------------------------------------------------ */
typedef struct cm_gnn_dropout {
    const uint8_t* keep[7]; /* [N,H] after bn1, bn2, bn3, fc_shared; [N,H/2] mask, instance, edge head hidden; NULL = keep all */
    float p_conv, p_fc;     /* 0.3, 0.2 */
} cm_gnn_dropout;
typedef struct cm_gnn_grads { /* one buffer per trainable tensor of cm_gnn_params, same shapes; overwritten */
    float *conv1_lin_weight, *conv1_att_src, *conv1_att_dst, *conv1_bias;
    float *conv_lin_weight[3], *conv_bias[3];
    float *bn_weight[4], *bn_bias[4];
    float *fc_shared_weight, *fc_shared_bias;
    float *head1_weight[3], *head1_bias[3], *head2_weight[3], *head2_bias[3];
} cm_gnn_grads;
size_t cm_gnn_train_workspace_bytes(const cm_gnn_params* p, int n_nodes, int n_edges);
/* Data-parallel BatchNorm: the reference normalises over ALL nodes of the batch (train.py:231,241-250), so
 * when the batch is sharded over ranks every BatchNorm layer must see every rank's (sum x, sum x^2, n), and its backward
 * every rank's (sum dy, sum dy xhat).  With a hook set, cm_gnn_train_forward / cm_gnn_backward copy those column sums
 * (float64, 2 * hidden + 1 / 2 * hidden values) to device_buf, call fn(user, count, stream) -- the host sums
 * device_buf[0..count) over the ranks on that stream (ncclAllReduce through torch.distributed) -- and read them back.
 * bn_weight / bn_bias gradients stay this rank's partial sums (the gradient all-reduce adds them like the rest).
 * fn == NULL clears the hook (single-process statistics).  One hook per process (one process per GPU). */
/* The dense building block of the training passes (also their unit-test hook): c [m,n] = a w^T (+ bias), fp32 in and out.
 * a is [m,k] row-major, or [k,m] with transposed_a (the weight-gradient access pattern); w is [n,k] (nn.Linear layout).
 * Large shapes run on the tcgen05 dense kernel with both operands split into 16-bit thirds (hi | lo | hi against
 * hi | hi | lo, fp32 accumulation; fp16 parts for the forward pattern, bf16 parts with transposed_a); small ones on fp32
 * CUDA-core tiles. */
int cm_train_gemm_f32(const float* a, const float* w, const float* bias, int m, int n, int k, int transposed_a, float* c, void* stream);
typedef void (*cm_bn_sync_fn)(void* user, int count, void* stream);
int cm_set_bn_sync(cm_bn_sync_fn fn, void* user, double* device_buf, int capacity);
/* Outputs: mask_out / instance_out [N,num_classes], edge_out [N,1], node_emb [N,hidden] (optional; the shared
 * embedding after dropout), bn_batch_mean / bn_batch_var [4,hidden] (optional; biased variance). */
int cm_gnn_train_forward(const cm_gnn_params* p, const float* x, const int32_t* csr_row_ptr, const int32_t* csr_col,
                         const float* csr_w, const float* self_w, int n_nodes, int n_edges, const cm_gnn_dropout* drop,
                         void* workspace, size_t workspace_bytes, float* mask_out, float* instance_out, float* edge_out,
                         float* node_emb, float* bn_batch_mean, float* bn_batch_var, uint32_t* status, void* stream);
/* d_* are the loss gradients with respect to the forward outputs (NULL = zero). */
int cm_gnn_backward(const cm_gnn_params* p, const float* x, const int32_t* csr_row_ptr, const int32_t* csr_col,
                    const float* csr_w, const float* self_w, int n_nodes, int n_edges, const cm_gnn_dropout* drop,
                    void* workspace, size_t workspace_bytes, const float* d_mask_out, const float* d_instance_out,
                    const float* d_edge_out, const float* d_node_emb, const cm_gnn_grads* grads, void* stream);

/* ------------------------------------------------------------------------------------------------
 * A10-A12: CrossAttentionFusion / LateFusion + the four detector heads
 * (models/multimodal/fusion_model.py:16-246), eval mode.  Raw parameters = reference state dict.
 * ------------------------------------------------------------------------------------------------ */
typedef struct cm_fusion_params {
    int rg_dim, kg_dim, hidden, heads, num_classes;
    int fusion_type; /* 0 = cross_attention, 1 = late */
    /* cross attention (NULL weight = nn.Identity for the two projections, fusion_model.py:29-30) */
    const float *rg_proj_w, *rg_proj_b, *kg_proj_w, *kg_proj_b;
    const float *r2k_in_w, *r2k_in_b, *r2k_out_w, *r2k_out_b; /* cross_attn_rg2kg */
    const float *k2r_in_w, *k2r_in_b, *k2r_out_w, *k2r_out_b; /* cross_attn_kg2rg */
    const float *ln_rg_w, *ln_rg_b, *ln_kg_w, *ln_kg_b;
    const float *ffn_rg_w1, *ffn_rg_b1, *ffn_rg_w2, *ffn_rg_b2;
    const float *ffn_kg_w1, *ffn_kg_b1, *ffn_kg_w2, *ffn_kg_b2;
    const float *fl_w1, *fl_b1, *fl_w2, *fl_b2; /* fusion_layer.0 / .3 */
    /* late fusion: fusion.fusion.{0,3,6} */
    const float *late_w[3], *late_b[3];
    /* heads: mask, instance, edge, score; layer .0 [final/2, final], layer .3 [c, final/2] */
    const float *head_w1[4], *head_b1[4], *head_w2[4], *head_b2[4];
} cm_fusion_params;

size_t cm_fusion_workspace_bytes(const cm_fusion_params* p, int n_graphs, int node_cap, int n_kg);
/* Tensor-core path for knowledge-graph rows shared by every image (the reference feeds the same category rows
 * to every sample -- "KG broadcast once"): cm_fusion_fold projects kg [n_kg, kg_dim] through the KG side of both
 * attention blocks once per (weights, kg) pair and stores fp16 copies of the remaining weights; the blob is then
 * passed to cm_fusion_forward, which runs the region-node side as tcgen05 GEMMs.  cm_fusion_folded_bytes returns
 * 0 when the shape has no folded form (late fusion, hidden_dim > 256, num_heads * slots(n_kg) > 128). */
size_t cm_fusion_folded_bytes(const cm_fusion_params* p, int n_kg);
int cm_fusion_fold(const cm_fusion_params* p, const float* kg, int n_kg, void* folded, size_t folded_bytes,
                   void* stream);
/* folded: blob of cm_fusion_fold (requires kg_graphs == 1; kg itself is then not read again) or NULL for the
 * general fp32 path (per-image kg rows, late fusion, large KG vocabularies).
 * rg [node_cap, rg_dim] ragged rows (node_ptr [n_graphs+1]); kg [kg_graphs, n_kg, kg_dim] with kg_graphs
 * either 1 (the knowledge-graph rows are shared by every image -- broadcast once) or n_graphs.
 * Outputs: fused [n_graphs, final]; mask/instance [n_graphs,num_classes]; edge/score [n_graphs,1];
 * attn_rg2kg [node_cap, n_kg] and attn_kg2rg (image b's block [n_kg, N_b] at offset n_kg*node_ptr[b])
 * are written only when non-NULL (MultimodalCamouflageDetector.forward(return_attention=...), :243-246). */
int cm_fusion_forward(const cm_fusion_params* p, const void* folded, const float* rg,
                      const int32_t* node_ptr, int n_graphs, int node_cap, const float* kg, int kg_graphs, int n_kg, void* workspace,
                      size_t workspace_bytes, float* fused, float* mask_out, float* instance_out,
                      float* edge_out, float* score_out, float* attn_rg2kg, float* attn_kg2rg,
                      void* stream);

/* Profiling aid: while buf (40 int64 on the device) is set, CTA 0 of the fused node kernel records a %globaltimer stamp at
 * every phase's MMA issue / epilogue start / epilogue end (tools/fusion_phase_profile.py).  NULL turns it off. */
void cm_debug_fusion_phase_stamps(long long* device_buf);

/* Fused QK^T -> softmax -> PV on the tcgen05 tensor cores for heads of 32 dims (the cross-attention of A10 when the
 * knowledge graph is large; cm_fusion_forward takes this path by itself for >= 128 shared KG rows when no attention
 * maps are requested).  q [mq, hidden], k [mk, hidden] row-major fp16; vt [hidden, ldv] = V transposed (ldv % 8 == 0,
 * columns >= mk finite); hidden % 64 == 0; heads = hidden / 32.  seg_ptr (device, n_seg + 1 ints) gives the key range
 * of each segment -- one set of output rows per segment: out [n_seg * mq, ldo] fp16 -- or NULL for one segment [0, mk).
 * scale multiplies the scores (1/sqrt(32) for nn.MultiheadAttention). */
int cm_flash_attention_f16(const void* q_f16, int mq, const void* k_f16, int mk, const void* vt_f16, int ldv, int hidden,
                           const int32_t* seg_ptr, int n_seg, float scale, void* out_f16, int ldo, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Training-mode CrossAttentionFusion + the four detector heads: forward under model.train() and the backward
 * pass (fusion_model.py:75-146, 229-246; the step is models/multimodal/train_multimodal.py:232-275).
 * Ragged batch as in cm_fusion_forward; kg is [n_graphs, n_kg, kg_dim] -- one copy of the KG rows per image,
 * which is what the reference's expand() feeds the model, so gradients that reach the KG side sum over images.
 * Dropout (nn.Dropout / attention dropout, all with the model's `dropout`) through caller-supplied keep masks:
 *   keep[0] rg->kg attention probabilities [M, heads, n_kg]      keep[1] ffn_rg hidden [M, 2*hidden]
 *   keep[2] kg->rg attention probabilities [heads, n_kg, M]      keep[3] ffn_kg hidden [n_graphs*n_kg, 2*hidden]
 *   keep[4] fusion_layer hidden [n_graphs, hidden]               keep[5..8] mask / instance / edge / score head hidden
 * NULL = keep all.  d_rg [M, rg_dim] and d_kg [n_graphs*n_kg, kg_dim] (optional) receive the input gradients, so the
 * node embeddings can be trained through (cm_gnn_backward's d_node_emb).
 * ------------------------------------------------------------------------------------------------ */
typedef struct cm_fusion_dropout {
    const uint8_t* keep[9];
    float p;
} cm_fusion_dropout;
typedef struct cm_fusion_grads { /* one buffer per trainable tensor of cm_fusion_params (cross attention), overwritten */
    float *rg_proj_w, *rg_proj_b, *kg_proj_w, *kg_proj_b;
    float *r2k_in_w, *r2k_in_b, *r2k_out_w, *r2k_out_b;
    float *k2r_in_w, *k2r_in_b, *k2r_out_w, *k2r_out_b;
    float *ln_rg_w, *ln_rg_b, *ln_kg_w, *ln_kg_b;
    float *ffn_rg_w1, *ffn_rg_b1, *ffn_rg_w2, *ffn_rg_b2;
    float *ffn_kg_w1, *ffn_kg_b1, *ffn_kg_w2, *ffn_kg_b2;
    float *fl_w1, *fl_b1, *fl_w2, *fl_b2;
    float *head_w1[4], *head_b1[4], *head_w2[4], *head_b2[4];   /* NULL-able together with cm_fusion_params' heads */
    float *late_w[3], *late_b[3];                                /* fusion_type 1 (LateFusion) */
} cm_fusion_grads;
size_t cm_fusion_train_workspace_bytes(const cm_fusion_params* p, int n_graphs, int n_rows, int n_kg);
int cm_fusion_train_forward(const cm_fusion_params* p, const float* rg, const int32_t* node_ptr, int n_graphs, int n_rows,
                            const float* kg, int n_kg, const cm_fusion_dropout* drop, void* workspace, size_t workspace_bytes,
                            float* fused, float* mask_out, float* instance_out, float* edge_out, float* score_out, void* stream);
int cm_fusion_backward(const cm_fusion_params* p, const float* rg, const int32_t* node_ptr, int n_graphs, int n_rows,
                       const float* kg, int n_kg, const cm_fusion_dropout* drop, void* workspace, size_t workspace_bytes,
                       const float* d_fused, const float* d_mask_out, const float* d_instance_out, const float* d_edge_out,
                       const float* d_score_out, const cm_fusion_grads* grads, float* d_rg, float* d_kg, void* stream);

/* ------------------------------------------------------------------------------------------------
 * the Canny edge map behind region feature 13 -- skimage.feature.canny(gray, sigma=2) with
 * gray = image . [0.2989, 0.5870, 0.1140] (extract_rg_embeddings.py:151-152): Gaussian (mode 'constant', bleed-over
 * corrected), scipy Sobel (mode 'reflect'), bilinear non-maximum suppression, hysteresis low/high on the float
 * magnitude (skimage defaults 0.1 / 0.2), all in float64 in scipy's operation order.  Parity unpinned (scikit-image
 * is not vendored by the reference): checked against the numpy/scipy restatement in oracle/canny.py.
 * image [B,H,W,3] float32 in [0,1] (image_is_u8 = 0) or uint8 (value = byte/255); edges [B,H,W] uint8 out (0/1).
 * ------------------------------------------------------------------------------------------------ */
size_t cm_canny_workspace_bytes(int batch, int h, int w);
int cm_canny(const void* image, int image_is_u8, int batch, int h, int w, double sigma, double low_threshold,
             double high_threshold, void* workspace, size_t workspace_bytes, uint8_t* edges, void* stream);

/* ------------------------------------------------------------------------------------------------
 * A8 + A9 (+A13): pixel-mask paint and the metric reductions.
 * Paint: models/region_graph/test.py:241-244 -- pixel with label l gets prob[l] when 0 <= l < N_b (the
 * reference indexes by RAW label); with id_of_label != NULL it gets prob[id_of_label[l]] instead
 * (label_offset_fix).  Metrics: utils/metrics.py:9-76 per image with strict `>` thresholds.
 * gt [B,H,W] uint8 (value/255.0 = the PIL 'L' convention, test.py:275).
 * counts [B,4] int64 out: tp, sum_pred, sum_gt, n_equal; abs_err [B] float64 out: sum |pred - gt|.
 * pred_mask [B,H,W] float32 out or NULL.
 * ------------------------------------------------------------------------------------------------ */
int cm_paint_metrics(const int32_t* labels, const uint8_t* gt, int batch, int h, int w,
                     const float* mask_prob, const int32_t* node_ptr, const int32_t* id_of_label,
                     int label_cap, float threshold, float* pred_mask, int64_t* counts, double* abs_err,
                     void* stream);
/* Binary masks that crossed PCIe one BIT per pixel (the Canny map and a {0,255} ground-truth mask are 2 of the 7 bytes
 * per pixel of the host -> device transfer; packed they are 0.25): bit i of byte j (little-endian bit order,
 * numpy.packbits(bitorder="little")) -> out[8 j + i] = one_value or 0.  out 8-byte aligned, 8 * n_bytes bytes. */
int cm_unpack_bits(const uint8_t* bits, int64_t n_bytes, int one_value, uint8_t* out, void* stream);
/* Label maps over PCIe at 10 bits per pixel (the hand-over of the host front end; skimage labels at
 * n_segments = 500 stay below 1024): packed = little-endian bit stream, label i in bits [10 i, 10 i + 10) (4 labels in 5 bytes;
 * pipeline.pack_labels10 on the host), n_labels a multiple of 16, out [n_labels] int16 (16-byte aligned). */
int cm_unpack_labels10(const uint8_t* packed, int64_t n_labels, int16_t* out, void* stream);
/* Host -> device staging of a batch by a kernel instead of a copy engine (the front end's hand-over; the
 * reference's `.to(device)`, models/multimodal/test_multimodal.py:96-99): host_src is PINNED host memory (cudaHostAlloc /
 * torch pin_memory: device-addressable under unified addressing), n_bytes a multiple of 16, both pointers 16-byte aligned;
 * `ctas` blocks of 256 threads (<= 0: 32) read it with 16-byte loads.  Being a kernel it can sit inside a captured step. */
int cm_ingest_copy(const void* host_src, void* dst, int64_t n_bytes, int ctas, void* stream);
/* Same reductions for a caller-supplied float prediction and float ground truth (the literal signature of
 * utils/metrics.py:evaluate_segmentation: two arrays, one threshold). */
int cm_metrics_dense(const float* pred, const float* gt, int batch, int64_t pixels, float threshold,
                     int64_t* counts, double* abs_err, void* stream);
/* per_image [B,7] float64 out: iou, dice, precision, recall, f1, mae, accuracy (utils/metrics.py:57-76);
 * partial [15] float64 out: n, sum m_k (7), sum m_k^2 (7) -- the vector that is all-reduced across ranks
 * and from which batch_evaluate's mean / population std follow (utils/metrics.py:78-92). */
int cm_metrics_finalize(const int64_t* counts, const double* abs_err, int batch, int64_t pixels,
                        double* per_image, double* partial, void* stream);

/* ------------------------------------------------------------------------------------------------
 * A9': S-measure (Fan et al. ICCV'17), E-measure (Fan et al. IJCAI'18), F-measure (beta^2 = 0.3) and weighted
 * F-measure (Margolin et al. CVPR'14).  NOT IN THE REFERENCE (utils/metrics.py:9-92 has none of them; BASELINE.json's
 * north_star names them): the published definitions, restated in oracle/sod_metrics.py -- parity unpinned.
 * pred [B,H,W] float32 in [0,1] (e.g. cm_paint_metrics' pred_mask), gt [B,H,W] uint8 (foreground = byte >= 128).
 * per_image [B,8] float64 out: s_measure, e_adaptive, e_mean, e_max, f_adaptive, f_mean, f_max, wf_measure
 * (adaptive threshold min(2 mean(pred), 1); mean / max over the image's own 256-threshold curve).
 * partial [528] float64 out: sum of the 8 keys, sum of their squares, then the batch sums of the F curve [256] and the
 * E curve [256] (entry i = threshold floor(255 pred) >= 255 - i) -- the vector to all-reduce next to
 * cm_metrics_finalize's: dataset-level max / mean F and E are taken over the dataset-AVERAGED curves.
 * ------------------------------------------------------------------------------------------------ */
size_t cm_sod_workspace_bytes(int batch, int h, int w);
/* with_weighted_f = 0 skips the weighted F-measure (it needs an exact Euclidean feature transform of the GT and a 7x7
 * filter per image -- several times the cost of everything else here); its slot in per_image / partial is then NaN. */
int cm_sod_metrics(const float* pred, const uint8_t* gt, int batch, int h, int w, int with_weighted_f, void* workspace,
                   size_t workspace_bytes, double* per_image, double* partial, void* stream);
/* No counterpart in the reference: added for callers that cannot allocate device memory
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  cm_sod_metrics without the
 * weighted F-measure (its slot is NaN) and without a workspace argument: the scratch is a 2 MiB array built into the library
 * (cm_sod_metrics_static_max_bytes; every GPU that loads the library holds it until exit; ~170 images of 256 x 256 per
 * call).  A batch that needs more is rejected with CM_ERR_WORKSPACE.  The call allocates nothing (no cudaMalloc / cudaFree /
 * allocating library routine).  Every call on a device shares the array: calls that overlap on the GPU (different streams
 * without ordering) race -- serialise them.  Inputs / outputs as in cm_sod_metrics. */
size_t cm_sod_metrics_static_max_bytes(void);
int cm_sod_metrics_static(const float* pred, const uint8_t* gt, int batch, int h, int w, double* per_image, double* partial,
                          void* stream);
/* The same on the PAINTED prediction of cm_paint_metrics (pixel with label l gets mask_prob[l], or
 * mask_prob[id_of_label[l]] with the offset fix; models/region_graph/test.py:241-244) without materialising the mask:
 * labels [B,H,W] int32 (label_bytes 4) or int16 (2); the other arguments as in cm_paint_metrics. */
int cm_sod_metrics_painted(const void* labels, int label_bytes, const uint8_t* gt, int batch, int h, int w, const float* mask_prob,
                           const int32_t* node_ptr, const int32_t* id_of_label, int label_cap, int with_weighted_f, void* workspace,
                           size_t workspace_bytes, double* per_image, double* partial, void* stream);
/* A8 + A9 + A9' on a painted mask WITHOUT a pixel pass behind the model.  A painted prediction is constant over a region, so
 * every pixel sum of utils/metrics.py:9-92 and of the S- / E- / F-measures factors through the contingency table
 * count[label][gt class][S-measure quadrant]; the quadrants are cut at the ground truth's centroid, so the table depends on
 * labels + gt only.
 *   cm_region_table_build    labels [B,H,W] int32 / int16 (label_bytes 4 / 2), gt [B,H,W] uint8, or bit-packed [B,H*W/8] with
 *                            gt_packed != 0 (bit i of byte j = pixel 8 j + i, a set bit = 255): builds the table in `workspace`.
 *                            Needs nothing from the model: run it next to the graph build.
 *   cm_region_table_metrics  the same workspace + the per-node probabilities: counts [B,4] / abs_err [B] / per_image [B,7] /
 *                            partial [15] exactly as cm_paint_metrics + cm_metrics_finalize give them, per_image_sod [B,8] /
 *                            partial_sod [528] as cm_sod_metrics_painted (weighted F = NaN).  gt_bytes: the byte ground truth
 *                            again, or NULL when it was bit-packed; |pred - gt| only factors for gt in {0, 255}, so for an
 *                            image where the build found other bytes that image's block walks its pixels for the A9 counts / MAE
 *                            the way cm_paint_metrics does (decided on the device, no synchronisation).  threshold in [0, 1).  label_cap <= 6000; labels outside [0, label_cap)
 *                            paint probability 0 as in cm_paint_metrics. */
size_t cm_region_table_workspace_bytes(int batch, int h, int w, int label_cap);
int cm_region_table_build(const void* labels, int label_bytes, const uint8_t* gt, int gt_packed, int batch, int h, int w, int label_cap,
                          void* workspace, size_t workspace_bytes, void* stream);
int cm_region_table_metrics(const void* labels, int label_bytes, const uint8_t* gt_bytes, int batch, int h, int w, const float* mask_prob,
                            const int32_t* node_ptr, const int32_t* id_of_label, int label_cap, float threshold, void* workspace,
                            size_t workspace_bytes, int64_t* counts, double* abs_err, double* per_image, double* partial,
                            double* per_image_sod, double* partial_sod, void* stream);
/* A13: image-level confusion counts tp, fp, fn, tn (models/multimodal/train_multimodal.py:197-220) from
 * predicted class = argmax(mask_out) and true class ids.  counts4 [4] int64 accumulated (caller zeroes). */
int cm_class_confusion(const float* mask_out, const int64_t* true_cls, int n, int num_classes,
                       int64_t* counts4, void* stream);

/* ------------------------------------------------------------------------------------------------
 * the training-loop pieces around the model passes (device-side, no host synchronisation).
 * ------------------------------------------------------------------------------------------------ */
/* Per-region ground-truth targets (models/region_graph/train.py:169-176): the reference thresholds the mean of the
 * ToTensor'ed masks (byte/255) over the region's pixels at 0.5 / 0.5 / 0.3; with byte masks that is the integer
 * test 2*S > 255*n (10*S > 765*n for the edge mask).  labels [B,H,W] int32, masks [B,H,W] uint8, node_* from
 * cm_rag_build.  y / instance_y [node_cap] int64 (what nn.CrossEntropyLoss takes), edge_y [node_cap] float32. */
int cm_region_targets(const int32_t* labels, const uint8_t* mask, const uint8_t* instance_mask, const uint8_t* edge_mask,
                      int batch, int h, int w, const int32_t* node_ptr, const int32_t* node_label, const int32_t* node_bbox,
                      int node_cap, int64_t* y, int64_t* instance_y, float* edge_y, void* stream);

/* Region-node multi-task loss (train.py:325-353): CrossEntropyLoss(weight=[1, mask_pos_weight]) * task_mask +
 * CrossEntropyLoss(weight=[1, instance_pos_weight]) * task_instance + BCEWithLogitsLoss(pos_weight) * task_edge.
 * Logits [node_cap, 2] / [node_cap, 2] / [node_cap]; *n_nodes_dev live rows.  loss_out [4] = total, mask, instance,
 * edge terms; d_* (optional) = d total / d logits (rows past the live count zeroed).  workspace: 40 bytes. */
typedef struct cm_rg_loss_config {
    float mask_pos_weight, instance_pos_weight, edge_pos_weight;   /* 5, 4, 3 in the reference */
    float task_mask, task_instance, task_edge;                     /* 2, 1, 0.5 */
} cm_rg_loss_config;
int cm_rg_node_loss(const float* mask_logits, const float* instance_logits, const float* edge_logits, const int64_t* y,
                    const int64_t* instance_y, const float* edge_y, const int32_t* n_nodes_dev, int node_cap,
                    const cm_rg_loss_config* cfg, void* workspace, size_t workspace_bytes, float* loss_out, float* d_mask,
                    float* d_instance, float* d_edge, void* stream);

/* Image-level loss of the multimodal detector (models/multimodal/train_multimodal.py:29-57, 253-268), summed over the
 * images of the batch (the reference calls backward once per sample):
 *   w_focal * AggressiveFocalLoss(alpha, gamma)(mask, label) + w_instance * CE(instance, label)
 *   + w_edge * BCEWithLogits(edge, edge_label) + w_score * (score - score_label)^2
 * mask / instance logits [G,2], edge logit [G], score [G] (after the sigmoid).  loss_out [1]; d_* optional. */
typedef struct cm_image_loss_config {
    float alpha, gamma;                            /* 0.75, 3.0 */
    float w_focal, w_instance, w_edge, w_score;    /* 3, 1, 0.5, 0.3 */
} cm_image_loss_config;
int cm_image_loss(const float* mask_logits, const float* instance_logits, const float* edge_logits, const float* score,
                  const int64_t* mask_label, const float* edge_label, const float* score_label, int n_graphs,
                  const cm_image_loss_config* cfg, float* loss_out, float* d_mask, float* d_instance, float* d_edge,
                  float* d_score, void* stream);

/* torch.nn.utils.clip_grad_norm_(params, max_norm) followed by torch.optim.AdamW.step() (train.py:318,356-357;
 * train_multimodal.py:276-277) over a table of tensors in two launches + one tiny one: params / grads / exp_avg /
 * exp_avg_sq are DEVICE arrays of n_tensors device pointers, numel the element counts (device), max_numel the
 * largest of them (host).  The gradients are scaled in place by min(1, max_norm / (total_norm + 1e-6)) (max_norm <= 0:
 * no clipping); norm_out [2] = total gradient norm before clipping, clip coefficient.  step counts from 1.
 * norm_workspace: n_tensors doubles. */
typedef struct cm_adamw_config {
    float lr, beta1, beta2, eps, weight_decay, max_norm;
} cm_adamw_config;
int cm_adamw_step(float* const* params, const float* const* grads, float* const* exp_avg, float* const* exp_avg_sq,
                  const int64_t* numel, int n_tensors, int64_t max_numel, const cm_adamw_config* cfg, int step,
                  double* norm_workspace, float* norm_out, void* stream);

/* ------------------------------------------------------------------------------------------------
 * SLIC superpixels -- skimage.segmentation.slic(img_u8, n_segments=500, compactness=10, sigma=1) as
 * called at models/region_graph/extract_rg_embeddings.py:143-144 (img_as_float -> Gaussian (reflect) -> rgb2lab ->
 * max_num_iter k-means rounds on the regular seed grid -> connectivity enforcement, labels from 1; label 0 can appear
 * where a first small segment has no labelled neighbour yet, as in skimage).  float64, sequential summation order.
 * image_u8 [B,H,W,3]; labels [B,H,W] int32 out.
 * ------------------------------------------------------------------------------------------------ */
size_t cm_slic_workspace_bytes(int batch, int h, int w, int n_segments);
int cm_slic(const uint8_t* image_u8, int batch, int h, int w, int n_segments, double compactness, double sigma,
            int max_num_iter, void* workspace, size_t workspace_bytes, int32_t* labels, void* stream);
/* The last stage on its own (skimage _enforce_label_connectivity_cython, 4-connected raster-order BFS): segments
 * smaller than min_size join the last labelled neighbour met, segments are cut at max_size pixels.
 * coords_workspace: batch * max_size int32. */
int cm_slic_enforce_connectivity(const int32_t* labels_in, int batch, int h, int w, int min_size, int max_size,
                                 int start_label, int32_t* coords_workspace, int32_t* labels_out, void* stream);

/* ------------------------------------------------------------------------------------------------
 * int16 label maps end to end (SLIC labels are < 32768; the host sends 2 bytes per pixel): the same three entry
 * points reading int16 directly, so no widened copy is written or re-read.  Semantics are those of cm_rag_build,
 * cm_node_features_u8 (region kernel: node_bbox required) and cm_paint_metrics.  cm_rag_build_i16 needs W % 8 == 0 and a
 * 16-byte aligned label buffer; otherwise widen with cm_labels_widen_i16 and use the int32 entry points.
 * ------------------------------------------------------------------------------------------------ */
int cm_rag_build_i16(const int16_t* labels, int batch, int h, int w, int label_cap, void* workspace, size_t workspace_bytes,
                     int32_t* id_of_label, int32_t* node_ptr, int32_t* edge_ptr, int32_t* node_label, int32_t* node_bbox,
                     int64_t* batch_vec, int32_t* csr_row_ptr, int32_t* csr_col, int64_t* edge_index, int local_ids,
                     int node_cap, int edge_cap, uint32_t* status, void* stream);
int cm_node_features_u8_i16(const uint8_t* image_u8, const int16_t* labels, const uint8_t* canny, int batch, int h, int w,
                            int label_cap, const int32_t* node_ptr, const int32_t* node_label, const int32_t* node_bbox,
                            float norm_side, float* x, int node_cap, void* stream);
int cm_paint_metrics_i16(const int16_t* labels, const uint8_t* gt, int batch, int h, int w, const float* mask_prob,
                         const int32_t* node_ptr, const int32_t* id_of_label, int label_cap, float threshold,
                         float* pred_mask, int64_t* counts, double* abs_err, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* CM_B200_H */
