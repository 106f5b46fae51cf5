// A10 + A11 on the tensor cores: CrossAttentionFusion + detector heads for a knowledge graph that is SHARED by
// every image of the batch ("KG embeddings broadcast once", the reference's use: the same 13 category rows for
// every sample, models/multimodal/train_multimodal.py / test_multimodal.py).
//
// With the KG rows fixed, everything on the KG side of both nn.MultiheadAttention blocks is a constant of the
// (weights, KG) pair and is folded once by cm_fusion_fold (fusion_model.py:108-131 restated):
//   rg -> kg   scores1[n,(h,k)] = q1[n,h,:] . k1[k,h,:] = r[n,:] . U1[(h,k),:] + c1[(h,k)]
//              U1[(h,k),c] = scale * sum_d Wq1[h*hd+d, c] k1[k,h*hd+d],  c1 = scale * bq1_h . k1[k,h,:]
//              attn_out1[n,:] = sum_(h,k) P1[n,(h,k)] VO[(h,k),:] + bo1,  VO[(h,k),c] = sum_d Wo1[c,h*hd+d] v1[k,h*hd+d]
//   kg -> rg   scores2[(h,k),n] = q2[k,h,:] . k2[n,h,:] = r[n,:] . U2[(h,k),:] (+ a constant in n, which the
//              softmax over n cancels),  U2[(h,k),c] = sum_d Wk2[h*hd+d, c] q2[k,h*hd+d]
// so per region node the two Q/K projections, QK^T of both directions, PV and the rg->kg out-projection collapse
// into two tensor-core GEMMs against [D -> 256] and [128 -> D] folded matrices.  (h,k) lives at column h*GS + k
// of a 128-column block, GS = 16 or 32 slots per head, so the fold needs heads * GS <= 128 and n_kg <= GS; other
// shapes (per-image KG, large KG vocabularies, late fusion) take the general path in fusion.cu.
#include <math.h>
#include <stdlib.h>

#include "dense_tc.cuh"
#include "fusion_kg.cuh"
#include "fusion_node.cuh"
#include "fusion_tc.cuh"
#include "linear.cuh"

namespace cm {

static size_t al(size_t v) { return (v + 255) / 256 * 256; }
long long* g_fusion_phase_dbg = nullptr;   // profiling aid (cm_debug_fusion_phase_stamps): 40 + 2 x grid device int64 or NULL

// shapes the single-kernel node side is written for (the reference's configs/multimodal_config.yaml defaults)
static bool fusion_node_fused_ok(const cm_fusion_params* p) { return p->hidden == 256 && p->rg_dim == 128 && p->rg_proj_w != nullptr; }

bool fusion_fold_supported(const cm_fusion_params* p, int n_kg) {
    if (p->fusion_type != 0) return false;
    const int D = p->hidden, H = p->heads;
    if (D % 32 != 0 || D > 256 || D < 32 || p->rg_dim % 8 != 0 || H < 1 || D % H != 0) return false;
    const int gs = n_kg <= 16 ? 16 : 32;
    return n_kg <= gs && H * gs <= 128;
}

FusionFolded fusion_folded_layout(const cm_fusion_params* p, int n_kg) {
    FusionFolded f;
    const size_t D = p->hidden, K = n_kg, dr = p->rg_dim;
    size_t o = 0;
    auto take = [&](size_t bytes) {
        size_t r = o;
        o += al(bytes);
        return r;
    };
    f.gs = n_kg <= 16 ? 16 : 32;
    f.n_kg = n_kg;
    f.w_rgproj = take(D * dr * 2);
    f.wcat = take(256 * D * 2);
    f.wv2 = take(D * D * 2);
    f.vot = take(D * 128 * 2);
    f.ffn1 = take(2 * D * D * 2);
    f.ffn2 = take(2 * D * D * 2);
    // KG-row chain (out-projection, FFN) on the tensor cores; the graph-level layers read the fp32 parameters as stored
    f.wo2_h = take(D * D * 2);
    f.kffn1_h = take(2 * D * D * 2);
    f.kffn2_h = take(2 * D * D * 2);
    f.ccat = take(256 * 4);
    f.kp = take(K * D * 4);
    f.k1v1 = take(K * 2 * D * 4);
    f.q2 = take(K * D * 4);
    f.total = o;
    return f;
}

__global__ void cast_weight_kernel(const float* __restrict__ in, op_t* __restrict__ out, size_t n) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        out[i] = f2op(in[i]);
}

static int cast_weight(const float* in, void* out, size_t n, cudaStream_t st) {
    {
        CM_TIMED("fusion_fold", st);
        cast_weight_kernel<<<ceil_div((long long)n, 256 * 4), 256, 0, st>>>(in, reinterpret_cast<op_t*>(out), n);
    }
    return post_launch("fusion_fold(cast)");
}

// One thread per folded element: wcat rows [0,128) = U1, [128,256) = U2 (slot h*gs+k, padding rows 0), ccat,
// vot [D,128] (VO transposed so it is an nn.Linear-style [out, in] weight for the P1 x VO product).
__global__ void fusion_fold_kernel(cm_fusion_params p, int n_kg, int gs, const float* __restrict__ k1v1, const float* __restrict__ q2,
                                   op_t* __restrict__ wcat, float* __restrict__ ccat, op_t* __restrict__ vot) {
    const int D = p.hidden, H = p.heads, hd = D / H;
    const float scale = 1.0f / sqrtf((float)hd);
    const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x, stride = (long long)gridDim.x * blockDim.x;
    for (long long i = tid; i < 256LL * D; i += stride) {
        const int row = (int)(i / D), c = (int)(i % D), slot = row & 127, h = slot / gs, k = slot % gs;
        float s = 0.f;
        if (h < H && k < n_kg) {
            if (row < 128) {
                for (int d = 0; d < hd; ++d) s = fmaf(p.r2k_in_w[(size_t)(h * hd + d) * D + c], k1v1[(size_t)k * 2 * D + h * hd + d], s);
                s *= scale;
            } else {
                for (int d = 0; d < hd; ++d) s = fmaf(p.k2r_in_w[(size_t)(D + h * hd + d) * D + c], q2[(size_t)k * D + h * hd + d], s);
            }
        }
        wcat[i] = f2op(s);
    }
    for (long long i = tid; i < 256; i += stride) {
        const int slot = (int)i & 127, h = slot / gs, k = slot % gs;
        float s = 0.f;
        if (i < 128 && h < H && k < n_kg) {
            for (int d = 0; d < hd; ++d) s = fmaf(p.r2k_in_b[h * hd + d], k1v1[(size_t)k * 2 * D + h * hd + d], s);
            s *= scale;
        }
        ccat[i] = s;
    }
    for (long long i = tid; i < 128LL * D; i += stride) {
        const int c = (int)(i / 128), slot = (int)(i % 128), h = slot / gs, k = slot % gs;
        float s = 0.f;
        if (h < H && k < n_kg)
            for (int d = 0; d < hd; ++d) s = fmaf(p.r2k_out_w[(size_t)c * D + h * hd + d], k1v1[(size_t)k * 2 * D + D + h * hd + d], s);
        vot[i] = f2op(s);
    }
}

int fusion_fold(const cm_fusion_params* p, const float* kg, int n_kg, void* folded, cudaStream_t st) {
    const FusionFolded f = fusion_folded_layout(p, n_kg);
    char* base = reinterpret_cast<char*>(folded);
    const int D = p->hidden, K = n_kg;
    int rc;
    float* kp = reinterpret_cast<float*>(base + f.kp);
    float* k1v1 = reinterpret_cast<float*>(base + f.k1v1);
    float* q2 = reinterpret_cast<float*>(base + f.q2);
    const float scale = 1.0f / sqrtf((float)(D / p->heads));
    Linear l;
    if (p->kg_proj_w) {
        l.M = K; l.A = kg; l.lda = p->kg_dim; l.K = p->kg_dim; l.W = p->kg_proj_w; l.bias = p->kg_proj_b; l.N = D; l.C = kp; l.ldc = D;
        if ((rc = launch_linear(l, st)) != CM_OK) return rc;
    } else {
        cudaMemcpyAsync(kp, kg, sizeof(float) * K * D, cudaMemcpyDeviceToDevice, st);
    }
    l = Linear(); l.M = K; l.A = kp; l.lda = D; l.K = D; l.W = p->r2k_in_w + (size_t)D * D; l.bias = p->r2k_in_b + D; l.N = 2 * D; l.C = k1v1; l.ldc = 2 * D;
    if ((rc = launch_linear(l, st)) != CM_OK) return rc;
    l = Linear(); l.M = K; l.A = kp; l.lda = D; l.K = D; l.W = p->k2r_in_w; l.bias = p->k2r_in_b; l.N = D; l.C = q2; l.ldc = D; l.alpha = scale;
    if ((rc = launch_linear(l, st)) != CM_OK) return rc;
    {
        CM_TIMED("fusion_fold", st);
        fusion_fold_kernel<<<256, 256, 0, st>>>(*p, K, f.gs, k1v1, q2, reinterpret_cast<op_t*>(base + f.wcat),
                                                reinterpret_cast<float*>(base + f.ccat), reinterpret_cast<op_t*>(base + f.vot));
    }
    if ((rc = post_launch("fusion_fold")) != CM_OK) return rc;
    if (p->rg_proj_w && (rc = cast_weight(p->rg_proj_w, base + f.w_rgproj, (size_t)D * p->rg_dim, st)) != CM_OK) return rc;
    if ((rc = cast_weight(p->k2r_in_w + (size_t)2 * D * D, base + f.wv2, (size_t)D * D, st)) != CM_OK) return rc;
    if ((rc = cast_weight(p->ffn_rg_w1, base + f.ffn1, (size_t)2 * D * D, st)) != CM_OK) return rc;
    if ((rc = cast_weight(p->ffn_rg_w2, base + f.ffn2, (size_t)2 * D * D, st)) != CM_OK) return rc;
    if ((rc = cast_weight(p->k2r_out_w, base + f.wo2_h, (size_t)D * D, st)) != CM_OK) return rc;
    if ((rc = cast_weight(p->ffn_kg_w1, base + f.kffn1_h, (size_t)2 * D * D, st)) != CM_OK) return rc;
    if ((rc = cast_weight(p->ffn_kg_w2, base + f.kffn2_h, (size_t)2 * D * D, st)) != CM_OK) return rc;
    return CM_OK;
}

struct FusionTcWs {
    op_t *rgb, *r, *p1, *v2, *y0, *hh, *ctx2, *z0, *hk;
    float *s2, *y, *part_m, *part_l, *part_acc, *part_y, *stat, *z, *both;
    size_t total;
};

static FusionTcWs carve_fusion_tc(void* ws, const cm_fusion_params* p, int G, int M, int Kk) {
    Carver c(ws);
    FusionTcWs w;
    const size_t D = p->hidden, m = M, gc = (size_t)G * KG_MAX_CHUNKS;
    const int gs = Kk <= 16 ? 16 : 32;
    w.rgb = c.take<op_t>(m * p->rg_dim);
    w.r = c.take<op_t>(m * D);
    w.p1 = c.take<op_t>(m * 128);
    w.v2 = c.take<op_t>(m * D);
    w.y0 = c.take<op_t>(m * D);
    w.hh = c.take<op_t>(m * 2 * D);
    w.s2 = c.take<float>(m * 128);
    w.y = c.take<float>(m * D);
    w.part_m = c.take<float>(gc * 128);
    w.part_l = c.take<float>(gc * 128);
    w.part_acc = c.take<float>(gc * gs * D);
    w.part_y = c.take<float>(gc * D);
    w.stat = c.take<float>((size_t)G * 256);
    const size_t gk = (size_t)G * Kk;
    w.ctx2 = c.take<op_t>(gk * D);
    w.z0 = c.take<op_t>(gk * D);
    w.hk = c.take<op_t>(gk * 2 * D);
    w.z = c.take<float>(gk * D);
    w.both = c.take<float>((size_t)G * 2 * D);
    w.total = align_up(c.off, 256);
    return w;
}

size_t fusion_tc_workspace_bytes(const cm_fusion_params* p, int G, int M, int Kk) { return carve_fusion_tc(nullptr, p, G, M, Kk).total; }

int fusion_forward_tc(const cm_fusion_params* p, const void* folded, const float* rg, const int32_t* node_ptr, int G, int M, int Kk,
                      void* workspace, float* fused_out, float* mask_out, float* instance_out, float* edge_out, float* score_out,
                      float* attn_rg2kg, float* attn_kg2rg, cudaStream_t st) {
    const FusionFolded f = fusion_folded_layout(p, Kk);
    const char* base = reinterpret_cast<const char*>(folded);
    auto W = [&](size_t off) { return reinterpret_cast<const op_t*>(base + off); };
    FusionTcWs w = carve_fusion_tc(workspace, p, G, M, Kk);
    const int D = p->hidden;
    const int* m_dev = node_ptr + G;
    int rc;
    Dense d;

    // With the single-kernel node side AND the fused KG-side tail the node kernel reduces what the KG side needs from the node
    // rows (kg->rg attention partials, column sums of y) per tile (fusion_node.cu MODE 2): S2 / V2 / y are not written and
    // kg2rg_partial does not run.  The kg->rg attention map needs S2, so asking for it keeps the unfused form.
    const bool fused_tail = D == 256 && Kk <= 16 && !getenv("CM_FUSION_UNFUSED_TAIL");
    const size_t part_slots = (size_t)G * KG_MAX_CHUNKS;
    const bool tile_partials = fusion_node_fused_ok(p) && fused_tail && p->heads == 8 && f.gs == 16 && !attn_kg2rg && !attn_rg2kg &&
                               (size_t)(M + 127) / 128 + (size_t)G + 1 <= part_slots && !getenv("CM_FUSION_NODE_UNFUSED_KG");
    if (!fusion_node_fused_ok(p) && (rc = launch_cast_f16(rg, p->rg_dim, M, p->rg_dim, m_dev, w.rgb, p->rg_dim, st)) != CM_OK) return rc;
    if (fusion_node_fused_ok(p)) {
        // the whole region-node side in one kernel (fusion_node.cu)
        FusionNodeArgs a;
        a.M = M; a.m_dev = m_dev;
        a.b_rgproj = p->rg_proj_b; a.ccat = reinterpret_cast<const float*>(base + f.ccat); a.b_v2 = p->k2r_in_b + 2 * D;
        a.b_out1 = p->r2k_out_b; a.ln_g = p->ln_rg_w; a.ln_b = p->ln_rg_b; a.b_ffn1 = p->ffn_rg_b1; a.b_ffn2 = p->ffn_rg_b2;
        a.sm_gs = f.gs; a.sm_valid = Kk;
        a.s2 = w.s2; a.v2 = w.v2; a.y = w.y; a.attn_mean = attn_rg2kg;
        a.x32 = rg;   // fp32 node embeddings are converted inside the kernel (rg_dim == 128 here)
        a.dbg = g_fusion_phase_dbg;
        if (tile_partials) {
            a.part_m = w.part_m; a.part_l = w.part_l; a.part_acc = w.part_acc; a.part_y = w.part_y;
            a.node_ptr = node_ptr; a.n_graphs = G;
        }
        if ((rc = launch_fusion_node(a, w.rgb, W(f.w_rgproj), W(f.wcat), W(f.wv2), W(f.vot), W(f.ffn1), W(f.ffn2), st)) != CM_OK) return rc;
    } else {
    const op_t* r = w.rgb;
    if (p->rg_proj_w) {   // fusion_model.py:108
        d = Dense(); d.A = w.rgb; d.lda = p->rg_dim; d.W = W(f.w_rgproj); d.M = M; d.N = D; d.K = p->rg_dim; d.m_dev = m_dev;
        d.bias = p->rg_proj_b; d.out_h = w.r; d.ldo_h = D;
        if ((rc = launch_dense(d, st)) != CM_OK) return rc;
        r = w.r;
    }
    // scores of both directions in one pass over r: P1 = softmax_k(r U1^T + c1), S2 = r U2^T
    d = Dense(); d.A = r; d.lda = D; d.W = W(f.wcat); d.M = M; d.N = 256; d.K = D; d.m_dev = m_dev;
    d.bias = reinterpret_cast<const float*>(base + f.ccat); d.out_h = w.p1; d.ldo_h = 128; d.out_f32 = w.s2; d.ldo_f32 = 128;
    d.sm_gs = f.gs; d.sm_valid = Kk; d.attn_mean = attn_rg2kg;
    if ((rc = launch_dense(d, st)) != CM_OK) return rc;
    // V of the kg -> rg direction (:123-129)
    d = Dense(); d.A = r; d.lda = D; d.W = W(f.wv2); d.M = M; d.N = D; d.K = D; d.m_dev = m_dev;
    d.bias = p->k2r_in_b + 2 * D; d.out_h = w.v2; d.ldo_h = D;
    if ((rc = launch_dense(d, st)) != CM_OK) return rc;
    // y0 = LN(r + P1 VO + bo1) (:119)
    d = Dense(); d.A = w.p1; d.lda = 128; d.W = W(f.vot); d.M = M; d.N = D; d.K = 128; d.m_dev = m_dev;
    d.bias = p->r2k_out_b; d.residual = r; d.ldr = D; d.ln_gamma = p->ln_rg_w; d.ln_beta = p->ln_rg_b; d.out_h = w.y0; d.ldo_h = D;
    if ((rc = launch_dense(d, st)) != CM_OK) return rc;
    // y = y0 + FFN(y0) (:120), mean over the image's nodes (:134)
    d = Dense(); d.A = w.y0; d.lda = D; d.W = W(f.ffn1); d.M = M; d.N = 2 * D; d.K = D; d.m_dev = m_dev;
    d.bias = p->ffn_rg_b1; d.relu = 1; d.out_h = w.hh; d.ldo_h = 2 * D;
    if ((rc = launch_dense(d, st)) != CM_OK) return rc;
    d = Dense(); d.A = w.hh; d.lda = 2 * D; d.W = W(f.ffn2); d.M = M; d.N = D; d.K = 2 * D; d.m_dev = m_dev;
    d.bias = p->ffn_rg_b2; d.residual = w.y0; d.ldr = D; d.out_f32 = w.y; d.ldo_f32 = D;
    if ((rc = launch_dense(d, st)) != CM_OK) return rc;
    }
    // ---- kg attends to rg (:123-131): chunk partials over the nodes, merged per image into ctx2 and the node mean (:134)
    KgSideArgs ks;
    ks.s2 = w.s2; ks.v2 = w.v2; ks.y = w.y; ks.part_m = w.part_m; ks.part_l = w.part_l; ks.part_acc = w.part_acc; ks.part_y = w.part_y;
    ks.attn_kg2rg = attn_kg2rg; ks.node_cap = M;
    if (!tile_partials && (rc = launch_kg2rg_partial(ks, D, p->heads, f.gs, node_ptr, G, st)) != CM_OK) return rc;
    // ---- mean over kg rows (:135), fusion layer (:138-139), heads (:239-242)
    GraphTailArgs gt;
    gt.D = D; gt.n_kg = Kk; gt.num_classes = p->num_classes;
    gt.want_heads = (mask_out || instance_out || edge_out || score_out) ? 1 : 0;
    gt.z = w.z; gt.both = w.both;
    gt.fl1_w = p->fl_w1; gt.fl1_b = p->fl_b1; gt.fl2_w = p->fl_w2; gt.fl2_b = p->fl_b2;
    for (int t = 0; t < 4; ++t) {
        gt.hw1[t] = p->head_w1[t]; gt.hb1[t] = p->head_b1[t]; gt.hw2[t] = p->head_w2[t]; gt.hb2[t] = p->head_b2[t];
        if (gt.want_heads) CM_REQUIRE(p->head_w1[t] && p->head_b1[t] && p->head_w2[t] && p->head_b2[t], "cm_fusion_forward: head %d parameters missing", t);
    }
    gt.fused = fused_out; gt.mask_out = mask_out; gt.inst_out = instance_out; gt.edge_out = edge_out; gt.score_out = score_out;
    const float* kp = reinterpret_cast<const float*>(base + f.kp);
    if (fused_tail) {
        // one CTA per image: merge + KG-row chain (mma.sync 16-row tiles) + fusion MLP + heads (fusion_kg.cu)
        KgTailArgs kt;
        kt.part_m = w.part_m; kt.part_l = w.part_l; kt.part_acc = w.part_acc; kt.part_y = w.part_y; kt.node_ptr = node_ptr;
        kt.tile_partials = tile_partials ? 1 : 0;
        kt.v2_bias = tile_partials ? p->k2r_in_b + 2 * D : nullptr;
        kt.n_kg = Kk; kt.gs = f.gs; kt.heads = p->heads;
        kt.wo2 = W(f.wo2_h); kt.w1 = W(f.kffn1_h); kt.w2 = W(f.kffn2_h);
        kt.kp = kp; kt.bo2 = p->k2r_out_b; kt.ln_g = p->ln_kg_w; kt.ln_b = p->ln_kg_b; kt.b1 = p->ffn_kg_b1; kt.b2 = p->ffn_kg_b2;
        kt.stat = w.stat; kt.gt = gt;
        if ((rc = launch_kg_tail(kt, ks, G, st)) != CM_OK) return rc;
        return post_launch("cm_fusion_forward(tc)");
    }
    if ((rc = launch_kg_merge(ks, D, p->heads, f.gs, Kk, node_ptr, G, w.ctx2, w.both, w.stat, st)) != CM_OK) return rc;
    // ---- KG rows: z0 = LN(kp + ctx2 Wo^T + bo) (:130), z = z0 + FFN(z0) (:131)
    const int GK = G * Kk;
    if (D == 256) {
        FusionNodeArgs a;
        a.M = GK; a.m_dev = nullptr;
        a.b_out1 = p->k2r_out_b; a.ln_g = p->ln_kg_w; a.ln_b = p->ln_kg_b; a.b_ffn1 = p->ffn_kg_b1; a.b_ffn2 = p->ffn_kg_b2;
        a.res_rows = kp; a.res_mod = Kk; a.y = w.z;
        if ((rc = launch_fusion_kgrows(a, w.ctx2, W(f.wo2_h), W(f.kffn1_h), W(f.kffn2_h), st)) != CM_OK) return rc;
    } else {
        d = Dense(); d.A = w.ctx2; d.lda = D; d.W = W(f.wo2_h); d.M = GK; d.N = D; d.K = D;
        d.bias = p->k2r_out_b; d.residual_f32 = kp; d.ldr = D; d.res_mod = Kk; d.ln_gamma = p->ln_kg_w; d.ln_beta = p->ln_kg_b;
        d.out_h = w.z0; d.ldo_h = D;
        if ((rc = launch_dense(d, st)) != CM_OK) return rc;
        d = Dense(); d.A = w.z0; d.lda = D; d.W = W(f.kffn1_h); d.M = GK; d.N = 2 * D; d.K = D;
        d.bias = p->ffn_kg_b1; d.relu = 1; d.out_h = w.hk; d.ldo_h = 2 * D;
        if ((rc = launch_dense(d, st)) != CM_OK) return rc;
        d = Dense(); d.A = w.hk; d.lda = 2 * D; d.W = W(f.kffn2_h); d.M = GK; d.N = D; d.K = 2 * D;
        d.bias = p->ffn_kg_b2; d.residual = w.z0; d.ldr = D; d.out_f32 = w.z; d.ldo_f32 = D;
        if ((rc = launch_dense(d, st)) != CM_OK) return rc;
    }
    if ((rc = launch_graph_tail(gt, G, st)) != CM_OK) return rc;
    return post_launch("cm_fusion_forward(tc)");
}

}  // namespace cm

// Profiling aid (tools/fusion_phase_profile.py): CTA 0 of fusion_node writes globaltimer stamps of every phase into buf
// and every CTA its start / end at [40 + 2 b], [41 + 2 b] (40 + 2 x grid int64 on the device) while set; NULL turns it off.  Not part of the product path.
extern "C" void cm_debug_fusion_phase_stamps(long long* device_buf) { cm::g_fusion_phase_dbg = device_buf; }

