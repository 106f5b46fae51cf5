// A10-A12: CrossAttentionFusion / LateFusion + detector heads (models/multimodal/fusion_model.py:16-246).
//
// Ragged batch: all images' region nodes form one [M, d] row block (node_ptr gives the segments), so the
// seven node-side linears run as M-row dense layers; knowledge-graph rows are shared by every image when
// kg_graphs == 1 (projected once per call -- "KG broadcast once").  nn.MultiheadAttention is written out:
// packed in-projection rows [q; k; v], q scaled by 1/sqrt(head_dim), softmax over keys, out-projection,
// head-averaged probabilities returned only when the caller asks for them (fusion_model.py:112-129, 243-246).
#include <math.h>

#include <algorithm>

#include "flash_tc.cuh"
#include "fusion_tc.cuh"
#include "linear.cuh"

namespace cm {

__device__ __forceinline__ int seg_of_row(const int* __restrict__ ptr, int n_seg, int r) {
    int lo = 0, hi = n_seg;
    while (hi - lo > 1) {
        int mid = (lo + hi) >> 1;
        if (ptr[mid] <= r) lo = mid; else hi = mid;
    }
    return lo;
}

// rg -> kg: one warp per region node.  Lane owns CPL = D/32 contiguous channels, so a head spans
// LPH = 32/heads lanes and every key costs log2(LPH) shuffles for all heads at once.
template <int CPL>
__global__ void __launch_bounds__(256) attn_rg2kg_kernel(const float* __restrict__ q, const float* __restrict__ kv, int D, int heads,
                                                         const int* __restrict__ node_ptr, int n_graphs, int cap, int n_kg,
                                                         int kg_shared, float* __restrict__ ctx, float* __restrict__ attn) {
    const int row = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    const int live = min(node_ptr[n_graphs], cap);
    if (row >= live) return;
    const int lph = 32 / heads;
    const int b = kg_shared ? 0 : seg_of_row(node_ptr, n_graphs, row);
    const float* kvb = kv + (size_t)b * n_kg * 2 * D;
    float qv[CPL], acc[CPL];
#pragma unroll
    for (int c = 0; c < CPL; ++c) {
        qv[c] = q[(size_t)row * D + lane * CPL + c];
        acc[c] = 0.f;
    }
    float m = -INFINITY, l = 0.f;
    for (int k = 0; k < n_kg; ++k) {
        const float* kr = kvb + (size_t)k * 2 * D + lane * CPL;
        float s = 0.f;
#pragma unroll
        for (int c = 0; c < CPL; ++c) s = fmaf(qv[c], kr[c], s);
        for (int o = 1; o < lph; o <<= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        const float mn = fmaxf(m, s);
        const float corr = expf(m - mn), p = expf(s - mn);
        l = l * corr + p;
        const float* vr = kr + D;
#pragma unroll
        for (int c = 0; c < CPL; ++c) acc[c] = fmaf(p, vr[c], acc[c] * corr);
        m = mn;
    }
    const float inv = 1.0f / l;
#pragma unroll
    for (int c = 0; c < CPL; ++c) ctx[(size_t)row * D + lane * CPL + c] = acc[c] * inv;
    if (attn) {
        for (int k = 0; k < n_kg; ++k) {
            const float* kr = kvb + (size_t)k * 2 * D + lane * CPL;
            float s = 0.f;
#pragma unroll
            for (int c = 0; c < CPL; ++c) s = fmaf(qv[c], kr[c], s);
            for (int o = 1; o < lph; o <<= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
            float p = expf(s - m) * inv;   // identical in the lph lanes of a head
            for (int o = lph; o < 32; o <<= 1) p += __shfl_xor_sync(0xffffffffu, p, o);   // sum over heads
            if (lane == 0) attn[(size_t)row * n_kg + k] = p / (float)heads;
        }
    }
}

// kg -> rg: one warp per (image, kg row, head).  Pass 1 finds the row maximum, pass 2 accumulates
// exp-weights and the context; (max, sum) are kept for the optional head-averaged weight pass.
__global__ void __launch_bounds__(256) attn_kg2rg_kernel(const float* __restrict__ q2, const float* __restrict__ kv, int D, int heads,
                                                         const int* __restrict__ node_ptr, int n_graphs, int n_kg, int kg_shared,
                                                         float* __restrict__ ctx, float* __restrict__ stat) {
    const long long wid = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (wid >= (long long)n_graphs * n_kg * heads) return;
    const int h = (int)(wid % heads);
    const int k = (int)((wid / heads) % n_kg);
    const int b = (int)(wid / ((long long)heads * n_kg));
    const int hd = D / heads;
    const int n0 = node_ptr[b], n1 = node_ptr[b + 1];
    const float* qr = q2 + ((size_t)(kg_shared ? 0 : b) * n_kg + k) * D + h * hd;
    float mx = -INFINITY;
    for (int n = n0 + lane; n < n1; n += 32) {
        const float* kr = kv + (size_t)n * 2 * D + h * hd;
        float s = 0.f;
        for (int c = 0; c < hd; ++c) s = fmaf(qr[c], kr[c], s);
        mx = fmaxf(mx, s);
    }
    mx = warp_max(mx);
    float lsum = 0.f;
    float acc[4] = {0.f, 0.f, 0.f, 0.f};   // channel d = lane + 32*t of this head (hd <= 128)
    for (int base = n0; base < n1; base += 32) {
        const int n = base + lane;
        float p = 0.f;
        if (n < n1) {
            const float* kr = kv + (size_t)n * 2 * D + h * hd;
            float s = 0.f;
            for (int c = 0; c < hd; ++c) s = fmaf(qr[c], kr[c], s);
            p = expf(s - mx);
        }
        lsum += p;
        const int cnt = min(32, n1 - base);
        for (int j = 0; j < cnt; ++j) {
            const float pj = __shfl_sync(0xffffffffu, p, j);
            const float* vr = kv + (size_t)(base + j) * 2 * D + D + h * hd;
#pragma unroll
            for (int t = 0; t < 4; ++t) {
                const int d = lane + 32 * t;
                if (d < hd) acc[t] = fmaf(pj, vr[d], acc[t]);
            }
        }
    }
    lsum = warp_sum(lsum);
    const float inv = n1 > n0 ? 1.0f / lsum : 0.f;
    float* o = ctx + ((size_t)b * n_kg + k) * D + h * hd;
#pragma unroll
    for (int t = 0; t < 4; ++t) {
        const int d = lane + 32 * t;
        if (d < hd) o[d] = acc[t] * inv;
    }
    if (stat && lane == 0) {
        stat[2 * wid] = mx;
        stat[2 * wid + 1] = inv;
    }
}

// head-averaged kg->rg probabilities: one warp per (image, kg row); block [n_kg, N_b] at n_kg*node_ptr[b]
__global__ void __launch_bounds__(256) attn_kg2rg_weights_kernel(const float* __restrict__ q2, const float* __restrict__ kv, int D,
                                                                 int heads, const int* __restrict__ node_ptr, int n_graphs, int n_kg,
                                                                 int kg_shared, const float* __restrict__ stat, float* __restrict__ attn) {
    const long long wid = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (wid >= (long long)n_graphs * n_kg) return;
    const int k = (int)(wid % n_kg), b = (int)(wid / n_kg);
    const int hd = D / heads;
    const int n0 = node_ptr[b], n1 = node_ptr[b + 1];
    const float* qr = q2 + ((size_t)(kg_shared ? 0 : b) * n_kg + k) * D;
    float* o = attn + (size_t)n_kg * n0 + (size_t)k * (n1 - n0);
    for (int n = n0 + lane; n < n1; n += 32) {
        const float* kr = kv + (size_t)n * 2 * D;
        float tot = 0.f;
        for (int h = 0; h < heads; ++h) {
            float s = 0.f;
            for (int c = 0; c < hd; ++c) s = fmaf(qr[h * hd + c], kr[h * hd + c], s);
            const float* st = stat + 2 * (wid * heads + h);
            tot += expf(s - st[0]) * st[1];
        }
        o[n - n0] = tot / (float)heads;
    }
}

}  // namespace cm

// ------------------------------------------------------------------------------------------- host side
namespace cm {

struct FusionWs {
    float *r, *q1, *ctx1, *t, *y0, *hh, *y, *k2v2;          // node-side [M, D] / [M, 2D]
    float *kp, *k1v1, *q2;                                  // kg-side, kg_graphs * n_kg rows
    float *ctx2, *c2, *z0, *hk, *z, *stat;                  // n_graphs * n_kg rows
    float *both, *f1, *fused, *hid, *pooled;                // n_graphs rows
    size_t total;
};

static FusionWs carve_fusion(void* ws, const cm_fusion_params* p, int G, int M, int Kk) {
    Carver c(ws);
    FusionWs w;
    const size_t D = p->hidden, m = M, gk = (size_t)G * Kk;
    w.r = c.take<float>(m * D);
    w.q1 = c.take<float>(m * D);
    w.ctx1 = c.take<float>(m * D);
    w.t = c.take<float>(m * D);
    w.y0 = c.take<float>(m * D);
    w.hh = c.take<float>(m * 2 * D);
    w.y = c.take<float>(m * D);
    w.k2v2 = c.take<float>(m * 2 * D);
    w.kp = c.take<float>(gk * D);
    w.k1v1 = c.take<float>(gk * 2 * D);
    w.q2 = c.take<float>(gk * D);
    w.ctx2 = c.take<float>(gk * D);
    w.c2 = c.take<float>(gk * D);
    w.z0 = c.take<float>(gk * D);
    w.hk = c.take<float>(gk * 2 * D);
    w.z = c.take<float>(gk * D);
    w.stat = c.take<float>(gk * p->heads * 2);
    w.both = c.take<float>((size_t)G * 2 * D);
    w.f1 = c.take<float>((size_t)G * D);
    w.fused = c.take<float>((size_t)G * D);
    w.hid = c.take<float>((size_t)G * 4 * D);
    w.pooled = c.take<float>((size_t)G * (p->rg_dim + p->kg_dim));
    w.total = align_up(c.off, 256);
    return w;
}

static int check_fusion(const cm_fusion_params* p) {
    CM_REQUIRE(p, "cm_fusion: NULL params");
    CM_REQUIRE(p->fusion_type == 0 || p->fusion_type == 1, "Unknown fusion_type: %d", p->fusion_type);
    CM_REQUIRE(p->rg_dim > 0 && p->kg_dim > 0 && p->hidden > 0 && p->num_classes >= 1, "cm_fusion: bad dims");
    if (p->fusion_type == 0) {
        const int D = p->hidden, hd = p->heads;
        CM_REQUIRE(hd >= 1 && hd <= 32 && (hd & (hd - 1)) == 0, "cm_fusion: num_heads=%d must be a power of two <= 32", hd);
        CM_REQUIRE(D % 64 == 0 && D <= 512 && D % hd == 0 && D / hd <= 128, "cm_fusion: hidden_dim=%d unsupported", D);
        CM_REQUIRE(p->rg_proj_w || p->rg_dim == D, "cm_fusion: rg_proj missing while rg_dim != hidden_dim");
        CM_REQUIRE(p->kg_proj_w || p->kg_dim == D, "cm_fusion: kg_proj missing while kg_dim != hidden_dim");
    } else {
        CM_REQUIRE(p->hidden % 2 == 0, "cm_fusion: late fusion needs an even hidden_dim");
    }
    return CM_OK;
}

template <int CPL>
static void launch_rg2kg(const float* q, const float* kv, int D, int heads, const int* node_ptr, int G, int M, int Kk, int shared,
                         float* ctx, float* attn, cudaStream_t st) {
    {
        CM_TIMED("attn_rg2kg", st);
        attn_rg2kg_kernel<CPL><<<ceil_div(M, 8), 256, 0, st>>>(q, kv, D, heads, node_ptr, G, M, Kk, shared, ctx, attn);
    }
}

static int detector_heads(const cm_fusion_params* p, const float* fused, int final_dim, int G, float* hid, float* mask_out,
                          float* instance_out, float* edge_out, float* score_out, cudaStream_t st) {
    const int H2 = final_dim / 2;
    int rc;
    if (!mask_out && !instance_out && !edge_out && !score_out) return CM_OK;   // fusion module used on its own
    for (int t = 0; t < 4; ++t)
        CM_REQUIRE(p->head_w1[t] && p->head_b1[t] && p->head_w2[t] && p->head_b2[t], "cm_fusion_forward: head %d parameters missing", t);
    Linear l;
    l.M = G; l.K = final_dim; l.A = fused; l.lda = final_dim; l.N = H2; l.ldc = 4 * H2; l.relu = 1;
    for (int t = 0; t < 4; ++t) {
        l.W = p->head_w1[t]; l.bias = p->head_b1[t]; l.C = hid + t * H2;
        if ((rc = launch_linear(l, st)) != CM_OK) return rc;
    }
    Heads2 h2;
    h2.n_heads = 4;
    h2.hdim = H2;
    float* outs[4] = {mask_out, instance_out, edge_out, score_out};
    for (int t = 0; t < 4; ++t) {
        h2.w2[t] = p->head_w2[t];
        h2.b2[t] = p->head_b2[t];
        h2.cdim[t] = t < 2 ? p->num_classes : 1;
        h2.out[t] = outs[t];
    }
    h2.sigmoid[3] = 1;   // score head ends in nn.Sigmoid (fusion_model.py:229-235)
    return launch_heads2(hid, 4 * H2, G, nullptr, h2, st);
}

}  // namespace cm

extern "C" size_t cm_fusion_workspace_bytes(const cm_fusion_params* p, int n_graphs, int node_cap, int n_kg) {
    if (!p || n_graphs <= 0 || node_cap <= 0 || n_kg <= 0 || cm::check_fusion(p) != CM_OK) return 0;
    size_t need = cm::carve_fusion(nullptr, p, n_graphs, node_cap, n_kg).total;
    if (cm::fusion_fold_supported(p, n_kg)) need = std::max(need, cm::fusion_tc_workspace_bytes(p, n_graphs, node_cap, n_kg));
    if (cm::fusion_flash_supported(p, n_kg)) need = std::max(need, cm::fusion_flash_workspace_bytes(p, n_graphs, node_cap, n_kg));
    return need;
}

extern "C" size_t cm_fusion_folded_bytes(const cm_fusion_params* p, int n_kg) {
    if (!p || n_kg <= 0 || cm::check_fusion(p) != CM_OK || !cm::fusion_fold_supported(p, n_kg)) return 0;
    return cm::fusion_folded_layout(p, n_kg).total;
}

extern "C" int cm_fusion_fold(const cm_fusion_params* p, const float* kg, int n_kg, void* folded, size_t folded_bytes, void* stream) {
    using namespace cm;
    CM_DEVICE_OR_RETURN();
    int rc = check_fusion(p);
    if (rc != CM_OK) return rc;
    CM_REQUIRE(kg && folded && n_kg > 0, "cm_fusion_fold: NULL pointer");
    CM_REQUIRE(fusion_fold_supported(p, n_kg),
               "cm_fusion_fold: needs cross_attention, hidden_dim <= 256 (multiple of 32), rg_dim %% 8 == 0 and num_heads * "
               "(16 or 32 slots >= n_kg) <= 128; got hidden=%d heads=%d n_kg=%d", p->hidden, p->heads, n_kg);
    const size_t need = fusion_folded_layout(p, n_kg).total;
    CM_REQUIRE(folded_bytes >= need, "cm_fusion_fold: blob %zu < %zu bytes", folded_bytes, need);
    return fusion_fold(p, kg, n_kg, folded, as_stream(stream));
}

extern "C" int cm_fusion_forward(const cm_fusion_params* p, const void* folded, const float* rg, const int32_t* node_ptr, int n_graphs, int node_cap,
                                 const float* kg, int kg_graphs, int n_kg, void* workspace, size_t workspace_bytes, float* fused_out,
                                 float* mask_out, float* instance_out, float* edge_out, float* score_out, float* attn_rg2kg,
                                 float* attn_kg2rg, void* stream) {
    using namespace cm;
    CM_DEVICE_OR_RETURN();
    int rc = check_fusion(p);
    if (rc != CM_OK) return rc;
    CM_REQUIRE(rg && node_ptr && kg && workspace, "cm_fusion_forward: NULL pointer");
    CM_REQUIRE(n_graphs > 0 && node_cap > 0 && n_kg > 0, "cm_fusion_forward: bad sizes");
    CM_REQUIRE(kg_graphs == 1 || kg_graphs == n_graphs, "cm_fusion_forward: kg_graphs must be 1 or n_graphs");
    const int G = n_graphs, M = node_cap, Kk = n_kg, D = p->hidden;
    FusionWs w = carve_fusion(workspace, p, G, M, Kk);
    if (workspace_bytes < w.total) {
        set_error("cm_fusion_forward: workspace %zu < %zu bytes", workspace_bytes, w.total);
        return CM_ERR_WORKSPACE;
    }
    cudaStream_t st = as_stream(stream);
    if (folded) {
        CM_REQUIRE(kg_graphs == 1 && fusion_fold_supported(p, Kk), "cm_fusion_forward: a folded blob needs shared kg rows (kg_graphs == 1)");
        const size_t need = fusion_tc_workspace_bytes(p, G, M, Kk);
        if (workspace_bytes < need) {
            set_error("cm_fusion_forward: workspace %zu < %zu bytes", workspace_bytes, need);
            return CM_ERR_WORKSPACE;
        }
        return fusion_forward_tc(p, folded, rg, node_ptr, G, M, Kk, workspace, fused_out, mask_out, instance_out, edge_out, score_out,
                                 attn_rg2kg, attn_kg2rg, st);
    }
    const int* m_dev = node_ptr + G;
    float* fused = fused_out ? fused_out : w.fused;
    Linear l;

    if (p->fusion_type == 1) {
        // LateFusion (fusion_model.py:164-171): mean over nodes, concat, Linear-ReLU-Linear-ReLU-Linear
        const int dr = p->rg_dim, dk = p->kg_dim, P = dr + dk;
        if ((rc = launch_segment_mean(rg, dr, node_ptr, 0, G, w.pooled, P, st)) != CM_OK) return rc;
        if (kg_graphs == 1) {
            // one shared mean row, replicated per graph through fixed_len = 0 segments is not expressible:
            // compute it once into row 0 then broadcast with a stride-0 dense layer input below
            if ((rc = launch_segment_mean(kg, dk, nullptr, Kk, 1, w.pooled + dr, P, st)) != CM_OK) return rc;
            for (int g = 1; g < G; ++g)
                cudaMemcpyAsync(w.pooled + (size_t)g * P + dr, w.pooled + dr, sizeof(float) * dk, cudaMemcpyDeviceToDevice, st);
        } else {
            if ((rc = launch_segment_mean(kg, dk, nullptr, Kk, G, w.pooled + dr, P, st)) != CM_OK) return rc;
        }
        l = Linear(); l.M = G; l.A = w.pooled; l.lda = P; l.K = P; l.W = p->late_w[0]; l.bias = p->late_b[0]; l.N = D; l.C = w.both; l.ldc = D; l.relu = 1;
        if ((rc = launch_linear(l, st)) != CM_OK) return rc;
        l.A = w.both; l.lda = D; l.K = D; l.W = p->late_w[1]; l.bias = p->late_b[1]; l.N = D / 2; l.C = w.f1; l.ldc = D / 2;
        if ((rc = launch_linear(l, st)) != CM_OK) return rc;
        l.A = w.f1; l.lda = D / 2; l.K = D / 2; l.W = p->late_w[2]; l.bias = p->late_b[2]; l.N = D / 2; l.C = fused; l.ldc = D / 2; l.relu = 0;
        if ((rc = launch_linear(l, st)) != CM_OK) return rc;
        if ((rc = detector_heads(p, fused, D / 2, G, w.hid, mask_out, instance_out, edge_out, score_out, st)) != CM_OK) return rc;
        return post_launch("cm_fusion_forward(late)");
    }

    // large shared KG vocabularies, attention maps not requested: fused QK^T.softmax.V on the tensor cores (flash_tc.cu)
    if (kg_graphs == 1 && !attn_rg2kg && !attn_kg2rg && fusion_flash_supported(p, Kk)) {
        const size_t need = fusion_flash_workspace_bytes(p, G, M, Kk);
        if (workspace_bytes < need) {
            set_error("cm_fusion_forward: workspace %zu < %zu bytes", workspace_bytes, need);
            return CM_ERR_WORKSPACE;
        }
        FlashTail t;
        if ((rc = fusion_forward_flash(p, rg, node_ptr, G, M, kg, Kk, workspace, &t, st)) != CM_OK) return rc;
        float* fz = fused_out ? fused_out : t.fused;
        l = Linear(); l.M = G; l.A = t.both; l.lda = 2 * D; l.K = 2 * D; l.W = p->fl_w1; l.bias = p->fl_b1; l.N = D; l.C = t.f1; l.ldc = D; l.relu = 1;
        if ((rc = launch_linear(l, st)) != CM_OK) return rc;
        l = Linear(); l.M = G; l.A = t.f1; l.lda = D; l.K = D; l.W = p->fl_w2; l.bias = p->fl_b2; l.N = D; l.C = fz; l.ldc = D;
        if ((rc = launch_linear(l, st)) != CM_OK) return rc;
        if ((rc = detector_heads(p, fz, D, G, t.hid, mask_out, instance_out, edge_out, score_out, st)) != CM_OK) return rc;
        return post_launch("cm_fusion_forward(flash)");
    }

    const int heads = p->heads, R = kg_graphs * Kk, GK = G * Kk, shared = kg_graphs == 1;
    const float scale = 1.0f / sqrtf((float)(D / heads));
    // projections (fusion_model.py:108-109)
    const float* r = rg;
    if (p->rg_proj_w) {
        l = Linear(); l.M = M; l.m_dev = m_dev; l.A = rg; l.lda = p->rg_dim; l.K = p->rg_dim; l.W = p->rg_proj_w; l.bias = p->rg_proj_b; l.N = D; l.C = w.r; l.ldc = D;
        if ((rc = launch_linear(l, st)) != CM_OK) return rc;
        r = w.r;
    }
    const float* kp = kg;
    if (p->kg_proj_w) {
        l = Linear(); l.M = R; l.A = kg; l.lda = p->kg_dim; l.K = p->kg_dim; l.W = p->kg_proj_w; l.bias = p->kg_proj_b; l.N = D; l.C = w.kp; l.ldc = D;
        if ((rc = launch_linear(l, st)) != CM_OK) return rc;
        kp = w.kp;
    }
    // ---- rg attends to kg (:112-120)
    l = Linear(); l.M = M; l.m_dev = m_dev; l.A = r; l.lda = D; l.K = D; l.W = p->r2k_in_w; l.bias = p->r2k_in_b; l.N = D; l.C = w.q1; l.ldc = D; l.alpha = scale;
    if ((rc = launch_linear(l, st)) != CM_OK) return rc;
    l = Linear(); l.M = R; l.A = kp; l.lda = D; l.K = D; l.W = p->r2k_in_w + (size_t)D * D; l.bias = p->r2k_in_b + D; l.N = 2 * D; l.C = w.k1v1; l.ldc = 2 * D;
    if ((rc = launch_linear(l, st)) != CM_OK) return rc;
    switch (D / 32) {
        case 2: launch_rg2kg<2>(w.q1, w.k1v1, D, heads, node_ptr, G, M, Kk, shared, w.ctx1, attn_rg2kg, st); break;
        case 4: launch_rg2kg<4>(w.q1, w.k1v1, D, heads, node_ptr, G, M, Kk, shared, w.ctx1, attn_rg2kg, st); break;
        case 6: launch_rg2kg<6>(w.q1, w.k1v1, D, heads, node_ptr, G, M, Kk, shared, w.ctx1, attn_rg2kg, st); break;
        case 8: launch_rg2kg<8>(w.q1, w.k1v1, D, heads, node_ptr, G, M, Kk, shared, w.ctx1, attn_rg2kg, st); break;
        case 10: launch_rg2kg<10>(w.q1, w.k1v1, D, heads, node_ptr, G, M, Kk, shared, w.ctx1, attn_rg2kg, st); break;
        case 12: launch_rg2kg<12>(w.q1, w.k1v1, D, heads, node_ptr, G, M, Kk, shared, w.ctx1, attn_rg2kg, st); break;
        case 14: launch_rg2kg<14>(w.q1, w.k1v1, D, heads, node_ptr, G, M, Kk, shared, w.ctx1, attn_rg2kg, st); break;
        default: launch_rg2kg<16>(w.q1, w.k1v1, D, heads, node_ptr, G, M, Kk, shared, w.ctx1, attn_rg2kg, st); break;
    }
    if ((rc = post_launch("attn_rg2kg")) != CM_OK) return rc;
    l = Linear(); l.M = M; l.m_dev = m_dev; l.A = w.ctx1; l.lda = D; l.K = D; l.W = p->r2k_out_w; l.bias = p->r2k_out_b; l.N = D; l.C = w.t; l.ldc = D; l.residual = r; l.ldr = D;
    if ((rc = launch_linear(l, st)) != CM_OK) return rc;
    if ((rc = launch_layernorm(w.t, w.y0, p->ln_rg_w, p->ln_rg_b, M, D, m_dev, st)) != CM_OK) return rc;
    l = Linear(); l.M = M; l.m_dev = m_dev; l.A = w.y0; l.lda = D; l.K = D; l.W = p->ffn_rg_w1; l.bias = p->ffn_rg_b1; l.N = 2 * D; l.C = w.hh; l.ldc = 2 * D; l.relu = 1;
    if ((rc = launch_linear(l, st)) != CM_OK) return rc;
    l = Linear(); l.M = M; l.m_dev = m_dev; l.A = w.hh; l.lda = 2 * D; l.K = 2 * D; l.W = p->ffn_rg_w2; l.bias = p->ffn_rg_b2; l.N = D; l.C = w.y; l.ldc = D; l.residual = w.y0; l.ldr = D;
    if ((rc = launch_linear(l, st)) != CM_OK) return rc;
    if ((rc = launch_segment_mean(w.y, D, node_ptr, 0, G, w.both, 2 * D, st)) != CM_OK) return rc;   // :134

    // ---- kg attends to rg (:123-131)
    l = Linear(); l.M = R; l.A = kp; l.lda = D; l.K = D; l.W = p->k2r_in_w; l.bias = p->k2r_in_b; l.N = D; l.C = w.q2; l.ldc = D; l.alpha = scale;
    if ((rc = launch_linear(l, st)) != CM_OK) return rc;
    l = Linear(); l.M = M; l.m_dev = m_dev; l.A = r; l.lda = D; l.K = D; l.W = p->k2r_in_w + (size_t)D * D; l.bias = p->k2r_in_b + D; l.N = 2 * D; l.C = w.k2v2; l.ldc = 2 * D;
    if ((rc = launch_linear(l, st)) != CM_OK) return rc;
    {
        CM_TIMED("attn_kg2rg", st);
        attn_kg2rg_kernel<<<ceil_div((long long)GK * heads, 8), 256, 0, st>>>(w.q2, w.k2v2, D, heads, node_ptr, G, Kk, shared, w.ctx2, w.stat);
    }
    if ((rc = post_launch("attn_kg2rg")) != CM_OK) return rc;
    if (attn_kg2rg) {
        {
            CM_TIMED("attn_kg2rg_weights", st);
            attn_kg2rg_weights_kernel<<<ceil_div(GK, 8), 256, 0, st>>>(w.q2, w.k2v2, D, heads, node_ptr, G, Kk, shared, w.stat, attn_kg2rg);
        }
        if ((rc = post_launch("attn_kg2rg_weights")) != CM_OK) return rc;
    }
    l = Linear(); l.M = GK; l.A = w.ctx2; l.lda = D; l.K = D; l.W = p->k2r_out_w; l.bias = p->k2r_out_b; l.N = D; l.C = w.c2; l.ldc = D; l.residual = kp; l.ldr = D;
    l.res_mod = shared ? Kk : 0;
    if ((rc = launch_linear(l, st)) != CM_OK) return rc;
    if ((rc = launch_layernorm(w.c2, w.z0, p->ln_kg_w, p->ln_kg_b, GK, D, nullptr, st)) != CM_OK) return rc;
    l = Linear(); l.M = GK; l.A = w.z0; l.lda = D; l.K = D; l.W = p->ffn_kg_w1; l.bias = p->ffn_kg_b1; l.N = 2 * D; l.C = w.hk; l.ldc = 2 * D; l.relu = 1;
    if ((rc = launch_linear(l, st)) != CM_OK) return rc;
    l = Linear(); l.M = GK; l.A = w.hk; l.lda = 2 * D; l.K = 2 * D; l.W = p->ffn_kg_w2; l.bias = p->ffn_kg_b2; l.N = D; l.C = w.z; l.ldc = D; l.residual = w.z0; l.ldr = D;
    if ((rc = launch_linear(l, st)) != CM_OK) return rc;
    if ((rc = launch_segment_mean(w.z, D, nullptr, Kk, G, w.both + D, 2 * D, st)) != CM_OK) return rc;   // :135

    // ---- fusion layer (:138-139) and the four heads (:239-242)
    l = Linear(); l.M = G; l.A = w.both; l.lda = 2 * D; l.K = 2 * D; l.W = p->fl_w1; l.bias = p->fl_b1; l.N = D; l.C = w.f1; l.ldc = D; l.relu = 1;
    if ((rc = launch_linear(l, st)) != CM_OK) return rc;
    l = Linear(); l.M = G; l.A = w.f1; l.lda = D; l.K = D; l.W = p->fl_w2; l.bias = p->fl_b2; l.N = D; l.C = fused; l.ldc = D;
    if ((rc = launch_linear(l, st)) != CM_OK) return rc;
    if ((rc = detector_heads(p, fused, D, G, w.hid, mask_out, instance_out, edge_out, score_out, st)) != CM_OK) return rc;
    return post_launch("cm_fusion_forward");
}
