// Training-mode CrossAttentionFusion + detector heads: forward with caller-supplied dropout masks and the full
// backward pass (models/multimodal/fusion_model.py:75-146, 208-246 under model.train(); the step is
// models/multimodal/train_multimodal.py:232-275).
//
// Ragged batch: image g owns region rows node_ptr[g] .. node_ptr[g+1]-1 and its own n_kg knowledge-graph rows
// (the caller expands a shared KG to [G, K, d], as the reference's expand() does, so that every gradient that
// reaches the KG side sums over images in the ordinary weight-gradient reductions).  Attention in both directions is
// written out per (row, head) with the probabilities saved for the backward; all reductions are gathers or
// fixed-order sums (no float atomics), dense layers are the fp32 GEMMs of train_common.cu.
#include <stdlib.h>

#include "train_common.cuh"

namespace cm {

constexpr int FT_MAX_DPL = 4;   // head_dim <= 128

struct FusionTrainWs {
    float *R, *Kp, *Q1, *KV1, *P1, *O1, *X1, *L1, *st1, *F1, *R2;
    float *Q2, *KV2, *P2, *O2, *Y1, *L2, *st2, *F2, *K2o;
    float *cat, *Fh, *fused, *hh[4], *score;
    // backward scratch
    float *dfused, *dhh, *dFh, *dcat, *dA, *dB, *dC, *dR, *dkA, *dkB, *dkC, *dKp, *dP1, *dP2, *prodM, *prodK, *dKV1;
    double* part_d;
    float* part_f;
    int splits;
    size_t total;
};

static FusionTrainWs carve_fusion_train(void* ws, const cm_fusion_params* p, int G, int M, int K) {
    Carver c(ws);
    FusionTrainWs w;
    const size_t m = (size_t)(M > 0 ? M : 1), g = (size_t)(G > 0 ? G : 1), gk = g * (size_t)(K > 0 ? K : 1), D = p->hidden, H = p->heads,
                 kk = (size_t)(K > 0 ? K : 1), D2 = D / 2;
    w.R = c.take<float>(m * D);
    w.Kp = c.take<float>(gk * D);
    w.Q1 = c.take<float>(m * D);
    w.KV1 = c.take<float>(gk * 2 * D);
    w.P1 = c.take<float>(m * H * kk);
    w.O1 = c.take<float>(m * D);
    w.X1 = c.take<float>(m * D);
    w.L1 = c.take<float>(m * D);
    w.st1 = c.take<float>(m * 2);
    w.F1 = c.take<float>(m * 2 * D);
    w.R2 = c.take<float>(m * D);
    w.Q2 = c.take<float>(gk * D);
    w.KV2 = c.take<float>(m * 2 * D);
    w.P2 = c.take<float>(H * kk * m);
    w.O2 = c.take<float>(gk * D);
    w.Y1 = c.take<float>(gk * D);
    w.L2 = c.take<float>(gk * D);
    w.st2 = c.take<float>(gk * 2);
    w.F2 = c.take<float>(gk * 2 * D);
    w.K2o = c.take<float>(gk * D);
    w.cat = c.take<float>(g * 2 * D);
    w.Fh = c.take<float>(g * D);
    w.fused = c.take<float>(g * D);
    for (int t = 0; t < 4; ++t) w.hh[t] = c.take<float>(g * D2);
    w.score = c.take<float>(g);
    w.dfused = c.take<float>(g * D);
    w.dhh = c.take<float>(g * D2);
    w.dFh = c.take<float>(g * D);
    w.dcat = c.take<float>(g * 2 * D);
    w.dA = c.take<float>(m * D);
    w.dB = c.take<float>(m * D);
    w.dC = c.take<float>(m * 2 * D);
    w.dR = c.take<float>(m * D);
    w.dkA = c.take<float>(gk * D);
    w.dkB = c.take<float>(gk * D);
    w.dkC = c.take<float>(gk * 2 * D);
    w.dKp = c.take<float>(gk * D);
    w.dP1 = c.take<float>(m * H * kk);
    w.dP2 = c.take<float>(H * kk * m);
    w.prodM = c.take<float>(m * D);
    w.prodK = c.take<float>(gk * D);
    w.dKV1 = c.take<float>(gk * 2 * D);
    const size_t rows = m > gk ? m : gk;
    w.part_d = c.take<double>((size_t)col_parts((int)rows) * 3 * D);
    w.splits = M >= 1024 ? (M / 512 < 64 ? M / 512 : 64) : 1;
    w.part_f = c.take<float>((size_t)w.splits * 3 * D * D);
    w.total = align_up(c.off, 256);
    return w;
}

__device__ __forceinline__ int ft_graph_of(const int* __restrict__ ptr, int G, int r) {
    int lo = 0, hi = G;
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (ptr[mid] <= r) lo = mid; else hi = mid;
    }
    return lo;
}

// ---------------------------------------------------------------- rg -> kg attention (queries = region rows)
// one warp per (row m, head h).  P [M,H,K] receives the softmax probabilities (before dropout); O [M,D] the context.
__global__ void __launch_bounds__(256) attn_r2k_fwd_kernel(const float* __restrict__ Q, const float* __restrict__ KV, const int* __restrict__ node_ptr,
                                                           int G, int M, int K, int D, int H, float scale, const uint8_t* __restrict__ keep,
                                                           float keep_scale, float* __restrict__ P, float* __restrict__ O) {
    const int wid = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (wid >= M * H) return;
    const int m = wid / H, h = wid - m * H, dh = D / H, g = ft_graph_of(node_ptr, G, m);
    const float* kv = KV + (size_t)g * K * 2 * D + h * dh;
    float q[FT_MAX_DPL], o[FT_MAX_DPL];
#pragma unroll
    for (int i = 0; i < FT_MAX_DPL; ++i) {
        const int d = lane + 32 * i;
        q[i] = d < dh ? Q[(size_t)m * D + h * dh + d] * scale : 0.f;
        o[i] = 0.f;
    }
    float* pr = P + ((size_t)m * H + h) * K;
    float mx = -INFINITY;
    for (int k = 0; k < K; ++k) {
        float s = 0.f;
#pragma unroll
        for (int i = 0; i < FT_MAX_DPL; ++i) {
            const int d = lane + 32 * i;
            if (d < dh) s = fmaf(q[i], kv[(size_t)k * 2 * D + d], s);
        }
        s = warp_sum(s);
        if (lane == 0) pr[k] = s;
        mx = fmaxf(mx, s);
    }
    __syncwarp();
    float sum = 0.f;
    for (int k = lane; k < K; k += 32) {
        const float e = expf(pr[k] - mx);
        pr[k] = e;
        sum += e;
    }
    sum = warp_sum(sum);
    const float inv = 1.0f / sum;
    for (int k = lane; k < K; k += 32) pr[k] *= inv;
    __syncwarp();
    const uint8_t* kp = keep ? keep + ((size_t)m * H + h) * K : nullptr;
    for (int k = 0; k < K; ++k) {
        float pk = pr[k];
        if (kp) pk = kp[k] ? pk * keep_scale : 0.f;
#pragma unroll
        for (int i = 0; i < FT_MAX_DPL; ++i) {
            const int d = lane + 32 * i;
            if (d < dh) o[i] = fmaf(pk, kv[(size_t)k * 2 * D + D + d], o[i]);
        }
    }
#pragma unroll
    for (int i = 0; i < FT_MAX_DPL; ++i) {
        const int d = lane + 32 * i;
        if (d < dh) O[(size_t)m * D + h * dh + d] = o[i];
    }
}

// dS [M,H,K] (softmax backward, scaled scores) and dQ [M,D]
__global__ void __launch_bounds__(256) attn_r2k_bwd_kernel(const float* __restrict__ KV, const float* __restrict__ P, const float* __restrict__ dO,
                                                           const int* __restrict__ node_ptr, int G, int M, int K, int D, int H, float scale,
                                                           const uint8_t* __restrict__ keep, float keep_scale, float* __restrict__ dS,
                                                           float* __restrict__ dQ) {
    const int wid = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (wid >= M * H) return;
    const int m = wid / H, h = wid - m * H, dh = D / H, g = ft_graph_of(node_ptr, G, m);
    const float* kv = KV + (size_t)g * K * 2 * D + h * dh;
    float go[FT_MAX_DPL], dq[FT_MAX_DPL];
#pragma unroll
    for (int i = 0; i < FT_MAX_DPL; ++i) {
        const int d = lane + 32 * i;
        go[i] = d < dh ? dO[(size_t)m * D + h * dh + d] : 0.f;
        dq[i] = 0.f;
    }
    const float* pr = P + ((size_t)m * H + h) * K;
    float* ds = dS + ((size_t)m * H + h) * K;
    const uint8_t* kp = keep ? keep + ((size_t)m * H + h) * K : nullptr;
    float S = 0.f;
    for (int k = 0; k < K; ++k) {
        float t = 0.f;
#pragma unroll
        for (int i = 0; i < FT_MAX_DPL; ++i) {
            const int d = lane + 32 * i;
            if (d < dh) t = fmaf(go[i], kv[(size_t)k * 2 * D + D + d], t);
        }
        t = warp_sum(t);
        if (kp) t = kp[k] ? t * keep_scale : 0.f;
        if (lane == 0) ds[k] = t;
        S = fmaf(pr[k], t, S);
    }
    __syncwarp();
    for (int k = lane; k < K; k += 32) ds[k] = pr[k] * (ds[k] - S);
    __syncwarp();
    for (int k = 0; k < K; ++k) {
        const float v = ds[k] * scale;
#pragma unroll
        for (int i = 0; i < FT_MAX_DPL; ++i) {
            const int d = lane + 32 * i;
            if (d < dh) dq[i] = fmaf(v, kv[(size_t)k * 2 * D + d], dq[i]);
        }
    }
#pragma unroll
    for (int i = 0; i < FT_MAX_DPL; ++i) {
        const int d = lane + 32 * i;
        if (d < dh) dQ[(size_t)m * D + h * dh + d] = dq[i];
    }
}

// dK1 / dV1 [G*K, 2D]: thread (key k, dim d) of graph g, head h sums over the graph's region rows
__global__ void attn_r2k_bwd_kv_kernel(const float* __restrict__ Q, const float* __restrict__ P, const float* __restrict__ dS,
                                       const float* __restrict__ dO, const int* __restrict__ node_ptr, int K, int D, int H, float scale,
                                       const uint8_t* __restrict__ keep, float keep_scale, float* __restrict__ dKV) {
    const int g = blockIdx.x, h = blockIdx.y, dh = D / H;
    const int m0 = node_ptr[g], m1 = node_ptr[g + 1];
    for (int i = threadIdx.x; i < K * dh; i += blockDim.x) {
        const int k = i / dh, d = i - k * dh;
        float dk = 0.f, dv = 0.f;
        for (int m = m0; m < m1; ++m) {
            const size_t pi = ((size_t)m * H + h) * K + k;
            float pt = P[pi];
            if (keep) pt = keep[pi] ? pt * keep_scale : 0.f;
            dk = fmaf(dS[pi], Q[(size_t)m * D + h * dh + d], dk);
            dv = fmaf(pt, dO[(size_t)m * D + h * dh + d], dv);
        }
        dKV[((size_t)g * K + k) * 2 * D + h * dh + d] = dk * scale;
        dKV[((size_t)g * K + k) * 2 * D + D + h * dh + d] = dv;
    }
}

// ---------------------------------------------------------------- kg -> rg attention (queries = KG rows)
// one warp per (graph g, key row k, head h); softmax over the graph's region rows.  P2 [H,K,M].
__global__ void __launch_bounds__(256) attn_k2r_fwd_kernel(const float* __restrict__ Q, const float* __restrict__ KV, const int* __restrict__ node_ptr,
                                                           int G, int M, int K, int D, int H, float scale, const uint8_t* __restrict__ keep,
                                                           float keep_scale, float* __restrict__ P, float* __restrict__ O) {
    const int wid = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (wid >= G * K * H) return;
    const int g = wid / (K * H), k = (wid / H) % K, h = wid % H, dh = D / H;
    const int m0 = node_ptr[g], m1 = node_ptr[g + 1];
    const float* q = Q + ((size_t)g * K + k) * D + h * dh;
    float* pr = P + ((size_t)h * K + k) * M;
    float mx = -INFINITY;
    for (int m = m0 + lane; m < m1; m += 32) {
        const float* kr = KV + (size_t)m * 2 * D + h * dh;
        float s = 0.f;
        for (int d = 0; d < dh; ++d) s = fmaf(q[d], kr[d], s);
        s *= scale;
        pr[m] = s;
        mx = fmaxf(mx, s);
    }
    mx = warp_max(mx);
    float sum = 0.f;
    for (int m = m0 + lane; m < m1; m += 32) {
        const float e = expf(pr[m] - mx);
        pr[m] = e;
        sum += e;
    }
    sum = warp_sum(sum);
    const float inv = 1.0f / sum;
    for (int m = m0 + lane; m < m1; m += 32) pr[m] *= inv;
    __syncwarp();
    const uint8_t* kp = keep ? keep + ((size_t)h * K + k) * M : nullptr;
    float o[FT_MAX_DPL];
#pragma unroll
    for (int i = 0; i < FT_MAX_DPL; ++i) o[i] = 0.f;
    for (int m = m0; m < m1; ++m) {
        float pm = pr[m];
        if (kp) pm = kp[m] ? pm * keep_scale : 0.f;
        const float* vr = KV + (size_t)m * 2 * D + D + h * dh;
#pragma unroll
        for (int i = 0; i < FT_MAX_DPL; ++i) {
            const int d = lane + 32 * i;
            if (d < dh) o[i] = fmaf(pm, vr[d], o[i]);
        }
    }
#pragma unroll
    for (int i = 0; i < FT_MAX_DPL; ++i) {
        const int d = lane + 32 * i;
        if (d < dh) O[((size_t)g * K + k) * D + h * dh + d] = o[i];
    }
}

// dS2 [H,K,M] and dQ2 [G*K, D]
__global__ void __launch_bounds__(256) attn_k2r_bwd_a_kernel(const float* __restrict__ KV, const float* __restrict__ P, const float* __restrict__ dO,
                                                             const int* __restrict__ node_ptr, int G, int M, int K, int D, int H, float scale,
                                                             const uint8_t* __restrict__ keep, float keep_scale, float* __restrict__ dS,
                                                             float* __restrict__ dQ) {
    const int wid = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (wid >= G * K * H) return;
    const int g = wid / (K * H), k = (wid / H) % K, h = wid % H, dh = D / H;
    const int m0 = node_ptr[g], m1 = node_ptr[g + 1];
    const float* go = dO + ((size_t)g * K + k) * D + h * dh;
    const float* pr = P + ((size_t)h * K + k) * M;
    float* ds = dS + ((size_t)h * K + k) * M;
    const uint8_t* kp = keep ? keep + ((size_t)h * K + k) * M : nullptr;
    float S = 0.f;
    for (int m = m0 + lane; m < m1; m += 32) {
        const float* vr = KV + (size_t)m * 2 * D + D + h * dh;
        float t = 0.f;
        for (int d = 0; d < dh; ++d) t = fmaf(go[d], vr[d], t);
        if (kp) t = kp[m] ? t * keep_scale : 0.f;
        ds[m] = t;
        S = fmaf(pr[m], t, S);
    }
    S = warp_sum(S);
    for (int m = m0 + lane; m < m1; m += 32) ds[m] = pr[m] * (ds[m] - S);
    __syncwarp();
    float dq[FT_MAX_DPL];
#pragma unroll
    for (int i = 0; i < FT_MAX_DPL; ++i) dq[i] = 0.f;
    for (int m = m0; m < m1; ++m) {
        const float v = ds[m] * scale;
        const float* kr = KV + (size_t)m * 2 * D + h * dh;
#pragma unroll
        for (int i = 0; i < FT_MAX_DPL; ++i) {
            const int d = lane + 32 * i;
            if (d < dh) dq[i] = fmaf(v, kr[d], dq[i]);
        }
    }
#pragma unroll
    for (int i = 0; i < FT_MAX_DPL; ++i) {
        const int d = lane + 32 * i;
        if (d < dh) dQ[((size_t)g * K + k) * D + h * dh + d] = dq[i];
    }
}

// dK2 / dV2 [M, 2D]: one warp per region row, gathers over the graph's K query rows
__global__ void __launch_bounds__(256) attn_k2r_bwd_b_kernel(const float* __restrict__ Q, const float* __restrict__ P, const float* __restrict__ dS,
                                                             const float* __restrict__ dO, const int* __restrict__ node_ptr, int G, int M, int K,
                                                             int D, int H, float scale, const uint8_t* __restrict__ keep, float keep_scale,
                                                             float* __restrict__ dKV) {
    const int m = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (m >= M) return;
    const int g = ft_graph_of(node_ptr, G, m), dh = D / H;
    for (int c = lane; c < D; c += 32) {
        const int h = c / dh;
        float dk = 0.f, dv = 0.f;
        for (int k = 0; k < K; ++k) {
            const size_t pi = ((size_t)h * K + k) * M + m;
            float pt = P[pi];
            if (keep) pt = keep[pi] ? pt * keep_scale : 0.f;
            dk = fmaf(dS[pi], Q[((size_t)g * K + k) * D + c], dk);
            dv = fmaf(pt, dO[((size_t)g * K + k) * D + c], dv);
        }
        dKV[(size_t)m * 2 * D + c] = dk * scale;
        dKV[(size_t)m * 2 * D + D + c] = dv;
    }
}

// ---------------------------------------------------------------- attention building blocks for head_dim 32, K <= 16
// The kernels above walk a graph's region rows one at a time per warp (kg->rg) or per thread (dK1 / dV1): at 480 rows per
// image those are 480-step dependent chains with strided row reads -- 0.93 ms of a 4.8 ms step.  Same sums, laid out for the
// machine: (i) scores of every region row against its graph's K query rows from a shared-memory tile, (ii) row operations
// (softmax / its backward) with a warp per (graph, query row, head) over contiguous memory, (iii) coefficient x row
// accumulation with a CTA per (graph, head): eight warps take 32 rows at a time, lane = channel, sixteen running sums each,
// combined in a fixed order (deterministic; the order differs from the sequential kernels, so results agree to rounding).
constexpr int FA_KMAX = 16;
static bool attn_fast_ok(int K, int D, int H) {
    static const bool off = getenv("CM_TRAIN_ATTN_SEQUENTIAL") != nullptr;   // the kernels above, for A/B runs
    return !off && H > 0 && D == 32 * H && K >= 1 && K <= FA_KMAX && D <= 256;
}

// out[(h K + k) M + m] = scale * <Qg[g(m) K + k, h 32 ..], X[m, xoff + h 32 ..]>; 32 region rows per CTA, thread = (row, head)
__global__ void __launch_bounds__(256) attn_scores32_kernel(const float* __restrict__ Qg, const float* __restrict__ X, int ldx, int xoff,
                                                            const int* __restrict__ node_ptr, int G, int M, int K, int D, int H, float scale,
                                                            float* __restrict__ out) {
    __shared__ float xs[32 * 257];
    const int mb = blockIdx.x * 32, ld = D + 1;
    for (int i = threadIdx.x; i < 32 * D; i += 256) {
        const int r = i / D, c = i - r * D;
        xs[r * ld + c] = mb + r < M ? X[(size_t)(mb + r) * ldx + xoff + c] : 0.f;
    }
    __syncthreads();
    const int r = threadIdx.x & 31, m = mb + r;
    if (m >= M) return;
    const int g = ft_graph_of(node_ptr, G, m);
    for (int h = threadIdx.x >> 5; h < H; h += 8) {
        float x[32];
#pragma unroll
        for (int d = 0; d < 32; ++d) x[d] = xs[r * ld + h * 32 + d];
        for (int k = 0; k < K; ++k) {
            const float4* q = reinterpret_cast<const float4*>(Qg + ((size_t)g * K + k) * D + h * 32);
            float s0 = 0.f, s1 = 0.f;
#pragma unroll
            for (int j = 0; j < 8; j += 2) {
                const float4 a = __ldg(q + j), b = __ldg(q + j + 1);
                s0 = fmaf(a.x, x[4 * j], s0); s0 = fmaf(a.y, x[4 * j + 1], s0); s0 = fmaf(a.z, x[4 * j + 2], s0); s0 = fmaf(a.w, x[4 * j + 3], s0);
                s1 = fmaf(b.x, x[4 * j + 4], s1); s1 = fmaf(b.y, x[4 * j + 5], s1); s1 = fmaf(b.z, x[4 * j + 6], s1); s1 = fmaf(b.w, x[4 * j + 7], s1);
            }
            out[((size_t)h * K + k) * M + m] = (s0 + s1) * scale;
        }
    }
}
// one warp per (graph, query row, head): P[(h K + k) M + m0 .. m1) raw scores -> softmax probabilities, in place
__global__ void __launch_bounds__(256) attn_softmax_rows_kernel(const int* __restrict__ node_ptr, int G, int M, int K, int H, float* __restrict__ P) {
    const int wid = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (wid >= G * K * H) return;
    const int g = wid / (K * H), k = (wid / H) % K, h = wid % H;
    const int m0 = node_ptr[g], m1 = node_ptr[g + 1];
    float* pr = P + ((size_t)h * K + k) * M;
    float mx = -INFINITY;
    for (int m = m0 + lane; m < m1; m += 32) mx = fmaxf(mx, pr[m]);
    mx = warp_max(mx);
    float sum = 0.f;
    for (int m = m0 + lane; m < m1; m += 32) {
        const float e = expf(pr[m] - mx);
        pr[m] = e;
        sum += e;
    }
    sum = warp_sum(sum);
    const float inv = 1.0f / sum;
    for (int m = m0 + lane; m < m1; m += 32) pr[m] *= inv;
}
// softmax backward over the same rows: T holds t = <dO, V> (raw); T <- P * (t' - sum P t'), t' = t after the dropout mask
__global__ void __launch_bounds__(256) attn_dsoftmax_rows_kernel(const float* __restrict__ P, const int* __restrict__ node_ptr, int G, int M, int K,
                                                                 int H, const uint8_t* __restrict__ keep, float keep_scale, float* __restrict__ T) {
    const int wid = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (wid >= G * K * H) return;
    const int g = wid / (K * H), k = (wid / H) % K, h = wid % H;
    const int m0 = node_ptr[g], m1 = node_ptr[g + 1];
    const size_t row = ((size_t)h * K + k) * M;
    float S = 0.f;
    for (int m = m0 + lane; m < m1; m += 32) {
        float t = T[row + m];
        if (keep) t = keep[row + m] ? t * keep_scale : 0.f;
        T[row + m] = t;
        S = fmaf(P[row + m], t, S);
    }
    S = warp_sum(S);
    for (int m = m0 + lane; m < m1; m += 32) T[row + m] = P[row + m] * (T[row + m] - S);
}

struct AccumPair {
    const float* coef;        // LAYOUT 0: [M, H, K]   LAYOUT 1: [H, K, M]
    const uint8_t* keep;      // dropout mask over coef (same layout) or NULL
    float keep_scale;
    const float* x;           // rows [M, ldx], channel h 32 + d at x[m ldx + h 32 + d]
    int ldx;
    float out_scale;
    float* out;               // out[(g K + k) ldo + h 32 + d]
    int ldo;
};
// out_p[g, k, h, d] = out_scale_p * sum over the graph's rows m of coef_p[m, h, k] * x_p[m, h, d]
template <int LAYOUT, int PAIRS>
__global__ void __launch_bounds__(256) attn_accum32_kernel(const AccumPair p0, const AccumPair p1, const int* __restrict__ node_ptr, int M, int K,
                                                           int H) {
    __shared__ float red[PAIRS][8][FA_KMAX][32];
    const int g = blockIdx.x, h = blockIdx.y, w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int m0 = node_ptr[g], m1 = node_ptr[g + 1];
    float acc[PAIRS][FA_KMAX];
#pragma unroll
    for (int p = 0; p < PAIRS; ++p)
#pragma unroll
        for (int k = 0; k < FA_KMAX; ++k) acc[p][k] = 0.f;
    for (int base = m0 + 32 * w; base < m1; base += 256) {
        const int mine = base + lane, cnt = min(32, m1 - base);
        float ck[PAIRS][FA_KMAX];
#pragma unroll
        for (int p = 0; p < PAIRS; ++p) {
            const AccumPair& a = p == 0 ? p0 : p1;
#pragma unroll
            for (int k = 0; k < FA_KMAX; ++k) {
                float c = 0.f;
                if (k < K && mine < m1) {
                    const size_t i = LAYOUT == 0 ? ((size_t)mine * H + h) * K + k : ((size_t)h * K + k) * M + mine;
                    c = a.coef[i];
                    if (a.keep) c = a.keep[i] ? c * a.keep_scale : 0.f;
                }
                ck[p][k] = c;
            }
        }
        for (int j0 = 0; j0 < cnt; j0 += 4) {
            float xv[PAIRS][4];
#pragma unroll
            for (int p = 0; p < PAIRS; ++p) {
                const AccumPair& a = p == 0 ? p0 : p1;
#pragma unroll
                for (int u = 0; u < 4; ++u) xv[p][u] = j0 + u < cnt ? a.x[(size_t)(base + j0 + u) * a.ldx + h * 32 + lane] : 0.f;
            }
#pragma unroll
            for (int u = 0; u < 4; ++u)
#pragma unroll
                for (int p = 0; p < PAIRS; ++p)
#pragma unroll
                    for (int k = 0; k < FA_KMAX; ++k)
                        acc[p][k] = fmaf(__shfl_sync(0xffffffffu, ck[p][k], (j0 + u) & 31), xv[p][u], acc[p][k]);
        }
    }
#pragma unroll
    for (int p = 0; p < PAIRS; ++p)
#pragma unroll
        for (int k = 0; k < FA_KMAX; ++k) red[p][w][k][lane] = acc[p][k];
    __syncthreads();
    for (int i = threadIdx.x; i < PAIRS * K * 32; i += 256) {
        const int p = i / (K * 32), k = (i / 32) % K, d = i & 31;
        float t = 0.f;
#pragma unroll
        for (int ww = 0; ww < 8; ++ww) t += red[p][ww][k][d];
        const AccumPair& a = p == 0 ? p0 : p1;
        a.out[((size_t)g * K + k) * a.ldo + h * 32 + d] = t * a.out_scale;
    }
}

// ---------------------------------------------------------------- pooling, sigmoid
// out[g, off .. off+D) = mean of rows seg[g] .. seg[g+1]-1 (seg == NULL: fixed `len` rows per graph)
__global__ void ft_mean_pool_kernel(const float* __restrict__ in, int D, const int* __restrict__ seg, int len, int G, float* __restrict__ out,
                                    int ldo, int off) {
    const int g = blockIdx.x;
    const int r0 = seg ? seg[g] : g * len, r1 = seg ? seg[g + 1] : (g + 1) * len;
    for (int c = threadIdx.x; c < D; c += blockDim.x) {
        float s = 0.f;
        for (int r = r0; r < r1; ++r) s += in[(size_t)r * D + c];
        out[(size_t)g * ldo + off + c] = r1 > r0 ? s / (float)(r1 - r0) : 0.f;
    }
}
// d[r, :] = dpool[g(r), off .. off+D) / rows(g)
__global__ void ft_mean_pool_bwd_kernel(const float* __restrict__ dpool, int ldo, int off, int D, const int* __restrict__ seg, int len, int G,
                                        int rows, float* __restrict__ d) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long)rows * D) return;
    const int r = (int)(i / D), c = (int)(i - (long)r * D);
    const int g = seg ? ft_graph_of(seg, G, r) : r / len;
    const int cnt = seg ? seg[g + 1] - seg[g] : len;
    d[i] = dpool[(size_t)g * ldo + off + c] / (float)cnt;
}
__global__ void ft_sigmoid_kernel(float* __restrict__ v, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) v[i] = 1.0f / (1.0f + expf(-v[i]));
}
__global__ void ft_sigmoid_bwd_kernel(const float* __restrict__ s, const float* __restrict__ ds, int n, float* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = ds[i] * s[i] * (1.0f - s[i]);
}

static float fdrop_scale(const cm_fusion_dropout* d, int site) { return d && d->keep[site] ? 1.f / (1.f - d->p) : 1.f; }
static const uint8_t* fdrop_keep(const cm_fusion_dropout* d, int site) { return d ? d->keep[site] : nullptr; }

// ------------------------------------------------------------------------------------------------ late fusion
// LateFusion under model.train() (fusion_model.py:149-171 + the detector heads :208-246 with final_dim = hidden / 2):
// mean over nodes / KG rows -> cat -> Linear ReLU Dropout -> Linear ReLU Dropout -> Linear -> heads.  Dropout sites:
// keep[0] [G, hidden], keep[1] [G, hidden/2], keep[5..8] the four head hidden layers [G, hidden/4].
struct LateTrainWs {
    float *cat, *h1, *h2, *fused, *hh[4], *score;
    float *dfused, *dhh, *dh2, *dh1, *dcat, *dsig;
    double* part_d;
    size_t total;
};
static LateTrainWs carve_late_train(void* ws, const cm_fusion_params* p, int G) {
    Carver c(ws);
    LateTrainWs w;
    const size_t g = (size_t)(G > 0 ? G : 1), D = p->hidden, FD = D / 2, HD = FD / 2, in = (size_t)p->rg_dim + p->kg_dim;
    w.cat = c.take<float>(g * in);
    w.h1 = c.take<float>(g * D);
    w.h2 = c.take<float>(g * FD);
    w.fused = c.take<float>(g * FD);
    for (int t = 0; t < 4; ++t) w.hh[t] = c.take<float>(g * HD);
    w.score = c.take<float>(g);
    w.dfused = c.take<float>(g * FD);
    w.dhh = c.take<float>(g * HD);
    w.dh2 = c.take<float>(g * FD);
    w.dh1 = c.take<float>(g * D);
    w.dcat = c.take<float>(g * in);
    w.dsig = c.take<float>(g);
    w.part_d = c.take<double>((size_t)col_parts((int)g) * (D > in ? D : in));
    w.total = align_up(c.off, 256);
    return w;
}
static int check_late_train(const cm_fusion_params* p, const char* who) {
    if (!p || p->fusion_type != 1 || p->hidden <= 0 || p->hidden % 4 || p->rg_dim <= 0 || p->kg_dim <= 0 || p->num_classes <= 0 ||
        !p->late_w[0] || !p->late_w[1] || !p->late_w[2]) {
        set_error("%s: late fusion needs hidden %% 4 == 0 and its three Linear layers", who);
        return CM_ERR_ARG;
    }
    return CM_OK;
}
// heads on `fused` [G, FD] (hidden layer FD / 2): shared by both fusion types
static void heads_train_forward(const cm_fusion_params* p, const cm_fusion_dropout* drop, const float* fusedp, int FD, int G, float* const* hh,
                                float* score_ws, float* mask_out, float* instance_out, float* edge_out, float* score_out, cudaStream_t st) {
    const int HD = FD / 2;
    float* outs[4] = {mask_out, instance_out, edge_out, score_out};
    const int cdim[4] = {p->num_classes, p->num_classes, 1, 1};
    for (int t = 0; t < 4; ++t) {
        if (!p->head_w1[t]) continue;   // a bare fusion module (CrossAttentionFusion / LateFusion on its own) has no heads
        Gemm g = gemm_nt(fusedp, FD, p->head_w1[t], hh[t], HD, G, HD, FD, p->head_b1[t], 1);
        g.keep = fdrop_keep(drop, 5 + t);
        g.keep_scale = fdrop_scale(drop, 5 + t);
        launch_gemm(g, st, "train_linear");
        float* o = t == 3 ? score_ws : outs[t];
        if (!o) continue;
        launch_gemm(gemm_nt(hh[t], HD, p->head_w2[t], o, cdim[t], G, cdim[t], HD, p->head_b2[t], 0), st, "train_linear");
        if (t == 3) {
            CM_TIMED("train_elementwise", st);
            ft_sigmoid_kernel<<<ceil_div(G, 128), 128, 0, st>>>(score_ws, G);   // :233
            if (score_out) cudaMemcpyAsync(score_out, score_ws, sizeof(float) * G, cudaMemcpyDeviceToDevice, st);
        }
    }
}
// accumulates the heads' contribution into dfused [G, FD] (pre-set by the caller) and writes the head gradients
static void heads_backward(const cm_fusion_params* p, const cm_fusion_dropout* drop, const cm_fusion_grads* gr, const float* fusedp, int FD, int G,
                           float* const* hh, const float* score_ws, const float* const* douts, float* dsig, float* dhh, double* part_d,
                           float* dfused, cudaStream_t st) {
    const int HD = FD / 2;
    const int cdim[4] = {p->num_classes, p->num_classes, 1, 1};
    for (int t = 0; t < 4; ++t) {
        if (!p->head_w1[t]) continue;
        if (!douts[t]) {
            cudaMemsetAsync(gr->head_w2[t], 0, sizeof(float) * cdim[t] * HD, st);
            cudaMemsetAsync(gr->head_b2[t], 0, sizeof(float) * cdim[t], st);
            cudaMemsetAsync(gr->head_w1[t], 0, sizeof(float) * HD * FD, st);
            cudaMemsetAsync(gr->head_b1[t], 0, sizeof(float) * HD, st);
            continue;
        }
        const float* dout = douts[t];
        if (t == 3) {   // through the sigmoid
            CM_TIMED("train_elementwise", st);
            ft_sigmoid_bwd_kernel<<<ceil_div(G, 128), 128, 0, st>>>(score_ws, douts[3], G, dsig);
            dout = dsig;
        }
        launch_gemm(gemm_tn(dout, cdim[t], hh[t], HD, gr->head_w2[t], G, cdim[t], HD, nullptr, 1), st, "train_wgrad");
        launch_colsum(dout, cdim[t], G, cdim[t], part_d, gr->head_b2[t], 0, st);
        launch_gemm(gemm_nn(dout, cdim[t], p->head_w2[t], dhh, HD, G, cdim[t], HD), st, "train_dgrad");
        launch_relu_drop_bwd(hh[t], fdrop_scale(drop, 5 + t), (long)G * HD, dhh, st);
        launch_gemm(gemm_tn(dhh, HD, fusedp, FD, gr->head_w1[t], G, HD, FD, nullptr, 1), st, "train_wgrad");
        launch_colsum(dhh, HD, G, HD, part_d, gr->head_b1[t], 0, st);
        Gemm g = gemm_nn(dhh, HD, p->head_w1[t], dfused, FD, G, HD, FD);
        g.accumulate = 1;
        launch_gemm(g, st, "train_dgrad");
    }
}

static int late_train_forward(const cm_fusion_params* p, const float* rg, const int32_t* node_ptr, int G, int M, const float* kg, int K,
                              const cm_fusion_dropout* drop, void* workspace, size_t workspace_bytes, float* fused, float* mask_out,
                              float* instance_out, float* edge_out, float* score_out, cudaStream_t st) {
    (void)M;
    const int D = p->hidden, FD = D / 2, in = p->rg_dim + p->kg_dim;
    LateTrainWs w = carve_late_train(workspace, p, G);
    if (workspace_bytes < w.total) {
        set_error("cm_fusion_train_forward: workspace %zu < %zu bytes", workspace_bytes, w.total);
        return CM_ERR_WORKSPACE;
    }
    {
        CM_TIMED("train_pool", st);
        ft_mean_pool_kernel<<<G, 256, 0, st>>>(rg, p->rg_dim, node_ptr, 0, G, w.cat, in, 0);          // :165-166
        ft_mean_pool_kernel<<<G, 256, 0, st>>>(kg, p->kg_dim, nullptr, K, G, w.cat, in, p->rg_dim);   // :167-168
    }
    {
        Gemm g = gemm_nt(w.cat, in, p->late_w[0], w.h1, D, G, D, in, p->late_b[0], 1);
        g.keep = fdrop_keep(drop, 0);
        g.keep_scale = fdrop_scale(drop, 0);
        launch_gemm(g, st, "train_linear");
    }
    {
        Gemm g = gemm_nt(w.h1, D, p->late_w[1], w.h2, FD, G, FD, D, p->late_b[1], 1);
        g.keep = fdrop_keep(drop, 1);
        g.keep_scale = fdrop_scale(drop, 1);
        launch_gemm(g, st, "train_linear");
    }
    launch_gemm(gemm_nt(w.h2, FD, p->late_w[2], w.fused, FD, G, FD, FD, p->late_b[2], 0), st, "train_linear");
    if (fused) cudaMemcpyAsync(fused, w.fused, sizeof(float) * G * FD, cudaMemcpyDeviceToDevice, st);
    heads_train_forward(p, drop, w.fused, FD, G, w.hh, w.score, mask_out, instance_out, edge_out, score_out, st);
    return post_launch("cm_fusion_train_forward");
}

static int late_backward(const cm_fusion_params* p, const float* rg, const int32_t* node_ptr, int G, int M, const float* kg, int K,
                         const cm_fusion_dropout* drop, void* workspace, size_t workspace_bytes, const float* d_fused, const float* const* douts,
                         const cm_fusion_grads* gr, float* d_rg, float* d_kg, cudaStream_t st) {
    (void)rg;
    (void)kg;
    const int D = p->hidden, FD = D / 2, in = p->rg_dim + p->kg_dim;
    CM_REQUIRE(gr->late_w[0] && gr->late_w[1] && gr->late_w[2] && gr->late_b[0] && gr->late_b[1] && gr->late_b[2],
               "cm_fusion_backward: late fusion needs the late_w / late_b gradient buffers");
    LateTrainWs w = carve_late_train(workspace, p, G);
    if (workspace_bytes < w.total) {
        set_error("cm_fusion_backward: workspace %zu < %zu bytes", workspace_bytes, w.total);
        return CM_ERR_WORKSPACE;
    }
    if (d_fused) cudaMemcpyAsync(w.dfused, d_fused, sizeof(float) * G * FD, cudaMemcpyDeviceToDevice, st);
    else cudaMemsetAsync(w.dfused, 0, sizeof(float) * G * FD, st);
    heads_backward(p, drop, gr, w.fused, FD, G, w.hh, w.score, douts, w.dsig, w.dhh, w.part_d, w.dfused, st);
    launch_gemm(gemm_tn(w.dfused, FD, w.h2, FD, gr->late_w[2], G, FD, FD, nullptr, 1), st, "train_wgrad");
    launch_colsum(w.dfused, FD, G, FD, w.part_d, gr->late_b[2], 0, st);
    launch_gemm(gemm_nn(w.dfused, FD, p->late_w[2], w.dh2, FD, G, FD, FD), st, "train_dgrad");
    launch_relu_drop_bwd(w.h2, fdrop_scale(drop, 1), (long)G * FD, w.dh2, st);
    launch_gemm(gemm_tn(w.dh2, FD, w.h1, D, gr->late_w[1], G, FD, D, nullptr, 1), st, "train_wgrad");
    launch_colsum(w.dh2, FD, G, FD, w.part_d, gr->late_b[1], 0, st);
    launch_gemm(gemm_nn(w.dh2, FD, p->late_w[1], w.dh1, D, G, FD, D), st, "train_dgrad");
    launch_relu_drop_bwd(w.h1, fdrop_scale(drop, 0), (long)G * D, w.dh1, st);
    launch_gemm(gemm_tn(w.dh1, D, w.cat, in, gr->late_w[0], G, D, in, nullptr, 1), st, "train_wgrad");
    launch_colsum(w.dh1, D, G, D, w.part_d, gr->late_b[0], 0, st);
    launch_gemm(gemm_nn(w.dh1, D, p->late_w[0], w.dcat, in, G, D, in), st, "train_dgrad");
    {
        CM_TIMED("train_pool", st);
        if (d_rg) ft_mean_pool_bwd_kernel<<<ceil_div((long)M * p->rg_dim, 256), 256, 0, st>>>(w.dcat, in, 0, p->rg_dim, node_ptr, 0, G, M, d_rg);
        if (d_kg)
            ft_mean_pool_bwd_kernel<<<ceil_div((long)G * K * p->kg_dim, 256), 256, 0, st>>>(w.dcat, in, p->rg_dim, p->kg_dim, nullptr, K, G, G * K, d_kg);
    }
    return post_launch("cm_fusion_backward");
}

static int check_fusion_train(const cm_fusion_params* p, const char* who) {
    if (!p || p->fusion_type != 0 || p->hidden <= 0 || p->heads <= 0 || p->hidden % p->heads || p->hidden / p->heads > 32 * FT_MAX_DPL ||
        p->hidden % 2 || p->num_classes <= 0) {
        set_error("%s: training path supports fusion_type cross_attention with head_dim <= %d", who, 32 * FT_MAX_DPL);
        return CM_ERR_ARG;
    }
    return CM_OK;
}

}  // namespace cm

extern "C" size_t cm_fusion_train_workspace_bytes(const cm_fusion_params* p, int n_graphs, int n_rows, int n_kg) {
    if (p && p->fusion_type == 1) {
        if (cm::check_late_train(p, "cm_fusion_train_workspace_bytes") != CM_OK || n_graphs <= 0) return 0;
        return cm::carve_late_train(nullptr, p, n_graphs).total;
    }
    if (cm::check_fusion_train(p, "cm_fusion_train_workspace_bytes") != CM_OK || n_graphs <= 0 || n_rows < 0 || n_kg <= 0) return 0;
    return cm::carve_fusion_train(nullptr, p, n_graphs, n_rows, n_kg).total;
}

extern "C" int cm_fusion_train_forward(const cm_fusion_params* p, const float* rg, const int32_t* node_ptr, int n_graphs, int n_rows,
                                       const float* kg, int n_kg, const cm_fusion_dropout* drop, void* workspace, size_t workspace_bytes,
                                       float* fused, float* mask_out, float* instance_out, float* edge_out, float* score_out, void* stream) {
    using namespace cm;
    CM_DEVICE_OR_RETURN();
    if (p && p->fusion_type == 1) {
        if (int rc = check_late_train(p, "cm_fusion_train_forward")) return rc;
        CM_REQUIRE(rg && node_ptr && kg && workspace && n_graphs > 0 && n_rows > 0 && n_kg > 0, "cm_fusion_train_forward: bad arguments");
        return late_train_forward(p, rg, node_ptr, n_graphs, n_rows, kg, n_kg, drop, workspace, workspace_bytes, fused, mask_out, instance_out,
                                  edge_out, score_out, as_stream(stream));
    }
    if (int rc = check_fusion_train(p, "cm_fusion_train_forward")) return rc;
    CM_REQUIRE(rg && node_ptr && kg && workspace && n_graphs > 0 && n_rows > 0 && n_kg > 0, "cm_fusion_train_forward: bad arguments");
    const int G = n_graphs, M = n_rows, K = n_kg, D = p->hidden, H = p->heads, GK = G * K;
    FusionTrainWs w = carve_fusion_train(workspace, p, G, M, K);
    if (workspace_bytes < w.total) {
        set_error("cm_fusion_train_forward: workspace %zu < %zu bytes", workspace_bytes, w.total);
        return CM_ERR_WORKSPACE;
    }
    cudaStream_t st = as_stream(stream);
    const float scale = 1.0f / sqrtf((float)(D / H));
    // projections (nn.Identity when the input width is already `hidden`, fusion_model.py:29-30)
    if (p->rg_proj_w) launch_gemm(gemm_nt(rg, p->rg_dim, p->rg_proj_w, w.R, D, M, D, p->rg_dim, p->rg_proj_b, 0), st, "train_linear");
    else cudaMemcpyAsync(w.R, rg, sizeof(float) * M * D, cudaMemcpyDeviceToDevice, st);
    if (p->kg_proj_w) launch_gemm(gemm_nt(kg, p->kg_dim, p->kg_proj_w, w.Kp, D, GK, D, p->kg_dim, p->kg_proj_b, 0), st, "train_linear");
    else cudaMemcpyAsync(w.Kp, kg, sizeof(float) * GK * D, cudaMemcpyDeviceToDevice, st);
    // ---- region rows attend to KG rows (:112-120)
    launch_gemm(gemm_nt(w.R, D, p->r2k_in_w, w.Q1, D, M, D, D, p->r2k_in_b, 0), st, "train_linear");
    launch_gemm(gemm_nt(w.Kp, D, p->r2k_in_w + (size_t)D * D, w.KV1, 2 * D, GK, 2 * D, D, p->r2k_in_b + D, 0), st, "train_linear");
    {
        CM_TIMED("train_attn_r2k_fwd", st);
        attn_r2k_fwd_kernel<<<ceil_div((long)M * H, 8), 256, 0, st>>>(w.Q1, w.KV1, node_ptr, G, M, K, D, H, scale, fdrop_keep(drop, 0),
                                                                      fdrop_scale(drop, 0), w.P1, w.O1);
    }
    launch_gemm(gemm_nt(w.O1, D, p->r2k_out_w, w.X1, D, M, D, D, p->r2k_out_b, 0), st, "train_linear");
    launch_add(w.X1, w.R, w.X1, (long)M * D, st);
    launch_ln_fwd(w.X1, p->ln_rg_w, p->ln_rg_b, M, D, w.L1, w.st1, st);
    {
        Gemm g = gemm_nt(w.L1, D, p->ffn_rg_w1, w.F1, 2 * D, M, 2 * D, D, p->ffn_rg_b1, 1);
        g.keep = fdrop_keep(drop, 1);
        g.keep_scale = fdrop_scale(drop, 1);
        launch_gemm(g, st, "train_linear");
    }
    launch_gemm(gemm_nt(w.F1, 2 * D, p->ffn_rg_w2, w.R2, D, M, D, 2 * D, p->ffn_rg_b2, 0), st, "train_linear");
    launch_add(w.R2, w.L1, w.R2, (long)M * D, st);
    // ---- KG rows attend to region rows (:123-131)
    launch_gemm(gemm_nt(w.Kp, D, p->k2r_in_w, w.Q2, D, GK, D, D, p->k2r_in_b, 0), st, "train_linear");
    launch_gemm(gemm_nt(w.R, D, p->k2r_in_w + (size_t)D * D, w.KV2, 2 * D, M, 2 * D, D, p->k2r_in_b + D, 0), st, "train_linear");
    {
        CM_TIMED("train_attn_k2r_fwd", st);
        if (attn_fast_ok(K, D, H)) {
            attn_scores32_kernel<<<ceil_div(M, 32), 256, 0, st>>>(w.Q2, w.KV2, 2 * D, 0, node_ptr, G, M, K, D, H, scale, w.P2);
            attn_softmax_rows_kernel<<<ceil_div((long)GK * H, 8), 256, 0, st>>>(node_ptr, G, M, K, H, w.P2);
            const AccumPair pv = {w.P2, fdrop_keep(drop, 2), fdrop_scale(drop, 2), w.KV2 + D, 2 * D, 1.f, w.O2, D};
            attn_accum32_kernel<1, 1><<<dim3(G, H), 256, 0, st>>>(pv, pv, node_ptr, M, K, H);
        } else {
            attn_k2r_fwd_kernel<<<ceil_div((long)GK * H, 8), 256, 0, st>>>(w.Q2, w.KV2, node_ptr, G, M, K, D, H, scale, fdrop_keep(drop, 2),
                                                                           fdrop_scale(drop, 2), w.P2, w.O2);
        }
    }
    launch_gemm(gemm_nt(w.O2, D, p->k2r_out_w, w.Y1, D, GK, D, D, p->k2r_out_b, 0), st, "train_linear");
    launch_add(w.Y1, w.Kp, w.Y1, (long)GK * D, st);
    launch_ln_fwd(w.Y1, p->ln_kg_w, p->ln_kg_b, GK, D, w.L2, w.st2, st);
    {
        Gemm g = gemm_nt(w.L2, D, p->ffn_kg_w1, w.F2, 2 * D, GK, 2 * D, D, p->ffn_kg_b1, 1);
        g.keep = fdrop_keep(drop, 3);
        g.keep_scale = fdrop_scale(drop, 3);
        launch_gemm(g, st, "train_linear");
    }
    launch_gemm(gemm_nt(w.F2, 2 * D, p->ffn_kg_w2, w.K2o, D, GK, D, 2 * D, p->ffn_kg_b2, 0), st, "train_linear");
    launch_add(w.K2o, w.L2, w.K2o, (long)GK * D, st);
    // ---- pooling, fusion MLP, heads (:134-139, 229-246)
    {
        CM_TIMED("train_pool", st);
        ft_mean_pool_kernel<<<G, 256, 0, st>>>(w.R2, D, node_ptr, 0, G, w.cat, 2 * D, 0);
        ft_mean_pool_kernel<<<G, 256, 0, st>>>(w.K2o, D, nullptr, K, G, w.cat, 2 * D, D);
    }
    {
        Gemm g = gemm_nt(w.cat, 2 * D, p->fl_w1, w.Fh, D, G, D, 2 * D, p->fl_b1, 1);
        g.keep = fdrop_keep(drop, 4);
        g.keep_scale = fdrop_scale(drop, 4);
        launch_gemm(g, st, "train_linear");
    }
    launch_gemm(gemm_nt(w.Fh, D, p->fl_w2, w.fused, D, G, D, D, p->fl_b2, 0), st, "train_linear");
    if (fused) cudaMemcpyAsync(fused, w.fused, sizeof(float) * G * D, cudaMemcpyDeviceToDevice, st);
    heads_train_forward(p, drop, w.fused, D, G, w.hh, w.score, mask_out, instance_out, edge_out, score_out, st);
    return post_launch("cm_fusion_train_forward");
}

extern "C" int cm_fusion_backward(const cm_fusion_params* p, const float* rg, const int32_t* node_ptr, int n_graphs, int n_rows,
                                  const float* kg, int n_kg, const cm_fusion_dropout* drop, void* workspace, size_t workspace_bytes,
                                  const float* d_fused, const float* d_mask_out, const float* d_instance_out, const float* d_edge_out,
                                  const float* d_score_out, const cm_fusion_grads* gr, float* d_rg, float* d_kg, void* stream) {
    using namespace cm;
    CM_DEVICE_OR_RETURN();
    const float* douts[4] = {d_mask_out, d_instance_out, d_edge_out, d_score_out};
    if (p && p->fusion_type == 1) {
        if (int rc = check_late_train(p, "cm_fusion_backward")) return rc;
        CM_REQUIRE(rg && node_ptr && kg && workspace && gr && n_graphs > 0 && n_rows > 0 && n_kg > 0, "cm_fusion_backward: bad arguments");
        return late_backward(p, rg, node_ptr, n_graphs, n_rows, kg, n_kg, drop, workspace, workspace_bytes, d_fused, douts, gr, d_rg, d_kg,
                             as_stream(stream));
    }
    if (int rc = check_fusion_train(p, "cm_fusion_backward")) return rc;
    CM_REQUIRE(rg && node_ptr && kg && workspace && gr && n_graphs > 0 && n_rows > 0 && n_kg > 0, "cm_fusion_backward: bad arguments");
    const int G = n_graphs, M = n_rows, K = n_kg, D = p->hidden, H = p->heads, GK = G * K;
    FusionTrainWs w = carve_fusion_train(workspace, p, G, M, K);
    if (workspace_bytes < w.total) {
        set_error("cm_fusion_backward: workspace %zu < %zu bytes", workspace_bytes, w.total);
        return CM_ERR_WORKSPACE;
    }
    cudaStream_t st = as_stream(stream);
    const float scale = 1.0f / sqrtf((float)(D / H));
    const int S = w.splits;
    const long MD = (long)M * D, KD = (long)GK * D;

    // ---- heads -> dfused
    if (d_fused) cudaMemcpyAsync(w.dfused, d_fused, sizeof(float) * G * D, cudaMemcpyDeviceToDevice, st);
    else cudaMemsetAsync(w.dfused, 0, sizeof(float) * G * D, st);
    heads_backward(p, drop, gr, w.fused, D, G, w.hh, w.score, douts, w.dcat, w.dhh, w.part_d, w.dfused, st);
    // ---- fusion_layer
    launch_gemm(gemm_tn(w.dfused, D, w.Fh, D, gr->fl_w2, G, D, D, nullptr, 1), st, "train_wgrad");
    launch_colsum(w.dfused, D, G, D, w.part_d, gr->fl_b2, 0, st);
    launch_gemm(gemm_nn(w.dfused, D, p->fl_w2, w.dFh, D, G, D, D), st, "train_dgrad");
    launch_relu_drop_bwd(w.Fh, fdrop_scale(drop, 4), (long)G * D, w.dFh, st);
    launch_gemm(gemm_tn(w.dFh, D, w.cat, 2 * D, gr->fl_w1, G, D, 2 * D, nullptr, 1), st, "train_wgrad");
    launch_colsum(w.dFh, D, G, D, w.part_d, gr->fl_b1, 0, st);
    launch_gemm(gemm_nn(w.dFh, D, p->fl_w1, w.dcat, 2 * D, G, D, 2 * D), st, "train_dgrad");
    {
        CM_TIMED("train_pool", st);
        ft_mean_pool_bwd_kernel<<<ceil_div(MD, 256), 256, 0, st>>>(w.dcat, 2 * D, 0, D, node_ptr, 0, G, M, w.dA);   // dA = dR2
        ft_mean_pool_bwd_kernel<<<ceil_div(KD, 256), 256, 0, st>>>(w.dcat, 2 * D, D, D, nullptr, K, G, GK, w.dkA);  // dkA = dK2o
    }

    // ================= KG stream: K2o = L2 + FFN_kg(L2); L2 = LN(Y1); Y1 = Kp + out_proj(O2)
    launch_gemm(gemm_tn(w.dkA, D, w.F2, 2 * D, gr->ffn_kg_w2, GK, D, 2 * D, nullptr, 1), st, "train_wgrad");
    launch_colsum(w.dkA, D, GK, D, w.part_d, gr->ffn_kg_b2, 0, st);
    launch_gemm(gemm_nn(w.dkA, D, p->ffn_kg_w2, w.dkC, 2 * D, GK, D, 2 * D), st, "train_dgrad");
    launch_relu_drop_bwd(w.F2, fdrop_scale(drop, 3), (long)GK * 2 * D, w.dkC, st);
    launch_gemm(gemm_tn(w.dkC, 2 * D, w.L2, D, gr->ffn_kg_w1, GK, 2 * D, D, nullptr, 1), st, "train_wgrad");
    launch_colsum(w.dkC, 2 * D, GK, 2 * D, w.part_d, gr->ffn_kg_b1, 0, st);
    {
        Gemm g = gemm_nn(w.dkC, 2 * D, p->ffn_kg_w1, w.dkA, D, GK, 2 * D, D);   // dL2 = dK2o + dF2pre W1
        g.accumulate = 1;
        launch_gemm(g, st, "train_dgrad");
    }
    launch_ln_bwd(w.Y1, w.st2, p->ln_kg_w, w.dkA, GK, D, w.dkB, w.prodK, st);   // dkB = dY1
    launch_colsum(w.prodK, D, GK, D, w.part_d, gr->ln_kg_w, 0, st);
    launch_colsum(w.dkA, D, GK, D, w.part_d, gr->ln_kg_b, 0, st);
    // dKp starts as the residual branch
    cudaMemcpyAsync(w.dKp, w.dkB, sizeof(float) * KD, cudaMemcpyDeviceToDevice, st);
    launch_gemm(gemm_tn(w.dkB, D, w.O2, D, gr->k2r_out_w, GK, D, D, nullptr, 1), st, "train_wgrad");
    launch_colsum(w.dkB, D, GK, D, w.part_d, gr->k2r_out_b, 0, st);
    launch_gemm(gemm_nn(w.dkB, D, p->k2r_out_w, w.dkA, D, GK, D, D), st, "train_dgrad");   // dkA = dO2
    {
        CM_TIMED("train_attn_k2r_bwd", st);
        if (attn_fast_ok(K, D, H)) {
            attn_scores32_kernel<<<ceil_div(M, 32), 256, 0, st>>>(w.dkA, w.KV2, 2 * D, D, node_ptr, G, M, K, D, H, 1.f, w.dP2);   // t = <dO2, V2>
            attn_dsoftmax_rows_kernel<<<ceil_div((long)GK * H, 8), 256, 0, st>>>(w.P2, node_ptr, G, M, K, H, fdrop_keep(drop, 2),
                                                                                 fdrop_scale(drop, 2), w.dP2);
            const AccumPair dq = {w.dP2, nullptr, 1.f, w.KV2, 2 * D, scale, w.dkB, D};                                            // dkB = dQ2
            attn_accum32_kernel<1, 1><<<dim3(G, H), 256, 0, st>>>(dq, dq, node_ptr, M, K, H);
        } else {
            attn_k2r_bwd_a_kernel<<<ceil_div((long)GK * H, 8), 256, 0, st>>>(w.KV2, w.P2, w.dkA, node_ptr, G, M, K, D, H, scale, fdrop_keep(drop, 2),
                                                                             fdrop_scale(drop, 2), w.dP2, w.dkB);   // dkB = dQ2
        }
        attn_k2r_bwd_b_kernel<<<ceil_div(M, 8), 256, 0, st>>>(w.Q2, w.P2, w.dP2, w.dkA, node_ptr, G, M, K, D, H, scale, fdrop_keep(drop, 2),
                                                              fdrop_scale(drop, 2), w.dC);                       // dC = dKV2 [M,2D]
    }
    // in_proj of kg->rg: q rows from Kp, k/v rows from R
    launch_gemm(gemm_tn(w.dkB, D, w.Kp, D, gr->k2r_in_w, GK, D, D, nullptr, 1), st, "train_wgrad");
    launch_colsum(w.dkB, D, GK, D, w.part_d, gr->k2r_in_b, 0, st);
    launch_gemm(gemm_tn(w.dC, 2 * D, w.R, D, gr->k2r_in_w + (size_t)D * D, M, 2 * D, D, w.part_f, S), st, "train_wgrad");
    launch_colsum(w.dC, 2 * D, M, 2 * D, w.part_d, gr->k2r_in_b + D, 0, st);
    {
        Gemm g = gemm_nn(w.dkB, D, p->k2r_in_w, w.dKp, D, GK, D, D);   // dKp += dQ2 Wq
        g.accumulate = 1;
        launch_gemm(g, st, "train_dgrad");
    }
    launch_gemm(gemm_nn(w.dC, 2 * D, p->k2r_in_w + (size_t)D * D, w.dR, D, M, 2 * D, D), st, "train_dgrad");   // dR = dKV2 Wkv

    // ================= region stream: R2 = L1 + FFN_rg(L1); L1 = LN(X1); X1 = R + out_proj(O1)
    launch_gemm(gemm_tn(w.dA, D, w.F1, 2 * D, gr->ffn_rg_w2, M, D, 2 * D, w.part_f, S), st, "train_wgrad");
    launch_colsum(w.dA, D, M, D, w.part_d, gr->ffn_rg_b2, 0, st);
    launch_gemm(gemm_nn(w.dA, D, p->ffn_rg_w2, w.dC, 2 * D, M, D, 2 * D), st, "train_dgrad");
    launch_relu_drop_bwd(w.F1, fdrop_scale(drop, 1), (long)M * 2 * D, w.dC, st);
    launch_gemm(gemm_tn(w.dC, 2 * D, w.L1, D, gr->ffn_rg_w1, M, 2 * D, D, w.part_f, S), st, "train_wgrad");
    launch_colsum(w.dC, 2 * D, M, 2 * D, w.part_d, gr->ffn_rg_b1, 0, st);
    {
        Gemm g = gemm_nn(w.dC, 2 * D, p->ffn_rg_w1, w.dA, D, M, 2 * D, D);   // dL1
        g.accumulate = 1;
        launch_gemm(g, st, "train_dgrad");
    }
    launch_ln_bwd(w.X1, w.st1, p->ln_rg_w, w.dA, M, D, w.dB, w.prodM, st);   // dB = dX1
    launch_colsum(w.prodM, D, M, D, w.part_d, gr->ln_rg_w, 0, st);
    launch_colsum(w.dA, D, M, D, w.part_d, gr->ln_rg_b, 0, st);
    launch_add(w.dR, w.dB, w.dR, MD, st);   // residual branch
    launch_gemm(gemm_tn(w.dB, D, w.O1, D, gr->r2k_out_w, M, D, D, w.part_f, S), st, "train_wgrad");
    launch_colsum(w.dB, D, M, D, w.part_d, gr->r2k_out_b, 0, st);
    launch_gemm(gemm_nn(w.dB, D, p->r2k_out_w, w.dA, D, M, D, D), st, "train_dgrad");   // dA = dO1
    {
        CM_TIMED("train_attn_r2k_bwd", st);
        attn_r2k_bwd_kernel<<<ceil_div((long)M * H, 8), 256, 0, st>>>(w.KV1, w.P1, w.dA, node_ptr, G, M, K, D, H, scale, fdrop_keep(drop, 0),
                                                                      fdrop_scale(drop, 0), w.dP1, w.dB);   // dB = dQ1
        if (attn_fast_ok(K, D, H)) {
            const AccumPair dk = {w.dP1, nullptr, 1.f, w.Q1, D, scale, w.dKV1, 2 * D};
            const AccumPair dv = {w.P1, fdrop_keep(drop, 0), fdrop_scale(drop, 0), w.dA, D, 1.f, w.dKV1 + D, 2 * D};
            attn_accum32_kernel<0, 2><<<dim3(G, H), 256, 0, st>>>(dk, dv, node_ptr, M, K, H);
        } else {
            attn_r2k_bwd_kv_kernel<<<dim3(G, H), 256, 0, st>>>(w.Q1, w.P1, w.dP1, w.dA, node_ptr, K, D, H, scale, fdrop_keep(drop, 0),
                                                               fdrop_scale(drop, 0), w.dKV1);
        }
    }
    launch_gemm(gemm_tn(w.dB, D, w.R, D, gr->r2k_in_w, M, D, D, w.part_f, S), st, "train_wgrad");
    launch_colsum(w.dB, D, M, D, w.part_d, gr->r2k_in_b, 0, st);
    launch_gemm(gemm_tn(w.dKV1, 2 * D, w.Kp, D, gr->r2k_in_w + (size_t)D * D, GK, 2 * D, D, nullptr, 1), st, "train_wgrad");
    launch_colsum(w.dKV1, 2 * D, GK, 2 * D, w.part_d, gr->r2k_in_b + D, 0, st);
    {
        Gemm g = gemm_nn(w.dB, D, p->r2k_in_w, w.dR, D, M, D, D);   // dR += dQ1 Wq
        g.accumulate = 1;
        launch_gemm(g, st, "train_dgrad");
    }
    {
        Gemm g = gemm_nn(w.dKV1, 2 * D, p->r2k_in_w + (size_t)D * D, w.dKp, D, GK, 2 * D, D);   // dKp += dKV1 Wkv
        g.accumulate = 1;
        launch_gemm(g, st, "train_dgrad");
    }
    // ---- input projections
    if (p->rg_proj_w) {
        launch_gemm(gemm_tn(w.dR, D, rg, p->rg_dim, gr->rg_proj_w, M, D, p->rg_dim, w.part_f, S), st, "train_wgrad");
        launch_colsum(w.dR, D, M, D, w.part_d, gr->rg_proj_b, 0, st);
        if (d_rg) launch_gemm(gemm_nn(w.dR, D, p->rg_proj_w, d_rg, p->rg_dim, M, D, p->rg_dim), st, "train_dgrad");
    } else if (d_rg) {
        cudaMemcpyAsync(d_rg, w.dR, sizeof(float) * MD, cudaMemcpyDeviceToDevice, st);
    }
    if (p->kg_proj_w) {
        launch_gemm(gemm_tn(w.dKp, D, kg, p->kg_dim, gr->kg_proj_w, GK, D, p->kg_dim, nullptr, 1), st, "train_wgrad");
        launch_colsum(w.dKp, D, GK, D, w.part_d, gr->kg_proj_b, 0, st);
        if (d_kg) launch_gemm(gemm_nn(w.dKp, D, p->kg_proj_w, d_kg, p->kg_dim, GK, D, p->kg_dim), st, "train_dgrad");
    } else if (d_kg) {
        cudaMemcpyAsync(d_kg, w.dKp, sizeof(float) * KD, cudaMemcpyDeviceToDevice, st);
    }
    return post_launch("cm_fusion_backward");
}
