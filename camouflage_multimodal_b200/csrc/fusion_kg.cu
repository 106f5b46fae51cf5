// KG side of the folded cross-attention fusion + the graph-level layers (fusion_model.py:123-139, 239-242), fp32:
//
//   kg2rg_partial  grid (image, node chunk): per folded key slot j = (head, kg row) the chunk's softmax statistics
//                  (max, sum of exp) of the raw scores S2[n, j], the un-normalised context sum_n e[n, j] V2[n, c]
//                  and the column sums of the rg stream y -- flash-attention style partials, no atomics, so the
//                  result does not depend on scheduling.
//   kg_tail        one CTA per image: merges the chunk partials into ctx2 [K, D] and mean_n y, then runs the 13-row
//                  KG stream (out-projection + residual + LayerNorm, FFN + residual, mean over KG rows), the fusion
//                  MLP and the four detector heads with all activations in shared memory.  Weights are read as
//                  transposed fp32 copies ([in, out], prepared by cm_fusion_fold) so consecutive threads read
//                  consecutive outputs.
//   kg2rg_weights  (only when the caller asked for attention maps) head-averaged probabilities from S2 and the
//                  merged statistics.
#include <math.h>

#include <algorithm>

#include "fusion_kg.cuh"

namespace cm {

constexpr int KG_SUB = 64;   // nodes per shared-memory sub-block (32 measured no faster: 25.3 vs 24.8 us)

template <int GS>
__global__ void __launch_bounds__(256) kg2rg_partial_kernel(const float* __restrict__ s2, const op_t* __restrict__ v2, const float* __restrict__ y,
                                                            int D, int heads, const int* __restrict__ node_ptr, int n_chunks,
                                                            float* __restrict__ part_m, float* __restrict__ part_l,
                                                            float* __restrict__ part_acc, float* __restrict__ part_y) {
    __shared__ float e[KG_SUB][128];
    __shared__ float s_f[128];
    const int b = blockIdx.x, ch = blockIdx.y, t = threadIdx.x;
    const int n0 = node_ptr[b], n1 = node_ptr[b + 1], nb = n1 - n0;
    // chunk length: a multiple of KG_SUB so that at most n_chunks chunks cover the image
    const int per = ((nb + n_chunks - 1) / n_chunks + KG_SUB - 1) / KG_SUB * KG_SUB;
    const int c_begin = n0 + ch * per, c_end = min(n1, c_begin + per);
    if (c_begin >= n1) return;
    const int hd = D / heads, my_h = t < D ? t / hd : 0;
    float acc[GS];
#pragma unroll
    for (int k = 0; k < GS; ++k) acc[k] = 0.f;
    float run_m = -INFINITY, run_l = 0.f, ysum = 0.f;   // run_m / run_l: slot t (threads < 128)
    for (int base = c_begin; base < c_end; base += KG_SUB) {
        const int cnt = min(KG_SUB, c_end - base);
        for (int i = t; i < KG_SUB * 32; i += 256) {     // float4 per thread, rows beyond the chunk -> -inf
            const int r = i >> 5, c4 = (i & 31) * 4;
            float4 v = make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
            if (r < cnt) v = *reinterpret_cast<const float4*>(s2 + (size_t)(base + r) * 128 + c4);
            *reinterpret_cast<float4*>(&e[r][c4]) = v;
        }
        __syncthreads();
        if (t < 128) {
            float m = run_m;
            for (int r = 0; r < cnt; ++r) m = fmaxf(m, e[r][t]);
            const float f = expf(run_m - m);   // 0 on the first sub-block (run_m = -inf)
            float l = 0.f;
            for (int r = 0; r < KG_SUB; ++r) {
                const float w = r < cnt ? expf(e[r][t] - m) : 0.f;
                e[r][t] = w;
                l += w;
            }
            run_l = run_l * f + l;
            run_m = m;
            s_f[t] = f;
        }
        __syncthreads();
        if (t < D) {
#pragma unroll
            for (int k = 0; k < GS; ++k) acc[k] *= s_f[my_h * GS + k];
            for (int r0 = 0; r0 < cnt; r0 += 8) {   // 16 independent global loads in flight per thread
                float v[8], yv[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const bool in = r0 + i < cnt;
                    v[i] = in ? op2f(v2[(size_t)(base + r0 + i) * D + t]) : 0.f;
                    yv[i] = in ? y[(size_t)(base + r0 + i) * D + t] : 0.f;
                }
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    ysum += yv[i];
#pragma unroll
                    for (int k = 0; k < GS; ++k) acc[k] = fmaf(e[r0 + i][my_h * GS + k], v[i], acc[k]);   // rows >= cnt hold e = 0
                }
            }
        }
        __syncthreads();
    }
    const size_t slot = (size_t)b * n_chunks + ch;
    if (t < 128) {
        part_m[slot * 128 + t] = run_m;
        part_l[slot * 128 + t] = run_l;
    }
    if (t < D) {
#pragma unroll
        for (int k = 0; k < GS; ++k) part_acc[(slot * GS + k) * D + t] = acc[k];
        part_y[slot * D + t] = ysum;
    }
}

// Merge the chunk partials of image b: ctx2 [K, D] = softmax_n(S2)^T V2 as fp16 rows (A operand of the KG-row
// chain) and the node mean of the rg stream -> both[b, 0:D].
__global__ void __launch_bounds__(1024) kg_merge_kernel(const float* __restrict__ part_m, const float* __restrict__ part_l,
                                                       const float* __restrict__ part_acc, const float* __restrict__ part_y,
                                                       const int* __restrict__ node_ptr, int n_chunks, int D, int heads, int gs, int n_kg,
                                                       op_t* __restrict__ ctx2, float* __restrict__ both, float* __restrict__ stat) {
    __shared__ float sc[KG_MAX_CHUNKS * 128];   // exp(m_ch - M) / L per chunk and folded key slot
    const int b = blockIdx.x, tid = threadIdx.x, nt = blockDim.x;
    const int n0 = node_ptr[b], nb = node_ptr[b + 1] - n0;
    const int per = ((nb + n_chunks - 1) / n_chunks + KG_SUB - 1) / KG_SUB * KG_SUB;
    const int nch = nb > 0 ? (nb + per - 1) / per : 0;
    const size_t slot0 = (size_t)b * n_chunks;
    if (tid < 128) {
        float pm[KG_MAX_CHUNKS], pl[KG_MAX_CHUNKS];
#pragma unroll
        for (int ch = 0; ch < KG_MAX_CHUNKS; ++ch) {
            pm[ch] = ch < nch ? part_m[(slot0 + ch) * 128 + tid] : -INFINITY;
            pl[ch] = ch < nch ? part_l[(slot0 + ch) * 128 + tid] : 0.f;
        }
        float M = -INFINITY;
#pragma unroll
        for (int ch = 0; ch < KG_MAX_CHUNKS; ++ch) M = fmaxf(M, pm[ch]);
        float L = 0.f;
#pragma unroll
        for (int ch = 0; ch < KG_MAX_CHUNKS; ++ch)
            if (ch < nch) L += pl[ch] * expf(pm[ch] - M);
        const float inv = L > 0.f ? 1.0f / L : 0.f;
#pragma unroll
        for (int ch = 0; ch < KG_MAX_CHUNKS; ++ch)
            if (ch < nch) sc[ch * 128 + tid] = expf(pm[ch] - M) * inv;
        if (stat) {
            stat[((size_t)b * 128 + tid) * 2] = M;
            stat[((size_t)b * 128 + tid) * 2 + 1] = inv;
        }
    }
    __syncthreads();
    const int hd = D / heads;
    for (int i = tid; i < n_kg * D; i += nt) {   // bv2 is already inside V2
        const int k = i / D, c = i - k * D, j = (c / hd) * gs + k;
        float pv[KG_MAX_CHUNKS];
#pragma unroll
        for (int ch = 0; ch < KG_MAX_CHUNKS; ++ch) pv[ch] = ch < nch ? part_acc[((slot0 + ch) * gs + k) * D + c] : 0.f;
        float s = 0.f;
#pragma unroll
        for (int ch = 0; ch < KG_MAX_CHUNKS; ++ch)
            if (ch < nch) s = fmaf(sc[ch * 128 + j], pv[ch], s);
        ctx2[((size_t)b * n_kg + k) * D + c] = f2op(s);
    }
    for (int c = tid; c < D; c += nt) {          // mean over the image's nodes of the rg stream (fusion_model.py:134)
        float pv[KG_MAX_CHUNKS];
#pragma unroll
        for (int ch = 0; ch < KG_MAX_CHUNKS; ++ch) pv[ch] = ch < nch ? part_y[(slot0 + ch) * D + c] : 0.f;
        float s = 0.f;
#pragma unroll
        for (int ch = 0; ch < KG_MAX_CHUNKS; ++ch) s += pv[ch];
        both[(size_t)b * 2 * D + c] = nb > 0 ? s / (float)nb : 0.f;
    }
}

// out[o] = act(sum_k a[k] w[o][k] + bias[o]): one warp per output, lanes stride K in float4 (w rows are contiguous:
// an nn.Linear weight as stored).  a, out in shared memory; K % 128 == 0 or any K % 4 == 0.
template <int KI>   // K == 128 * KI: every weight load of a warp pass (8 outputs x KI float4 per lane) is issued before the first FMA
__device__ __forceinline__ void gemv_warp_t(const float* a, const float* const* __restrict__ wrows, const float* const* __restrict__ brows,
                                            int rows_per_mat, float* out, int N, bool relu) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
    constexpr int K = 128 * KI;
    for (int o0 = warp * 8; o0 < N; o0 += nwarps * 8) {
        float4 wv[8][KI];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int o = min(o0 + i, N - 1), m = o / rows_per_mat, r = o - m * rows_per_mat;
            const float* wr = wrows[m] + (size_t)r * K;
#pragma unroll
            for (int q = 0; q < KI; ++q) wv[i][q] = __ldg(reinterpret_cast<const float4*>(wr + q * 128 + lane * 4));
        }
        float s[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) s[i] = 0.f;
#pragma unroll
        for (int q = 0; q < KI; ++q) {
            const float4 av = *reinterpret_cast<const float4*>(a + q * 128 + lane * 4);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                s[i] = fmaf(wv[i][q].x, av.x, s[i]);
                s[i] = fmaf(wv[i][q].y, av.y, s[i]);
                s[i] = fmaf(wv[i][q].z, av.z, s[i]);
                s[i] = fmaf(wv[i][q].w, av.w, s[i]);
            }
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const float t = warp_sum(s[i]);
            const int o = o0 + i;
            if (lane == 0 && o < N) {
                const int m = o / rows_per_mat, r = o - m * rows_per_mat;
                const float v = t + (brows[m] ? brows[m][r] : 0.f);
                out[o] = relu ? fmaxf(v, 0.f) : v;
            }
        }
    }
}

// out[o] = act(sum_k a[k] w[o][k] + bias[o]) for the rows of n_mat stacked nn.Linear weights ([rows_per_mat, K] each, as
// stored); one warp per 8 outputs, lanes stride K in float4.  a, out in shared memory; K % 4 == 0.
__device__ __forceinline__ void gemv_warp(const float* a, const float* const* wrows, const float* const* brows, int rows_per_mat, float* out,
                                          int N, int K, bool relu) {
    if (K == 512) return gemv_warp_t<4>(a, wrows, brows, rows_per_mat, out, N, relu);
    if (K == 256) return gemv_warp_t<2>(a, wrows, brows, rows_per_mat, out, N, relu);
    if (K == 128) return gemv_warp_t<1>(a, wrows, brows, rows_per_mat, out, N, relu);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
    for (int o = warp; o < N; o += nwarps) {
        const int m = o / rows_per_mat, r = o - m * rows_per_mat;
        const float* wr = wrows[m] + (size_t)r * K;
        float s = 0.f;
        for (int k = lane * 4; k < K; k += 128) {
            const float4 wv = __ldg(reinterpret_cast<const float4*>(wr + k));
            const float4 av = *reinterpret_cast<const float4*>(a + k);
            s = fmaf(wv.x, av.x, s);
            s = fmaf(wv.y, av.y, s);
            s = fmaf(wv.z, av.z, s);
            s = fmaf(wv.w, av.w, s);
        }
        s = warp_sum(s);
        if (lane == 0) {
            const float v = s + (brows[m] ? brows[m][r] : 0.f);
            out[o] = relu ? fmaxf(v, 0.f) : v;
        }
    }
}

// Per image: mean over the KG rows of z (:135), fusion layer (:138-139), the four heads (:239-242); fp32.
__global__ void __launch_bounds__(1024) graph_tail_kernel(const GraphTailArgs a) {
    __shared__ __align__(16) float both[512], f1[256], fz[256], hid[512];
    const int D = a.D, K = a.n_kg, tid = threadIdx.x, b = blockIdx.x;
    const int warp = tid >> 5, lane = tid & 31, nwarps = blockDim.x >> 5;
    for (int c = tid; c < D; c += blockDim.x) {
        float s = 0.f;
#pragma unroll 8
        for (int r = 0; r < K; ++r) s += a.z[((size_t)b * K + r) * D + c];
        both[D + c] = s / (float)K;
        both[c] = a.both[(size_t)b * 2 * D + c];
    }
    __syncthreads();
    gemv_warp(both, &a.fl1_w, &a.fl1_b, D, f1, D, 2 * D, true);
    __syncthreads();
    gemv_warp(f1, &a.fl2_w, &a.fl2_b, D, fz, D, D, false);
    __syncthreads();
    if (a.fused)
        for (int c = tid; c < D; c += blockDim.x) a.fused[(size_t)b * D + c] = fz[c];
    if (!a.want_heads) return;
    const int H2 = D / 2;
    gemv_warp(fz, a.hw1, a.hb1, H2, hid, 4 * H2, D, true);   // first layers of the four heads as one stacked matrix
    __syncthreads();
    int o = 0;
    for (int t = 0; t < 4; ++t) {
        const int cdim = t < 2 ? a.num_classes : 1;
        for (int j = 0; j < cdim; ++j, ++o) {
            if (o % nwarps != warp) continue;
            const float* wrow = a.hw2[t] + (size_t)j * H2;
            float s = 0.f;
            for (int k = lane; k < H2; k += 32) s = fmaf(hid[t * H2 + k], wrow[k], s);
            s = warp_sum(s) + a.hb2[t][j];
            if (t == 3) s = 1.0f / (1.0f + expf(-s));   // score head ends in nn.Sigmoid (fusion_model.py:229-235)
            float* out = t == 0 ? a.mask_out : t == 1 ? a.inst_out : t == 2 ? a.edge_out : a.score_out;
            if (lane == 0 && out) out[(size_t)b * cdim + j] = s;
        }
    }
}

// attn block of image b: [n_kg, N_b] at n_kg * node_ptr[b]; one thread per (node, kg row)
__global__ void kg2rg_weights_kernel(const float* __restrict__ s2, const float* __restrict__ stat, const int* __restrict__ node_ptr, int n_graphs,
                                     int n_kg, int heads, int gs, float* __restrict__ attn) {
    const int total = node_ptr[n_graphs];
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)total * n_kg) return;
    const int n = (int)(i / n_kg), k = (int)(i % n_kg);
    int lo = 0, hi = n_graphs;
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (node_ptr[mid] <= n) lo = mid; else hi = mid;
    }
    const int n0 = node_ptr[lo], nb = node_ptr[lo + 1] - n0;
    float tot = 0.f;
    for (int h = 0; h < heads; ++h) {
        const int j = h * gs + k;
        tot += expf(s2[(size_t)n * 128 + j] - stat[((size_t)lo * 128 + j) * 2]) * stat[((size_t)lo * 128 + j) * 2 + 1];
    }
    attn[(size_t)n_kg * n0 + (size_t)k * nb + (n - n0)] = tot / (float)heads;
}

int launch_kg2rg_partial(const KgSideArgs& k, int D, int heads, int gs, const int* node_ptr, int n_graphs, cudaStream_t st) {
    dim3 grid(n_graphs, KG_MAX_CHUNKS);
    {
        CM_TIMED("kg2rg_partial", st);
        if (gs == 16)
            kg2rg_partial_kernel<16><<<grid, 256, 0, st>>>(k.s2, k.v2, k.y, D, heads, node_ptr, KG_MAX_CHUNKS, k.part_m, k.part_l, k.part_acc, k.part_y);
        else
            kg2rg_partial_kernel<32><<<grid, 256, 0, st>>>(k.s2, k.v2, k.y, D, heads, node_ptr, KG_MAX_CHUNKS, k.part_m, k.part_l, k.part_acc, k.part_y);
    }
    return post_launch("kg2rg_partial");
}

int launch_kg_merge(const KgSideArgs& k, int D, int heads, int gs, int n_kg, const int* node_ptr, int n_graphs, op_t* ctx2, float* both, float* stat,
                    cudaStream_t st) {
    {
        CM_TIMED("kg_merge", st);
        kg_merge_kernel<<<n_graphs, 1024, 0, st>>>(k.part_m, k.part_l, k.part_acc, k.part_y, node_ptr, KG_MAX_CHUNKS, D, heads, gs, n_kg, ctx2, both, stat);
    }
    int rc = post_launch("kg_merge");
    if (rc != CM_OK || !k.attn_kg2rg) return rc;
    {
        CM_TIMED("kg2rg_weights", st);
        kg2rg_weights_kernel<<<ceil_div((long long)k.node_cap * n_kg, 256), 256, 0, st>>>(k.s2, stat, node_ptr, n_graphs, n_kg, heads, gs, k.attn_kg2rg);
    }
    return post_launch("kg2rg_weights");
}

int launch_graph_tail(const GraphTailArgs& a, int n_graphs, cudaStream_t st) {
    {
        CM_TIMED("graph_tail", st);
        graph_tail_kernel<<<n_graphs, 1024, 0, st>>>(a);
    }
    return post_launch("graph_tail");
}

}  // namespace cm
