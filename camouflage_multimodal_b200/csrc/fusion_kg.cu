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
#include <cooperative_groups.h>
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

// ------------------------------------------------------------------------------------------------ fused KG-side tail
// A cluster of 4 CTAs per image for everything after the node-side kernel: merge of the kg->rg chunk partials -> ctx2
// [K,256]; z0 = LN(kp + ctx2 Wo^T + bo); z = z0 + W2 relu(W1 z0 + b1) + b2 (fusion_model.py:130-131); mean over the K
// rows (:135); fusion MLP (:138-139); the four heads (:239-242).  Each CTA owns a quarter of every layer's OUTPUT columns
// (so a quarter of its weights: the layer chain is bound by how fast one SM can pull weights out of L2) and broadcasts
// its slice of the activations into the shared memory of its three peers (DSMEM), one cluster barrier per layer.
// The three K-row GEMMs are 16-row problems (K <= 16 rows, zero padded): warp-level mma.sync m16n8k16 tiles -- a
// 128-row tcgen05 tile would be 90 % padding -- with the activations split into fp16 hi + lo parts (two MMAs against
// the same fp16 weight fragment, fp32 accumulation) and the weight fragments read straight from the folded fp16 copies
// in L2.  Replaces kg_merge + fusion_kgrows + graph_tail (19 + 37 + 30 us per batch of 32).
namespace cg = cooperative_groups;
constexpr int KT_CL = 4, KT_THREADS = 256, KT_LDA = 256 + 8, KT_LDH = 512 + 8;   // +8 halves: rows land on different banks
struct KtSmem {
    float sc[KG_MAX_CHUNKS * 128];
    float z[16 * 256];                 // pre-LN rows, then z0 (fp32), then this CTA's columns of z
    __half a_hi[16 * KT_LDA], a_lo[16 * KT_LDA];
    __half h_hi[16 * KT_LDH], h_lo[16 * KT_LDH];
    float both[512], f1[256], fz[256], hid[128];
};

__device__ __forceinline__ void mma16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void split2(float v0, float v1, __half2* hi, __half2* lo) {
    const __half h0 = __float2half_rn(v0), h1 = __float2half_rn(v1);
    *hi = __halves2half2(h0, h1);
    *lo = __halves2half2(__float2half_rn(v0 - __half2float(h0)), __float2half_rn(v1 - __half2float(h1)));
}
// C[16, n) += A[16, KD] W[n, KD]^T for the n-tiles nt0 + warp, nt0 + warp + nwarps, ... < nt1; A = hi + lo in shared
// memory (row stride lda halves), W fp16 [N, KD] row-major in global memory.  epi(row, col, v0, v1) gets column pairs.
template <int KD, typename Epi>
__device__ __forceinline__ void kt_gemm16(const __half* a_hi, const __half* a_lo, int lda, const op_t* __restrict__ W, int nt0, int nt1, Epi epi) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5, g = lane >> 2, t = lane & 3;
    for (int nt = nt0 + warp; nt < nt1; nt += nwarps) {
        const uint32_t* wrow = reinterpret_cast<const uint32_t*>(W + (size_t)(nt * 8 + g) * KD);
        float c[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int kb = 0; kb < KD / 256; ++kb) {   // 256 columns of K at a time: 32 weight fragments in flight before the first MMA
            uint32_t b[16][2];
#pragma unroll
            for (int ks = 0; ks < 16; ++ks) {
                b[ks][0] = __ldg(wrow + (kb * 16 + ks) * 8 + t);
                b[ks][1] = __ldg(wrow + (kb * 16 + ks) * 8 + 4 + t);
            }
#pragma unroll
            for (int ks = 0; ks < 16; ++ks) {
                const int k0 = (kb * 16 + ks) * 16 + 2 * t;
                uint32_t ah[4], al[4];
                ah[0] = *reinterpret_cast<const uint32_t*>(a_hi + g * lda + k0);
                ah[1] = *reinterpret_cast<const uint32_t*>(a_hi + (g + 8) * lda + k0);
                ah[2] = *reinterpret_cast<const uint32_t*>(a_hi + g * lda + k0 + 8);
                ah[3] = *reinterpret_cast<const uint32_t*>(a_hi + (g + 8) * lda + k0 + 8);
                al[0] = *reinterpret_cast<const uint32_t*>(a_lo + g * lda + k0);
                al[1] = *reinterpret_cast<const uint32_t*>(a_lo + (g + 8) * lda + k0);
                al[2] = *reinterpret_cast<const uint32_t*>(a_lo + g * lda + k0 + 8);
                al[3] = *reinterpret_cast<const uint32_t*>(a_lo + (g + 8) * lda + k0 + 8);
                mma16816(c, ah, b[ks][0], b[ks][1]);
                mma16816(c, al, b[ks][0], b[ks][1]);
            }
        }
        epi(g, nt * 8 + 2 * t, c[0], c[1]);
        epi(g + 8, nt * 8 + 2 * t, c[2], c[3]);
    }
}

__global__ void __cluster_dims__(KT_CL, 1, 1) __launch_bounds__(KT_THREADS) kg_tail_kernel(const KgTailArgs a) {
    extern __shared__ __align__(16) uint8_t kt_raw[];
    KtSmem& sm = *reinterpret_cast<KtSmem*>(kt_raw);
    cg::cluster_group cluster = cg::this_cluster();
    constexpr int D = 256, QD = D / KT_CL;   // this CTA's 64 columns of every 256-wide layer
    const int q = (int)cluster.block_rank(), b = blockIdx.y, tid = threadIdx.x, nt = blockDim.x, warp = tid >> 5, lane = tid & 31;
    KtSmem* peer[KT_CL];
#pragma unroll
    for (int r = 0; r < KT_CL; ++r) peer[r] = cluster.map_shared_rank(&sm, r);
    const int K = a.n_kg, gs = a.gs, heads = a.heads, n_chunks = KG_MAX_CHUNKS;
    const int n0 = a.node_ptr[b], nb = a.node_ptr[b + 1] - n0;
    // the image's partials: slots slot0 .. slot0 + nch - 1
    size_t slot0;
    int nch;
    if (a.tile_partials) {
        slot0 = (size_t)(n0 >> 7) + b;
        nch = nb > 0 ? ((n0 + nb - 1) >> 7) - (n0 >> 7) + 1 : 0;
    } else {
        const int per = ((nb + n_chunks - 1) / n_chunks + KG_SUB - 1) / KG_SUB * KG_SUB;
        slot0 = (size_t)b * n_chunks;
        nch = nb > 0 ? (nb + per - 1) / per : 0;
    }
    // ---- merge statistics of the partials per folded key slot (every CTA: 128 threads); groups of KG_MAX_CHUNKS partials,
    // one group unless an image spans more than 16 tiles
    float M = -INFINITY, inv = 0.f;
    if (tid < 128) {
        for (int g0 = 0; g0 < nch; g0 += KG_MAX_CHUNKS) {
            float pm[KG_MAX_CHUNKS];
#pragma unroll
            for (int ch = 0; ch < KG_MAX_CHUNKS; ++ch) pm[ch] = g0 + ch < nch ? a.part_m[(slot0 + g0 + ch) * 128 + tid] : -INFINITY;
#pragma unroll
            for (int ch = 0; ch < KG_MAX_CHUNKS; ++ch) M = fmaxf(M, pm[ch]);
        }
        float L = 0.f;
        for (int g0 = 0; g0 < nch; g0 += KG_MAX_CHUNKS) {
            float pm[KG_MAX_CHUNKS], pl[KG_MAX_CHUNKS];
#pragma unroll
            for (int ch = 0; ch < KG_MAX_CHUNKS; ++ch) {
                pm[ch] = g0 + ch < nch ? a.part_m[(slot0 + g0 + ch) * 128 + tid] : -INFINITY;
                pl[ch] = g0 + ch < nch ? a.part_l[(slot0 + g0 + ch) * 128 + tid] : 0.f;
            }
#pragma unroll
            for (int ch = 0; ch < KG_MAX_CHUNKS; ++ch)
                if (g0 + ch < nch) L += pl[ch] * expf(pm[ch] - M);
        }
        inv = L > 0.f ? 1.0f / L : 0.f;
        if (a.stat && q == 0) {
            a.stat[((size_t)b * 128 + tid) * 2] = M;
            a.stat[((size_t)b * 128 + tid) * 2 + 1] = inv;
        }
    }
    cluster.sync();   // every CTA of the cluster is running before the first write into a peer's shared memory
    // ---- this CTA's 64 columns of ctx2 [16, 256] (rows >= K zero) as fp16 hi / lo, broadcast; its columns of mean_n y
    const int hd = D / heads;
    constexpr int ITEMS = 16 * (QD / 2) / KT_THREADS;   // (row, column pair) items per thread
    float s0[ITEMS], s1[ITEMS];
#pragma unroll
    for (int it = 0; it < ITEMS; ++it) s0[it] = s1[it] = 0.f;
    for (int g0 = 0; g0 < nch || g0 == 0; g0 += KG_MAX_CHUNKS) {
        if (g0 > 0) __syncthreads();              // the previous group's factors have been consumed
        if (tid < 128) {
#pragma unroll
            for (int ch = 0; ch < KG_MAX_CHUNKS; ++ch)
                sm.sc[ch * 128 + tid] = g0 + ch < nch ? expf(a.part_m[(slot0 + g0 + ch) * 128 + tid] - M) * inv : 0.f;
        }
        __syncthreads();
#pragma unroll
        for (int it = 0; it < ITEMS; ++it) {
            const int i = tid + it * KT_THREADS, k = i / (QD / 2), c = q * QD + (i - k * (QD / 2)) * 2;
            if (k < K) {
                const int jj = (c / hd) * gs + k;   // c and c + 1 lie in the same head (hd is even)
                float2 pv[KG_MAX_CHUNKS];
#pragma unroll
                for (int ch = 0; ch < KG_MAX_CHUNKS; ++ch)
                    pv[ch] = g0 + ch < nch ? *reinterpret_cast<const float2*>(a.part_acc + ((slot0 + g0 + ch) * gs + k) * D + c) : make_float2(0.f, 0.f);
#pragma unroll
                for (int ch = 0; ch < KG_MAX_CHUNKS; ++ch) {
                    s0[it] = fmaf(sm.sc[ch * 128 + jj], pv[ch].x, s0[it]);
                    s1[it] = fmaf(sm.sc[ch * 128 + jj], pv[ch].y, s1[it]);
                }
            }
        }
    }
#pragma unroll
    for (int it = 0; it < ITEMS; ++it) {
        const int i = tid + it * KT_THREADS, k = i / (QD / 2), c = q * QD + (i - k * (QD / 2)) * 2;
        if (a.v2_bias && k < K && nb > 0) {   // softmax weights sum to one: the bias of V passes through the attention
            s0[it] += a.v2_bias[c];
            s1[it] += a.v2_bias[c + 1];
        }
        __half2 hi, lo;
        split2(s0[it], s1[it], &hi, &lo);
#pragma unroll
        for (int r = 0; r < KT_CL; ++r) {
            *reinterpret_cast<__half2*>(peer[r]->a_hi + k * KT_LDA + c) = hi;
            *reinterpret_cast<__half2*>(peer[r]->a_lo + k * KT_LDA + c) = lo;
        }
    }
    if (tid < QD) {
        const int c = q * QD + tid;
        float s = 0.f;
        for (int g0 = 0; g0 < nch; g0 += KG_MAX_CHUNKS) {
            float pv[KG_MAX_CHUNKS];
#pragma unroll
            for (int ch = 0; ch < KG_MAX_CHUNKS; ++ch) pv[ch] = g0 + ch < nch ? a.part_y[(slot0 + g0 + ch) * D + c] : 0.f;
#pragma unroll
            for (int ch = 0; ch < KG_MAX_CHUNKS; ++ch) s += pv[ch];
        }
        s = nb > 0 ? s / (float)nb : 0.f;
#pragma unroll
        for (int r = 0; r < KT_CL; ++r) peer[r]->both[c] = s;
    }
    cluster.sync();
    // ---- pre-LN rows kp + ctx2 Wo^T + bo: this CTA's 64 columns, broadcast
    kt_gemm16<256>(sm.a_hi, sm.a_lo, KT_LDA, a.wo2, q * (QD / 8), (q + 1) * (QD / 8), [&](int row, int col, float v0, float v1) {
        const float2 bo = *reinterpret_cast<const float2*>(a.bo2 + col);
        float2 kp = make_float2(0.f, 0.f);
        if (row < K) kp = *reinterpret_cast<const float2*>(a.kp + (size_t)row * D + col);
        const float2 v = make_float2(v0 + bo.x + kp.x, v1 + bo.y + kp.y);
#pragma unroll
        for (int r = 0; r < KT_CL; ++r) *reinterpret_cast<float2*>(peer[r]->z + row * D + col) = v;
    });
    cluster.sync();
    // ---- z0 = LayerNorm(row) (eps 1e-5, biased variance): every CTA normalises all 16 rows for itself, one warp per row
    for (int row = warp; row < 16; row += nt >> 5) {
        float x[8], s1 = 0.f, s2 = 0.f;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            x[j] = sm.z[row * D + lane * 8 + j];
            s1 += x[j];
        }
        s1 = warp_sum(s1);
        const float mean = s1 * (1.0f / 256.f);
#pragma unroll
        for (int j = 0; j < 8; ++j) s2 = fmaf(x[j] - mean, x[j] - mean, s2);
        s2 = warp_sum(s2);
        const float rstd = rsqrtf(s2 * (1.0f / 256.f) + 1e-5f);
#pragma unroll
        for (int j = 0; j < 8; j += 2) {
            const int c = lane * 8 + j;
            float v0 = 0.f, v1 = 0.f;
            if (row < K) {
                v0 = (x[j] - mean) * rstd * a.ln_g[c] + a.ln_b[c];
                v1 = (x[j + 1] - mean) * rstd * a.ln_g[c + 1] + a.ln_b[c + 1];
            }
            sm.z[row * D + c] = v0;
            sm.z[row * D + c + 1] = v1;
            split2(v0, v1, reinterpret_cast<__half2*>(sm.a_hi + row * KT_LDA + c), reinterpret_cast<__half2*>(sm.a_lo + row * KT_LDA + c));
        }
    }
    __syncthreads();
    // ---- hk = relu(z0 W1^T + b1) [16, 512]: this CTA's 128 columns, broadcast as hi / lo
    kt_gemm16<256>(sm.a_hi, sm.a_lo, KT_LDA, a.w1, q * (2 * QD / 8), (q + 1) * (2 * QD / 8), [&](int row, int col, float v0, float v1) {
        const float2 b1 = *reinterpret_cast<const float2*>(a.b1 + col);
        __half2 hi, lo;
        split2(fmaxf(v0 + b1.x, 0.f), fmaxf(v1 + b1.y, 0.f), &hi, &lo);
#pragma unroll
        for (int r = 0; r < KT_CL; ++r) {
            *reinterpret_cast<__half2*>(peer[r]->h_hi + row * KT_LDH + col) = hi;
            *reinterpret_cast<__half2*>(peer[r]->h_lo + row * KT_LDH + col) = lo;
        }
    });
    cluster.sync();
    // ---- z = z0 + hk W2^T + b2, this CTA's 64 columns (in place: each cell is read and written by the same thread)
    kt_gemm16<512>(sm.h_hi, sm.h_lo, KT_LDH, a.w2, q * (QD / 8), (q + 1) * (QD / 8), [&](int row, int col, float v0, float v1) {
        const float2 b2 = *reinterpret_cast<const float2*>(a.b2 + col);
        float2* cell = reinterpret_cast<float2*>(sm.z + row * D + col);
        const float2 z0 = *cell;
        *cell = make_float2(z0.x + v0 + b2.x, z0.y + v1 + b2.y);
    });
    __syncthreads();
    if (tid < QD) {   // mean over the KG rows (:135) of this CTA's columns, broadcast
        const int c = q * QD + tid;
        float s = 0.f;
        for (int r = 0; r < K; ++r) s += sm.z[r * D + c];
        s /= (float)K;
#pragma unroll
        for (int r = 0; r < KT_CL; ++r) peer[r]->both[D + c] = s;
    }
    cluster.sync();
    // ---- fusion layer (:138-139): this CTA's 64 outputs of each Linear, fp32 as stored, broadcast
    const GraphTailArgs& gta = a.gt;
    {
        const float* wq = gta.fl1_w + (size_t)q * QD * 2 * D;
        const float* bq = gta.fl1_b + q * QD;
        gemv_warp(sm.both, &wq, &bq, QD, sm.hid, QD, 2 * D, true);
        __syncthreads();
        if (tid < QD) {
#pragma unroll
            for (int r = 0; r < KT_CL; ++r) peer[r]->f1[q * QD + tid] = sm.hid[tid];
        }
    }
    cluster.sync();
    {
        const float* wq = gta.fl2_w + (size_t)q * QD * D;
        const float* bq = gta.fl2_b + q * QD;
        gemv_warp(sm.f1, &wq, &bq, QD, sm.hid, QD, D, false);
        __syncthreads();
        if (tid < QD) {
            const float v = sm.hid[tid];
            if (gta.fused) gta.fused[(size_t)b * D + q * QD + tid] = v;
#pragma unroll
            for (int r = 0; r < KT_CL; ++r) peer[r]->fz[q * QD + tid] = v;
        }
    }
    cluster.sync();   // the last remote writes: after this barrier a CTA only touches its own shared memory
    if (!gta.want_heads) return;
    // ---- head q (:239-242): mask, instance, edge, score -- one head per CTA of the cluster
    const int H2 = D / 2, nwarps = nt >> 5;
    gemv_warp(sm.fz, &gta.hw1[q], &gta.hb1[q], H2, sm.hid, H2, D, true);
    __syncthreads();
    const int cdim = q < 2 ? gta.num_classes : 1;
    for (int j = warp; j < cdim; j += nwarps) {
        const float* wrow = gta.hw2[q] + (size_t)j * H2;
        float s = 0.f;
        for (int k = lane; k < H2; k += 32) s = fmaf(sm.hid[k], wrow[k], s);
        s = warp_sum(s) + gta.hb2[q][j];
        if (q == 3) s = 1.0f / (1.0f + expf(-s));   // score head ends in nn.Sigmoid (fusion_model.py:229-235)
        float* out = q == 0 ? gta.mask_out : q == 1 ? gta.inst_out : q == 2 ? gta.edge_out : gta.score_out;
        if (lane == 0 && out) out[(size_t)b * cdim + j] = s;
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

int launch_kg_tail(const KgTailArgs& a, const KgSideArgs& k, int n_graphs, cudaStream_t st) {
    static PerDeviceOnce attr;
    if (attr.first()) cudaFuncSetAttribute(kg_tail_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(KtSmem));
    {
        CM_TIMED("kg_tail", st);
        kg_tail_kernel<<<dim3(KT_CL, n_graphs), KT_THREADS, sizeof(KtSmem), st>>>(a);
    }
    int rc = post_launch("kg_tail");
    if (rc != CM_OK || !k.attn_kg2rg) return rc;
    {
        CM_TIMED("kg2rg_weights", st);
        kg2rg_weights_kernel<<<ceil_div((long long)k.node_cap * a.n_kg, 256), 256, 0, st>>>(k.s2, a.stat, a.node_ptr, n_graphs, a.n_kg, a.heads, a.gs,
                                                                                           k.attn_kg2rg);
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
