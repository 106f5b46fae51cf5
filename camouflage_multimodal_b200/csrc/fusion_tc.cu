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

#include "dense_tc.cuh"
#include "fusion_tc.cuh"
#include "linear.cuh"

namespace cm {

static size_t al(size_t v) { return (v + 255) / 256 * 256; }

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
    // KG-side and graph-level layers run in split precision: weights stored [hi | hi | lo] along K (3x)
    f.wo2 = take(3 * D * D * 2);
    f.kffn1 = take(3 * 2 * D * D * 2);
    f.kffn2 = take(3 * 2 * D * D * 2);
    f.fl1 = take(3 * 2 * D * D * 2);
    f.fl2 = take(3 * D * D * 2);
    f.hw1 = take(3 * 4 * (D / 2) * D * 2);
    f.kp_h = take(K * D * 2);
    f.ccat = take(256 * 4);
    f.hb1 = take(4 * (D / 2) * 4);
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

__global__ void concat_heads1_kernel(cm_fusion_params p, op_t* __restrict__ hw1, float* __restrict__ hb1) {
    const int D = p.hidden, H2 = D / 2;
    const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x, stride = (long long)gridDim.x * blockDim.x;
    for (long long i = tid; i < 4LL * H2 * D; i += stride) {
        const int t = (int)(i / ((long long)H2 * D)), o = (int)(i / D), c = (int)(i % D);
        const float v = p.head_w1[t][i - (long long)t * H2 * D];
        const op_t hi = f2op(v);
        op_t* w = hw1 + (size_t)o * 3 * D + c;
        w[0] = hi;
        w[D] = hi;
        w[2 * D] = f2op(v - op2f(hi));
    }
    for (long long i = tid; i < 4LL * H2; i += stride) hb1[i] = p.head_b1[(int)(i / H2)][i % H2];
}

// kg -> rg attention on the folded scores: grid (n_graphs, n_kg), one thread per channel c (head h = c / hd).
//   p2[h][n] = softmax over the image's nodes n of s2[n, h*gs + k];  ctx2[b*K + k, c] = sum_n p2[h(c)][n] v2[n, c]
// attn (optional) block [n_kg, N_b] at n_kg*node_ptr[b]: mean over heads of p2.
__global__ void __launch_bounds__(256) attn_kg2rg_fold_kernel(const float* __restrict__ s2, const op_t* __restrict__ v2, int D,
                                                              int heads, int gs, const int* __restrict__ node_ptr, int n_kg,
                                                              op_t* __restrict__ ctx2, float* __restrict__ attn) {
    __shared__ float s_max[32], s_inv[32];
    __shared__ float s_p[8][33];
    const int b = blockIdx.x, k = blockIdx.y, c = threadIdx.x;
    const int warp = c >> 5, lane = c & 31, nwarps = blockDim.x >> 5;
    const int n0 = node_ptr[b], n1 = node_ptr[b + 1];
    const int hd = D / heads;
    for (int h = warp; h < heads; h += nwarps) {
        const float* col = s2 + h * gs + k;
        float mx = -INFINITY;
        for (int n = n0 + lane; n < n1; n += 32) mx = fmaxf(mx, col[(size_t)n * 128]);
        mx = warp_max(mx);
        float sum = 0.f;
        for (int n = n0 + lane; n < n1; n += 32) sum += expf(col[(size_t)n * 128] - mx);
        sum = warp_sum(sum);
        if (lane == 0) {
            s_max[h] = mx;
            s_inv[h] = n1 > n0 ? 1.0f / sum : 0.f;
        }
    }
    __syncthreads();
    const int my_h = c / hd;
    float acc = 0.f;
    for (int base = n0; base < n1; base += 32) {
        for (int t = c; t < heads * 32; t += blockDim.x) {
            const int h = t >> 5, i = t & 31, n = base + i;
            s_p[h][i] = n < n1 ? expf(s2[(size_t)n * 128 + h * gs + k] - s_max[h]) * s_inv[h] : 0.f;
        }
        __syncthreads();
        const int cnt = min(32, n1 - base);
        if (c < D)
            for (int i = 0; i < cnt; ++i) acc = fmaf(s_p[my_h][i], op2f(v2[(size_t)(base + i) * D + c]), acc);
        if (attn && c < cnt) {
            float tot = 0.f;
            for (int h = 0; h < heads; ++h) tot += s_p[h][c];
            attn[(size_t)n_kg * n0 + (size_t)k * (n1 - n0) + (base - n0) + c] = tot / (float)heads;
        }
        __syncthreads();
    }
    if (c < D) {   // split row [hi | lo | hi]
        op_t* o = ctx2 + ((size_t)b * n_kg + k) * 3 * D + c;
        const op_t hi = f2op(acc);
        o[0] = hi;
        o[D] = f2op(acc - op2f(hi));
        o[2 * D] = hi;
    }
}

// out[g, c] = mean over rows seg_ptr[g] .. (or fixed_len rows) of in, written as a split row: thirds of `third`
// columns [hi | lo | hi], this segment mean landing at column offset col0 of each third
__global__ void __launch_bounds__(128) segment_mean_split_kernel(const float* __restrict__ in, int dim, const int* __restrict__ seg_ptr,
                                                                 int fixed_len, op_t* __restrict__ out, int third, int col0) {
    const int g = blockIdx.x, c = blockIdx.y * 128 + threadIdx.x;
    if (c >= dim) return;
    const int a = seg_ptr ? seg_ptr[g] : g * fixed_len;
    const int z = seg_ptr ? seg_ptr[g + 1] : a + fixed_len;
    float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
    int r = a;
    for (; r + 3 < z; r += 4) {
        s0 += in[(size_t)r * dim + c];
        s1 += in[(size_t)(r + 1) * dim + c];
        s2 += in[(size_t)(r + 2) * dim + c];
        s3 += in[(size_t)(r + 3) * dim + c];
    }
    for (; r < z; ++r) s0 += in[(size_t)r * dim + c];
    const int n = z - a;
    const float v = n > 0 ? ((s0 + s1) + (s2 + s3)) / (float)n : 0.f;
    op_t* o = out + (size_t)g * 3 * third + col0 + c;
    const op_t hi = f2op(v);
    o[0] = hi;
    o[third] = f2op(v - op2f(hi));
    o[2 * third] = hi;
}

static int seg_mean_split(const float* in, int dim, const int* seg_ptr, int fixed_len, int n_seg, op_t* out, int third, int col0,
                          cudaStream_t st) {
    dim3 grid(n_seg, ceil_div(dim, 128));
    {
        CM_TIMED("segment_mean", st);
        segment_mean_split_kernel<<<grid, 128, 0, st>>>(in, dim, seg_ptr, fixed_len, out, third, col0);
    }
    return post_launch("segment_mean");
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
    auto SW = [&](const float* w, int rows, int cols, size_t off) {
        return launch_split_weight(w, rows, cols, cols, nullptr, reinterpret_cast<op_t*>(base + off), st);
    };
    if ((rc = SW(p->k2r_out_w, D, D, f.wo2)) != CM_OK) return rc;
    if ((rc = SW(p->ffn_kg_w1, 2 * D, D, f.kffn1)) != CM_OK) return rc;
    if ((rc = SW(p->ffn_kg_w2, D, 2 * D, f.kffn2)) != CM_OK) return rc;
    if ((rc = SW(p->fl_w1, D, 2 * D, f.fl1)) != CM_OK) return rc;
    if ((rc = SW(p->fl_w2, D, D, f.fl2)) != CM_OK) return rc;
    if ((rc = cast_weight(kp, base + f.kp_h, (size_t)K * D, st)) != CM_OK) return rc;
    if (p->head_w1[0]) {
        {
            CM_TIMED("fusion_fold", st);
            concat_heads1_kernel<<<64, 256, 0, st>>>(*p, reinterpret_cast<op_t*>(base + f.hw1), reinterpret_cast<float*>(base + f.hb1));
        }
        if ((rc = post_launch("fusion_fold(heads)")) != CM_OK) return rc;
    }
    return CM_OK;
}

struct FusionTcWs {
    op_t *rgb, *r, *p1, *v2, *y0, *hh, *ctx2, *z0, *hk, *both, *f1, *fusedb;
    float *s2, *y, *z, *z0f, *fused, *hid;
    size_t total;
};

static FusionTcWs carve_fusion_tc(void* ws, const cm_fusion_params* p, int G, int M, int Kk) {
    Carver c(ws);
    FusionTcWs w;
    const size_t D = p->hidden, m = M, gk = (size_t)G * Kk;
    w.rgb = c.take<op_t>(m * p->rg_dim);
    w.r = c.take<op_t>(m * D);
    w.p1 = c.take<op_t>(m * 128);
    w.v2 = c.take<op_t>(m * D);
    w.y0 = c.take<op_t>(m * D);
    w.hh = c.take<op_t>(m * 2 * D);
    w.ctx2 = c.take<op_t>(gk * 3 * D);
    w.z0 = c.take<op_t>(gk * 3 * D);
    w.hk = c.take<op_t>(gk * 3 * 2 * D);
    w.both = c.take<op_t>((size_t)G * 3 * 2 * D);
    w.f1 = c.take<op_t>((size_t)G * 3 * D);
    w.fusedb = c.take<op_t>((size_t)G * 3 * D);
    w.z0f = c.take<float>(gk * D);
    w.s2 = c.take<float>(m * 128);
    w.y = c.take<float>(m * D);
    w.z = c.take<float>(gk * D);
    w.fused = c.take<float>((size_t)G * D);
    w.hid = c.take<float>((size_t)G * 4 * (D / 2));
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
    const int D = p->hidden, GK = G * Kk;
    const int* m_dev = node_ptr + G;
    int rc;
    Dense d;

    if ((rc = launch_cast_f16(rg, p->rg_dim, M, p->rg_dim, m_dev, w.rgb, p->rg_dim, st)) != CM_OK) return rc;
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
    if ((rc = seg_mean_split(w.y, D, node_ptr, 0, G, w.both, 2 * D, 0, st)) != CM_OK) return rc;

    // ---- kg attends to rg (:123-131).  From here on rows are per (image, kg row) or per image: few rows whose
    // rounding errors are not averaged over hundreds of nodes, so these layers run in split precision
    // (activations [hi | lo | hi] x weights [hi | hi | lo], K tripled).
    {
        CM_TIMED("attn_kg2rg", st);
        attn_kg2rg_fold_kernel<<<dim3(G, Kk), D, 0, st>>>(w.s2, w.v2, D, p->heads, f.gs, node_ptr, Kk, w.ctx2, attn_kg2rg);
    }
    if ((rc = post_launch("attn_kg2rg")) != CM_OK) return rc;
    d = Dense(); d.A = w.ctx2; d.lda = 3 * D; d.W = W(f.wo2); d.M = GK; d.N = D; d.K = 3 * D;
    d.bias = p->k2r_out_b; d.residual_f32 = reinterpret_cast<const float*>(base + f.kp); d.ldr = D; d.res_mod = Kk;
    d.ln_gamma = p->ln_kg_w; d.ln_beta = p->ln_kg_b; d.out_h = w.z0; d.ldo_h = 3 * D; d.split_out = 1; d.out_f32 = w.z0f; d.ldo_f32 = D;
    if ((rc = launch_dense(d, st)) != CM_OK) return rc;
    d = Dense(); d.A = w.z0; d.lda = 3 * D; d.W = W(f.kffn1); d.M = GK; d.N = 2 * D; d.K = 3 * D;
    d.bias = p->ffn_kg_b1; d.relu = 1; d.out_h = w.hk; d.ldo_h = 6 * D; d.split_out = 1;
    if ((rc = launch_dense(d, st)) != CM_OK) return rc;
    d = Dense(); d.A = w.hk; d.lda = 6 * D; d.W = W(f.kffn2); d.M = GK; d.N = D; d.K = 6 * D;
    d.bias = p->ffn_kg_b2; d.residual_f32 = w.z0f; d.ldr = D; d.out_f32 = w.z; d.ldo_f32 = D;
    if ((rc = launch_dense(d, st)) != CM_OK) return rc;
    if ((rc = seg_mean_split(w.z, D, nullptr, Kk, G, w.both, 2 * D, D, st)) != CM_OK) return rc;   // :135

    // ---- fusion layer (:138-139) and the four heads (:239-242)
    float* fused = fused_out ? fused_out : w.fused;
    d = Dense(); d.A = w.both; d.lda = 6 * D; d.W = W(f.fl1); d.M = G; d.N = D; d.K = 6 * D;
    d.bias = p->fl_b1; d.relu = 1; d.out_h = w.f1; d.ldo_h = 3 * D; d.split_out = 1;
    if ((rc = launch_dense(d, st)) != CM_OK) return rc;
    d = Dense(); d.A = w.f1; d.lda = 3 * D; d.W = W(f.fl2); d.M = G; d.N = D; d.K = 3 * D;
    d.bias = p->fl_b2; d.out_h = w.fusedb; d.ldo_h = 3 * D; d.split_out = 1; d.out_f32 = fused; d.ldo_f32 = D;
    if ((rc = launch_dense(d, st)) != CM_OK) return rc;
    if (mask_out || instance_out || edge_out || score_out) {
        const int H2 = D / 2;
        d = Dense(); d.A = w.fusedb; d.lda = 3 * D; d.W = W(f.hw1); d.M = G; d.N = 4 * H2; d.K = 3 * D;
        d.bias = reinterpret_cast<const float*>(base + f.hb1); d.relu = 1; d.out_f32 = w.hid; d.ldo_f32 = 4 * H2;
        if ((rc = launch_dense(d, st)) != CM_OK) return rc;
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
        if ((rc = launch_heads2(w.hid, 4 * H2, G, nullptr, h2, st)) != CM_OK) return rc;
    }
    return post_launch("cm_fusion_forward(tc)");
}

}  // namespace cm
