// Training-mode RegionGraphGNN: forward with BatchNorm batch statistics and caller-supplied dropout masks, and the
// full backward pass (models/region_graph/extract_rg_embeddings.py:54-90 under model.train(), driven by
// models/region_graph/train.py:345-357: out = model(batch); loss.backward()).
//
// Message passing keeps the shape of the inference path: warp-per-target-node gathers over the CSR rows, the GAT
// layer aggregated in the 15-d input space (z = sum_j alpha_ij x_j, then one [60 -> 128] linear), GCN aggregation
// applied before the layer's linear.  Backward of an aggregation is the same gather over the REVERSE edges: the
// region adjacency is symmetric, so the out-edges of node j are the entries of row j and `rev[e]` (built once per
// graph) names the slot of the opposite direction, whose weight / attention coefficient is the one needed.  No
// float atomics anywhere: every gradient is a gather or a fixed-order reduction, so a step is bit-reproducible.
// Dense layers are fp32 (train_common.cu); saved activations live in the caller's workspace between the two calls.
#include <type_traits>

#include "train_common.cuh"

namespace cm {

struct GnnTrainWs {
    float *u_src, *u_dst, *wf;
    float *asrc, *adst, *gmax, *gden, *z;
    float *y[4], *a[4], *g[3], *dis, *e, *t[3];
    float *mean, *invstd;   // [4][H]
    double *bn_sums, *bn_n;  // [2H + 1] column sums (+ row count) of the layer in flight; [4] global row count per layer
    unsigned* bn_counter;    // [ceil(H / 128)] arrival counters of bn_stats_kernel (zeroed by the forward, self-resetting)
    int* rev;
    // backward scratch
    float *dA, *dB, *dT, *dz, *dpre_edge, *dpre_self, *dasrc, *dadst, *du, *dwf, *s1, *s2;
    double* part_d;
    float* part_f;
    int splits;
    size_t total;
};

static int gnn_train_splits(int n) { return n >= 1024 ? (n / 512 < 64 ? n / 512 : 64) : 1; }

static GnnTrainWs carve_gnn_train(void* ws, const cm_gnn_params* p, int n, int e) {
    Carver c(ws);
    GnnTrainWs w;
    const size_t N = (size_t)(n > 0 ? n : 1), E = (size_t)(e > 0 ? e : 1), H = p->hidden, C = p->in_channels, hd = p->heads, H2 = H / 2;
    w.u_src = c.take<float>(hd * C);
    w.u_dst = c.take<float>(hd * C);
    w.wf = c.take<float>(H * hd * C);
    w.asrc = c.take<float>(N * hd);
    w.adst = c.take<float>(N * hd);
    w.gmax = c.take<float>(N * hd);
    w.gden = c.take<float>(N * hd);
    w.z = c.take<float>(N * hd * C);
    for (int k = 0; k < 4; ++k) w.y[k] = c.take<float>(N * H);
    for (int k = 0; k < 4; ++k) w.a[k] = c.take<float>(N * H);
    for (int k = 0; k < 3; ++k) w.g[k] = c.take<float>(N * H);
    w.dis = c.take<float>(N);
    w.e = c.take<float>(N * H);
    for (int k = 0; k < 3; ++k) w.t[k] = c.take<float>(N * H2);
    w.mean = c.take<float>(4 * H);
    w.invstd = c.take<float>(4 * H);
    w.bn_sums = c.take<double>(2 * H + 1);
    w.bn_n = c.take<double>(4);
    w.bn_counter = c.take<unsigned>((H + 127) / 128);
    w.rev = c.take<int>(E);
    w.dA = c.take<float>(N * H);
    w.dB = c.take<float>(N * H);
    w.dT = c.take<float>(N * H2);
    w.dz = c.take<float>(N * hd * C);
    w.dpre_edge = c.take<float>(E * hd);
    w.dpre_self = c.take<float>(N * hd);
    w.dasrc = c.take<float>(N * hd);
    w.dadst = c.take<float>(N * hd);
    w.du = c.take<float>(2 * hd * C);
    w.dwf = c.take<float>(H * hd * C);
    w.s1 = c.take<float>(H);
    w.s2 = c.take<float>(H);
    w.part_d = c.take<double>((size_t)col_parts(n) * 2 * H);
    w.splits = gnn_train_splits(n);
    w.part_f = c.take<float>((size_t)w.splits * H * H);
    w.total = align_up(c.off, 256);
    return w;
}

// ---------------------------------------------------------------- GAT
// u_src/u_dst [heads, C] = W_h^T att_h; wf [H, heads*C] = W[h*H+o, c] / heads (mean over heads folded in)
__global__ void gat_fold_train_kernel(const float* __restrict__ W, const float* __restrict__ att_src, const float* __restrict__ att_dst, int H,
                                      int C, int hd, float* __restrict__ u_src, float* __restrict__ u_dst, float* __restrict__ wf) {
    const int tid = blockIdx.x * blockDim.x + threadIdx.x, stride = gridDim.x * blockDim.x;
    for (int i = tid; i < hd * C; i += stride) {
        const int h = i / C, c = i % C;
        float s = 0.f, d = 0.f;
        for (int o = 0; o < H; ++o) {
            const float w = W[(size_t)(h * H + o) * C + c];
            s = fmaf(w, att_src[h * H + o], s);
            d = fmaf(w, att_dst[h * H + o], d);
        }
        u_src[i] = s;
        u_dst[i] = d;
    }
    for (int i = tid; i < H * hd * C; i += stride) {
        const int o = i / (hd * C), hc = i % (hd * C), h = hc / C, c = hc % C;
        wf[i] = W[(size_t)(h * H + o) * C + c] / (float)hd;
    }
}

__global__ void gat_logits_train_kernel(const float* __restrict__ x, int C, int hd, const float* __restrict__ u_src,
                                        const float* __restrict__ u_dst, int n, float* __restrict__ a_src, float* __restrict__ a_dst) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n * hd) return;
    const int g = i / hd, h = i - g * hd;
    const float* xr = x + (size_t)g * C;
    float s = 0.f, d = 0.f;
    for (int c = 0; c < C; ++c) {
        s = fmaf(xr[c], u_src[h * C + c], s);
        d = fmaf(xr[c], u_dst[h * C + c], d);
    }
    a_src[i] = s;
    a_dst[i] = d;
}

__device__ __forceinline__ float lrelu02(float v) { return v > 0.f ? v : 0.2f * v; }

// one warp per target node; lane c < C owns input feature c for all heads.  Saves the softmax max and denominator.
template <int HEADS>
__global__ void __launch_bounds__(256) gat_agg_train_fwd_kernel(const float* __restrict__ x, int C, const int* __restrict__ row_ptr,
                                                                const int* __restrict__ col, const float* __restrict__ a_src,
                                                                const float* __restrict__ a_dst, int n, float* __restrict__ z,
                                                                float* __restrict__ gmax, float* __restrict__ gden) {
    const int g = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (g >= n) return;
    const int e0 = row_ptr[g], e1 = row_ptr[g + 1];
    float ad[HEADS], mx[HEADS], den[HEADS], acc[HEADS];
#pragma unroll
    for (int h = 0; h < HEADS; ++h) {
        ad[h] = a_dst[(size_t)g * HEADS + h];
        mx[h] = lrelu02(a_src[(size_t)g * HEADS + h] + ad[h]);   // the added self loop
        den[h] = 0.f;
        acc[h] = 0.f;
    }
    for (int e = e0; e < e1; ++e) {
        const int j = col[e];
        if (j == g) continue;   // PyG removes existing self loops first
#pragma unroll
        for (int h = 0; h < HEADS; ++h) mx[h] = fmaxf(mx[h], lrelu02(a_src[(size_t)j * HEADS + h] + ad[h]));
    }
    for (int e = e0; e <= e1; ++e) {   // e == e1 is the self loop
        const int j = e < e1 ? col[e] : g;
        if (e < e1 && j == g) continue;
        const float xv = lane < C ? x[(size_t)j * C + lane] : 0.f;
#pragma unroll
        for (int h = 0; h < HEADS; ++h) {
            const float w = expf(lrelu02(a_src[(size_t)j * HEADS + h] + ad[h]) - mx[h]);
            den[h] += w;
            acc[h] = fmaf(w, xv, acc[h]);
        }
    }
    if (lane < C) {
#pragma unroll
        for (int h = 0; h < HEADS; ++h) z[(size_t)g * HEADS * C + h * C + lane] = acc[h] / (den[h] + 1e-16f);
    }
    if (lane < HEADS) {
        float m = 0.f, d = 0.f;
#pragma unroll
        for (int h = 0; h < HEADS; ++h)
            if (lane == h) { m = mx[h]; d = den[h] + 1e-16f; }
        gmax[(size_t)g * HEADS + lane] = m;
        gden[(size_t)g * HEADS + lane] = d;
    }
}

// Backward of the attention softmax for target node g (one warp, lanes over the incoming edges incl. the self loop):
//   dalpha_j = <dz[g,h,:], x_j>;  ds_j = alpha_j (dalpha_j - sum_k alpha_k dalpha_k);  dpre_j = ds_j * LeakyReLU'(pre_j)
// dpre of edge slot e goes to dpre_edge[e,h], of the self loop to dpre_self[g,h]; d a_dst[g,h] = sum_j dpre_j.
template <int HEADS>
__global__ void __launch_bounds__(256) gat_agg_train_bwd_kernel(const float* __restrict__ x, int C, const int* __restrict__ row_ptr,
                                                                const int* __restrict__ col, const float* __restrict__ a_src,
                                                                const float* __restrict__ a_dst, const float* __restrict__ gmax,
                                                                const float* __restrict__ gden, const float* __restrict__ dz, int n,
                                                                float* __restrict__ dpre_edge, float* __restrict__ dpre_self,
                                                                float* __restrict__ dadst) {
    extern __shared__ float gat_dz_s[];   // [8 warps][HEADS*C]
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = blockIdx.x * 8 + warp;
    if (g >= n) return;
    float* dzs = gat_dz_s + warp * HEADS * C;
    for (int i = lane; i < HEADS * C; i += 32) dzs[i] = dz[(size_t)g * HEADS * C + i];
    __syncwarp();
    const int e0 = row_ptr[g], e1 = row_ptr[g + 1];
    float ad[HEADS], mx[HEADS], rden[HEADS], S[HEADS], sum_dpre[HEADS];
#pragma unroll
    for (int h = 0; h < HEADS; ++h) {
        ad[h] = a_dst[(size_t)g * HEADS + h];
        mx[h] = gmax[(size_t)g * HEADS + h];
        rden[h] = 1.0f / gden[(size_t)g * HEADS + h];
        S[h] = 0.f;
        sum_dpre[h] = 0.f;
    }
    for (int pass = 0; pass < 2; ++pass) {
        for (int eb = e0; eb <= e1; eb += 32) {
            const int e = eb + lane;
            const bool self = e == e1;
            const int j = e < e1 ? col[e] : g;
            const bool live = e <= e1 && (self || j != g);
            float alpha[HEADS], dal[HEADS], pre[HEADS];
#pragma unroll
            for (int h = 0; h < HEADS; ++h) alpha[h] = dal[h] = pre[h] = 0.f;
            if (live) {
                const float* xj = x + (size_t)j * C;
#pragma unroll
                for (int h = 0; h < HEADS; ++h) {
                    pre[h] = a_src[(size_t)j * HEADS + h] + ad[h];
                    alpha[h] = expf(lrelu02(pre[h]) - mx[h]) * rden[h];
                }
                for (int c = 0; c < C; ++c) {
                    const float xv = xj[c];
#pragma unroll
                    for (int h = 0; h < HEADS; ++h) dal[h] = fmaf(dzs[h * C + c], xv, dal[h]);
                }
            }
            if (pass == 0) {
#pragma unroll
                for (int h = 0; h < HEADS; ++h) S[h] += warp_sum(alpha[h] * dal[h]);
            } else {
#pragma unroll
                for (int h = 0; h < HEADS; ++h) {
                    const float dp = alpha[h] * (dal[h] - S[h]) * (pre[h] > 0.f ? 1.0f : 0.2f);
                    if (live) {
                        if (self) dpre_self[(size_t)g * HEADS + h] = dp;
                        else dpre_edge[(size_t)e * HEADS + h] = dp;
                    }
                    sum_dpre[h] += warp_sum(live ? dp : 0.f);
                }
            }
        }
    }
    if (lane == 0) {
#pragma unroll
        for (int h = 0; h < HEADS; ++h) dadst[(size_t)g * HEADS + h] = sum_dpre[h];
    }
    // existing self-loop slots (removed by PyG) carry no gradient
    for (int e = e0 + lane; e < e1; e += 32)
        if (col[e] == g)
            for (int h = 0; h < HEADS; ++h) dpre_edge[(size_t)e * HEADS + h] = 0.f;
}

// d a_src[j,h] = dpre_self[j,h] + sum over the out-edges of j (= reverse slots of row j)
__global__ void gat_src_gather_kernel(const int* __restrict__ row_ptr, const int* __restrict__ rev, const float* __restrict__ dpre_edge,
                                      const float* __restrict__ dpre_self, int n, int hd, float* __restrict__ dasrc) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n * hd) return;
    const int j = i / hd, h = i - j * hd;
    float s = dpre_self[i];
    for (int e = row_ptr[j]; e < row_ptr[j + 1]; ++e) s += dpre_edge[(size_t)rev[e] * hd + h];
    dasrc[i] = s;
}

// dW[h*H+o, c] = dwf[o, h*C+c] / heads + att_src[h,o] du_src[h,c] + att_dst[h,o] du_dst[h,c];
// datt_src[h,o] = sum_c W[h*H+o, c] du_src[h,c] (same for dst)
__global__ void gat_param_grads_kernel(const float* __restrict__ W, const float* __restrict__ att_src, const float* __restrict__ att_dst,
                                       const float* __restrict__ dwf, const float* __restrict__ du_src, const float* __restrict__ du_dst,
                                       int H, int C, int hd, float* __restrict__ dW, float* __restrict__ datt_src,
                                       float* __restrict__ datt_dst) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= hd * H) return;
    const int h = i / H, o = i - h * H;
    float ds = 0.f, dd = 0.f;
    for (int c = 0; c < C; ++c) {
        const float w = W[(size_t)i * C + c], us = du_src[h * C + c], ud = du_dst[h * C + c];
        ds = fmaf(w, us, ds);
        dd = fmaf(w, ud, dd);
        dW[(size_t)i * C + c] = dwf[(size_t)o * hd * C + h * C + c] / (float)hd + att_src[i] * us + att_dst[i] * ud;
    }
    datt_src[i] = ds;
    datt_dst[i] = dd;
}

// rev[e] for e = (row i, neighbour j): the slot of i in row j; CM_STATUS_EDGE_OVERFLOW if the graph is not symmetric
__global__ void csr_reverse_kernel(const int* __restrict__ row_ptr, const int* __restrict__ col, int n, int* __restrict__ rev,
                                   uint32_t* status) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    for (int e = row_ptr[i]; e < row_ptr[i + 1]; ++e) {
        const int j = col[e];
        int r = -1;
        if (j >= 0 && j < n)
            for (int f = row_ptr[j]; f < row_ptr[j + 1]; ++f)
                if (col[f] == i) { r = f; break; }
        if (r < 0) {
            if (status) atomicOr(status, CM_STATUS_EDGE_OVERFLOW);
            r = e;
        }
        rev[e] = r;
    }
}

// ---------------------------------------------------------------- GCN aggregation
__global__ void gcn_dis_train_kernel(const int* __restrict__ row_ptr, const int* __restrict__ col, const float* __restrict__ w,
                                     const float* __restrict__ self_w, int n, float* __restrict__ dis) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float d = 0.f;
    for (int e = row_ptr[i]; e < row_ptr[i + 1]; ++e)
        if (col[e] != i) d += w ? w[e] : 1.0f;
    d += self_w ? self_w[i] : 1.0f;
    const float r = 1.0f / sqrtf(d);
    dis[i] = isinf(r) ? 0.f : r;
}

// out_i = dis_i * (sum_{e in row i} w[perm ? perm[e] : e] * dis_j * in_j + self_w_i * dis_i * in_i); one warp per node.
// perm == NULL: forward (weights of the edges INTO i); perm == rev: transpose (weights of the edges OUT of i).
__global__ void __launch_bounds__(256) gcn_agg_train_kernel(const float* __restrict__ in, const int* __restrict__ row_ptr,
                                                            const int* __restrict__ col, const float* __restrict__ w,
                                                            const int* __restrict__ perm, const float* __restrict__ self_w,
                                                            const float* __restrict__ dis, int n, int H, float* __restrict__ out) {
    const int g = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (g >= n) return;
    const float di = dis[g], sw = self_w ? self_w[g] : 1.0f;
    const int e0 = row_ptr[g], e1 = row_ptr[g + 1];
    for (int c = lane * 4; c < H; c += 128) {
        const float4 s = *reinterpret_cast<const float4*>(in + (size_t)g * H + c);
        const float ws = sw * di;
        float4 acc = make_float4(ws * s.x, ws * s.y, ws * s.z, ws * s.w);
        for (int e = e0; e < e1; ++e) {
            const int j = col[e];
            if (j == g) continue;
            const float we = (w ? w[perm ? perm[e] : e] : 1.0f) * dis[j];
            const float4 v = *reinterpret_cast<const float4*>(in + (size_t)j * H + c);
            acc.x = fmaf(we, v.x, acc.x); acc.y = fmaf(we, v.y, acc.y); acc.z = fmaf(we, v.z, acc.z); acc.w = fmaf(we, v.w, acc.w);
        }
        *reinterpret_cast<float4*>(out + (size_t)g * H + c) = make_float4(di * acc.x, di * acc.y, di * acc.z, di * acc.w);
    }
}

// ---------------------------------------------------------------- data-parallel BatchNorm hook
// One process per GPU: when set, every BatchNorm layer's column sums are copied to `buf`, the hook is called (the host
// all-reduces buf[0..count) over the ranks ON THE GIVEN STREAM) and the statistics are taken from the reduced sums.
static cm_bn_sync_fn g_bn_sync = nullptr;
static void* g_bn_sync_user = nullptr;
static double* g_bn_sync_buf = nullptr;
static int g_bn_sync_cap = 0;
static int bn_sync(double* sums, int count, cudaStream_t st) {
    if (!g_bn_sync) return CM_OK;
    if (count > g_bn_sync_cap) {
        set_error("BatchNorm sync buffer holds %d doubles, layer needs %d", g_bn_sync_cap, count);
        return CM_ERR_ARG;
    }
    cudaMemcpyAsync(g_bn_sync_buf, sums, sizeof(double) * count, cudaMemcpyDeviceToDevice, st);
    g_bn_sync(g_bn_sync_user, count, (void*)st);
    cudaMemcpyAsync(sums, g_bn_sync_buf, sizeof(double) * count, cudaMemcpyDeviceToDevice, st);
    return CM_OK;
}

// ---------------------------------------------------------------- BatchNorm (batch statistics) + ReLU + dropout
constexpr int BN_ROWS = 64;
__global__ void bn_fwd_part_kernel(const float* __restrict__ y, int n, int H, double* __restrict__ part) {
    const int c = blockIdx.y * blockDim.x + threadIdx.x;
    if (c >= H) return;
    const int r0 = blockIdx.x * BN_ROWS, r1 = min(n, r0 + BN_ROWS);
    double s = 0.0, q = 0.0;
    for (int r = r0; r < r1; ++r) {
        const double v = y[(size_t)r * H + c];
        s += v;
        q += v * v;
    }
    part[((size_t)blockIdx.x * 2) * H + c] = s;
    part[((size_t)blockIdx.x * 2 + 1) * H + c] = q;
}
// sums [2H + 1] = column sums of the parts (fixed order) and, in the last slot, the number of rows
__global__ void bn_reduce_parts_kernel(const double* __restrict__ part, int parts, int n, int H, double* __restrict__ sums) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c == 0) sums[2 * H] = (double)n;
    if (c >= H) return;
    double s = 0.0, q = 0.0;
    for (int p = 0; p < parts; ++p) {
        s += part[((size_t)p * 2) * H + c];
        q += part[((size_t)p * 2 + 1) * H + c];
    }
    sums[c] = s;
    sums[H + c] = q;
}
// sums = (sum x, sum x^2, n) of ALL rows of the batch -- this rank's, or every rank's after the data-parallel hook
__global__ void bn_fwd_final_kernel(const double* __restrict__ sums, int H, float eps, float* __restrict__ mean, float* __restrict__ invstd,
                                    float* __restrict__ batch_mean, float* __restrict__ batch_var, double* __restrict__ n_out) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= H) return;
    const double n = sums[2 * H];
    if (c == 0) *n_out = n;
    const double m = sums[c] / n;
    double v = sums[H + c] / n - m * m;   // biased variance (BatchNorm1d training normalisation)
    v = v > 0.0 ? v : 0.0;
    mean[c] = (float)m;
    invstd[c] = (float)(1.0 / sqrt(v + (double)eps));
    if (batch_mean) batch_mean[c] = (float)m;
    if (batch_var) batch_var[c] = (float)v;
}
__global__ void bn_fwd_apply_kernel(const float* __restrict__ y, const float* __restrict__ mean, const float* __restrict__ invstd,
                                    const float* __restrict__ gamma, const float* __restrict__ beta, const uint8_t* __restrict__ keep,
                                    float scale, long count, int H, float* __restrict__ a) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    const int c = (int)(i % H);
    float v = fmaxf((y[i] - mean[c]) * invstd[c] * gamma[c] + beta[c], 0.f);
    if (keep) v = keep[i] ? v * scale : 0.f;
    a[i] = v;
}
// d (in: grad wrt the layer output a; out: grad wrt the BatchNorm output) and the two column sums BN backward needs
__global__ void bn_bwd_part_kernel(const float* __restrict__ y, const float* __restrict__ mean, const float* __restrict__ invstd,
                                   const float* __restrict__ gamma, const float* __restrict__ beta, const uint8_t* __restrict__ keep,
                                   float scale, float* __restrict__ d, int n, int H, double* __restrict__ part) {
    const int c = blockIdx.y * blockDim.x + threadIdx.x;
    if (c >= H) return;
    const int r0 = blockIdx.x * BN_ROWS, r1 = min(n, r0 + BN_ROWS);
    const float m = mean[c], is = invstd[c], ga = gamma[c], be = beta[c];
    double s = 0.0, q = 0.0;
    for (int r = r0; r < r1; ++r) {
        const size_t i = (size_t)r * H + c;
        const float xh = (y[i] - m) * is;
        float dv = d[i];
        if (keep) dv = keep[i] ? dv * scale : 0.f;
        if (!(xh * ga + be > 0.f)) dv = 0.f;
        d[i] = dv;
        s += dv;
        q += (double)dv * xh;
    }
    part[((size_t)blockIdx.x * 2) * H + c] = s;
    part[((size_t)blockIdx.x * 2 + 1) * H + c] = q;
}
// The statistics of a BatchNorm layer in ONE launch (H % 4 == 0): 64 rows x 128 columns per CTA (8 row lanes x 32 float4
// lanes, eight independent loads per thread, float64 sums combined in a fixed order), the parts written to `part`; the CTA
// that arrives last at the column tile's counter adds the parts up (fixed order: the result does not depend on which CTA
// it is), and -- when no data-parallel hook sits between the sums and their use (`fuse_final`) -- derives what the apply
// kernel needs.  FWD: sums = (sum x, sum x^2, n) -> mean / invstd / batch statistics.  BWD: also rewrites d (dropout +
// ReLU gate), sums = (sum dy, sum dy xhat) -> dgamma / dbeta (this rank's sums, always) and s1 / s2.
struct BnStats {
    const float* y;
    int n, H;
    double *part, *sums;
    unsigned* counter;
    int fuse_final;
    // BWD inputs
    const float *mean, *invstd, *gamma, *beta;
    const uint8_t* keep;
    float scale;
    float* d;
    // FWD final
    float eps;
    float *mean_out, *invstd_out, *batch_mean, *batch_var;
    double* n_out;
    // BWD final
    float *dgamma, *dbeta, *s1, *s2;
    const double* n_glob;
};
template <bool BWD>
__global__ void __launch_bounds__(256) bn_stats_kernel(const BnStats a) {
    __shared__ double sh_s[8][128], sh_q[8][128];
    __shared__ int sh_last;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5, H = a.H;
    const int cb = blockIdx.y * 128, c0 = cb + tx * 4;
    const int r0 = blockIdx.x * BN_ROWS;
    double s[4] = {0.0, 0.0, 0.0, 0.0}, q[4] = {0.0, 0.0, 0.0, 0.0};
    if (c0 < H) {
        float4 v[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const int r = r0 + ty + 8 * k;
            v[k] = r < a.n ? *reinterpret_cast<const float4*>(a.y + (size_t)r * H + c0) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
        if (!BWD) {
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                const float f[4] = {v[k].x, v[k].y, v[k].z, v[k].w};
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const double x = f[j];
                    s[j] += x;
                    q[j] += x * x;
                }
            }
        } else {
            const float4 m4 = *reinterpret_cast<const float4*>(a.mean + c0), i4 = *reinterpret_cast<const float4*>(a.invstd + c0);
            const float m[4] = {m4.x, m4.y, m4.z, m4.w}, is[4] = {i4.x, i4.y, i4.z, i4.w};
            const float ga[4] = {a.gamma[c0], a.gamma[c0 + 1], a.gamma[c0 + 2], a.gamma[c0 + 3]};   // parameters: any 4-byte alignment
            const float be[4] = {a.beta[c0], a.beta[c0 + 1], a.beta[c0 + 2], a.beta[c0 + 3]};
            float4 dv4[8];
            uchar4 kp[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                const int r = r0 + ty + 8 * k;
                const size_t i = (size_t)r * H + c0;
                dv4[k] = r < a.n ? *reinterpret_cast<const float4*>(a.d + i) : make_float4(0.f, 0.f, 0.f, 0.f);
                kp[k] = (a.keep && r < a.n) ? *reinterpret_cast<const uchar4*>(a.keep + i) : make_uchar4(1, 1, 1, 1);
            }
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                const int r = r0 + ty + 8 * k;
                if (r >= a.n) continue;
                const float f[4] = {v[k].x, v[k].y, v[k].z, v[k].w};
                float dv[4] = {dv4[k].x, dv4[k].y, dv4[k].z, dv4[k].w};
                const unsigned char kk[4] = {kp[k].x, kp[k].y, kp[k].z, kp[k].w};
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float xh = (f[j] - m[j]) * is[j];
                    if (a.keep) dv[j] = kk[j] ? dv[j] * a.scale : 0.f;
                    if (!(xh * ga[j] + be[j] > 0.f)) dv[j] = 0.f;
                    s[j] += dv[j];
                    q[j] += (double)dv[j] * xh;
                }
                *reinterpret_cast<float4*>(a.d + (size_t)r * H + c0) = make_float4(dv[0], dv[1], dv[2], dv[3]);
            }
        }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        sh_s[ty][tx * 4 + j] = s[j];
        sh_q[ty][tx * 4 + j] = q[j];
    }
    __syncthreads();
    const int col = threadIdx.x & 127, c = cb + col;
    const bool second = threadIdx.x >= 128;   // threads 0..127 own the first sum of a column, 128..255 the second
    {
        double t = 0.0;
#pragma unroll
        for (int k = 0; k < 8; ++k) t += second ? sh_q[k][col] : sh_s[k][col];
        if (c < H) a.part[((size_t)blockIdx.x * 2 + (second ? 1 : 0)) * H + c] = t;
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) sh_last = atomicAdd(a.counter + blockIdx.y, 1u) == gridDim.x - 1;
    __syncthreads();
    if (!sh_last) return;
    __threadfence();
    double t0 = 0.0, t1 = 0.0, t2 = 0.0, t3 = 0.0;
    if (c < H) {
        const double* pp = a.part + (second ? H : 0) + c;
        const int parts = gridDim.x;
        int p = 0;
        for (; p + 4 <= parts; p += 4) {
            t0 += __ldcg(pp + (size_t)(p + 0) * 2 * H);
            t1 += __ldcg(pp + (size_t)(p + 1) * 2 * H);
            t2 += __ldcg(pp + (size_t)(p + 2) * 2 * H);
            t3 += __ldcg(pp + (size_t)(p + 3) * 2 * H);
        }
        for (; p < parts; ++p) t0 += __ldcg(pp + (size_t)p * 2 * H);
    }
    const double tot = (t0 + t1) + (t2 + t3);
    if (c < H) a.sums[(second ? H : 0) + c] = tot;
    if (!BWD && threadIdx.x == 0 && blockIdx.y == 0) a.sums[2 * H] = (double)a.n;
    double* sh_tot = &sh_s[0][0];   // [256]: both sums of the tile's columns, for the finishing step
    __syncthreads();
    sh_tot[threadIdx.x] = tot;
    __syncthreads();
    if (threadIdx.x == 0) a.counter[blockIdx.y] = 0;   // ready for the next layer / pass
    if (second || c >= H) return;
    const double sum0 = sh_tot[col], sum1 = sh_tot[128 + col];
    if (BWD) {
        a.dbeta[c] = (float)sum0;
        a.dgamma[c] = (float)sum1;
        if (a.fuse_final) {
            const double n = *a.n_glob;
            a.s1[c] = (float)(sum0 / n);
            a.s2[c] = (float)(sum1 / n);
        }
    } else if (a.fuse_final) {
        const double n = (double)a.n;
        if (c == 0) *a.n_out = n;
        const double m = sum0 / n;
        double v = sum1 / n - m * m;   // biased variance (BatchNorm1d training normalisation)
        v = v > 0.0 ? v : 0.0;
        a.mean_out[c] = (float)m;
        a.invstd_out[c] = (float)(1.0 / sqrt(v + (double)a.eps));
        if (a.batch_mean) a.batch_mean[c] = (float)m;
        if (a.batch_var) a.batch_var[c] = (float)v;
    }
}
// dgamma / dbeta are THIS rank's sums (the gradient all-reduce adds the ranks up like every other gradient)
__global__ void bn_bwd_grads_kernel(const double* __restrict__ sums, int H, float* __restrict__ dgamma, float* __restrict__ dbeta) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= H) return;
    dbeta[c] = (float)sums[c];
    dgamma[c] = (float)sums[H + c];
}
// s1, s2 = the two batch means of the BatchNorm backward over ALL rows (after the data-parallel hook: every rank's)
__global__ void bn_bwd_final_kernel(const double* __restrict__ sums, const double* __restrict__ n_glob, int H, float* __restrict__ s1,
                                    float* __restrict__ s2) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= H) return;
    const double n = *n_glob;
    s1[c] = (float)(sums[c] / n);
    s2[c] = (float)(sums[H + c] / n);
}
__global__ void bn_bwd_apply_kernel(const float* __restrict__ y, const float* __restrict__ mean, const float* __restrict__ invstd,
                                    const float* __restrict__ gamma, const float* __restrict__ dpre, const float* __restrict__ s1,
                                    const float* __restrict__ s2, long count, int H, float* __restrict__ dy) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    const int c = (int)(i % H);
    const float xh = (y[i] - mean[c]) * invstd[c];
    dy[i] = gamma[c] * invstd[c] * (dpre[i] - s1[c] - xh * s2[c]);
}
// d = (act > 0) ? d * scale : 0, act = relu(v) * keep * scale (so act > 0 <=> v > 0 and kept)
__global__ void relu_drop_bwd_kernel(const float* __restrict__ act, float scale, long count, float* __restrict__ d) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    d[i] = act[i] > 0.f ? d[i] * scale : 0.f;
}
__global__ void fill_or_copy_kernel(const float* __restrict__ src, long count, float* __restrict__ dst) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < count) dst[i] = src ? src[i] : 0.f;
}

static float drop_scale(const cm_gnn_dropout* d, int site) {
    if (!d || !d->keep[site]) return 1.f;
    const float p = site < 3 ? d->p_conv : d->p_fc;
    return 1.f / (1.f - p);
}
static const uint8_t* drop_keep(const cm_gnn_dropout* d, int site) { return d ? d->keep[site] : nullptr; }

template <typename F>
static int gat_dispatch(int heads, F f) {
    switch (heads) {
        case 1: f(std::integral_constant<int, 1>()); return CM_OK;
        case 2: f(std::integral_constant<int, 2>()); return CM_OK;
        case 4: f(std::integral_constant<int, 4>()); return CM_OK;
        case 8: f(std::integral_constant<int, 8>()); return CM_OK;
    }
    set_error("GAT heads must be 1, 2, 4 or 8 (got %d)", heads);
    return CM_ERR_ARG;
}

static int check_train_params(const cm_gnn_params* p, const char* who) {
    if (!p || p->hidden <= 0 || p->hidden % 4 || p->in_channels <= 0 || p->in_channels > 32 || p->heads <= 0 || p->num_classes <= 0) {
        set_error("%s: bad model sizes (hidden %% 4 == 0, in_channels <= 32)", who);
        return CM_ERR_ARG;
    }
    return CM_OK;
}

}  // namespace cm

extern "C" int cm_set_bn_sync(cm_bn_sync_fn fn, void* user, double* device_buf, int capacity) {
    if (fn && (!device_buf || capacity <= 0)) {
        cm::set_error("cm_set_bn_sync: a hook needs a device buffer");
        return CM_ERR_ARG;
    }
    cm::g_bn_sync = fn;
    cm::g_bn_sync_user = user;
    cm::g_bn_sync_buf = fn ? device_buf : nullptr;
    cm::g_bn_sync_cap = fn ? capacity : 0;
    return CM_OK;
}

extern "C" size_t cm_gnn_train_workspace_bytes(const cm_gnn_params* p, int n_nodes, int n_edges) {
    if (!p || n_nodes < 0 || n_edges < 0 || cm::check_train_params(p, "cm_gnn_train_workspace_bytes") != CM_OK) return 0;
    return cm::carve_gnn_train(nullptr, p, n_nodes, n_edges).total;
}

extern "C" int cm_gnn_train_forward(const cm_gnn_params* p, const float* x, const int32_t* csr_row_ptr, const int32_t* csr_col,
                                    const float* csr_w, const float* self_w, int n_nodes, int n_edges, const cm_gnn_dropout* drop,
                                    void* workspace, size_t workspace_bytes, float* mask_out, float* instance_out, float* edge_out,
                                    float* node_emb, float* bn_batch_mean, float* bn_batch_var, uint32_t* status, void* stream) {
    using namespace cm;
    CM_DEVICE_OR_RETURN();
    if (int rc = check_train_params(p, "cm_gnn_train_forward")) return rc;
    CM_REQUIRE(x && csr_row_ptr && workspace && n_nodes > 0 && n_edges >= 0 && (n_edges == 0 || csr_col), "cm_gnn_train_forward: bad arguments");
    GnnTrainWs w = carve_gnn_train(workspace, p, n_nodes, n_edges);
    if (workspace_bytes < w.total) {
        set_error("cm_gnn_train_forward: workspace %zu < %zu bytes", workspace_bytes, w.total);
        return CM_ERR_WORKSPACE;
    }
    cudaStream_t st = as_stream(stream);
    const int N = n_nodes, H = p->hidden, C = p->in_channels, hd = p->heads, H2 = H / 2, KC = hd * C;
    const long NH = (long)N * H;
    cudaMemsetAsync(w.bn_counter, 0, sizeof(unsigned) * ((H + 127) / 128), st);
    {
        CM_TIMED("train_gat_fold", st);
        gat_fold_train_kernel<<<ceil_div(H * KC, 256), 256, 0, st>>>(p->conv1_lin_weight, p->conv1_att_src, p->conv1_att_dst, H, C, hd, w.u_src,
                                                                     w.u_dst, w.wf);
    }
    if (n_edges > 0) {
        CM_TIMED("train_csr_reverse", st);
        csr_reverse_kernel<<<ceil_div(N, 128), 128, 0, st>>>(csr_row_ptr, csr_col, N, w.rev, status);
    }
    {
        CM_TIMED("train_gat_logits", st);
        gat_logits_train_kernel<<<ceil_div((long)N * hd, 256), 256, 0, st>>>(x, C, hd, w.u_src, w.u_dst, N, w.asrc, w.adst);
    }
    if (int rc = gat_dispatch(hd, [&](auto hc) {
            CM_TIMED("train_gat_aggregate", st);
            gat_agg_train_fwd_kernel<decltype(hc)::value><<<ceil_div(N, 8), 256, 0, st>>>(x, C, csr_row_ptr, csr_col, w.asrc, w.adst, N, w.z,
                                                                                          w.gmax, w.gden);
        }))
        return rc;
    launch_gemm(gemm_nt(w.z, KC, w.wf, w.y[0], H, N, H, KC, p->conv1_bias, 0), st, "train_linear");
    {
        CM_TIMED("train_gcn_norm", st);
        gcn_dis_train_kernel<<<ceil_div(N, 256), 256, 0, st>>>(csr_row_ptr, csr_col, csr_w, self_w, N, w.dis);
    }
    const int parts = ceil_div(N, BN_ROWS);
    for (int k = 0; k < 4; ++k) {
        if (k > 0) {
            {
                CM_TIMED("train_gcn_aggregate", st);
                gcn_agg_train_kernel<<<ceil_div(N, 8), 256, 0, st>>>(w.a[k - 1], csr_row_ptr, csr_col, csr_w, nullptr, self_w, w.dis, N, H,
                                                                    w.g[k - 1]);
            }
            launch_gemm(gemm_nt(w.g[k - 1], H, p->conv_lin_weight[k - 1], w.y[k], H, N, H, H, p->conv_bias[k - 1], 0), st, "train_linear");
        }
        {
            CM_TIMED("train_bn", st);
            float* bmean = bn_batch_mean ? bn_batch_mean + k * H : nullptr;
            float* bvar = bn_batch_var ? bn_batch_var + k * H : nullptr;
            if (H % 4 == 0) {
                BnStats b = {};
                b.y = w.y[k]; b.n = N; b.H = H; b.part = w.part_d; b.sums = w.bn_sums; b.counter = w.bn_counter; b.fuse_final = !g_bn_sync;
                b.eps = p->bn_eps; b.mean_out = w.mean + k * H; b.invstd_out = w.invstd + k * H; b.batch_mean = bmean; b.batch_var = bvar;
                b.n_out = w.bn_n + k;
                bn_stats_kernel<false><<<dim3(parts, ceil_div(H, 128)), 256, 0, st>>>(b);
            } else {
                bn_fwd_part_kernel<<<dim3(parts, ceil_div(H, 128)), 128, 0, st>>>(w.y[k], N, H, w.part_d);
                bn_reduce_parts_kernel<<<ceil_div(H, 128), 128, 0, st>>>(w.part_d, parts, N, H, w.bn_sums);
            }
            if (g_bn_sync || H % 4 != 0) {
                if (int rc = bn_sync(w.bn_sums, 2 * H + 1, st)) return rc;   // data parallel: (sum x, sum x^2, n) over every rank's nodes
                bn_fwd_final_kernel<<<ceil_div(H, 128), 128, 0, st>>>(w.bn_sums, H, p->bn_eps, w.mean + k * H, w.invstd + k * H, bmean, bvar,
                                                                      w.bn_n + k);
            }
            bn_fwd_apply_kernel<<<ceil_div(NH, 256), 256, 0, st>>>(w.y[k], w.mean + k * H, w.invstd + k * H, p->bn_weight[k], p->bn_bias[k],
                                                                   k < 3 ? drop_keep(drop, k) : nullptr, k < 3 ? drop_scale(drop, k) : 1.f,
                                                                   NH, H, w.a[k]);
        }
    }
    {
        Gemm g = gemm_nt(w.a[3], H, p->fc_shared_weight, w.e, H, N, H, H, p->fc_shared_bias, 1);
        g.keep = drop_keep(drop, 3);
        g.keep_scale = drop_scale(drop, 3);
        launch_gemm(g, st, "train_linear");
    }
    if (node_emb) cudaMemcpyAsync(node_emb, w.e, sizeof(float) * NH, cudaMemcpyDeviceToDevice, st);
    float* outs[3] = {mask_out, instance_out, edge_out};
    const int cdim[3] = {p->num_classes, p->num_classes, 1};
    for (int t = 0; t < 3; ++t) {
        Gemm g = gemm_nt(w.e, H, p->head1_weight[t], w.t[t], H2, N, H2, H, p->head1_bias[t], 1);
        g.keep = drop_keep(drop, 4 + t);
        g.keep_scale = drop_scale(drop, 4 + t);
        launch_gemm(g, st, "train_linear");
        if (outs[t]) launch_gemm(gemm_nt(w.t[t], H2, p->head2_weight[t], outs[t], cdim[t], N, cdim[t], H2, p->head2_bias[t], 0), st, "train_linear");
    }
    return post_launch("cm_gnn_train_forward");
}

extern "C" int cm_gnn_backward(const cm_gnn_params* p, const float* x, const int32_t* csr_row_ptr, const int32_t* csr_col,
                               const float* csr_w, const float* self_w, int n_nodes, int n_edges, const cm_gnn_dropout* drop,
                               void* workspace, size_t workspace_bytes, const float* d_mask_out, const float* d_instance_out,
                               const float* d_edge_out, const float* d_node_emb, const cm_gnn_grads* gr, void* stream) {
    using namespace cm;
    CM_DEVICE_OR_RETURN();
    if (int rc = check_train_params(p, "cm_gnn_backward")) return rc;
    CM_REQUIRE(x && csr_row_ptr && workspace && gr && n_nodes > 0 && n_edges >= 0, "cm_gnn_backward: bad arguments");
    GnnTrainWs w = carve_gnn_train(workspace, p, n_nodes, n_edges);
    if (workspace_bytes < w.total) {
        set_error("cm_gnn_backward: workspace %zu < %zu bytes", workspace_bytes, w.total);
        return CM_ERR_WORKSPACE;
    }
    cudaStream_t st = as_stream(stream);
    const int N = n_nodes, H = p->hidden, C = p->in_channels, hd = p->heads, H2 = H / 2, KC = hd * C, S = w.splits;
    const long NH = (long)N * H, NH2 = (long)N * H2;
    const int parts = ceil_div(N, BN_ROWS);

    // ---- heads and fc_shared: de accumulates in dA
    {
        CM_TIMED("train_elementwise", st);
        fill_or_copy_kernel<<<ceil_div(NH, 256), 256, 0, st>>>(d_node_emb, NH, w.dA);
    }
    const float* douts[3] = {d_mask_out, d_instance_out, d_edge_out};
    const int cdim[3] = {p->num_classes, p->num_classes, 1};
    for (int t = 0; t < 3; ++t) {
        if (!douts[t]) {   // no loss on this head: zero parameter gradients
            cudaMemsetAsync(gr->head2_weight[t], 0, sizeof(float) * cdim[t] * H2, st);
            cudaMemsetAsync(gr->head2_bias[t], 0, sizeof(float) * cdim[t], st);
            cudaMemsetAsync(gr->head1_weight[t], 0, sizeof(float) * H2 * H, st);
            cudaMemsetAsync(gr->head1_bias[t], 0, sizeof(float) * H2, st);
            continue;
        }
        launch_gemm(gemm_tn(douts[t], cdim[t], w.t[t], H2, gr->head2_weight[t], N, cdim[t], H2, w.part_f, S), st, "train_wgrad");
        launch_colsum(douts[t], cdim[t], N, cdim[t], w.part_d, gr->head2_bias[t], 0, st);
        launch_gemm(gemm_nn(douts[t], cdim[t], p->head2_weight[t], w.dT, H2, N, cdim[t], H2), st, "train_dgrad");
        {
            CM_TIMED("train_elementwise", st);
            relu_drop_bwd_kernel<<<ceil_div(NH2, 256), 256, 0, st>>>(w.t[t], drop_scale(drop, 4 + t), NH2, w.dT);
        }
        launch_gemm(gemm_tn(w.dT, H2, w.e, H, gr->head1_weight[t], N, H2, H, w.part_f, S), st, "train_wgrad");
        launch_colsum(w.dT, H2, N, H2, w.part_d, gr->head1_bias[t], 0, st);
        Gemm g = gemm_nn(w.dT, H2, p->head1_weight[t], w.dA, H, N, H2, H);   // de += dT W1
        g.accumulate = 1;
        launch_gemm(g, st, "train_dgrad");
    }
    {
        CM_TIMED("train_elementwise", st);
        relu_drop_bwd_kernel<<<ceil_div(NH, 256), 256, 0, st>>>(w.e, drop_scale(drop, 3), NH, w.dA);
    }
    launch_gemm(gemm_tn(w.dA, H, w.a[3], H, gr->fc_shared_weight, N, H, H, w.part_f, S), st, "train_wgrad");
    launch_colsum(w.dA, H, N, H, w.part_d, gr->fc_shared_bias, 0, st);
    launch_gemm(gemm_nn(w.dA, H, p->fc_shared_weight, w.dB, H, N, H, H), st, "train_dgrad");
    float *dcur = w.dB, *dtmp = w.dA;   // dcur = grad wrt a[3]

    for (int k = 3; k >= 0; --k) {
        {
            CM_TIMED("train_bn", st);
            const uint8_t* keep_k = k < 3 ? drop_keep(drop, k) : nullptr;
            const float scale_k = k < 3 ? drop_scale(drop, k) : 1.f;
            if (H % 4 == 0) {
                BnStats b = {};
                b.y = w.y[k]; b.n = N; b.H = H; b.part = w.part_d; b.sums = w.bn_sums; b.counter = w.bn_counter; b.fuse_final = !g_bn_sync;
                b.mean = w.mean + k * H; b.invstd = w.invstd + k * H; b.gamma = p->bn_weight[k]; b.beta = p->bn_bias[k];
                b.keep = keep_k; b.scale = scale_k; b.d = dcur;
                b.dgamma = gr->bn_weight[k]; b.dbeta = gr->bn_bias[k]; b.s1 = w.s1; b.s2 = w.s2; b.n_glob = w.bn_n + k;
                bn_stats_kernel<true><<<dim3(parts, ceil_div(H, 128)), 256, 0, st>>>(b);
            } else {
                bn_bwd_part_kernel<<<dim3(parts, ceil_div(H, 128)), 128, 0, st>>>(w.y[k], w.mean + k * H, w.invstd + k * H, p->bn_weight[k],
                                                                                  p->bn_bias[k], keep_k, scale_k, dcur, N, H, w.part_d);
                bn_reduce_parts_kernel<<<ceil_div(H, 128), 128, 0, st>>>(w.part_d, parts, N, H, w.bn_sums);
                bn_bwd_grads_kernel<<<ceil_div(H, 128), 128, 0, st>>>(w.bn_sums, H, gr->bn_weight[k], gr->bn_bias[k]);
            }
            if (g_bn_sync || H % 4 != 0) {
                if (int rc = bn_sync(w.bn_sums, 2 * H, st)) return rc;       // (sum dy, sum dy xhat) over every rank's nodes
                bn_bwd_final_kernel<<<ceil_div(H, 128), 128, 0, st>>>(w.bn_sums, w.bn_n + k, H, w.s1, w.s2);
            }
            bn_bwd_apply_kernel<<<ceil_div(NH, 256), 256, 0, st>>>(w.y[k], w.mean + k * H, w.invstd + k * H, p->bn_weight[k], dcur, w.s1, w.s2,
                                                                   NH, H, dtmp);
        }
        // dtmp = dy_k
        if (k > 0) {
            launch_gemm(gemm_tn(dtmp, H, w.g[k - 1], H, gr->conv_lin_weight[k - 1], N, H, H, w.part_f, S), st, "train_wgrad");
            launch_colsum(dtmp, H, N, H, w.part_d, gr->conv_bias[k - 1], 0, st);
            launch_gemm(gemm_nn(dtmp, H, p->conv_lin_weight[k - 1], dcur, H, N, H, H), st, "train_dgrad");   // dcur = dg
            {
                CM_TIMED("train_gcn_aggregate", st);
                gcn_agg_train_kernel<<<ceil_div(N, 8), 256, 0, st>>>(dcur, csr_row_ptr, csr_col, csr_w, n_edges > 0 ? w.rev : nullptr, self_w,
                                                                    w.dis, N, H, dtmp);   // dtmp = grad wrt a[k-1]
            }
            float* s = dcur;
            dcur = dtmp;
            dtmp = s;
        } else {
            launch_gemm(gemm_tn(dtmp, H, w.z, KC, w.dwf, N, H, KC, w.part_f, S), st, "train_wgrad");
            launch_colsum(dtmp, H, N, H, w.part_d, gr->conv1_bias, 0, st);
            launch_gemm(gemm_nn(dtmp, H, w.wf, w.dz, KC, N, H, KC), st, "train_dgrad");
        }
    }
    // ---- GAT attention backward
    if (int rc = gat_dispatch(hd, [&](auto hc) {
            CM_TIMED("train_gat_aggregate", st);
            constexpr int HD = decltype(hc)::value;
            gat_agg_train_bwd_kernel<HD><<<ceil_div(N, 8), 256, 8 * HD * C * sizeof(float), st>>>(
                x, C, csr_row_ptr, csr_col, w.asrc, w.adst, w.gmax, w.gden, w.dz, N, w.dpre_edge, w.dpre_self, w.dadst);
        }))
        return rc;
    {
        CM_TIMED("train_gat_logits", st);
        gat_src_gather_kernel<<<ceil_div((long)N * hd, 256), 256, 0, st>>>(csr_row_ptr, w.rev, w.dpre_edge, w.dpre_self, N, hd, w.dasrc);
    }
    launch_gemm(gemm_tn(w.dasrc, hd, x, C, w.du, N, hd, C, w.part_f, S), st, "train_wgrad");
    launch_gemm(gemm_tn(w.dadst, hd, x, C, w.du + KC, N, hd, C, w.part_f, S), st, "train_wgrad");
    {
        CM_TIMED("train_gat_fold", st);
        gat_param_grads_kernel<<<ceil_div(hd * H, 128), 128, 0, st>>>(p->conv1_lin_weight, p->conv1_att_src, p->conv1_att_dst, w.dwf, w.du,
                                                                      w.du + KC, H, C, hd, gr->conv1_lin_weight, gr->conv1_att_src,
                                                                      gr->conv1_att_dst);
    }
    return post_launch("cm_gnn_backward");
}
