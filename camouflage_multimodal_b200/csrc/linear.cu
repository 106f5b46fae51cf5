// Dense layers, LayerNorm, segment mean and small-head epilogues (fp32 path).
#include "linear.cuh"

namespace cm {

constexpr int BM = 64, BN = 64, BK = 16;

// 256 threads, 4x4 register tile per thread, operands staged transposed in shared memory.
template <bool VEC>
__global__ void __launch_bounds__(256) linear_kernel(Linear a) {
    __shared__ __align__(16) float As[BK][BM + 4];
    __shared__ __align__(16) float Ws[BK][BN + 4];
    const int m_live = a.m_dev ? min(*a.m_dev, a.M) : a.M;
    const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
    if (m0 >= m_live) return;
    const int tid = threadIdx.x;
    const int lr = tid >> 2, lk = (tid & 3) * 4;   // loader: row within tile, k offset
    const int ty = tid >> 4, tx = tid & 15;
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
    const int ar = m0 + lr, wr = n0 + lr;
    const float* ap = a.A + (size_t)ar * a.lda;
    const float* wp = a.W + (size_t)wr * a.K;
    for (int k0 = 0; k0 < a.K; k0 += BK) {
        float av[4] = {0.f, 0.f, 0.f, 0.f}, wv[4] = {0.f, 0.f, 0.f, 0.f};
        const int k = k0 + lk;
        if (VEC) {
            if (ar < m_live && k < a.K) {
                const float4 t = *reinterpret_cast<const float4*>(ap + k);
                av[0] = t.x; av[1] = t.y; av[2] = t.z; av[3] = t.w;
            }
            if (wr < a.N && k < a.K) {
                const float4 t = *reinterpret_cast<const float4*>(wp + k);
                wv[0] = t.x; wv[1] = t.y; wv[2] = t.z; wv[3] = t.w;
            }
        } else {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                if (ar < m_live && k + j < a.K) av[j] = ap[k + j];
                if (wr < a.N && k + j < a.K) wv[j] = wp[k + j];
            }
        }
        __syncthreads();
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            As[lk + j][lr] = av[j];
            Ws[lk + j][lr] = wv[j];
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < BK; ++kk) {
            const float4 x = *reinterpret_cast<const float4*>(&As[kk][ty * 4]);
            const float4 y = *reinterpret_cast<const float4*>(&Ws[kk][tx * 4]);
            const float xa[4] = {x.x, x.y, x.z, x.w}, ya[4] = {y.x, y.y, y.z, y.w};
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(xa[i], ya[j], acc[i][j]);
        }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int r = m0 + ty * 4 + i;
        if (r >= m_live) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int c = n0 + tx * 4 + j;
            if (c >= a.N) continue;
            float v = acc[i][j];
            if (a.bias) v += a.bias[c];
            v *= a.alpha;
            if (a.residual) v += a.residual[(size_t)(a.res_mod > 0 ? r % a.res_mod : r) * a.ldr + c];
            if (a.relu) v = fmaxf(v, 0.f);
            a.C[(size_t)r * a.ldc + c] = v;
        }
    }
}

int launch_linear(const Linear& a, cudaStream_t st) {
    if (a.M <= 0 || a.N <= 0 || a.K <= 0) return CM_OK;
    dim3 grid(ceil_div(a.N, BN), ceil_div(a.M, BM));
    const bool vec = (a.K % 4 == 0) && (a.lda % 4 == 0) && ((reinterpret_cast<uintptr_t>(a.A) & 15) == 0) &&
                     ((reinterpret_cast<uintptr_t>(a.W) & 15) == 0);
    if (vec) {
        CM_TIMED("linear", st);
        linear_kernel<true><<<grid, 256, 0, st>>>(a);
    }
    else {
        CM_TIMED("linear", st);
        linear_kernel<false><<<grid, 256, 0, st>>>(a);
    }
    return post_launch("linear");
}

// one warp per row
__global__ void __launch_bounds__(256) layernorm_kernel(const float* __restrict__ in, float* __restrict__ out, const float* __restrict__ gamma,
                                                        const float* __restrict__ beta, int rows, int dim, const int* rows_dev) {
    const int r = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    const int live = rows_dev ? min(*rows_dev, rows) : rows;
    if (r >= live) return;
    const float* x = in + (size_t)r * dim;
    float s = 0.f;
    for (int c = lane; c < dim; c += 32) s += x[c];
    const float mean = warp_sum(s) / (float)dim;
    float v = 0.f;
    for (int c = lane; c < dim; c += 32) {
        const float d = x[c] - mean;
        v += d * d;
    }
    const float rstd = 1.0f / sqrtf(warp_sum(v) / (float)dim + 1e-5f);
    float* o = out + (size_t)r * dim;
    for (int c = lane; c < dim; c += 32) o[c] = (x[c] - mean) * rstd * gamma[c] + beta[c];
}

int launch_layernorm(const float* in, float* out, const float* gamma, const float* beta, int rows, int dim,
                     const int* rows_dev, cudaStream_t st) {
    if (rows <= 0) return CM_OK;
    {
        CM_TIMED("layernorm", st);
        layernorm_kernel<<<ceil_div(rows, 8), 256, 0, st>>>(in, out, gamma, beta, rows, dim, rows_dev);
    }
    return post_launch("layernorm");
}

// grid (n_seg, ceil(dim/128)), block 128: thread per column, 4 row-partitions would need smem; rows are ~500
__global__ void __launch_bounds__(128) segment_mean_kernel(const float* __restrict__ in, int dim, const int* __restrict__ seg_ptr,
                                                           int fixed_len, float* __restrict__ out, int ldo) {
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
    out[(size_t)g * ldo + c] = n > 0 ? ((s0 + s1) + (s2 + s3)) / (float)n : 0.f;
}

int launch_segment_mean(const float* in, int dim, const int* seg_ptr, int fixed_len, int n_seg, float* out, int ldo,
                        cudaStream_t st) {
    if (n_seg <= 0) return CM_OK;
    dim3 grid(n_seg, ceil_div(dim, 128));
    {
        CM_TIMED("segment_mean", st);
        segment_mean_kernel<<<grid, 128, 0, st>>>(in, dim, seg_ptr, fixed_len, out, ldo);
    }
    return post_launch("segment_mean");
}

// one warp per row
__global__ void __launch_bounds__(256) heads2_kernel(const float* __restrict__ hid, int ld, int rows, const int* rows_dev, Heads2 h) {
    const int r = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    const int live = rows_dev ? min(*rows_dev, rows) : rows;
    if (r >= live) return;
    float first[2] = {0.f, 0.f};
    for (int t = 0; t < h.n_heads; ++t) {
        const float* x = hid + (size_t)r * ld + t * h.hdim;
        for (int j = 0; j < h.cdim[t]; ++j) {
            const float* wrow = h.w2[t] + (size_t)j * h.hdim;
            float s = 0.f;
            for (int k = lane; k < h.hdim; k += 32) s = fmaf(x[k], wrow[k], s);
            s = warp_sum(s) + h.b2[t][j];
            if (h.sigmoid[t]) s = 1.0f / (1.0f + expf(-s));
            if (t == 0 && j < 2) first[j] = s;
            if (lane == 0 && h.out[t]) h.out[t][(size_t)r * h.cdim[t] + j] = s;
        }
    }
    if (h.prob_out && lane == 0) {
        // softmax over two logits, class 1 (models/region_graph/test.py:238)
        const float m = fmaxf(first[0], first[1]);
        const float e0 = expf(first[0] - m), e1 = expf(first[1] - m);
        h.prob_out[r] = e1 / (e0 + e1);
    }
}

int launch_heads2(const float* hid, int ld, int rows, const int* rows_dev, const Heads2& h, cudaStream_t st) {
    if (rows <= 0) return CM_OK;
    {
        CM_TIMED("heads2", st);
        heads2_kernel<<<ceil_div(rows, 8), 256, 0, st>>>(hid, ld, rows, rows_dev, h);
    }
    return post_launch("heads2");
}

}  // namespace cm
