// Strided fp32 GEMM with deterministic split-K, and column sums with fp64 partials: the dense building blocks of
// the training-mode passes (see train_common.cuh).
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>

#include <cuda_bf16.h>

#include "dense_tc.cuh"
#include "train_common.cuh"

namespace cm {

constexpr int GB = 64, GK = 16;

// ---------------------------------------------------------------- tensor-core path of launch_gemm
// The training GEMMs (forward linears, dgrad, wgrad of both models) on the tcgen05 dense kernel (dense_tc.cu) with
// near-fp32 accuracy: both operands are rewritten as 16-bit thirds along the reduction dimension,
//   A' = [A_hi | A_lo | A_hi],  B' = [B_hi | B_hi | B_lo]   (hi = round16(v), lo = round16(v - hi)),
// so that one 16-bit GEMM with fp32 accumulation over 3 K computes A_hi B_hi + A_lo B_hi + A_hi B_lo.  Forward layers use
// fp16 parts (22 mantissa bits per operand: 5e-7 of the output scale, like the inference GNN layers), the backward GEMMs
// bf16 parts (16 bits, fp32's exponent range: gradients of a mean loss sit at 1e-6 .. 1e-9; 5e-6 .. 3e-5 of the scale).  Operands of any stride pattern (NT forward,
// NN dgrad, TN wgrad) are brought to K-major by the split kernels; wgrad runs split-K with a fixed-order reduction.
// k contiguous in the source: out[r, t * kp + c], thread per (r, c)
// hi / lo 16-bit parts of v as raw bits: fp16 (22 mantissa bits together, fp16's exponent range) or bf16 (16 bits, fp32's range)
__device__ __forceinline__ void split_parts(float v, int bf16, unsigned short* hi, unsigned short* lo) {
    if (bf16) {
        const __nv_bfloat16 h = __float2bfloat16_rn(v), l = __float2bfloat16_rn(v - __bfloat162float(h));
        *hi = __bfloat16_as_ushort(h);
        *lo = __bfloat16_as_ushort(l);
    } else {
        const op_t h = f2op(v), l = f2op(v - op2f(h));
        *hi = __half_as_ushort(h);
        *lo = __half_as_ushort(l);
    }
}
__global__ void __launch_bounds__(256) split_rows_kernel(const float* __restrict__ src, long s_row, int rows, int K, int kp, int b_pattern, int bf16,
                                                         unsigned short* __restrict__ out) {
    const long total = (long)rows * kp;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const int r = (int)(i / kp), c = (int)(i - (long)r * kp);
        const float v = c < K ? src[(long)r * s_row + c] : 0.f;
        unsigned short hi, lo;
        split_parts(v, bf16, &hi, &lo);
        unsigned short* o = out + (size_t)r * 3 * kp + c;
        o[0] = hi;
        o[kp] = b_pattern ? hi : lo;
        o[2 * kp] = b_pattern ? lo : hi;
    }
}
// rows contiguous in the source (element (r, k) at src[k * s_k + r]): 32 x 32 tiles through shared memory
__global__ void __launch_bounds__(256) split_cols_kernel(const float* __restrict__ src, long s_k, int rows, int K, int kp, int b_pattern, int bf16,
                                                         unsigned short* __restrict__ out) {
    __shared__ float t[32][33];
    const int r0 = blockIdx.x * 32, k0 = blockIdx.y * 32, tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    for (int j = ty; j < 32; j += 8) {
        const int k = k0 + j, r = r0 + tx;
        t[j][tx] = (k < K && r < rows) ? src[(long)k * s_k + r] : 0.f;
    }
    __syncthreads();
    for (int j = ty; j < 32; j += 8) {
        const int r = r0 + j, k = k0 + tx;
        if (r >= rows || k >= kp) continue;
        const float v = t[tx][j];
        unsigned short hi, lo;
        split_parts(v, bf16, &hi, &lo);
        unsigned short* o = out + (size_t)r * 3 * kp + k;
        o[0] = hi;
        o[kp] = b_pattern ? hi : lo;
        o[2 * kp] = b_pattern ? lo : hi;
    }
}
// vector forms of the two kernels above (same arithmetic, so the same bits): 8 consecutive k per thread, 16-byte stores
__device__ __forceinline__ void split8_store(const float (&v)[8], int b_pattern, int bf16, unsigned short* o, int kp) {
    unsigned short hi[8], lo[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) split_parts(v[q], bf16, &hi[q], &lo[q]);
    uint4 H, L;
    H.x = hi[0] | ((unsigned)hi[1] << 16); H.y = hi[2] | ((unsigned)hi[3] << 16); H.z = hi[4] | ((unsigned)hi[5] << 16); H.w = hi[6] | ((unsigned)hi[7] << 16);
    L.x = lo[0] | ((unsigned)lo[1] << 16); L.y = lo[2] | ((unsigned)lo[3] << 16); L.z = lo[4] | ((unsigned)lo[5] << 16); L.w = lo[6] | ((unsigned)lo[7] << 16);
    *reinterpret_cast<uint4*>(o) = H;
    *reinterpret_cast<uint4*>(o + kp) = b_pattern ? H : L;
    *reinterpret_cast<uint4*>(o + 2 * kp) = b_pattern ? L : H;
}
// One operand of a split launch.  mode 0: k contiguous, source rows 16-byte aligned (stride % 4 == 0) -- thread per (r, 8 k);
// mode 1: rows contiguous in the source (element (r, k) at src[k * stride + r]) -- a 32-row x 256-k tile through shared memory.
struct SplitOp {
    const float* src;
    unsigned short* out;
    long stride;
    int rows, K, mode, b_pattern, blocks, nbx;
};
constexpr int SC_K = 256;
__device__ __forceinline__ void split_rows_body(const SplitOp& o, int kp, int bf16, int bid) {
    const int kp8 = kp >> 3;
    const long total = (long)o.rows * kp8;
    for (long i = (long)bid * 256 + threadIdx.x; i < total; i += (long)o.blocks * 256) {
        const int r = (int)(i / kp8), c = (int)(i - (long)r * kp8) * 8;
        const float* s = o.src + (long)r * o.stride + c;
        float v[8];
        if (c + 8 <= o.K) {
            const float4 a = __ldg(reinterpret_cast<const float4*>(s)), b = __ldg(reinterpret_cast<const float4*>(s) + 1);
            v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
        } else {
#pragma unroll
            for (int q = 0; q < 8; ++q) v[q] = c + q < o.K ? s[q] : 0.f;
        }
        split8_store(v, o.b_pattern, bf16, o.out + (size_t)r * 3 * kp + c, kp);
    }
}
__device__ __forceinline__ void split_cols_body(const SplitOp& o, int kp, int bf16, int bid, float (*t)[33]) {
    const int r0 = (bid % o.nbx) * 32, k0 = (bid / o.nbx) * SC_K, tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int kn = min(SC_K, kp - k0);   // a multiple of 8
#pragma unroll 8
    for (int j = ty; j < kn; j += 8) {
        const int k = k0 + j, r = r0 + tx;
        t[j][tx] = (k < o.K && r < o.rows) ? __ldg(o.src + (long)k * o.stride + r) : 0.f;
    }
    __syncthreads();
    // a warp takes 8 rows x 4 octets at a time: lane -> (row, octet) so that the 32 shared-memory reads of a step fall in 32
    // banks ((8 octet + row) mod 32), and each row receives 64 contiguous bytes per part
    const int kgroups = (kn + 31) >> 5;
    for (int u = ty; u < 4 * kgroups; u += 8) {
        const int j = (u & 3) * 8 + (tx & 7), c = ((u >> 2) * 4 + (tx >> 3)) * 8, r = r0 + j;
        if (r >= o.rows || c >= kn) continue;
        float v[8];
#pragma unroll
        for (int q = 0; q < 8; ++q) v[q] = t[c + q][j];
        split8_store(v, o.b_pattern, bf16, o.out + (size_t)r * 3 * kp + k0 + c, kp);
    }
}
// both operands of a GEMM in one launch: blocks [0, a.blocks) work on A, the rest on B
__global__ void __launch_bounds__(256) split_pair_kernel(const SplitOp a, const SplitOp b, int kp, int bf16) {
    __shared__ float t[SC_K][33];
    const bool first = (int)blockIdx.x < a.blocks;
    const SplitOp& o = first ? a : b;
    const int bid = first ? blockIdx.x : blockIdx.x - a.blocks;
    if (o.mode == 0) split_rows_body(o, kp, bf16, bid);
    else split_cols_body(o, kp, bf16, bid, t);
}
// -> false when the operand needs the one-element kernels (unaligned k-contiguous rows)
static bool split_op_make(const float* src, long s_row, long s_k, int rows, int K, int kp, int b_pattern, unsigned short* out, SplitOp* o) {
    o->src = src; o->out = out; o->rows = rows; o->K = K; o->b_pattern = b_pattern;
    if (s_k == 1) {
        if ((s_row & 3) != 0 || (reinterpret_cast<uintptr_t>(src) & 15) != 0) return false;
        o->mode = 0; o->stride = s_row; o->nbx = 1;
        o->blocks = (int)std::min<long>(((long)rows * (kp >> 3) + 255) / 256, 148L * 8);
    } else {
        o->mode = 1; o->stride = s_k; o->nbx = ceil_div(rows, 32);
        o->blocks = o->nbx * ceil_div(kp, SC_K);
    }
    return true;
}
static void split_operand_scalar(const float* src, long s_row, long s_k, int rows, int K, int kp, int b_pattern, int bf16, unsigned short* out, cudaStream_t st) {
    if (s_k == 1) {
        const long total = (long)rows * kp;
        split_rows_kernel<<<(int)std::min<long>((total + 255) / 256, 148L * 16), 256, 0, st>>>(src, s_row, rows, K, kp, b_pattern, bf16, out);
    } else {
        split_cols_kernel<<<dim3(ceil_div(rows, 32), ceil_div(kp, 32)), 256, 0, st>>>(src, s_k, rows, K, kp, b_pattern, bf16, out);
    }
}
// A' = [hi | lo | hi] and B' = [hi | hi | lo] of one GEMM (header of this section)
static void split_operands(const Gemm& g, int kp, int bf16, unsigned short* a3, unsigned short* b3, cudaStream_t st) {
    static const bool scalar = getenv("CM_TRAIN_SPLIT_SCALAR") != nullptr;   // the one-element kernels, for A/B runs
    SplitOp a, b;
    if (!scalar && split_op_make(g.A, g.sam, g.sak, g.M, g.K, kp, 0, a3, &a) && split_op_make(g.B, g.sbn, g.sbk, g.N, g.K, kp, 1, b3, &b)) {
        split_pair_kernel<<<a.blocks + b.blocks, 256, 0, st>>>(a, b, kp, bf16);
        return;
    }
    split_operand_scalar(g.A, g.sam, g.sak, g.M, g.K, kp, 0, bf16, a3, st);
    split_operand_scalar(g.B, g.sbn, g.sbk, g.N, g.K, kp, 1, bf16, b3, st);
}
__global__ void gemm_epilogue_kernel(const float* __restrict__ raw, int M, int N, float* __restrict__ C, long ldc, const float* __restrict__ bias,
                                     int relu, const uint8_t* __restrict__ keep, float keep_scale, float alpha, int accumulate) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long)M * N) return;
    const int m = (int)(i / N), n = (int)(i - (long)m * N);
    float v = raw[i] * alpha;
    if (bias) v += bias[n];
    if (relu) v = fmaxf(v, 0.f);
    if (keep) v = keep[i] ? v * keep_scale : 0.f;
    float* c = C + (long)m * ldc + n;
    *c = accumulate ? *c + v : v;
}
__global__ void splitk_reduce_kernel(const float* __restrict__ part, int splits, int M, int N, float* __restrict__ C, long ldc,
                                     const float* __restrict__ bias, int relu, float alpha, int accumulate);

static bool tc_gemm_enabled() {
    static int on = -1;
    if (on < 0) {
        const char* e = getenv("CM_TRAIN_GEMM");
        on = !(e && e[0] == 'f');   // CM_TRAIN_GEMM=fp32 keeps the CUDA-core tiles
    }
    return on == 1;
}
static long tc_gemm_threshold() {
    static long t = -1;
    if (t < 0) {
        const char* e = getenv("CM_TRAIN_GEMM_MIN");   // multiply-adds
        t = e ? atol(e) : (1L << 24);
    }
    return t;
}
static void* scratch_alloc(size_t bytes, cudaStream_t st) {
    static PerDeviceOnce pool_once;
    if (pool_once.first()) {   // keep freed blocks in the stream-ordered pool instead of returning them to the driver
        int dev = 0;
        cudaGetDevice(&dev);
        cudaMemPool_t pool;
        if (cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess) {
            unsigned long long keep = ~0ull;
            cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
        }
    }
    void* p = nullptr;
    if (cudaMallocAsync(&p, bytes, st) != cudaSuccess) {
        cudaGetLastError();
        return nullptr;
    }
    return p;
}

// -> CM_OK when the GEMM ran on the tensor cores, 1 when the shape is not worth it (caller falls back), < 0 on error
static int launch_gemm_tc(const Gemm& g, cudaStream_t st, const char* name) {
    // below ~2^24 multiply-adds the fp32 tile kernel (one launch) is as fast as split + tcgen05 (three launches): KG-row and
    // graph-level layers stay there, the region-node layers of both models and every weight gradient over them come here
    if (!tc_gemm_enabled() || g.K < 64 || g.N < 16 || (long)g.M * g.N * g.K < tc_gemm_threshold()) return 1;
    if (g.sak != 1 && g.sam != 1) return 1;
    if (g.sbk != 1 && g.sbn != 1) return 1;
    if (const char* skip = getenv("CM_TC_SKIP")) {   // debugging aid: keep one access pattern on the fp32 tiles
        if (strstr(skip, "nn") && g.sak == 1 && g.sbk != 1) return 1;
        if (strstr(skip, "tn") && g.sak != 1) return 1;
        if (strstr(skip, "nt") && g.sak == 1 && g.sbk == 1) return 1;
        if (strstr(skip, "keep") && (g.keep || g.accumulate)) return 1;
        if (strstr(skip, "split") && g.splits > 1) return 1;
    }
    const int kp = (g.K + 7) / 8 * 8, K3 = 3 * kp;
    const size_t a_bytes = align_up((size_t)g.M * K3 * sizeof(op_t), 256), b_bytes = align_up((size_t)g.N * K3 * sizeof(op_t), 256);
    // split-K only where the caller provided the partial buffer (wgrad); every slice gets at least four k-blocks
    int splits = 1;
    if (g.splits > 1 && g.part) splits = std::max(1, std::min(g.splits, (K3 / 64) / 4));
    const bool direct = splits == 1 && !g.keep && !g.accumulate;
    const size_t raw_bytes = (splits == 1 && !direct) ? align_up((size_t)g.M * g.N * sizeof(float), 256) : 0;
    char* buf = reinterpret_cast<char*>(scratch_alloc(a_bytes + b_bytes + raw_bytes, st));
    if (!buf) return 1;
    unsigned short* a3 = reinterpret_cast<unsigned short*>(buf);
    unsigned short* b3 = reinterpret_cast<unsigned short*>(buf + a_bytes);
    // forward layers (both operands k-contiguous: activations x weights, O(1) values) keep fp16 thirds: 22 mantissa bits,
    // so ReLU gates and BatchNorm statistics match the fp32 reference; dgrad / wgrad take gradients (1e-6 .. 1e-9 under a
    // mean loss, below fp16's range) and use bf16 thirds
    const int bf16 = !(g.sak == 1 && g.sbk == 1);
    float* raw = reinterpret_cast<float*>(buf + a_bytes + b_bytes);
    int rc;
    {
        CM_TIMED("train_split_operand", st);
        split_operands(g, kp, bf16, a3, b3, st);
        rc = post_launch("split_operand");
    }
    if (rc == CM_OK) {
        Dense d;
        d.A = reinterpret_cast<const op_t*>(a3); d.lda = K3; d.W = reinterpret_cast<const op_t*>(b3); d.M = g.M; d.N = g.N; d.K = K3;
        d.bf16_operands = bf16;
        if (splits > 1) {
            d.out_f32 = g.part; d.ldo_f32 = g.N; d.k_splits = splits; d.split_stride = (size_t)g.M * g.N;
        } else if (direct) {
            d.out_f32 = g.C; d.ldo_f32 = (int)g.ldc; d.bias = g.bias; d.relu = g.relu; d.alpha = g.alpha;
        } else {
            d.out_f32 = raw; d.ldo_f32 = g.N;
        }
        rc = launch_dense(d, st);
    }
    if (rc == CM_OK && splits > 1) {
        CM_TIMED("splitk_reduce", st);
        splitk_reduce_kernel<<<ceil_div((long)g.M * g.N, 256), 256, 0, st>>>(g.part, splits, g.M, g.N, g.C, g.ldc, g.bias, g.relu, g.alpha, g.accumulate);
    } else if (rc == CM_OK && !direct) {
        CM_TIMED("train_elementwise", st);
        gemm_epilogue_kernel<<<ceil_div((long)g.M * g.N, 256), 256, 0, st>>>(raw, g.M, g.N, g.C, g.ldc, g.bias, g.relu, g.keep, g.keep_scale, g.alpha,
                                                                            g.accumulate);
    }
    cudaFreeAsync(buf, st);
    if (rc != CM_OK) return rc;
    return post_launch(name);
}

template <bool AKC, bool BKC>
__global__ void __launch_bounds__(256) sgemm_kernel(const float* __restrict__ A, long sam, long sak, const float* __restrict__ B, long sbk,
                                                    long sbn, float* __restrict__ C, long ldc, int M, int N, int K, int k_chunk,
                                                    long split_stride, const float* __restrict__ bias, int relu,
                                                    const uint8_t* __restrict__ keep, float keep_scale, float alpha, int accumulate) {
    __shared__ float As[GK][GB + 4], Bs[GK][GB + 4];
    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    const int m0 = blockIdx.y * GB, n0 = blockIdx.x * GB;
    const int kb = blockIdx.z * k_chunk, ke = min(K, kb + k_chunk);
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
    // the next k-tile's global loads are issued before the current tile's arithmetic (registers ra / rb): the small GEMMs that
    // run here (KG rows, per-image layers: 16 .. 64 CTAs) are bound by the latency of that load, not by the FMAs
    float ra[4], rb[4];
    auto load_tiles = [&](int k0) {
        if (AKC) {
            const int m = tid >> 2, kk = (tid & 3) * 4;
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int k = k0 + kk + q;
                ra[q] = (m0 + m < M && k < ke) ? A[(long)(m0 + m) * sam + (long)k * sak] : 0.f;
            }
        } else {
            const int kk = tid >> 4, m = (tid & 15) * 4, k = k0 + kk;
#pragma unroll
            for (int q = 0; q < 4; ++q) ra[q] = (m0 + m + q < M && k < ke) ? A[(long)(m0 + m + q) * sam + (long)k * sak] : 0.f;
        }
        if (BKC) {
            const int n = tid >> 2, kk = (tid & 3) * 4;
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int k = k0 + kk + q;
                rb[q] = (n0 + n < N && k < ke) ? B[(long)k * sbk + (long)(n0 + n) * sbn] : 0.f;
            }
        } else {
            const int kk = tid >> 4, n = (tid & 15) * 4, k = k0 + kk;
#pragma unroll
            for (int q = 0; q < 4; ++q) rb[q] = (n0 + n + q < N && k < ke) ? B[(long)k * sbk + (long)(n0 + n + q) * sbn] : 0.f;
        }
    };
    auto store_tiles = [&]() {
        if (AKC) {
            const int m = tid >> 2, kk = (tid & 3) * 4;
#pragma unroll
            for (int q = 0; q < 4; ++q) As[kk + q][m] = ra[q];
        } else {
            const int kk = tid >> 4, m = (tid & 15) * 4;
#pragma unroll
            for (int q = 0; q < 4; ++q) As[kk][m + q] = ra[q];
        }
        if (BKC) {
            const int n = tid >> 2, kk = (tid & 3) * 4;
#pragma unroll
            for (int q = 0; q < 4; ++q) Bs[kk + q][n] = rb[q];
        } else {
            const int kk = tid >> 4, n = (tid & 15) * 4;
#pragma unroll
            for (int q = 0; q < 4; ++q) Bs[kk][n + q] = rb[q];
        }
    };
    if (kb < ke) load_tiles(kb);
    for (int k0 = kb; k0 < ke; k0 += GK) {
        store_tiles();
        __syncthreads();
        if (k0 + GK < ke) load_tiles(k0 + GK);
#pragma unroll
        for (int kk = 0; kk < GK; ++kk) {
            const float4 a = *reinterpret_cast<const float4*>(&As[kk][ty * 4]);
            const float4 b = *reinterpret_cast<const float4*>(&Bs[kk][tx * 4]);
            const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
        }
        __syncthreads();
    }
    float* Cz = C + (long)blockIdx.z * split_stride;
    const bool final = split_stride == 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int m = m0 + ty * 4 + i;
        if (m >= M) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int n = n0 + tx * 4 + j;
            if (n >= N) continue;
            float v = acc[i][j];
            if (final) {
                v *= alpha;
                if (bias) v += bias[n];
                if (relu) v = fmaxf(v, 0.f);
                if (keep) v = keep[(long)m * N + n] ? v * keep_scale : 0.f;
                if (accumulate) v += Cz[(long)m * ldc + n];
            }
            Cz[(long)m * ldc + n] = v;
        }
    }
}

__global__ void splitk_reduce_kernel(const float* __restrict__ part, int splits, int M, int N, float* __restrict__ C, long ldc,
                                     const float* __restrict__ bias, int relu, float alpha, int accumulate) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long)M * N) return;
    float s = 0.f;
    for (int z = 0; z < splits; ++z) s += part[(long)z * M * N + i];
    const int m = (int)(i / N), n = (int)(i - (long)m * N);
    s *= alpha;
    if (bias) s += bias[n];
    if (relu) s = fmaxf(s, 0.f);
    float* c = C + (long)m * ldc + n;
    *c = accumulate ? *c + s : s;
}

int launch_gemm(const Gemm& g, cudaStream_t st, const char* name) {
    if (g.M <= 0 || g.N <= 0) return CM_OK;
    {
        const int rc = launch_gemm_tc(g, st, name);   // tcgen05 (split fp16 thirds); 1 = shape not worth it / unsupported strides
        if (rc <= 0) return rc;
    }
    const int splits = g.splits > 1 && g.part ? g.splits : 1;
    const int k_chunk = splits > 1 ? ((g.K + splits - 1) / splits + GK - 1) / GK * GK : (g.K > 0 ? g.K : 1);
    dim3 grid(ceil_div(g.N, GB), ceil_div(g.M, GB), splits);
    float* out = splits > 1 ? g.part : g.C;
    const long ldc = splits > 1 ? g.N : g.ldc, ss = splits > 1 ? (long)g.M * g.N : 0;
    {
        CM_TIMED(name, st);
#define CM_SGEMM(a, b)                                                                                                              \
    sgemm_kernel<a, b><<<grid, 256, 0, st>>>(g.A, g.sam, g.sak, g.B, g.sbk, g.sbn, out, ldc, g.M, g.N, g.K, k_chunk, ss, g.bias, g.relu, \
                                             g.keep, g.keep_scale, g.alpha, splits > 1 ? 0 : g.accumulate)
        if (g.a_k_contig && g.b_k_contig) CM_SGEMM(true, true);
        else if (g.a_k_contig) CM_SGEMM(true, false);
        else if (g.b_k_contig) CM_SGEMM(false, true);
        else CM_SGEMM(false, false);
#undef CM_SGEMM
    }
    if (splits > 1) {
        CM_TIMED("splitk_reduce", st);
        splitk_reduce_kernel<<<ceil_div((long)g.M * g.N, 256), 256, 0, st>>>(g.part, splits, g.M, g.N, g.C, g.ldc, g.bias, g.relu, g.alpha,
                                                                            g.accumulate);
    }
    return post_launch(name);
}

// ---------------------------------------------------------------- column sums
constexpr int CS_ROWS = 64;   // rows per CTA
int col_parts(int rows) { return rows > 0 ? ceil_div(rows, CS_ROWS) : 1; }

// 32 columns x 8 row lanes per CTA over CS_ROWS rows: eight independent loads in flight per thread, fixed-order tree
__global__ void __launch_bounds__(256) colsum_part_kernel(const float* __restrict__ a, int lda, int rows, int cols, double* __restrict__ part) {
    __shared__ double sh[8][33];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int c = blockIdx.y * 32 + tx;
    const int r0 = blockIdx.x * CS_ROWS, r1 = min(rows, r0 + CS_ROWS);
    double s = 0.0;
    if (c < cols) {
        float v[CS_ROWS / 8];
#pragma unroll
        for (int k = 0; k < CS_ROWS / 8; ++k) {
            const int r = r0 + ty + 8 * k;
            v[k] = r < r1 ? a[(long)r * lda + c] : 0.f;
        }
#pragma unroll
        for (int k = 0; k < CS_ROWS / 8; ++k) s += (double)v[k];
    }
    sh[ty][tx] = s;
    __syncthreads();
    if (ty == 0 && c < cols) {
#pragma unroll
        for (int k = 1; k < 8; ++k) s += sh[k][tx];
        part[(long)blockIdx.x * cols + c] = s;
    }
}
// 32 columns x 8 part lanes per CTA: coalesced reads of the partial rows, fixed-order tree over the 8 lanes (deterministic)
__global__ void __launch_bounds__(256) colsum_final_kernel(const double* __restrict__ part, int parts, int cols, float* __restrict__ out,
                                                           int accumulate) {
    __shared__ double sh[8][33];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int c = blockIdx.x * 32 + tx;
    double s = 0.0;
    if (c < cols)
        for (int p = ty; p < parts; p += 8) s += part[(long)p * cols + c];
    sh[ty][tx] = s;
    __syncthreads();
    if (ty == 0 && c < cols) {
#pragma unroll
        for (int k = 1; k < 8; ++k) s += sh[k][tx];
        out[c] = accumulate ? out[c] + (float)s : (float)s;
    }
}

// up to 8 row parts (512 rows: the per-image and KG-row matrices) in one launch: the same partial sums in the same order as
// colsum_part_kernel + colsum_final_kernel (the final's eight lanes hold one part each), so the same bits
__global__ void __launch_bounds__(256) colsum_small_kernel(const float* __restrict__ a, int lda, int rows, int cols, float* __restrict__ out,
                                                           int accumulate) {
    __shared__ double sh[8][33];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int c = blockIdx.x * 32 + tx;
    double total = 0.0;
    for (int r0 = 0; r0 < rows; r0 += CS_ROWS) {
        const int r1 = min(rows, r0 + CS_ROWS);
        double s = 0.0;
        if (c < cols) {
            float v[CS_ROWS / 8];
#pragma unroll
            for (int k = 0; k < CS_ROWS / 8; ++k) {
                const int r = r0 + ty + 8 * k;
                v[k] = r < r1 ? a[(long)r * lda + c] : 0.f;
            }
#pragma unroll
            for (int k = 0; k < CS_ROWS / 8; ++k) s += (double)v[k];
        }
        __syncthreads();
        sh[ty][tx] = s;
        __syncthreads();
        if (ty == 0) {
#pragma unroll
            for (int k = 1; k < 8; ++k) s += sh[k][tx];
            total += s;
        }
    }
    if (ty == 0 && c < cols) out[c] = accumulate ? out[c] + (float)total : (float)total;
}

int launch_colsum(const float* a, int lda, int rows, int cols, double* part, float* out, int accumulate, cudaStream_t st) {
    const int parts = col_parts(rows);
    if (rows <= 8 * CS_ROWS) {
        CM_TIMED("colsum", st);
        colsum_small_kernel<<<ceil_div(cols, 32), 256, 0, st>>>(a, lda, rows, cols, out, accumulate);
        return post_launch("colsum");
    }
    {
        CM_TIMED("colsum", st);
        if (rows > 0) colsum_part_kernel<<<dim3(parts, ceil_div(cols, 32)), 256, 0, st>>>(a, lda, rows, cols, part);
        colsum_final_kernel<<<ceil_div(cols, 32), 256, 0, st>>>(part, rows > 0 ? parts : 0, cols, out, accumulate);
    }
    return post_launch("colsum");
}

// ---------------------------------------------------------------- elementwise + LayerNorm
__global__ void tc_add_kernel(const float* __restrict__ a, const float* __restrict__ b, float* __restrict__ out, long count) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < count) out[i] = a[i] + (b ? b[i] : 0.f);
}
int launch_add(const float* a, const float* b, float* out, long count, cudaStream_t st) {
    if (count <= 0) return CM_OK;
    CM_TIMED("train_elementwise", st);
    tc_add_kernel<<<ceil_div(count, 256), 256, 0, st>>>(a, b, out, count);
    return post_launch("add");
}
__global__ void tc_relu_drop_bwd_kernel(const float* __restrict__ act, float scale, long count, float* __restrict__ d) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < count) d[i] = act[i] > 0.f ? d[i] * scale : 0.f;
}
int launch_relu_drop_bwd(const float* act, float scale, long count, float* d, cudaStream_t st) {
    if (count <= 0) return CM_OK;
    CM_TIMED("train_elementwise", st);
    tc_relu_drop_bwd_kernel<<<ceil_div(count, 256), 256, 0, st>>>(act, scale, count, d);
    return post_launch("relu_drop_bwd");
}

__global__ void __launch_bounds__(256) ln_fwd_kernel(const float* __restrict__ x, const float* __restrict__ gamma,
                                                     const float* __restrict__ beta, int rows, int dim, float* __restrict__ y,
                                                     float* __restrict__ stats) {
    const int r = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (r >= rows) return;
    const float* xr = x + (size_t)r * dim;
    float s = 0.f;
    for (int c = lane; c < dim; c += 32) s += xr[c];
    const float mean = warp_sum(s) / dim;
    float q = 0.f;
    for (int c = lane; c < dim; c += 32) {
        const float d = xr[c] - mean;
        q = fmaf(d, d, q);
    }
    const float rstd = rsqrtf(warp_sum(q) / dim + 1e-5f);
    for (int c = lane; c < dim; c += 32) y[(size_t)r * dim + c] = (xr[c] - mean) * rstd * gamma[c] + beta[c];
    if (lane == 0) {
        stats[2 * r] = mean;
        stats[2 * r + 1] = rstd;
    }
}
int launch_ln_fwd(const float* x, const float* gamma, const float* beta, int rows, int dim, float* y, float* stats, cudaStream_t st) {
    if (rows <= 0) return CM_OK;
    CM_TIMED("train_layernorm", st);
    ln_fwd_kernel<<<ceil_div(rows, 8), 256, 0, st>>>(x, gamma, beta, rows, dim, y, stats);
    return post_launch("ln_fwd");
}
__global__ void __launch_bounds__(256) ln_bwd_kernel(const float* __restrict__ x, const float* __restrict__ stats,
                                                     const float* __restrict__ gamma, const float* __restrict__ dy, int rows, int dim,
                                                     float* __restrict__ dx, float* __restrict__ prod) {
    const int r = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (r >= rows) return;
    const float mean = stats[2 * r], rstd = stats[2 * r + 1];
    const float* xr = x + (size_t)r * dim;
    const float* dr = dy + (size_t)r * dim;
    float a = 0.f, b = 0.f;
    for (int c = lane; c < dim; c += 32) {
        const float xh = (xr[c] - mean) * rstd, g = dr[c] * gamma[c];
        a += g;
        b = fmaf(g, xh, b);
    }
    a = warp_sum(a) / dim;
    b = warp_sum(b) / dim;
    for (int c = lane; c < dim; c += 32) {
        const float xh = (xr[c] - mean) * rstd, d = dr[c];
        prod[(size_t)r * dim + c] = d * xh;
        dx[(size_t)r * dim + c] = rstd * (d * gamma[c] - a - xh * b);
    }
}
int launch_ln_bwd(const float* x, const float* stats, const float* gamma, const float* dy, int rows, int dim, float* dx, float* prod,
                  cudaStream_t st) {
    if (rows <= 0) return CM_OK;
    CM_TIMED("train_layernorm", st);
    ln_bwd_kernel<<<ceil_div(rows, 8), 256, 0, st>>>(x, stats, gamma, dy, rows, dim, dx, prod);
    return post_launch("ln_bwd");
}

}  // namespace cm

// Exported building block / unit-test hook of the training GEMM (include/cm_b200.h): C[M,N] = A[M,K] W[N,K]^T (+ bias).
extern "C" int cm_train_gemm_f32(const float* a, const float* w, const float* bias, int m, int n, int k, int transposed_a, float* c, void* stream) {
    using namespace cm;
    CM_DEVICE_OR_RETURN();
    CM_REQUIRE(a && w && c && m > 0 && n > 0 && k > 0, "cm_train_gemm_f32: bad arguments");
    Gemm g = gemm_nt(a, k, w, c, n, m, n, k, bias, 0);
    if (transposed_a) {   // a is [K, M] (the wgrad pattern: reduction over the rows of the stored matrix)
        g.sam = 1;
        g.sak = m;
        g.a_k_contig = false;
    }
    return launch_gemm(g, as_stream(stream), "train_linear");
}

