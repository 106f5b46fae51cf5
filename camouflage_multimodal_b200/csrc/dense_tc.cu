// tcgen05 / TMEM / TMA dense layer (see dense_tc.cuh).  One 128 x BN output tile per CTA:
//   warp 0   TMEM allocation, then TMA producer (one elected lane)
//   warp 1   mbarrier init, then tcgen05.mma issuer (one elected lane), commits free the smem stages
//   warps 2-5  epilogue: TMEM lane quarter (warp & 3), one output row per thread
// Shared memory: 4 stages x (A 128x64 fp16 = 16 KB + W BNx64 fp16) in the canonical K-major SWIZZLE_128B
// layout that both the TMA tensor map and the UMMA shared-memory descriptor describe.
#include <cuda.h>
#include <cudaTypedefs.h>

#include <algorithm>
#include <map>
#include <tuple>

#include "dense_tc.cuh"
#include "tc_ptx.cuh"

namespace cm {

constexpr int TC_BM = 128, TC_BK = 64, TC_STAGES = 4, TC_THREADS = 320, TC_EPI_THREADS = 256;

template <int BN>
struct TcSmem {
    static constexpr int A_BYTES = TC_BM * TC_BK * 2;
    static constexpr int B_BYTES = BN * TC_BK * 2;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int BAR_OFFSET = TC_STAGES * STAGE_BYTES;
    static constexpr int VEC_OFFSET = BAR_OFFSET + 128;          // bias / gamma / beta of the tile's columns (3 x 256 fp32)
    static constexpr int STAT_OFFSET = VEC_OFFSET + 3 * 1024;    // LayerNorm partial sums: [2 halves][128 rows] float2
    static constexpr int TOTAL = STAT_OFFSET + 2048 + 1024;      // + slack for 1 KB alignment
};

template <int BN>
__global__ void __launch_bounds__(TC_THREADS, 1)
dense_tc_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_w, Dense p) {
    using S = TcSmem<BN>;
    extern __shared__ uint8_t smem_raw[];
    const int m_live = p.m_dev ? min(*p.m_dev, p.M) : p.M;
    const int m0 = blockIdx.y * TC_BM, n0 = blockIdx.x * BN;
    if (m0 >= m_live) return;
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + S::BAR_OFFSET);
    uint64_t* empty = full + TC_STAGES;
    uint64_t* acc_ready = empty + TC_STAGES;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_ready + 1);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int all_kb = (p.K + TC_BK - 1) / TC_BK;
    const int kb_per = p.k_splits > 1 ? (all_kb + p.k_splits - 1) / p.k_splits : all_kb;
    const int kb_first = p.k_splits > 1 ? blockIdx.z * kb_per : 0;
    const int num_kb = max(min(all_kb, kb_first + kb_per) - kb_first, 0);   // k-blocks of this CTA's slice (>= 1: the launcher sizes k_splits)
    if (p.k_splits > 1) p.out_f32 += (size_t)blockIdx.z * p.split_stride;

    if (warp == 1 && lane == 0) {
        for (int s = 0; s < TC_STAGES; ++s) {
            mbar_init(full + s, 1);
            mbar_init(empty + s, 1);
        }
        mbar_init(acc_ready, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"((uint32_t)BN)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ---------------------------------------------------------------- TMA producer
        for (int kb = 0; kb < num_kb; ++kb) {
            if (lane == 0) {
                const int s = kb % TC_STAGES, n = kb / TC_STAGES;
                if (n > 0) mbar_wait(empty + s, (n - 1) & 1);
                mbar_expect_tx(full + s, S::STAGE_BYTES);
                uint8_t* a_dst = smem + s * S::STAGE_BYTES;
                tma_load_2d(a_dst, &map_a, full + s, (kb_first + kb) * TC_BK, m0);
                tma_load_2d(a_dst + S::A_BYTES, &map_w, full + s, (kb_first + kb) * TC_BK, n0);
            }
            __syncwarp();
        }
    } else if (warp == 1) {
        // ---------------------------------------------------------------- MMA issuer
        const uint32_t idesc = umma_idesc_f16(TC_BM, BN) | (p.bf16_operands ? ((1u << 7) | (1u << 10)) : 0u);   // A / B format: 0 fp16, 1 bf16
        for (int kb = 0; kb < num_kb; ++kb) {
            if (lane == 0) {
                const int s = kb % TC_STAGES, n = kb / TC_STAGES;
                mbar_wait(full + s, n & 1);
                tc_fence_after();
                const uint32_t a_addr = smem_u32(smem + s * S::STAGE_BYTES);
                const uint64_t adesc = umma_desc_sw128(a_addr), bdesc = umma_desc_sw128(a_addr + S::A_BYTES);
#pragma unroll
                for (int k = 0; k < TC_BK / 16; ++k)   // +32 bytes (2 x 16 B units) per UMMA_K = 16 step
                    tc_mma_f16(tmem_base, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) != 0);
                tc_commit(empty + s);                  // frees the stage once these MMAs retire
                if (kb == num_kb - 1) tc_commit(acc_ready);
            }
            __syncwarp();
        }
    } else {
        // ---------------------------------------------------------------- epilogue (warps 2..9)
        // TMEM lane quarter q = warp % 4 (hardware rule); the two warps of a quarter split the tile's 32-column
        // chunks between them (half 0: even chunks, half 1: odd chunks; scores mode: [0,128) / [128,256)).
        const int q = warp & 3, half = (warp - 2) >> 2;
        const int et = threadIdx.x - 64;
        const int row = m0 + q * 32 + lane;
        const bool row_ok = row < m_live;
        const int n_cols = min(BN, p.N - n0);
        float* s_bias = reinterpret_cast<float*>(smem + S::VEC_OFFSET);
        float* s_gamma = s_bias + 256;
        float* s_beta = s_gamma + 256;
        float2* s_stat = reinterpret_cast<float2*>(smem + S::STAT_OFFSET);
        for (int c = et; c < BN; c += TC_EPI_THREADS) {
            const bool in = c < n_cols;
            s_bias[c] = (in && p.bias) ? p.bias[n0 + c] : 0.f;
            if (p.ln_gamma) {
                s_gamma[c] = in ? p.ln_gamma[n0 + c] : 0.f;
                s_beta[c] = in ? p.ln_beta[n0 + c] : 0.f;
            }
        }
        asm volatile("bar.sync 1, %0;" ::"n"(TC_EPI_THREADS) : "memory");
        mbar_wait(acc_ready, 0);
        __syncwarp();
        tc_fence_after();
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16);
        const size_t res_row = (size_t)(p.res_mod > 0 ? row % p.res_mod : row) * p.ldr;
        const op_t* res = (p.residual && row_ok) ? p.residual + res_row + n0 : nullptr;
        const float* resf = (p.residual_f32 && row_ok) ? p.residual_f32 + res_row + n0 : nullptr;
        const bool res_vec = res && (p.ldr & 7) == 0 && (reinterpret_cast<uintptr_t>(p.residual + n0) & 15) == 0;
        const bool resf_vec = resf && (p.ldr & 3) == 0 && (reinterpret_cast<uintptr_t>(p.residual_f32 + n0) & 15) == 0;
        const int c_begin = p.sm_gs ? half * 128 : half * 32, c_step = p.sm_gs ? 32 : 64;
        const int c_end = p.sm_gs ? min(n_cols, c_begin + 128) : n_cols;

        // x[j] = relu?( alpha * (acc + bias) + residual ) for the 32 columns starting at c0 (all lanes take part
        // in the TMEM load; global loads only for live rows)
        auto load_chunk = [&](int c0, float (&x)[32]) {
            uint32_t v[32];
            tc_ld32(taddr + c0, v);
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
                const float4 b4 = *reinterpret_cast<const float4*>(s_bias + c0 + j);
                x[j] = __uint_as_float(v[j]) + b4.x;
                x[j + 1] = __uint_as_float(v[j + 1]) + b4.y;
                x[j + 2] = __uint_as_float(v[j + 2]) + b4.z;
                x[j + 3] = __uint_as_float(v[j + 3]) + b4.w;
            }
            if (p.alpha != 1.0f) {
#pragma unroll
                for (int j = 0; j < 32; ++j) x[j] *= p.alpha;
            }
            const int valid = n_cols - c0;
            if (res) {
                if (res_vec && valid >= 32) {
#pragma unroll
                    for (int j = 0; j < 32; j += 8) {
                        const uint4 r8 = *reinterpret_cast<const uint4*>(res + c0 + j);
                        const uint32_t w[4] = {r8.x, r8.y, r8.z, r8.w};
#pragma unroll
                        for (int t = 0; t < 4; ++t) {
                            const float2 f = op22f(*reinterpret_cast<const op2_t*>(&w[t]));
                            x[j + 2 * t] += f.x;
                            x[j + 2 * t + 1] += f.y;
                        }
                    }
                } else {
#pragma unroll
                    for (int j = 0; j < 32; ++j)
                        if (j < valid) x[j] += op2f(res[c0 + j]);
                }
            }
            if (resf) {
                if (resf_vec && valid >= 32) {
#pragma unroll
                    for (int j = 0; j < 32; j += 4) {
                        const float4 r4 = *reinterpret_cast<const float4*>(resf + c0 + j);
                        x[j] += r4.x;
                        x[j + 1] += r4.y;
                        x[j + 2] += r4.z;
                        x[j + 3] += r4.w;
                    }
                } else {
#pragma unroll
                    for (int j = 0; j < 32; ++j)
                        if (j < valid) x[j] += resf[c0 + j];
                }
            }
            if (p.relu) {
#pragma unroll
                for (int j = 0; j < 32; ++j) x[j] = fmaxf(x[j], 0.f);
            }
        };

        float mean = 0.f, rstd = 1.f;
        if (p.ln_gamma) {   // pass 1: row statistics over the whole tile (the two halves exchange partial sums)
            float s1 = 0.f, s2 = 0.f;
            for (int c0 = c_begin; c0 < c_end; c0 += c_step) {
                float x[32];
                load_chunk(c0, x);
                const int valid = n_cols - c0;
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    const float t = (valid >= 32 || j < valid) ? x[j] : 0.f;
                    s1 += t;
                    s2 = fmaf(t, t, s2);
                }
            }
            s_stat[half * 128 + q * 32 + lane] = make_float2(s1, s2);
            asm volatile("bar.sync 1, %0;" ::"n"(TC_EPI_THREADS) : "memory");
            const float2 o = s_stat[(half ^ 1) * 128 + q * 32 + lane];
            s1 += o.x;
            s2 += o.y;
            mean = s1 / (float)n_cols;
            rstd = rsqrtf(fmaxf(s2 / (float)n_cols - mean * mean, 0.f) + 1e-5f);
        }
        float am[32];   // scores mode: head-summed probabilities of this row
#pragma unroll
        for (int j = 0; j < 32; ++j) am[j] = 0.f;
        for (int c0 = c_begin; c0 < c_end; c0 += c_step) {
            float x[32];
            load_chunk(c0, x);
            if (p.ln_gamma) {
#pragma unroll
                for (int j = 0; j < 32; j += 4) {
                    const float4 g4 = *reinterpret_cast<const float4*>(s_gamma + c0 + j), b4 = *reinterpret_cast<const float4*>(s_beta + c0 + j);
                    x[j] = (x[j] - mean) * rstd * g4.x + b4.x;
                    x[j + 1] = (x[j + 1] - mean) * rstd * g4.y + b4.y;
                    x[j + 2] = (x[j + 2] - mean) * rstd * g4.z + b4.z;
                    x[j + 3] = (x[j + 3] - mean) * rstd * g4.w + b4.w;
                }
            }
            const int valid = n_cols - c0;
            const bool full32 = valid >= 32;
            if (p.sm_gs) {
                // scores mode (N == 256): columns [0,128) hold per-head groups of sm_gs slots (sm_valid live keys
                // each) -> softmax within the group, fp16 out [M,128]; columns [128,256) -> raw fp32 scores [M,128]
                const int c = n0 + c0;
                if (c < 128) {
                    if (p.sm_gs == 16) softmax_groups<16>(x, p.sm_valid, am); else softmax_groups<32>(x, p.sm_valid, am);
                    if (row_ok) {
                        op_t* o = p.out_h + (size_t)row * 128 + c;
#pragma unroll
                        for (int j = 0; j < 32; j += 8) *reinterpret_cast<uint4*>(o + j) = pack8(x + j);
                        if (c == 96 && p.attn_mean) {
                            const float inv = (float)p.sm_gs / 128.f;   // 1 / heads
#pragma unroll
                            for (int j = 0; j < 32; ++j)
                                if (j < p.sm_valid) p.attn_mean[(size_t)row * p.sm_valid + j] = am[j] * inv;
                        }
                    }
                } else if (row_ok) {
                    float* o = p.out_f32 + (size_t)row * 128 + (c - 128);
#pragma unroll
                    for (int j = 0; j < 32; j += 4) *reinterpret_cast<float4*>(o + j) = make_float4(x[j], x[j + 1], x[j + 2], x[j + 3]);
                }
            } else if (row_ok) {
                if (p.out_h && p.split_out) {
                    // [hi | lo | hi] thirds of a 3N-wide row: the next layer multiplies them with [W_hi | W_hi | W_lo]
                    op_t* o = p.out_h + (size_t)row * p.ldo_h + n0 + c0;
                    float lo[32];
#pragma unroll
                    for (int j = 0; j < 32; ++j) lo[j] = x[j] - op2f(f2op(x[j]));
                    if (full32 && ((reinterpret_cast<uintptr_t>(o) & 15) == 0) && (p.N & 7) == 0) {
#pragma unroll
                        for (int j = 0; j < 32; j += 8) {
                            const uint4 h8 = pack8(x + j);
                            *reinterpret_cast<uint4*>(o + j) = h8;
                            *reinterpret_cast<uint4*>(o + p.N + j) = pack8(lo + j);
                            *reinterpret_cast<uint4*>(o + 2 * p.N + j) = h8;
                        }
                    } else {
#pragma unroll
                        for (int j = 0; j < 32; ++j)
                            if (j < valid) {
                                o[j] = o[2 * p.N + j] = f2op(x[j]);
                                o[p.N + j] = f2op(lo[j]);
                            }
                    }
                } else if (p.out_h) {
                    op_t* o = p.out_h + (size_t)row * p.ldo_h + n0 + c0;
                    if (full32 && ((reinterpret_cast<uintptr_t>(o) & 15) == 0)) {
#pragma unroll
                        for (int j = 0; j < 32; j += 8) *reinterpret_cast<uint4*>(o + j) = pack8(x + j);
                    } else {
#pragma unroll
                        for (int j = 0; j < 32; ++j)
                            if (j < valid) o[j] = f2op(x[j]);
                    }
                }
                if (p.out_f32) {
                    float* o = p.out_f32 + (size_t)row * p.ldo_f32 + n0 + c0;
                    if (full32 && ((reinterpret_cast<uintptr_t>(o) & 15) == 0)) {
#pragma unroll
                        for (int j = 0; j < 32; j += 4) *reinterpret_cast<float4*>(o + j) = make_float4(x[j], x[j + 1], x[j + 2], x[j + 3]);
                    } else {
#pragma unroll
                        for (int j = 0; j < 32; ++j)
                            if (j < valid) o[j] = x[j];
                    }
                }
            }
            __syncwarp();
        }
        tc_fence_before();
    }
    __syncthreads();
    if (warp == 0) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)BN) : "memory");
    }
}

// ------------------------------------------------------------------------------------------- host side
static PFN_cuTensorMapEncodeTiled_v12000 g_encode = nullptr;

static int get_encoder() {
    if (g_encode) return CM_OK;
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
    if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !fn) {
        set_error("cuTensorMapEncodeTiled unavailable: %s", cudaGetErrorString(e));
        cudaGetLastError();
        return CM_ERR_DEVICE;
    }
    g_encode = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(fn);
    return CM_OK;
}

// 2-D fp16 tensor map: inner dim = cols (K), outer = rows; box = 64 x box_rows, 128-byte swizzle, OOB -> 0.
int make_map(CUtensorMap* out, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows) {
    using Key = std::tuple<const void*, uint64_t, uint64_t, uint64_t, uint32_t>;
    static std::map<Key, CUtensorMap> cache;
    Key key{ptr, rows, cols, ld, box_rows};
    auto it = cache.find(key);
    if (it != cache.end()) {
        *out = it->second;
        return CM_OK;
    }
    int rc = get_encoder();
    if (rc != CM_OK) return rc;
    cuuint64_t gdim[2] = {cols, rows};
    cuuint64_t gstride[1] = {ld * 2};
    cuuint32_t box[2] = {(cuuint32_t)TC_BK, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = g_encode(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void*>(ptr), gdim, gstride, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled failed (%d) for ptr=%p rows=%llu cols=%llu ld=%llu", (int)r, ptr,
                  (unsigned long long)rows, (unsigned long long)cols, (unsigned long long)ld);
        return CM_ERR_LAUNCH;
    }
    if (cache.size() > 4096) cache.clear();
    cache[key] = *out;
    return CM_OK;
}

template <int BN>
static int launch_dense_bn(const Dense& d, cudaStream_t st) {
    CUtensorMap ma, mw;
    int rc = make_map(&ma, d.A, (uint64_t)d.M, (uint64_t)d.K, (uint64_t)d.lda, TC_BM);
    if (rc != CM_OK) return rc;
    rc = make_map(&mw, d.W, (uint64_t)d.N, (uint64_t)d.K, (uint64_t)d.K, BN);
    if (rc != CM_OK) return rc;
    static PerDeviceOnce attr;
    if (attr.first()) {
        cudaFuncSetAttribute(dense_tc_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, TcSmem<BN>::TOTAL);
    }
    dim3 grid(ceil_div(d.N, BN), ceil_div(d.M, TC_BM), d.k_splits > 1 ? d.k_splits : 1);
    {
        CM_TIMED("dense_tc", st);
        dense_tc_kernel<BN><<<grid, TC_THREADS, TcSmem<BN>::TOTAL, st>>>(ma, mw, d);
    }
    return post_launch("dense_tc");
}

int launch_dense(const Dense& d, cudaStream_t st) {
    if (d.M <= 0 || d.N <= 0 || d.K <= 0) return CM_OK;
    CM_REQUIRE(d.A && d.W && (d.out_h || d.out_f32), "dense: NULL operand");
    CM_REQUIRE(d.K % 8 == 0 && d.lda % 8 == 0 && d.lda >= d.K, "dense: K=%d / lda=%d must be multiples of 8", d.K, d.lda);
    CM_REQUIRE((reinterpret_cast<uintptr_t>(d.A) & 15) == 0 && (reinterpret_cast<uintptr_t>(d.W) & 15) == 0, "dense: operands must be 16-byte aligned");
    CM_REQUIRE(!d.ln_gamma || d.N <= 256, "dense: fused LayerNorm needs N <= 256 (N=%d)", d.N);
    CM_REQUIRE(!d.bf16_operands || (!d.out_h && !d.residual), "dense: bf16 operands go with fp32 outputs only");
    if (d.k_splits > 1) {
        const int all_kb = (d.K + TC_BK - 1) / TC_BK, per = (all_kb + d.k_splits - 1) / d.k_splits;
        CM_REQUIRE(d.out_f32 && !d.out_h && !d.bias && !d.residual && !d.residual_f32 && !d.ln_gamma && !d.relu && !d.sm_gs && d.alpha == 1.f &&
                       !d.m_dev && (long long)per * (d.k_splits - 1) < all_kb,
                   "dense: split-K writes raw fp32 partial tiles only, every slice non-empty");
    }
    CM_REQUIRE(!d.sm_gs || ((d.sm_gs == 16 || d.sm_gs == 32) && d.N == 256 && d.out_h && d.out_f32 && !d.ln_gamma &&
                            d.sm_valid >= 1 && d.sm_valid <= d.sm_gs),
               "dense: scores mode needs N == 256, group size 16/32, fp16 + f32 outputs");
    if (d.N > 128) return launch_dense_bn<256>(d, st);
    return launch_dense_bn<128>(d, st);
}

__global__ void cast_f16_kernel(const float* __restrict__ in, int ld_in, int rows, int cols, const int* rows_dev,
                                 op_t* __restrict__ out, int ld_out) {
    const int live = rows_dev ? min(*rows_dev, rows) : rows;
    const long long total = (long long)live * ld_out;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int r = (int)(i / ld_out), c = (int)(i - (long long)r * ld_out);
        out[i] = f2op(c < cols ? in[(size_t)r * ld_in + c] : 0.f);
    }
}

// dense rows (ld_in == ld_out == cols, cols % 8 == 0, 16-byte aligned): eight values per thread per trip
__global__ void __launch_bounds__(256) cast_f16_flat_kernel(const float* __restrict__ in, int rows, int cols, const int* rows_dev,
                                                             op_t* __restrict__ out) {
    const int live = rows_dev ? min(*rows_dev, rows) : rows;
    const long long n8 = (long long)live * cols / 8;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (long long)gridDim.x * blockDim.x) {
        const float4 a = reinterpret_cast<const float4*>(in)[2 * i], b = reinterpret_cast<const float4*>(in)[2 * i + 1];
        const float x[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
        reinterpret_cast<uint4*>(out)[i] = pack8(x);
    }
}

int launch_cast_f16(const float* in, int ld_in, int rows, int cols, const int* rows_dev, op_t* out, int ld_out,
                     cudaStream_t st) {
    if (rows <= 0) return CM_OK;
    const long long total = (long long)rows * ld_out;
    if (ld_in == cols && ld_out == cols && (cols & 7) == 0 && ((reinterpret_cast<uintptr_t>(in) | reinterpret_cast<uintptr_t>(out)) & 15) == 0) {
        const int blocks = (int)std::min<long long>((total / 8 + 255) / 256, 148LL * 8);
        CM_TIMED("cast_f16", st);
        cast_f16_flat_kernel<<<blocks, 256, 0, st>>>(in, rows, cols, rows_dev, out);
        return post_launch("cast_f16");
    }
    int blocks = (int)std::min<long long>((total + 255) / 256, 148LL * 16);
    {
        CM_TIMED("cast_f16", st);
        cast_f16_kernel<<<blocks, 256, 0, st>>>(in, ld_in, rows, cols, rows_dev, out, ld_out);
    }
    return post_launch("cast_f16");
}

// fp32 [rows, cols] (optionally scaled per row) -> fp16 [rows, 3*kpad] = [hi | hi | lo], each third zero-padded to kpad
__global__ void split_weight_kernel(const float* __restrict__ in, int rows, int cols, int kpad, const float* __restrict__ row_scale,
                                    op_t* __restrict__ out) {
    const long long total = (long long)rows * kpad;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int r = (int)(i / kpad), c = (int)(i - (long long)r * kpad);
        float v = c < cols ? in[(size_t)r * cols + c] : 0.f;
        if (row_scale) v *= row_scale[r];
        const op_t hi = f2op(v);
        op_t* o = out + (size_t)r * 3 * kpad + c;
        o[0] = hi;
        o[kpad] = hi;
        o[2 * kpad] = f2op(v - op2f(hi));
    }
}

int launch_split_weight(const float* in, int rows, int cols, int kpad, const float* row_scale, op_t* out, cudaStream_t st) {
    if (rows <= 0) return CM_OK;
    {
        CM_TIMED("fold_weights", st);
        split_weight_kernel<<<ceil_div((long long)rows * kpad, 256), 256, 0, st>>>(in, rows, cols, kpad, row_scale, out);
    }
    return post_launch("split_weight");
}

}  // namespace cm

// Exported building block (also the unit-test hook for the tensor-core path): see include/cm_b200.h.
extern "C" int cm_linear_f16(const void* a_f16, int lda, const void* w_f16, int m, int n, int k, const int32_t* m_dev,
                              const float* bias, float alpha, int relu, const void* residual_h, int ldr,
                              const float* ln_gamma, const float* ln_beta, void* out_h, float* out_f32, void* stream) {
    using namespace cm;
    CM_DEVICE_OR_RETURN();
    Dense d;
    d.A = reinterpret_cast<const op_t*>(a_f16);
    d.lda = lda;
    d.W = reinterpret_cast<const op_t*>(w_f16);
    d.M = m; d.N = n; d.K = k;
    d.m_dev = m_dev;
    d.bias = bias; d.alpha = alpha; d.relu = relu;
    d.residual = reinterpret_cast<const op_t*>(residual_h);
    d.ldr = ldr;
    d.ln_gamma = ln_gamma; d.ln_beta = ln_beta;
    d.out_h = reinterpret_cast<op_t*>(out_h);
    d.ldo_h = n;
    d.out_f32 = out_f32;
    d.ldo_f32 = n;
    CM_REQUIRE(m > 0 && n > 0 && k > 0, "cm_linear_f16: bad shape");
    CM_REQUIRE(!ln_gamma || ln_beta, "cm_linear_f16: ln_beta missing");
    return launch_dense(d, as_stream(stream));
}
