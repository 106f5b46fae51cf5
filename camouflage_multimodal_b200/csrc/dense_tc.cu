// tcgen05 / TMEM / TMA dense layer (see dense_tc.cuh).  One 128 x BN output tile per CTA:
//   warp 0   TMEM allocation, then TMA producer (one elected lane)
//   warp 1   mbarrier init, then tcgen05.mma issuer (one elected lane), commits free the smem stages
//   warps 2-5  epilogue: TMEM lane quarter (warp & 3), one output row per thread
// Shared memory: 4 stages x (A 128x64 bf16 = 16 KB + W BNx64 bf16) in the canonical K-major SWIZZLE_128B
// layout that both the TMA tensor map and the UMMA shared-memory descriptor describe.
#include <cuda.h>
#include <cudaTypedefs.h>

#include <algorithm>
#include <map>
#include <tuple>

#include "dense_tc.cuh"

namespace cm {

constexpr int TC_BM = 128, TC_BK = 64, TC_STAGES = 4, TC_THREADS = 192;

// ------------------------------------------------------------------------------------------ PTX wrappers
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
// Bounded wait: a broken pipeline traps (launch error) instead of hanging the GPU.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    const uint32_t addr = smem_u32(bar);
    for (uint32_t spin = 0;; ++spin) {
        uint32_t done;
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(addr), "r"(parity)
            : "memory");
        if (done) return;
        if (spin > (1u << 26)) __trap();
    }
}
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
            smem_u32(smem_dst)),
        "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// D[tmem] (+)= A[smem] * B[smem]^T, bf16 x bf16 -> fp32, issued by one thread for the CTA
__device__ __forceinline__ void tc_mma_bf16(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// 32 lanes x 32 columns of fp32: thread t of the warp gets row (lane quarter base + t), columns c..c+31
__device__ __forceinline__ void tc_ld32(uint32_t taddr, uint32_t (&v)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
          "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
          "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
          "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// UMMA shared-memory descriptor, K-major, SWIZZLE_128B: 8-row groups 1024 B apart (SBO), LBO unused (1),
// version 1 (sm_100), layout type 2.  Start address in 16-byte units.
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t smem_addr) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr & 0x3FFFFu) >> 4);
    d |= (uint64_t)1 << 16;
    d |= (uint64_t)(1024 >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}
// Instruction descriptor, kind::f16: D fp32 (bits 4-5 = 1), A = B = bf16 (bits 7-9, 10-12 = 1), both K-major,
// N >> 3 at bits 17-22, M >> 4 at bits 24-28.
__host__ __device__ constexpr uint32_t umma_idesc_bf16(int m, int n) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}

template <int BN>
struct TcSmem {
    static constexpr int A_BYTES = TC_BM * TC_BK * 2;
    static constexpr int B_BYTES = BN * TC_BK * 2;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int BAR_OFFSET = TC_STAGES * STAGE_BYTES;
    static constexpr int TOTAL = BAR_OFFSET + 128 + 1024;   // barriers + tmem slot, + slack for 1 KB alignment
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
    const int num_kb = (p.K + TC_BK - 1) / TC_BK;

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
                tma_load_2d(a_dst, &map_a, full + s, kb * TC_BK, m0);
                tma_load_2d(a_dst + S::A_BYTES, &map_w, full + s, kb * TC_BK, n0);
            }
            __syncwarp();
        }
    } else if (warp == 1) {
        // ---------------------------------------------------------------- MMA issuer
        constexpr uint32_t idesc = umma_idesc_bf16(TC_BM, BN);
        for (int kb = 0; kb < num_kb; ++kb) {
            if (lane == 0) {
                const int s = kb % TC_STAGES, n = kb / TC_STAGES;
                mbar_wait(full + s, n & 1);
                tc_fence_after();
                const uint32_t a_addr = smem_u32(smem + s * S::STAGE_BYTES);
                const uint64_t adesc = umma_desc_sw128(a_addr), bdesc = umma_desc_sw128(a_addr + S::A_BYTES);
#pragma unroll
                for (int k = 0; k < TC_BK / 16; ++k)   // +32 bytes (2 x 16 B units) per UMMA_K = 16 step
                    tc_mma_bf16(tmem_base, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) != 0);
                tc_commit(empty + s);                  // frees the stage once these MMAs retire
                if (kb == num_kb - 1) tc_commit(acc_ready);
            }
            __syncwarp();
        }
    } else {
        // ---------------------------------------------------------------- epilogue (warps 2..5)
        const int q = warp & 3;                       // TMEM lane quarter this warp may read
        const int row = m0 + q * 32 + lane;
        const bool row_ok = row < m_live;
        mbar_wait(acc_ready, 0);
        __syncwarp();
        tc_fence_after();
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16);
        const __nv_bfloat16* res = nullptr;
        if (p.residual && row_ok) res = p.residual + (size_t)(p.res_mod > 0 ? row % p.res_mod : row) * p.ldr;
        const int n_cols = min(BN, p.N - n0);
        float mean = 0.f, rstd = 1.f;
        if (p.ln_gamma) {
            float s1 = 0.f, s2 = 0.f;
            for (int c0 = 0; c0 < n_cols; c0 += 32) {
                uint32_t v[32];
                tc_ld32(taddr + c0, v);
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    const int c = n0 + c0 + j;
                    if (c0 + j < n_cols) {
                        float x = __uint_as_float(v[j]);
                        if (p.bias) x += p.bias[c];
                        x *= p.alpha;
                        if (res) x += __bfloat162float(res[c]);
                        if (p.relu) x = fmaxf(x, 0.f);
                        s1 += x;
                        s2 = fmaf(x, x, s2);
                    }
                }
            }
            mean = s1 / (float)n_cols;
            const float var = fmaxf(s2 / (float)n_cols - mean * mean, 0.f);
            rstd = rsqrtf(var + 1e-5f);
        }
        for (int c0 = 0; c0 < n_cols; c0 += 32) {
            uint32_t v[32];
            tc_ld32(taddr + c0, v);   // warp-collective: every lane takes part, stores are per row
            if (row_ok) {
            float x[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) {
                const int c = n0 + c0 + j;
                float t = __uint_as_float(v[j]);
                if (c0 + j < n_cols) {
                    if (p.bias) t += p.bias[c];
                    t *= p.alpha;
                    if (res) t += __bfloat162float(res[c]);
                    if (p.relu) t = fmaxf(t, 0.f);
                    if (p.ln_gamma) t = (t - mean) * rstd * p.ln_gamma[c] + p.ln_beta[c];
                }
                x[j] = t;
            }
            const bool full32 = c0 + 32 <= n_cols;
            if (p.out_bf16) {
                __nv_bfloat16* o = p.out_bf16 + (size_t)row * p.ldo_bf16 + n0 + c0;
                if (full32 && ((reinterpret_cast<uintptr_t>(o) & 15) == 0)) {
#pragma unroll
                    for (int j = 0; j < 32; j += 8) {
                        uint4 pk;
                        __nv_bfloat162 b0 = __floats2bfloat162_rn(x[j], x[j + 1]), b1 = __floats2bfloat162_rn(x[j + 2], x[j + 3]);
                        __nv_bfloat162 b2 = __floats2bfloat162_rn(x[j + 4], x[j + 5]), b3 = __floats2bfloat162_rn(x[j + 6], x[j + 7]);
                        pk.x = *reinterpret_cast<uint32_t*>(&b0);
                        pk.y = *reinterpret_cast<uint32_t*>(&b1);
                        pk.z = *reinterpret_cast<uint32_t*>(&b2);
                        pk.w = *reinterpret_cast<uint32_t*>(&b3);
                        *reinterpret_cast<uint4*>(o + j) = pk;
                    }
                } else {
                    for (int j = 0; j < 32 && c0 + j < n_cols; ++j) o[j] = __float2bfloat16(x[j]);
                }
            }
            if (p.out_f32) {
                float* o = p.out_f32 + (size_t)row * p.ldo_f32 + n0 + c0;
                if (full32 && ((reinterpret_cast<uintptr_t>(o) & 15) == 0)) {
#pragma unroll
                    for (int j = 0; j < 32; j += 4) *reinterpret_cast<float4*>(o + j) = make_float4(x[j], x[j + 1], x[j + 2], x[j + 3]);
                } else {
                    for (int j = 0; j < 32 && c0 + j < n_cols; ++j) o[j] = x[j];
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

// 2-D bf16 tensor map: inner dim = cols (K), outer = rows; box = 64 x box_rows, 128-byte swizzle, OOB -> 0.
static int make_map(CUtensorMap* out, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows) {
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
    CUresult r = g_encode(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), gdim, gstride, box, estr,
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
    static bool attr = false;
    if (!attr) {
        cudaFuncSetAttribute(dense_tc_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, TcSmem<BN>::TOTAL);
        attr = true;
    }
    dim3 grid(ceil_div(d.N, BN), ceil_div(d.M, TC_BM));
    {
        CM_TIMED("dense_tc", st);
        dense_tc_kernel<BN><<<grid, TC_THREADS, TcSmem<BN>::TOTAL, st>>>(ma, mw, d);
    }
    return post_launch("dense_tc");
}

int launch_dense(const Dense& d, cudaStream_t st) {
    if (d.M <= 0 || d.N <= 0 || d.K <= 0) return CM_OK;
    CM_REQUIRE(d.A && d.W && (d.out_bf16 || d.out_f32), "dense: NULL operand");
    CM_REQUIRE(d.K % 8 == 0 && d.lda % 8 == 0 && d.lda >= d.K, "dense: K=%d / lda=%d must be multiples of 8", d.K, d.lda);
    CM_REQUIRE((reinterpret_cast<uintptr_t>(d.A) & 15) == 0 && (reinterpret_cast<uintptr_t>(d.W) & 15) == 0, "dense: operands must be 16-byte aligned");
    CM_REQUIRE(!d.ln_gamma || d.N <= 256, "dense: fused LayerNorm needs N <= 256 (N=%d)", d.N);
    if (d.N > 128) return launch_dense_bn<256>(d, st);
    return launch_dense_bn<128>(d, st);
}

__global__ void cast_bf16_kernel(const float* __restrict__ in, int ld_in, int rows, int cols, const int* rows_dev,
                                 __nv_bfloat16* __restrict__ out, int ld_out) {
    const int live = rows_dev ? min(*rows_dev, rows) : rows;
    const long long total = (long long)live * ld_out;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int r = (int)(i / ld_out), c = (int)(i - (long long)r * ld_out);
        out[i] = __float2bfloat16(c < cols ? in[(size_t)r * ld_in + c] : 0.f);
    }
}

int launch_cast_bf16(const float* in, int ld_in, int rows, int cols, const int* rows_dev, __nv_bfloat16* out, int ld_out,
                     cudaStream_t st) {
    if (rows <= 0) return CM_OK;
    const long long total = (long long)rows * ld_out;
    int blocks = (int)std::min<long long>((total + 255) / 256, 148LL * 16);
    {
        CM_TIMED("cast_bf16", st);
        cast_bf16_kernel<<<blocks, 256, 0, st>>>(in, ld_in, rows, cols, rows_dev, out, ld_out);
    }
    return post_launch("cast_bf16");
}

}  // namespace cm

// Exported building block (also the unit-test hook for the tensor-core path): see include/cm_b200.h.
extern "C" int cm_linear_bf16(const void* a_bf16, int lda, const void* w_bf16, int m, int n, int k, const int32_t* m_dev,
                              const float* bias, float alpha, int relu, const void* residual_bf16, int ldr,
                              const float* ln_gamma, const float* ln_beta, void* out_bf16, float* out_f32, void* stream) {
    using namespace cm;
    CM_DEVICE_OR_RETURN();
    Dense d;
    d.A = reinterpret_cast<const __nv_bfloat16*>(a_bf16);
    d.lda = lda;
    d.W = reinterpret_cast<const __nv_bfloat16*>(w_bf16);
    d.M = m; d.N = n; d.K = k;
    d.m_dev = m_dev;
    d.bias = bias; d.alpha = alpha; d.relu = relu;
    d.residual = reinterpret_cast<const __nv_bfloat16*>(residual_bf16);
    d.ldr = ldr;
    d.ln_gamma = ln_gamma; d.ln_beta = ln_beta;
    d.out_bf16 = reinterpret_cast<__nv_bfloat16*>(out_bf16);
    d.ldo_bf16 = n;
    d.out_f32 = out_f32;
    d.ldo_f32 = n;
    CM_REQUIRE(m > 0 && n > 0 && k > 0, "cm_linear_bf16: bad shape");
    CM_REQUIRE(!ln_gamma || ln_beta, "cm_linear_bf16: ln_beta missing");
    return launch_dense(d, as_stream(stream));
}
