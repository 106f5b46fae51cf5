// Cross-attention for LARGE knowledge-graph vocabularies (BASELINE configs[4]: 1k .. 64k KG rows, up to 4096 regions
// per image): fused QK^T -> online softmax -> PV on the tcgen05 tensor cores, scores and the partial context in
// TMEM, operands staged by TMA, plus the host orchestration that runs the whole CrossAttentionFusion
// (models/multimodal/fusion_model.py:75-146) around it on the tcgen05 dense layer (dense_tc.cu).
//
// Kernel (head_dim 32, two heads per CTA so that one 64-column TMA box feeds both):
//   CTA = 128 query rows x one head pair x one key segment; 4 warps, thread t owns query row t (= TMEM lane t).
//   per block of 128 keys:  TMA K tile [128 keys x 64] + V^T tiles [64 dims x 128 keys]  (128-byte swizzle)
//     per head:  S[128x128] = Q_h K_h^T   tcgen05.mma M128 N128 K32 (2 k-steps at +0/+32 B inside the head's 64 B slice)
//                row max / exp2 / row sum from TMEM (tcgen05.ld), P -> fp16 -> shared memory in the UMMA K-major layout
//                PV[128x32] = P V_h       tcgen05.mma M128 N32 K128 (8 k-steps), fresh accumulator
//                o = o * exp2(m_old - m_new) + PV   in registers (32 floats per head per thread)
//   no S or P ever reaches HBM; the softmax scale and log2(e) are folded into one multiply; keys past the segment end
//   are masked by index (select), so ragged segments and zero-filled TMA tails are exact.
// Two CTAs fit per SM (81 KB shared memory, 256 TMEM columns each), so one CTA's TMA / MMA phases overlap the
// other's exponentials -- the MUFU unit is the bound at head_dim 32 (one exp per 64 tensor FLOPs).
#include <cuda.h>

#include <algorithm>

#include "dense_tc.cuh"
#include "flash_tc.cuh"
#include "linear.cuh"
#include "tc_ptx.cuh"

namespace cm {

// 64-key blocks, K / V^T double-buffered (the next block is in flight while this one is computed): 66 KB of shared
// memory and 128 TMEM columns per CTA, three CTAs per SM whose tensor, exp and TMA phases overlap
constexpr int FA_BM = 128, FA_BN = 64, FA_THREADS = 128;
constexpr int FA_Q = 0, FA_K = 16384, FA_V = 24576, FA_KV_STRIDE = 16384, FA_P = 49152, FA_BAR = 65536, FA_SMEM = FA_BAR + 128 + 1024;   // K/V double-buffered

__device__ __forceinline__ float fa_ex2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

__global__ void __launch_bounds__(FA_THREADS, 3) flash_attn_kernel(const __grid_constant__ CUtensorMap map_q, const __grid_constant__ CUtensorMap map_k,
                                                               const __grid_constant__ CUtensorMap map_vt, int Mq, const int* __restrict__ seg_ptr,
                                                               int n_keys, float scale_log2, op_t* __restrict__ out, int ldo) {
    extern __shared__ uint8_t fa_smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(fa_smem_raw) + 1023) & ~(uintptr_t)1023);
    const int tid = threadIdx.x, warp = tid >> 5;
    const int q0 = blockIdx.x * FA_BM, pair = blockIdx.y, seg = blockIdx.z;
    const int key0 = seg_ptr ? seg_ptr[seg] : 0, key1 = seg_ptr ? seg_ptr[seg + 1] : n_keys;
    // the V^T tiles have the keys on TMA's inner dimension, whose start must be 16-byte aligned: blocks are laid out from
    // the segment start rounded down to 8 keys, and the (at most 7) keys in front of the segment are masked like its tail
    const int key0_al = key0 & ~7;
    const int nk = key1 - key0, nblk = (key1 - key0_al + FA_BN - 1) / FA_BN;
    const int row = q0 + tid;
    op_t* orow = out + ((size_t)seg * Mq + row) * ldo + pair * 64;
    if (nk <= 0) {   // empty key segment: the reference would produce NaN (softmax over nothing); write zeros
        if (row < Mq)
            for (int i = 0; i < 8; ++i) reinterpret_cast<uint4*>(orow)[i] = make_uint4(0, 0, 0, 0);
        return;
    }
    uint64_t* bar_q = reinterpret_cast<uint64_t*>(smem + FA_BAR);
    uint64_t *bar_kv = bar_q + 1, *bar_s = bar_q + 3, *bar_o = bar_q + 4;   // bar_kv[2]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_q + 5);
    if (tid == 32) {
        mbar_init(bar_q, 1);
        mbar_init(bar_kv, 1);
        mbar_init(bar_kv + 1, 1);
        mbar_init(bar_s, 1);
        mbar_init(bar_o, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(128u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t t_s = tmem_base + ((uint32_t)(warp * 32) << 16), t_o = t_s + FA_BN;   // this warp's lane quarter
    constexpr uint32_t idesc_s = umma_idesc_f16(FA_BM, FA_BN), idesc_o = umma_idesc_f16(FA_BM, 32);
    const uint32_t aq = smem_u32(smem + FA_Q), ap = smem_u32(smem + FA_P);
    auto load_kv = [&](int blk) {   // thread 0: K rows and V^T columns of key block blk into buffer blk & 1
        const int kb = key0_al + blk * FA_BN, buf = blk & 1;
        mbar_expect_tx(bar_kv + buf, 16384);
        tma_load_2d(smem + FA_K + buf * FA_KV_STRIDE, &map_k, bar_kv + buf, pair * 64, kb);
        tma_load_2d(smem + FA_V + buf * FA_KV_STRIDE, &map_vt, bar_kv + buf, kb, pair * 64);
    };

    if (tid == 0) {
        mbar_expect_tx(bar_q, 16384);
        tma_load_2d(smem + FA_Q, &map_q, bar_q, pair * 64, q0);
        load_kv(0);
    }
    float m_run[2] = {-INFINITY, -INFINITY}, l_run[2] = {0.f, 0.f}, o[2][32];
#pragma unroll
    for (int i = 0; i < 32; ++i) o[0][i] = o[1][i] = 0.f;
    mbar_wait(bar_q, 0);
    uint32_t ph_s = 0, ph_o = 0;

    for (int blk = 0; blk < nblk; ++blk) {
        const int kb = key0_al + blk * FA_BN;
        // the other buffer was last read by block blk - 1, whose products retired before the barrier that ended it
        if (tid == 0 && blk + 1 < nblk) load_kv(blk + 1);
        mbar_wait(bar_kv + (blk & 1), (blk >> 1) & 1);
        const uint32_t ak = smem_u32(smem + FA_K + (blk & 1) * FA_KV_STRIDE), av = smem_u32(smem + FA_V + (blk & 1) * FA_KV_STRIDE);
        const int n_valid = min(FA_BN, key1 - kb), n_lo = max(0, key0 - kb);   // columns [n_lo, n_valid) are keys of the segment
#pragma unroll
        for (int hd = 0; hd < 2; ++hd) {
            if (tid == 0) {
                tc_fence_after();
                const uint64_t da = umma_desc_sw128(aq) + 4 * hd, db = umma_desc_sw128(ak) + 4 * hd;
                tc_mma_f16(tmem_base, da, db, idesc_s, 0);
                tc_mma_f16(tmem_base, da + 2, db + 2, idesc_s, 1);
                tc_commit(bar_s);
            }
            mbar_wait(bar_s, ph_s);
            ph_s ^= 1;
            tc_fence_after();
            // ---- pass 1: block maximum of this row (log2 domain)
            float mx = -INFINITY;
            uint32_t v[32];
#pragma unroll
            for (int c = 0; c < FA_BN / 32; ++c) {
                tc_ld32(t_s + c * 32, v);
#pragma unroll
                for (int j = 0; j < 32; ++j)
                    if (c * 32 + j < n_valid && c * 32 + j >= n_lo) mx = fmaxf(mx, __uint_as_float(v[j]) * scale_log2);
            }
            const float m_new = fmaxf(m_run[hd], mx);   // every block holds at least one key of the segment, so m_new is finite
            const float alpha = fa_ex2(m_run[hd] - m_new);   // exp2(-inf) = 0 on the first block
            // ---- pass 2: probabilities -> fp16 -> shared memory (two [128 x 64] K-major swizzled sub-tiles)
            float sum = 0.f;
#pragma unroll
            for (int c = 0; c < FA_BN / 32; ++c) {
                tc_ld32(t_s + c * 32, v);
                float p[32];
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    p[j] = (c * 32 + j < n_valid && c * 32 + j >= n_lo) ? fa_ex2(fmaf(__uint_as_float(v[j]), scale_log2, -m_new)) : 0.f;
                    sum += p[j];
                }
                uint8_t* pt = smem + FA_P;   // one [128 x 64] K-major swizzled tile
#pragma unroll
                for (int j = 0; j < 4; ++j) *reinterpret_cast<uint4*>(pt + sw128_offset(tid, (c & 1) * 4 + j)) = pack8(p + 8 * j);
            }
            l_run[hd] = l_run[hd] * alpha + sum;
            m_run[hd] = m_new;
            fence_proxy_async_smem();
            tc_fence_before();
            __syncthreads();
            if (tid == 0) {
                tc_fence_after();
#pragma unroll
                for (int k = 0; k < FA_BN / 16; ++k) {
                    const uint64_t da = umma_desc_sw128(ap) + 2 * k;
                    const uint64_t db = umma_desc_sw128(av + hd * 4096) + 2 * k;
                    tc_mma_f16(tmem_base + FA_BN, da, db, idesc_o, k != 0);
                }
                tc_commit(bar_o);
            }
            mbar_wait(bar_o, ph_o);
            ph_o ^= 1;
            tc_fence_after();
            tc_ld32(t_o, v);
#pragma unroll
            for (int i = 0; i < 32; ++i) o[hd][i] = fmaf(o[hd][i], alpha, __uint_as_float(v[i]));
            tc_fence_before();
            __syncthreads();   // S, P and (after the second head) K / V^T may be overwritten
        }
    }
    if (row < Mq) {
#pragma unroll
        for (int hd = 0; hd < 2; ++hd) {
            const float inv = l_run[hd] > 0.f ? 1.0f / l_run[hd] : 0.f;
            float t[32];
#pragma unroll
            for (int i = 0; i < 32; ++i) t[i] = o[hd][i] * inv;
#pragma unroll
            for (int j = 0; j < 4; ++j) reinterpret_cast<uint4*>(orow + hd * 32)[j] = pack8(t + 8 * j);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(128u) : "memory");
    }
}

// out [Mq * n_seg, ldo] fp16 = softmax(Q K^T * scale) V per head of 32 dims.  Q [Mq, D], K [Mk, D] row-major fp16;
// Vt [D, ldv] = V transposed (ldv % 8 == 0, columns >= Mk zero or finite).  seg_ptr (device, n_seg + 1 ints) gives the
// key range of every segment (NULL: one segment [0, Mk)); segment s writes rows s*Mq .. s*Mq + Mq - 1.
int launch_flash(const op_t* Q, int Mq, const op_t* K, int Mk, const op_t* Vt, int ldv, int D, const int* seg_ptr, int n_seg, float scale,
                 op_t* out, int ldo, cudaStream_t st) {
    if (Mq <= 0 || Mk <= 0) return CM_OK;
    CM_REQUIRE(D % 64 == 0 && ldv % 8 == 0 && ldo % 8 == 0, "flash attention: hidden %% 64 == 0 and 16-byte rows required");
    CUtensorMap mq, mk, mv;
    int rc;
    if ((rc = make_map(&mq, Q, (uint64_t)Mq, (uint64_t)D, (uint64_t)D, FA_BM)) != CM_OK) return rc;
    if ((rc = make_map(&mk, K, (uint64_t)Mk, (uint64_t)D, (uint64_t)D, FA_BN)) != CM_OK) return rc;
    if ((rc = make_map(&mv, Vt, (uint64_t)D, (uint64_t)Mk, (uint64_t)ldv, 64)) != CM_OK) return rc;
    static PerDeviceOnce attr;
    if (attr.first()) {
        cudaFuncSetAttribute(flash_attn_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FA_SMEM);
    }
    const int segs = seg_ptr ? n_seg : 1;
    CM_REQUIRE(segs >= 1 && segs <= 65535, "flash attention: 1..65535 key segments");
    dim3 grid(ceil_div(Mq, FA_BM), D / 64, segs);
    {
        CM_TIMED("flash_attention", st);
        flash_attn_kernel<<<grid, FA_THREADS, FA_SMEM, st>>>(mq, mk, mv, Mq, seg_ptr, Mk, scale * 1.4426950408889634f, out, ldo);
    }
    return post_launch("flash_attention");
}

// out [cols, ldo] = in [rows, ld_in]^T (fp16); rows >= *rows_dev (if given) and the padding columns are written as zeros
__global__ void transpose_f16_kernel(const op_t* __restrict__ in, int ld_in, int rows, int cols, const int* rows_dev, op_t* __restrict__ out,
                                     int ldo) {
    __shared__ op_t tile[32][33];
    const int live = rows_dev ? min(*rows_dev, rows) : rows;
    const int r0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        const int r = r0 + i, c = c0 + threadIdx.x;
        tile[i][threadIdx.x] = (r < live && c < cols) ? in[(size_t)r * ld_in + c] : f2op(0.f);
    }
    __syncthreads();
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        const int c = c0 + i, r = r0 + threadIdx.x;
        if (c < cols && r < ldo) out[(size_t)c * ldo + r] = tile[threadIdx.x][i];
    }
}
static int launch_transpose(const op_t* in, int ld_in, int rows, int cols, const int* rows_dev, op_t* out, int ldo, cudaStream_t st) {
    {
        CM_TIMED("transpose_f16", st);
        transpose_f16_kernel<<<dim3(ceil_div(ldo, 32), ceil_div(cols, 32)), dim3(32, 8), 0, st>>>(in, ld_in, rows, cols, rows_dev, out, ldo);
    }
    return post_launch("transpose_f16");
}

// ------------------------------------------------------------------------------------------- fusion around it
bool fusion_flash_supported(const cm_fusion_params* p, int n_kg) {
    return p->fusion_type == 0 && p->hidden % 64 == 0 && p->hidden <= 256 && p->heads > 0 && p->hidden / p->heads == 32 &&
           p->hidden % p->heads == 0 && p->rg_dim % 8 == 0 && p->kg_dim % 8 == 0 && n_kg >= FUSION_FLASH_MIN_KG;
}

struct FlashWs {
    op_t *wr, *wk, *in1, *o1, *in2, *o2, *f1a, *f1b, *f2a, *f2b;
    op_t *rg16, *kg16, *R, *Kp, *Q1, *K1, *V1, *V1t, *O1, *L1, *F1, *Q2, *K2, *V2, *V2t, *O2, *L2, *F2;
    float *R2f, *K2of, *both, *f1, *fused, *hid;
    int kpad, mpad;
    size_t total;
};
static FlashWs carve_flash(void* ws, const cm_fusion_params* p, int G, int M, int K) {
    Carver c(ws);
    FlashWs w;
    const size_t D = p->hidden, m = M, k = K, gk = (size_t)G * K;
    w.kpad = (K + 7) / 8 * 8;
    w.mpad = (M + 7) / 8 * 8;
    w.wr = c.take<op_t>(D * p->rg_dim);
    w.wk = c.take<op_t>(D * p->kg_dim);
    w.in1 = c.take<op_t>(3 * D * D);
    w.o1 = c.take<op_t>(D * D);
    w.in2 = c.take<op_t>(3 * D * D);
    w.o2 = c.take<op_t>(D * D);
    w.f1a = c.take<op_t>(2 * D * D);
    w.f1b = c.take<op_t>(2 * D * D);
    w.f2a = c.take<op_t>(2 * D * D);
    w.f2b = c.take<op_t>(2 * D * D);
    w.rg16 = c.take<op_t>(m * p->rg_dim);
    w.kg16 = c.take<op_t>(k * p->kg_dim);
    w.R = c.take<op_t>(m * D);
    w.Kp = c.take<op_t>(k * D);
    w.Q1 = c.take<op_t>(m * D);
    w.K1 = c.take<op_t>(k * D);
    w.V1 = c.take<op_t>(k * D);
    w.V1t = c.take<op_t>(D * (size_t)w.kpad);
    w.O1 = c.take<op_t>(m * D);
    w.L1 = c.take<op_t>(m * D);
    w.F1 = c.take<op_t>(m * 2 * D);
    w.Q2 = c.take<op_t>(k * D);
    w.K2 = c.take<op_t>(m * D);
    w.V2 = c.take<op_t>(m * D);
    w.V2t = c.take<op_t>(D * (size_t)w.mpad);
    w.O2 = c.take<op_t>(gk * D);
    w.L2 = c.take<op_t>(gk * D);
    w.F2 = c.take<op_t>(gk * 2 * D);
    w.R2f = c.take<float>(m * D);
    w.K2of = c.take<float>(gk * D);
    w.both = c.take<float>((size_t)G * 2 * D);
    w.f1 = c.take<float>((size_t)G * D);
    w.fused = c.take<float>((size_t)G * D);
    w.hid = c.take<float>((size_t)G * 4 * D);
    w.total = align_up(c.off, 256);
    return w;
}
size_t fusion_flash_workspace_bytes(const cm_fusion_params* p, int n_graphs, int node_cap, int n_kg) {
    return carve_flash(nullptr, p, n_graphs, node_cap, n_kg).total;
}

static Dense dense_of(const op_t* A, int lda, const op_t* W, int M, int N, int K, const int* m_dev, const float* bias) {
    Dense d;
    d.A = A; d.lda = lda; d.W = W; d.M = M; d.N = N; d.K = K; d.m_dev = m_dev; d.bias = bias;
    return d;
}

int fusion_forward_flash(const cm_fusion_params* p, const float* rg, const int32_t* node_ptr, int G, int M, const float* kg, int K,
                         void* workspace, FlashTail* tail, cudaStream_t st) {
    FlashWs w = carve_flash(workspace, p, G, M, K);
    tail->both = w.both; tail->f1 = w.f1; tail->fused = w.fused; tail->hid = w.hid;
    float* both_out = w.both;
    const int D = p->hidden, GK = G * K;
    const int* m_dev = node_ptr + G;
    const float scale = 1.0f / sqrtf((float)(D / p->heads));
    int rc;
#define CM_TRY(x) \
    if ((rc = (x)) != CM_OK) return rc
    // fp16 copies of the weights and inputs (1.3 M parameters: a few microseconds next to the attention)
    if (p->rg_proj_w) CM_TRY(launch_cast_f16(p->rg_proj_w, p->rg_dim, D, p->rg_dim, nullptr, w.wr, p->rg_dim, st));
    if (p->kg_proj_w) CM_TRY(launch_cast_f16(p->kg_proj_w, p->kg_dim, D, p->kg_dim, nullptr, w.wk, p->kg_dim, st));
    CM_TRY(launch_cast_f16(p->r2k_in_w, D, 3 * D, D, nullptr, w.in1, D, st));
    CM_TRY(launch_cast_f16(p->r2k_out_w, D, D, D, nullptr, w.o1, D, st));
    CM_TRY(launch_cast_f16(p->k2r_in_w, D, 3 * D, D, nullptr, w.in2, D, st));
    CM_TRY(launch_cast_f16(p->k2r_out_w, D, D, D, nullptr, w.o2, D, st));
    CM_TRY(launch_cast_f16(p->ffn_rg_w1, D, 2 * D, D, nullptr, w.f1a, D, st));
    CM_TRY(launch_cast_f16(p->ffn_rg_w2, 2 * D, D, 2 * D, nullptr, w.f1b, 2 * D, st));
    CM_TRY(launch_cast_f16(p->ffn_kg_w1, D, 2 * D, D, nullptr, w.f2a, D, st));
    CM_TRY(launch_cast_f16(p->ffn_kg_w2, 2 * D, D, 2 * D, nullptr, w.f2b, 2 * D, st));
    CM_TRY(launch_cast_f16(rg, p->rg_dim, M, p->rg_dim, m_dev, w.rg16, p->rg_dim, st));
    CM_TRY(launch_cast_f16(kg, p->kg_dim, K, p->kg_dim, nullptr, w.kg16, p->kg_dim, st));
    Dense d;
    // projections (:108-109)
    const op_t* R = w.rg16;
    if (p->rg_proj_w) {
        d = dense_of(w.rg16, p->rg_dim, w.wr, M, D, p->rg_dim, m_dev, p->rg_proj_b);
        d.out_h = w.R; d.ldo_h = D;
        CM_TRY(launch_dense(d, st));
        R = w.R;
    }
    const op_t* Kp = w.kg16;
    if (p->kg_proj_w) {
        d = dense_of(w.kg16, p->kg_dim, w.wk, K, D, p->kg_dim, nullptr, p->kg_proj_b);
        d.out_h = w.Kp; d.ldo_h = D;
        CM_TRY(launch_dense(d, st));
        Kp = w.Kp;
    }
    // ---- region rows attend to the KG rows (:112-120)
    d = dense_of(R, D, w.in1, M, D, D, m_dev, p->r2k_in_b);
    d.out_h = w.Q1; d.ldo_h = D;
    CM_TRY(launch_dense(d, st));
    d = dense_of(Kp, D, w.in1 + (size_t)D * D, K, D, D, nullptr, p->r2k_in_b + D);
    d.out_h = w.K1; d.ldo_h = D;
    CM_TRY(launch_dense(d, st));
    d = dense_of(Kp, D, w.in1 + (size_t)2 * D * D, K, D, D, nullptr, p->r2k_in_b + 2 * D);
    d.out_h = w.V1; d.ldo_h = D;
    CM_TRY(launch_dense(d, st));
    CM_TRY(launch_transpose(w.V1, D, K, D, nullptr, w.V1t, w.kpad, st));
    CM_TRY(launch_flash(w.Q1, M, w.K1, K, w.V1t, w.kpad, D, nullptr, 1, scale, w.O1, D, st));
    d = dense_of(w.O1, D, w.o1, M, D, D, m_dev, p->r2k_out_b);
    d.residual = R; d.ldr = D; d.ln_gamma = p->ln_rg_w; d.ln_beta = p->ln_rg_b; d.out_h = w.L1; d.ldo_h = D;
    CM_TRY(launch_dense(d, st));
    d = dense_of(w.L1, D, w.f1a, M, 2 * D, D, m_dev, p->ffn_rg_b1);
    d.relu = 1; d.out_h = w.F1; d.ldo_h = 2 * D;
    CM_TRY(launch_dense(d, st));
    d = dense_of(w.F1, 2 * D, w.f1b, M, D, 2 * D, m_dev, p->ffn_rg_b2);
    d.residual = w.L1; d.ldr = D; d.out_f32 = w.R2f; d.ldo_f32 = D;
    CM_TRY(launch_dense(d, st));
    CM_TRY(launch_segment_mean(w.R2f, D, node_ptr, 0, G, both_out, 2 * D, st));   // :134
    // ---- KG rows attend to each image's region rows (:123-131)
    d = dense_of(Kp, D, w.in2, K, D, D, nullptr, p->k2r_in_b);
    d.out_h = w.Q2; d.ldo_h = D;
    CM_TRY(launch_dense(d, st));
    d = dense_of(R, D, w.in2 + (size_t)D * D, M, D, D, m_dev, p->k2r_in_b + D);
    d.out_h = w.K2; d.ldo_h = D;
    CM_TRY(launch_dense(d, st));
    d = dense_of(R, D, w.in2 + (size_t)2 * D * D, M, D, D, m_dev, p->k2r_in_b + 2 * D);
    d.out_h = w.V2; d.ldo_h = D;
    CM_TRY(launch_dense(d, st));
    CM_TRY(launch_transpose(w.V2, D, M, D, m_dev, w.V2t, w.mpad, st));
    CM_TRY(launch_flash(w.Q2, K, w.K2, M, w.V2t, w.mpad, D, node_ptr, G, scale, w.O2, D, st));
    d = dense_of(w.O2, D, w.o2, GK, D, D, nullptr, p->k2r_out_b);
    d.residual = Kp; d.ldr = D; d.res_mod = K; d.ln_gamma = p->ln_kg_w; d.ln_beta = p->ln_kg_b; d.out_h = w.L2; d.ldo_h = D;
    CM_TRY(launch_dense(d, st));
    d = dense_of(w.L2, D, w.f2a, GK, 2 * D, D, nullptr, p->ffn_kg_b1);
    d.relu = 1; d.out_h = w.F2; d.ldo_h = 2 * D;
    CM_TRY(launch_dense(d, st));
    d = dense_of(w.F2, 2 * D, w.f2b, GK, D, 2 * D, nullptr, p->ffn_kg_b2);
    d.residual = w.L2; d.ldr = D; d.out_f32 = w.K2of; d.ldo_f32 = D;
    CM_TRY(launch_dense(d, st));
    CM_TRY(launch_segment_mean(w.K2of, D, nullptr, K, G, both_out + D, 2 * D, st));   // :135
#undef CM_TRY
    return CM_OK;
}

}  // namespace cm

// Exported building block (and the unit-test hook of the flash path): see include/cm_b200.h.
extern "C" int cm_flash_attention_f16(const void* q_f16, int mq, const void* k_f16, int mk, const void* vt_f16, int ldv, int hidden,
                                      const int32_t* seg_ptr, int n_seg, float scale, void* out_f16, int ldo, void* stream) {
    using namespace cm;
    CM_DEVICE_OR_RETURN();
    CM_REQUIRE(q_f16 && k_f16 && vt_f16 && out_f16 && mq > 0 && mk > 0 && hidden > 0, "cm_flash_attention_f16: bad arguments");
    CM_REQUIRE(ldv >= mk, "cm_flash_attention_f16: ldv < mk");
    return launch_flash(reinterpret_cast<const op_t*>(q_f16), mq, reinterpret_cast<const op_t*>(k_f16), mk,
                        reinterpret_cast<const op_t*>(vt_f16), ldv, hidden, seg_ptr, n_seg, scale, reinterpret_cast<op_t*>(out_f16), ldo,
                        as_stream(stream));
}
