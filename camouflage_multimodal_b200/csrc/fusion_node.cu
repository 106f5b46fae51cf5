// Region-node side of the KG-folded cross-attention fusion as ONE kernel (fusion_model.py:108-120 + the K/V side of
// :123-129), one CTA per tile of 128 region nodes, all activations resident in shared memory / TMEM:
//
//   ph0  r   = x Wr^T + br                         x [128,128] by TMA            -> A1 (smem, fp16, SW128 K-major)
//   ph1  S   = r [U1;U2]^T + [c1;0]                softmax per head -> P1        -> A2 (smem);  S2 raw -> global fp32
//   ph2  V2  = r Wv2^T + bv2                                                     -> global fp16
//   ph3  y0  = LN(r + P1 VO^T + bo1)                                             -> A1 (in place)
//   ph4  hh0 = relu(y0 W1[0:256]^T + b1)                                         -> A3 (smem)
//   ph5  Y   = hh0 W2[:,0:256]^T
//   ph6  hh1 = relu(y0 W1[256:512]^T + b1)                                       -> A3
//   ph7  Y  += hh1 W2[:,256:512]^T;  y = Y + b2 + y0                             -> global fp32 (mean over nodes follows)
//
// MODE 2 (the evaluation path when no kg->rg attention map is asked for): the KG side needs the node rows only through sums
// over each image's nodes -- the flash-style partials of the kg->rg attention (per folded key slot: max, sum of exp, sum of
// exp x V2) and the column sums of y -- so the tile reduces them itself and S2 / V2 / y never reach global memory (31 MB per
// batch of 32 written here and read back by kg2rg_partial, which disappears):
//   epi1, upper half  raw kg->rg scores -> column maxima of the tile (warp butterfly + 4-warp combine) -> E = exp(s - max) as
//                     fp16 [128 x 128] in the (dead) x tile, sum of E per slot and image -> part_m / part_l
//   after epi3        V2 (still in T0) -> fp16 [128 x 128] halves in the (dead) P1 tile; thread (slot octet, column) adds
//                     E[n, slot] V2[n, column] up over the rows of each image -> part_acc            (T0 is then released to ph4)
//   epi7              y -> fp32 tile in the (dead) hh / weight-ring space -> column sums per image -> part_y
// Partials are indexed by tile + image (rows are sorted by image, so (tile, image) pairs form a staircase and the sum is a
// unique slot); the KG-side tail merges the slots of an image.
//
// Every GEMM is 128 x 256 x K on tcgen05 (fp16 operands, fp32 accumulators in the two 256-column halves T0 / T1 of
// TMEM), its weight tiles [256 x 64] streamed through a 3-stage TMA ring; the epilogue of phase i (8 warps, one
// TMEM lane = one node row per thread) overlaps the MMAs of the phases that do not depend on it.
//   warp 0: TMA producer   warp 1: MMA issuer   warps 2-9: epilogues
// Dependencies: MMA(ph) waits for the epilogue that produced its A operand and for the epilogue that last read its
// TMEM half; epilogue(ph) waits for acc_ready[ph] only -- tcgen05.commit covers every earlier MMA, which is also
// what makes the in-place / aliased shared-memory buffers safe (see the notes at each phase).
#include <cuda.h>
#include <math.h>

#include "fusion_node.cuh"
#include "tc_ptx.cuh"

namespace cm {

constexpr int FN_THREADS = 320, FN_EPI_THREADS = 256, FN_STAGES = 3, FN_PHASES = 8;
constexpr int FN_A1 = 0, FN_A3 = 65536, FN_X = FN_A3, FN_P1 = FN_A3 + 32768, FN_RING = 131072, FN_STAGE_BYTES = 32768;
constexpr int FN_BAR = FN_RING + FN_STAGES * FN_STAGE_BYTES;   // 229376
constexpr int FN_SCR = FN_BAR + 256;                           // MODE 2: tile maxima [128] f32 + image of every row [128] i32
constexpr int FN_SMEM = FN_BAR + 256 + 1024 + 1024;            // + barriers, + scratch, + 1 KB alignment slack
constexpr int FN_KB_TILE = 128 * 64 * 2;                       // one [128 x 64] fp16 k-block of an activation tile

// GEMM schedule: A operand (smem offset, k-blocks), weight map index + K offset, TMEM column half, accumulate,
// epilogue the MMA warp must wait for before issuing (-1: none; phase 0 waits for the x tile instead).
struct FnPhase {
    int a_off, a_kb, wmap, w_row0, w_k0, tmem_col, accumulate, wait_epi;
};
__constant__ FnPhase fn_phases[3][FN_PHASES] = {
    {
        // MODE 0: region-node chain
        {FN_X, 2, 1, 0, 0, 0, 0, -1},      // r
        {FN_A1, 4, 2, 0, 0, 256, 0, 0},    // scores        (A1 = r ready after epi0)
        {FN_A1, 4, 3, 0, 0, 0, 0, 0},      // V2            (T0 drained by epi0)
        {FN_P1, 2, 4, 0, 0, 256, 0, 1},    // P1 VO^T       (P1 ready, T1 drained after epi1)
        {FN_A1, 4, 5, 0, 0, 0, 0, 3},      // hh0           (A1 = y0 after epi3; T0 drained by epi2 < epi3)
        {FN_A3, 4, 6, 0, 0, 256, 0, 4},    // Y  = hh0 ...  (A3 = hh0 after epi4; T1 drained by epi3)
        {FN_A1, 4, 5, 256, 0, 0, 0, 4},    // hh1           (T0 drained by epi4)
        {FN_A3, 4, 6, 0, 256, 256, 1, 6},  // Y += hh1 ...  (A3 = hh1 after epi6)
    },
    {
        // MODE 1: KG-row chain (fusion_model.py:130-131): x = ctx2 tile [128,256] in A3
        {FN_A3, 4, 1, 0, 0, 256, 0, -1},   // ctx2 Wo2^T    -> LN(kp + . + bo2) = z0 -> A1
        {FN_A1, 4, 2, 0, 0, 0, 0, 0},      // hh0           (A1 = z0 after epi0)
        {FN_A3, 4, 3, 0, 0, 256, 0, 1},    // Y  = hh0 ...  (A3 = hh0 after epi1; T1 drained by epi0)
        {FN_A1, 4, 2, 256, 0, 0, 0, 1},    // hh1           (T0 drained by epi1)
        {FN_A3, 4, 3, 0, 256, 256, 1, 3},  // Y += hh1 ...  (A3 = hh1 after epi3)
        {0, 0, 0, 0, 0, 0, 0, -1}, {0, 0, 0, 0, 0, 0, 0, -1}, {0, 0, 0, 0, 0, 0, 0, -1},
    },
    {
        // MODE 2: the region-node chain again (ph4 additionally waits for V2 to leave T0: epi_done[2])
        {FN_X, 2, 1, 0, 0, 0, 0, -1},
        {FN_A1, 4, 2, 0, 0, 256, 0, 0},
        {FN_A1, 4, 3, 0, 0, 0, 0, 0},
        {FN_P1, 2, 4, 0, 0, 256, 0, 1},
        {FN_A1, 4, 5, 0, 0, 0, 0, 3},
        {FN_A3, 4, 6, 0, 0, 256, 0, 4},
        {FN_A1, 4, 5, 256, 0, 0, 0, 4},
        {FN_A3, 4, 6, 0, 256, 256, 1, 6},
    },
};

__device__ __forceinline__ void st_smem_v4(uint32_t addr, uint4 v) {
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ uint4 ld_smem_v4(uint32_t addr) {
    uint4 v;
    asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr) : "memory");
    return v;
}
// 256-bit global stores (sm_100): a lane writes one full 32-byte sector per instruction.  The row-per-thread epilogues
// store 128 B (fp32) / 64 B (fp16) pieces at a 512 B / 1 KB stride, so with 16-byte stores every sector was written twice.
__device__ __forceinline__ void st_global_f32x8(float* p, const float* x) {
    asm volatile("st.global.v8.f32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(p), "f"(x[0]), "f"(x[1]), "f"(x[2]), "f"(x[3]), "f"(x[4]),
                 "f"(x[5]), "f"(x[6]), "f"(x[7])
                 : "memory");
}
__device__ __forceinline__ void st_global_b32x8(void* p, uint4 a, uint4 b) {
    asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(p), "r"(a.x), "r"(a.y), "r"(a.z), "r"(a.w), "r"(b.x), "r"(b.y),
                 "r"(b.z), "r"(b.w)
                 : "memory");
}
// 32 fp32 values of row `row`, columns c0..c0+31, into a K-major SW128 activation tile set starting at `base`
__device__ __forceinline__ void store_chunk_smem(uint32_t base, int row, int c0, const float* x) {
    const uint32_t tile = base + (uint32_t)(c0 >> 6) * FN_KB_TILE;
    const int ch0 = (c0 & 63) >> 3;
#pragma unroll
    for (int t = 0; t < 4; ++t) st_smem_v4(tile + sw128_offset(row, ch0 + t), pack8(x + 8 * t));
}
__device__ __forceinline__ void add_chunk_smem(uint32_t base, int row, int c0, float* x) {
    const uint32_t tile = base + (uint32_t)(c0 >> 6) * FN_KB_TILE;
    const int ch0 = (c0 & 63) >> 3;
#pragma unroll
    for (int t = 0; t < 4; ++t) {
        const uint4 v = ld_smem_v4(tile + sw128_offset(row, ch0 + t));
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const float2 f = op22f(*reinterpret_cast<const op2_t*>(&w[u]));
            x[8 * t + 2 * u] += f.x;
            x[8 * t + 2 * u + 1] += f.y;
        }
    }
}
// Per-column vectors (biases, LayerNorm gamma / beta) without a memory access on the epilogue's critical path: a thread's
// four 32-column chunks take the same columns in every lane of the warp (one row per lane), so lane l fetches element
// c0 + l of each chunk BEFORE the wait for the accumulator (the loads fly during the MMA) and the chunk code broadcasts
// element j from lane j by shuffle.  (The 10 KB of vectors have no room in shared memory next to the activations and
// the weight ring; with __ldg in the chunk loop every epilogue spent ~1.5 us of its ~2.2 us waiting for L1 / L2.)
struct VecSlice {
    float r[4];
};
__device__ __forceinline__ VecSlice vec_slice(const float* __restrict__ vec, int cbase, int cstep, int lane) {
    VecSlice s;
#pragma unroll
    for (int i = 0; i < 4; ++i) s.r[i] = __ldg(vec + cbase + i * cstep + lane);
    return s;
}
// x[j] = acc[j] + vec[c0 + j], the vector element coming from lane j's slice register
__device__ __forceinline__ void load_acc_bias_s(uint32_t taddr, int c0, float breg, float* x) {
    uint32_t v[32];
    tc_ld32(taddr + c0, v);
#pragma unroll
    for (int j = 0; j < 32; ++j) x[j] = __uint_as_float(v[j]) + __shfl_sync(0xffffffffu, breg, j);
}

// MODE 2 helpers.  Butterfly transpose-reduce over the warp's 32 rows (one row per lane, v = 32 consecutive columns of it):
// lane j ends up with column j's maximum / sum; 31 shuffles, fixed order.
template <bool MAX>
__device__ __forceinline__ float warp_col_reduce(float (&v)[32], int lane) {
    const unsigned full = 0xffffffffu;
#pragma unroll
    for (int w = 16; w >= 1; w >>= 1) {
        const bool up = (lane & w) != 0;
#pragma unroll
        for (int i = 0; i < w; ++i) {
            const float send = up ? v[i] : v[i + w], keep = up ? v[i + w] : v[i], got = __shfl_xor_sync(full, send, w);
            v[i] = MAX ? fmaxf(keep, got) : keep + got;
        }
    }
    return v[0];
}
// [128 rows x 128 columns] fp16 tile, 256 bytes per row, the 16-byte chunk index XORed with row % 16: row-per-lane 16-byte
// stores and column-per-lane 2-byte loads are both free of bank conflicts
__device__ __forceinline__ uint32_t kgt_off(int row, int chunk) { return (uint32_t)(row * 256 + ((chunk ^ (row & 15)) << 4)); }

__device__ __forceinline__ void ldsm_x4_trans(uint32_t addr, uint32_t (&r)[4]) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0, %1, %2, %3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void mma16816_f16(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// Measured and dropped: the per-column vectors as 7.5 KB of kernel parameters (host copies taken at fold time), read from the
// constant bank with a warp-uniform index instead of lane slices + shuffles: LDC.64 with a register offset, no local copy --
// and the same tile time (29.8 against 29.9 us, bench 199 k against 202 k img/s): the broadcast is not what the epilogues wait for.
// Measured and dropped: a node-pooled FFN.  The rg stream is used only through its mean over the image's nodes
// (fusion_model.py:134) and the mean commutes with the FFN's second layer, so ph5 / ph7, the y epilogue and the 1 KB per node
// of y can go: the y0 / hh epilogues add their columns up per image (warp transpose-reduce) and the KG-side tail applies W2 to
// one 512-vector per image.  Built and measured (the parity suite was not run on it): this kernel 41.0 -> 37.6 us (the column sums cost ~0.35 us per 32-column
// chunk, eating half of what the two GEMM phases and the 15.6 MB store saved), kg2rg_partial unchanged (it is latency-bound,
// not bound by the y bytes it no longer read), kg_tail 40.6 -> 45.5 us (the GEMV and the block sums sit on its critical
// path): nothing gained over the step, so the full FFN stayed.
// Measured and dropped: four more warps (448 threads) for the two epilogues that only feed global memory (raw kg->rg scores,
// V2), the softmax split over both halves.  The chain gets its accumulator halves back sooner, but the tile was slower both
// with row-per-thread st.global in the side warps (26.5 -> 30.2 us) and with 16 KB staged tiles handed to TMA stores
// (30.7 us); with the side warps only waiting it was 27.7 us: the softmax / LayerNorm epilogues of the chain warps stretch
// by what the side work would have saved (the register file allows 128 registers per thread at 448 threads).
// Measured and dropped: keeping the next chunk's TMEM load in flight while the current chunk is processed (tc_ld32_issue /
// tc_ld_wait in tc_ptx.cuh) needs 32 more live registers; a 320-thread CTA is capped at 168 (three warps per 16 K-register
// sub-partition), and the spills cost more than the hidden latency (43 -> 58 us).  Tried again for the plain-bias epilogues once
// the kernel was down to ~125 registers (two buffers, no spills): no gain either (epi0 1.03 us as before, the ReLU epilogues
// 1.5 instead of 1.2-1.4 us) -- with two warps per scheduler the epilogues are bound by instruction issue (~2.7 instructions
// per element: bias broadcast, add, pack, store), not by the TMEM load latency.
template <int MODE>
__global__ void __launch_bounds__(FN_THREADS, 1) fusion_node_kernel(const __grid_constant__ FusionNodeArgs a) {
    constexpr int NPH = MODE == 1 ? 5 : 8, XKB = MODE == 1 ? 4 : 2;
    extern __shared__ uint8_t smem_raw[];
    if (a.dbg && threadIdx.x == 0) {   // every CTA: first instruction / last (the buffer holds 40 + 2 x grid stamps)
        long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        a.dbg[40 + 2 * blockIdx.x] = t;
    }
    const int m_live = a.m_dev ? min(*a.m_dev, a.M) : a.M;
    const int m0 = blockIdx.x * 128;
    if (m0 >= m_live) return;
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    const uint32_t sbase = smem_u32(smem);
    uint64_t* w_full = reinterpret_cast<uint64_t*>(smem + FN_BAR);
    uint64_t* w_empty = w_full + FN_STAGES;
    uint64_t* x_full = w_empty + FN_STAGES;
    uint64_t* acc_ready = x_full + 1;
    uint64_t* epi_done = acc_ready + FN_PHASES;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(epi_done + FN_PHASES);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (warp == 1 && lane == 0) {
        for (int s = 0; s < FN_STAGES; ++s) {
            mbar_init(w_full + s, 1);
            mbar_init(w_empty + s, 1);
        }
        mbar_init(x_full, (MODE != 1 && a.x32) ? 8 : 1);   // one TMA transaction, or one arrival per converting warp
        for (int i = 0; i < FN_PHASES; ++i) {
            mbar_init(acc_ready + i, 1);
            mbar_init(epi_done + i, 8);   // one arrival per epilogue warp
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    if (warp >= 2) {
        // the per-column vectors the epilogues add (biases, LayerNorm gamma / beta: 10 KB; shared memory is full) are pulled
        // into L1 now, so the __ldg of every epilogue chunk hits L1 instead of stalling the phase chain on L2
        const int t = threadIdx.x - 64;   // 0..255: one 128-byte line each
        auto pf = [&](const float* v, int n_floats) {
            if (v && t * 32 < n_floats) asm volatile("prefetch.global.L1 [%0];" ::"l"(v + t * 32));
        };
        if (MODE != 1) {
            pf(a.b_rgproj, 256);
            pf(a.ccat, 256);
            pf(a.b_v2, 256);
        }
        pf(a.b_out1, 256);
        pf(a.ln_g, 256);
        pf(a.ln_b, 256);
        pf(a.b_ffn1, 512);
        pf(a.b_ffn2, 256);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const bool stamp = a.dbg != nullptr && blockIdx.x == 0;
    auto now = [] {
        long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        return t;
    };
    if (stamp && threadIdx.x == 0) a.dbg[32] = now();

    if (warp == 0) {
        // ------------------------------------------------------------------------------ TMA producer
        if (lane == 0) {
            if (!(MODE != 1 && a.x32)) {
                mbar_expect_tx(x_full, XKB * FN_KB_TILE);
                for (int kb = 0; kb < XKB; ++kb) tma_load_2d(smem + FN_X + kb * FN_KB_TILE, &a.maps[0], x_full, kb * 64, m0);
            }
            int t = 0;
            for (int ph = 0; ph < NPH; ++ph) {
                const FnPhase p = fn_phases[MODE][ph];
                for (int kb = 0; kb < p.a_kb; ++kb, ++t) {
                    const int s = t % FN_STAGES, n = t / FN_STAGES;
                    if (n > 0) mbar_wait(w_empty + s, (n - 1) & 1);
                    mbar_expect_tx(w_full + s, FN_STAGE_BYTES);
                    tma_load_2d(smem + FN_RING + s * FN_STAGE_BYTES, &a.maps[p.wmap], w_full + s, p.w_k0 + kb * 64, p.w_row0);
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ------------------------------------------------------------------------------ MMA issuer
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_f16(128, 256);
            int t = 0;
            for (int ph = 0; ph < NPH; ++ph) {
                const FnPhase p = fn_phases[MODE][ph];
                if (stamp) a.dbg[ph] = now();
                if (ph == 0) mbar_wait(x_full, 0);
                else if (p.wait_epi >= 0) mbar_wait(epi_done + p.wait_epi, 0);
                if (MODE == 2 && ph == 4) mbar_wait(epi_done + 2, 0);   // V2 has left T0 (it stays there until the kg->rg pass)
                tc_fence_after();
                if (stamp) a.dbg[8 + ph] = now();
                for (int kb = 0; kb < p.a_kb; ++kb, ++t) {
                    const int s = t % FN_STAGES, n = t / FN_STAGES;
                    mbar_wait(w_full + s, n & 1);
                    tc_fence_after();
                    const uint64_t adesc = umma_desc_sw128(sbase + p.a_off + kb * FN_KB_TILE);
                    const uint64_t bdesc = umma_desc_sw128(sbase + FN_RING + s * FN_STAGE_BYTES);
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        tc_mma_f16(tmem_base + p.tmem_col, adesc + 2 * k, bdesc + 2 * k, idesc, (p.accumulate | kb | k) != 0);
                    tc_commit(w_empty + s);
                }
                tc_commit(acc_ready + ph);
            }
        }
        __syncwarp();
    } else {
        // ------------------------------------------------------------------------------ epilogues (warps 2..9)
        const int q = warp & 3, half = (warp - 2) >> 2;
        const int lrow = q * 32 + lane, row = m0 + lrow;
        const bool row_ok = row < m_live;
        const uint32_t t0 = tmem_base + ((uint32_t)(q * 32) << 16), t1 = t0 + 256;
        auto done = [&](int ph) {   // smem writes -> async proxy, TMEM reads retired, then one arrival per warp
            fence_proxy_async_smem();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(epi_done + ph);
            if (stamp && threadIdx.x == 64) a.dbg[24 + ph] = now();
        };
        auto wait_acc = [&](int ph) {
            mbar_wait(acc_ready + ph, 0);
            __syncwarp();
            tc_fence_after();
            if (stamp && threadIdx.x == 64) a.dbg[16 + ph] = now();
        };
        float x[32];

        if (MODE != 1 && a.x32) {
            // fp32 rows -> fp16 A operand of phase 0 (SWIZZLE_128B, two 64-column K blocks): coalesced 16-byte loads,
            // eight in flight per thread, 8-byte shared-memory stores
            const int et = threadIdx.x - 64;
#pragma unroll
            for (int j0 = 0; j0 < 16; j0 += 8) {
                float4 v[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const int i = et + 256 * (j0 + j), r = i >> 5, c4 = i & 31;
                    v[j] = m0 + r < a.M ? *reinterpret_cast<const float4*>(a.x32 + (size_t)(m0 + r) * 128 + c4 * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
                }
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const int i = et + 256 * (j0 + j), r = i >> 5, c4 = i & 31;
                    const op2_t lo = f2op2(v[j].x, v[j].y), hi = f2op2(v[j].z, v[j].w);
                    uint2 pk;
                    pk.x = *reinterpret_cast<const uint32_t*>(&lo);
                    pk.y = *reinterpret_cast<const uint32_t*>(&hi);
                    *reinterpret_cast<uint2*>(smem + FN_X + (c4 >> 4) * FN_KB_TILE + sw128_offset(r, (c4 & 15) >> 1) + (c4 & 1) * 8) = pk;
                }
            }
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) mbar_arrive(x_full);
        }

        // MODE 2: the image of this thread's row (node_ptr[bi] <= row < node_ptr[bi + 1]; -1 past the end) for every thread of the
        // CTA to see.  Images hold about the same number of nodes, so the proportional guess is right or one off: two loads
        // instead of a bisection's dependent chain.
        const int nlive = min(128, m_live - m0);   // live rows of the tile (>= 1)
        float* s_m = reinterpret_cast<float*>(smem + FN_SCR);
        int* s_img = reinterpret_cast<int*>(smem + FN_SCR + 512);
        auto epi_sync = [] { asm volatile("bar.sync 1, %0;" ::"n"(FN_EPI_THREADS) : "memory"); };
        if (MODE == 2) {
            if (half == 0) {
                int bi = -1;
                if (row_ok) {
                    int b = min(__float2int_rz((float)row * (float)a.n_graphs / (float)m_live), a.n_graphs - 1);
                    int lo = a.node_ptr[b], hi = a.node_ptr[b + 1];
                    while (lo > row) {
                        hi = lo;
                        lo = a.node_ptr[--b];
                    }
                    while (hi <= row) {
                        lo = hi;
                        hi = a.node_ptr[++b + 1];
                    }
                    bi = b;
                }
                s_img[lrow] = bi;
            }
            epi_sync();
        }
        // row range [lo, hi) of image b inside this tile; the first three boundaries are fetched here, once (a tile of SLIC-sized
        // images holds two images at most), so the per-image loops below do not wait for global memory
        int npc[3] = {0, 0, 0};
        if (MODE == 2) {
            const int b0 = s_img[0];
#pragma unroll
            for (int k = 0; k < 3; ++k) npc[k] = a.node_ptr[min(b0 + k, a.n_graphs)];
        }
        auto seg_range = [&](int b, int& lo, int& hi) {
            const int k = b - s_img[0];
            const int p0 = k == 0 ? npc[0] : k == 1 ? npc[1] : a.node_ptr[b], p1 = k == 0 ? npc[1] : k == 1 ? npc[2] : a.node_ptr[b + 1];
            lo = max(p0 - m0, 0);
            hi = min(p1 - m0, nlive);
        };

        // y0 = LN(acc + bias + residual) -> A1.  Residual: A1 itself (in place: each thread re-reads and rewrites only
        // its own (row, chunk) cells, and every MMA that read A1 has completed because acc_ready[ph] covers them) or
        // fp32 rows from global memory (row % res_mod).  The statistics scratch is the x tile, dead since phase 0.
        auto epi_ln = [&](int ph, uint32_t taddr, const float* bias, const float* res_rows) {
            const VecSlice bs = vec_slice(bias, half * 32, 64, lane), gs = vec_slice(a.ln_g, half * 32, 64, lane),
                           es = vec_slice(a.ln_b, half * 32, 64, lane);
            wait_acc(ph);
            const float* res = res_rows ? res_rows + (size_t)(row % a.res_mod) * 256 : nullptr;
            auto load = [&](int c0) {
                load_acc_bias_s(taddr, c0, bs.r[c0 >> 6], x);
                if (res) {
#pragma unroll
                    for (int j = 0; j < 32; j += 4) {
                        const float4 r4 = __ldg(reinterpret_cast<const float4*>(res + c0 + j));
                        x[j] += r4.x;
                        x[j + 1] += r4.y;
                        x[j + 2] += r4.z;
                        x[j + 3] += r4.w;
                    }
                } else {
                    add_chunk_smem(sbase + FN_A1, lrow, c0, x);
                }
            };
            float s1 = 0.f, s2 = 0.f;
            for (int c0 = half * 32; c0 < 256; c0 += 64) {
                load(c0);
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    s1 += x[j];
                    s2 = fmaf(x[j], x[j], s2);
                }
            }
            float2* stat = reinterpret_cast<float2*>(smem + (MODE == 2 ? FN_P1 : FN_X));   // MODE 2: the x tile holds E; P1 is dead after acc_ready[ph]
            stat[half * 128 + lrow] = make_float2(s1, s2);
            asm volatile("bar.sync 1, %0;" ::"n"(FN_EPI_THREADS) : "memory");
            const float2 o = stat[(half ^ 1) * 128 + lrow];
            s1 += o.x;
            s2 += o.y;
            const float mean = s1 * (1.0f / 256.f);
            const float rstd = rsqrtf(fmaxf(s2 * (1.0f / 256.f) - mean * mean, 0.f) + 1e-5f);
            for (int c0 = half * 32; c0 < 256; c0 += 64) {
                load(c0);
                const float gr = gs.r[c0 >> 6], er = es.r[c0 >> 6];
#pragma unroll
                for (int j = 0; j < 32; ++j)
                    x[j] = (x[j] - mean) * rstd * __shfl_sync(0xffffffffu, gr, j) + __shfl_sync(0xffffffffu, er, j);
                store_chunk_smem(sbase + FN_A1, lrow, c0, x);
            }
            done(ph);   // the scratch is not rewritten before every warp has arrived here (next A3 writer waits on it)
        };
        // hh half = relu(acc + b1) -> A3
        auto epi_relu = [&](int ph, const float* b1) {
            const VecSlice bs = vec_slice(b1, half * 32, 64, lane);
            wait_acc(ph);
            for (int c0 = half * 32; c0 < 256; c0 += 64) {
                load_acc_bias_s(t0, c0, bs.r[c0 >> 6], x);
#pragma unroll
                for (int j = 0; j < 32; ++j) x[j] = fmaxf(x[j], 0.f);
                store_chunk_smem(sbase + FN_A3, lrow, c0, x);
            }
            done(ph);
        };
        // y = Y + b2 + A1 -> global fp32
        auto epi_final = [&](int ph) {
            const VecSlice bs = vec_slice(a.b_ffn2, half * 32, 64, lane);
            wait_acc(ph);
            for (int c0 = half * 32; c0 < 256; c0 += 64) {
                load_acc_bias_s(t1, c0, bs.r[c0 >> 6], x);
                add_chunk_smem(sbase + FN_A1, lrow, c0, x);
                if (row_ok) {
                    float* o = a.y + (size_t)row * 256 + c0;
#pragma unroll
                    for (int j = 0; j < 32; j += 8) st_global_f32x8(o + j, x + j);
                }
            }
        };
        // MODE 2: y = Y + b2 + A1 -> fp32 tile [128][260] over the dead hh buffer + weight ring -> column sums per image
        auto epi_final_sum = [&](int ph) {
            const VecSlice bs = vec_slice(a.b_ffn2, half * 32, 64, lane);
            wait_acc(ph);
            for (int c0 = half * 32; c0 < 256; c0 += 64) {
                load_acc_bias_s(t1, c0, bs.r[c0 >> 6], x);
                add_chunk_smem(sbase + FN_A1, lrow, c0, x);
#pragma unroll
                for (int j = 0; j < 32; j += 4)
                    st_smem_v4(sbase + FN_A3 + (uint32_t)(lrow * 260 + c0 + j) * 4,
                               make_uint4(__float_as_uint(x[j]), __float_as_uint(x[j + 1]), __float_as_uint(x[j + 2]), __float_as_uint(x[j + 3])));
            }
            if (stamp && threadIdx.x == 64) a.dbg[39] = now();
            epi_sync();
            const float* yt = reinterpret_cast<const float*>(smem + FN_A3);
            const int c = threadIdx.x - 64;
            for (int b = s_img[0]; b <= s_img[nlive - 1]; ++b) {
                int lo, hi;
                seg_range(b, lo, hi);
                if (hi <= lo) continue;
                float s4[4] = {0.f, 0.f, 0.f, 0.f};   // rows n with n % 4 == u, then combined in a fixed order
                int n = lo;
                for (; n + 8 <= hi; n += 8) {
                    float v[8];
#pragma unroll
                    for (int u = 0; u < 8; ++u) v[u] = yt[(n + u) * 260 + c];
#pragma unroll
                    for (int u = 0; u < 8; ++u) s4[u & 3] += v[u];
                }
                for (; n < hi; ++n) s4[(n - lo) & 3] += yt[n * 260 + c];
                a.part_y[(size_t)(blockIdx.x + b) * 256 + c] = (s4[0] + s4[1]) + (s4[2] + s4[3]);
            }
        };

        if (MODE == 1) {
            epi_ln(0, t1, a.b_out1, a.res_rows);
            epi_relu(1, a.b_ffn1);
            epi_relu(3, a.b_ffn1 + 256);
            epi_final(4);
        } else {
        // ---- epi0: r -> A1
        const VecSlice b0s = vec_slice(a.b_rgproj, half * 32, 64, lane);
        wait_acc(0);
        for (int c0 = half * 32; c0 < 256; c0 += 64) {
            load_acc_bias_s(t0, c0, b0s.r[c0 >> 6], x);
            store_chunk_smem(sbase + FN_A1, lrow, c0, x);
        }
        done(0);

        // ---- epi1: half 0 -> P1 = per-head softmax of the rg->kg scores (A2); half 1 -> raw kg->rg scores (global)
        if (MODE == 2) {
            // Each half takes two of the four 32-column chunks of BOTH score blocks: the softmax of the rg->kg scores -> P1, and
            // the raw kg->rg scores -> flash-style partial statistics (nothing of them leaves the SM but 1 KB per image).
            // (the kg->rg block has no additive term: a per-slot constant cancels in the softmax over the nodes, ccat[128..] = 0)
            float cs[2];
#pragma unroll
            for (int u = 0; u < 2; ++u) cs[u] = __ldg(a.ccat + 64 * half + 32 * u + lane);
            auto load_raw = [&](int c0) {
                uint32_t v[32];
                tc_ld32(t1 + c0, v);
#pragma unroll
                for (int j = 0; j < 32; ++j) x[j] = __uint_as_float(v[j]);
            };
            wait_acc(1);
            float am[32];   // unused head means (the rg->kg attention map is not produced in this mode)
#pragma unroll
            for (int u = 0; u < 2; ++u) {
                const int c0 = 64 * half + 32 * u;
                load_acc_bias_s(t1, c0, cs[u], x);
                if (a.sm_gs == 16) softmax_groups<16, false>(x, a.sm_valid, am); else softmax_groups<32, false>(x, a.sm_valid, am);
                store_chunk_smem(sbase + FN_P1, lrow, c0, x);
            }
            if (stamp && threadIdx.x == 64) a.dbg[33] = now();
            auto kg_sync = [&] {   // the four warps of this half
                if (half == 0) asm volatile("bar.sync 2, 128;" ::: "memory"); else asm volatile("bar.sync 3, 128;" ::: "memory");
            };
            float* cm = reinterpret_cast<float*>(smem + FN_X);     // [4 warps][128] column maxima (E goes over them below)
#pragma unroll 1
            for (int u = 0; u < 2; ++u) {
                load_raw(128 + 64 * half + 32 * u);
                if (!row_ok) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) x[j] = -INFINITY;
                }
                cm[q * 128 + 64 * half + 32 * u + lane] = warp_col_reduce<true>(x, lane);
            }
            kg_sync();
            const int sj = 64 * half + (lrow & 63);   // threads lrow < 64 of the half: one slot each
            float mj = 0.f;
            if (lrow < 64) mj = fmaxf(fmaxf(cm[sj], cm[128 + sj]), fmaxf(cm[256 + sj], cm[384 + sj]));   // finite: row 0 of a tile is live
            epi_sync();                                  // both halves are done with cm: E goes over it
            if (lrow < 64) s_m[sj] = mj;
            kg_sync();
#pragma unroll 1
            for (int u = 0; u < 2; ++u) {
                const int i = 2 * half + u;
                load_raw(128 + 32 * i);
#pragma unroll
                for (int j = 0; j < 32; ++j) x[j] = row_ok ? __expf(x[j] - s_m[32 * i + j]) : 0.f;
#pragma unroll
                for (int v = 0; v < 4; ++v) st_smem_v4(sbase + FN_X + kgt_off(lrow, 4 * i + v), pack8(x + 8 * v));
            }
            // the tile's maxima for every image that has rows in the tile (the sums of E come out of the product pass below)
            if (lrow < 64)
                for (int b = s_img[0]; b <= s_img[nlive - 1]; ++b) a.part_m[(size_t)(blockIdx.x + b) * 128 + sj] = mj;
            if (stamp && threadIdx.x == 192) a.dbg[34] = now();
        } else {
            const VecSlice c1s = vec_slice(a.ccat, half * 128, 32, lane);
            wait_acc(1);
            float am[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) am[j] = 0.f;
            for (int c0 = half * 128; c0 < half * 128 + 128; c0 += 32) {
                load_acc_bias_s(t1, c0, c1s.r[(c0 >> 5) & 3], x);
                if (stamp && threadIdx.x == 64 && c0 == 0) a.dbg[35] = now();
                if (half == 0) {
                    if (a.attn_mean) {
                        if (a.sm_gs == 16) softmax_groups<16, true>(x, a.sm_valid, am); else softmax_groups<32, true>(x, a.sm_valid, am);
                    } else {
                        if (a.sm_gs == 16) softmax_groups<16, false>(x, a.sm_valid, am); else softmax_groups<32, false>(x, a.sm_valid, am);
                    }
                    if (stamp && threadIdx.x == 64 && c0 == 0) a.dbg[36] = now();
                    store_chunk_smem(sbase + FN_P1, lrow, c0, x);
                    if (stamp && threadIdx.x == 64 && c0 == 0) a.dbg[37] = now();
                } else if (row_ok) {
                    float* o = a.s2 + (size_t)row * 128 + (c0 - 128);
#pragma unroll
                    for (int j = 0; j < 32; j += 8) st_global_f32x8(o + j, x + j);
                }
            }
            if (stamp && threadIdx.x == 64) a.dbg[33] = now();
            if (stamp && threadIdx.x == 192) a.dbg[34] = now();
            if (half == 0 && a.attn_mean && row_ok) {
                const float inv = (float)a.sm_gs / 128.f;   // 1 / heads
#pragma unroll
                for (int j = 0; j < 32; ++j)
                    if (j < a.sm_valid) a.attn_mean[(size_t)row * a.sm_valid + j] = am[j] * inv;
            }
        }
        done(1);

        if (MODE == 0) {
            // ---- epi2: V2 -> global fp16
            const VecSlice b2s = vec_slice(a.b_v2, half * 32, 64, lane);
            wait_acc(2);
            for (int c0 = half * 32; c0 < 256; c0 += 64) {
                load_acc_bias_s(t0, c0, b2s.r[c0 >> 6], x);
                if (row_ok) {
                    op_t* o = a.v2 + (size_t)row * 256 + c0;
#pragma unroll
                    for (int j = 0; j < 32; j += 16) st_global_b32x8(o + j, pack8(x + j), pack8(x + j + 8));
                }
            }
            done(2);
        }

        epi_ln(3, t1, a.b_out1, nullptr);     // y0 = LN(r + P1 VO^T + bo1)

        if (MODE == 2) {
            // ---- kg->rg partial context: V2 (in T0 since ph2) x E, 128 columns at a time through the dead P1 tile.  Thread
            // (slot octet kh, column cc) owns 8 slots of the column's head; rows are walked once per image of the tile.
            // (V2's bias is not added here: the attention weights of a slot sum to one, so the KG-side tail adds it to the
            // normalised context instead)
            wait_acc(2);
            epi_sync();                      // every warp is out of the LayerNorm scratch (P1 space); E is complete
            // Per pass (128 columns = 4 heads): warp (hw = head of the pass, n0 = column half of the head) computes the 16 slots x
            // 16 columns block acc = E_h^T V2 as mma.sync m16n8k16 tiles, both operands by ldmatrix.trans from the row-major
            // tiles; rows outside the image's range are masked out of the A fragment.  A third accumulator against a
            // constant-one B fragment gives the sums of E (warps with n0 == 0).
            const int ew = warp - 2, hw = ew >> 1, n0 = ew & 1, g = lane >> 2, t = lane & 3, lg = lane >> 3, lr = lane & 7;
            const int b_first = s_img[0], b_last = s_img[nlive - 1];
            // V2 leaves T0 at once -- the first 128 columns into the tile, the other 128 packed into registers until the first
            // pass is through -- so that ph4 can start while the products are formed
            uint4 keep[2][4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int p = i >> 1, u = i & 1, c0 = 128 * p + 64 * half + 32 * u;
                {
                    uint32_t v[32];
                    tc_ld32(t0 + c0, v);
#pragma unroll
                    for (int j = 0; j < 32; ++j) x[j] = row_ok ? __uint_as_float(v[j]) : 0.f;   // rows past the end: keep NaN out of the sums
                }
#pragma unroll
                for (int v = 0; v < 4; ++v) {
                    if (p == 0) st_smem_v4(sbase + FN_P1 + kgt_off(lrow, ((c0 & 127) >> 3) + v), pack8(x + 8 * v));
                    else keep[u][v] = pack8(x + 8 * v);
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(epi_done + 2);   // T0 is free for ph4
            if (stamp && threadIdx.x == 64) a.dbg[35] = now();
#pragma unroll 1
            for (int p = 0; p < 2; ++p) {
                if (p == 1) {
#pragma unroll
                    for (int u = 0; u < 2; ++u)
#pragma unroll
                        for (int v = 0; v < 4; ++v) st_smem_v4(sbase + FN_P1 + kgt_off(lrow, ((64 * half + 32 * u) >> 3) + v), keep[u][v]);
                    if (stamp && threadIdx.x == 64) a.dbg[37] = now();
                }
                epi_sync();
                const int h = 4 * p + hw;                       // head of this warp's block
                const int ech = 2 * h, vch = 4 * hw + 2 * n0;   // 16-byte chunk of E (slots 16 h ..) / of the V2 tile (columns of the pass)
                for (int b = b_first; b <= b_last; ++b) {
                    int lo, hi;
                    seg_range(b, lo, hi);
                    if (hi <= lo) continue;
                    float c0[4] = {0.f, 0.f, 0.f, 0.f}, c1[4] = {0.f, 0.f, 0.f, 0.f}, cl[4] = {0.f, 0.f, 0.f, 0.f};
                    for (int ks = lo >> 4; ks <= (hi - 1) >> 4; ++ks) {
                        uint32_t af[4], bf[4];
                        ldsm_x4_trans(sbase + FN_X + kgt_off(16 * ks + (lg >> 1) * 8 + lr, ech + (lg & 1)), af);
                        ldsm_x4_trans(sbase + FN_P1 + kgt_off(16 * ks + (lg & 1) * 8 + lr, vch + (lg >> 1)), bf);
                        // af[0], af[1]: rows 16 ks + 2 t (+1), slots g / g + 8; af[2], af[3]: rows + 8
                        const int r0 = 16 * ks + 2 * t;
                        const uint32_t m_lo = (r0 >= lo && r0 < hi ? 0x0000ffffu : 0u) | (r0 + 1 >= lo && r0 + 1 < hi ? 0xffff0000u : 0u);
                        const uint32_t m_hi = (r0 + 8 >= lo && r0 + 8 < hi ? 0x0000ffffu : 0u) | (r0 + 9 >= lo && r0 + 9 < hi ? 0xffff0000u : 0u);
                        // mma A fragment order: (m = g, k lo), (m = g + 8, k lo), (m = g, k hi), (m = g + 8, k hi)
                        const uint32_t am[4] = {af[0] & m_lo, af[1] & m_lo, af[2] & m_hi, af[3] & m_hi};
                        mma16816_f16(c0, am, bf[0], bf[1]);
                        mma16816_f16(c1, am, bf[2], bf[3]);
                        if (n0 == 0) mma16816_f16(cl, am, 0x3c003c00u, 0x3c003c00u);
                    }
                    const size_t slot = (size_t)blockIdx.x + b;
                    float* o = a.part_acc + (slot * 16) * 256 + 128 * p + 32 * hw + 16 * n0 + 2 * t;
                    *reinterpret_cast<float2*>(o + (size_t)g * 256) = make_float2(c0[0], c0[1]);
                    *reinterpret_cast<float2*>(o + (size_t)(g + 8) * 256) = make_float2(c0[2], c0[3]);
                    *reinterpret_cast<float2*>(o + (size_t)g * 256 + 8) = make_float2(c1[0], c1[1]);
                    *reinterpret_cast<float2*>(o + (size_t)(g + 8) * 256 + 8) = make_float2(c1[2], c1[3]);
                    if (n0 == 0 && t == 0) {
                        a.part_l[slot * 128 + 16 * h + g] = cl[0];
                        a.part_l[slot * 128 + 16 * h + g + 8] = cl[2];
                    }
                }
                if (stamp && threadIdx.x == 64 && p == 0) a.dbg[36] = now();
                epi_sync();                  // the V2 tile (and after the second pass E) may be overwritten
            }
            if (stamp && threadIdx.x == 64) a.dbg[24 + 2] = now();
        }

        epi_relu(4, a.b_ffn1);                // phase 5 has no epilogue: its accumulator stays in T1
        epi_relu(6, a.b_ffn1 + 256);
        if (MODE == 0) epi_final(7);          // y = Y + b2 + y0
        else epi_final_sum(7);
        }
        tc_fence_before();
    }
    __syncthreads();
    if (warp == 0) {
        if (stamp && lane == 0) a.dbg[38] = now();
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
        if (a.dbg && lane == 0) a.dbg[41 + 2 * blockIdx.x] = now();
    }
}

static void fn_set_attr() {
    static PerDeviceOnce attr;
    if (attr.first()) {
        cudaFuncSetAttribute(fusion_node_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, FN_SMEM);
        cudaFuncSetAttribute(fusion_node_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, FN_SMEM);
        cudaFuncSetAttribute(fusion_node_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, FN_SMEM);
    }
}

int launch_fusion_node(FusionNodeArgs& a, const void* x, const void* w_rgproj, const void* wcat, const void* wv2, const void* vot,
                       const void* ffn1, const void* ffn2, cudaStream_t st) {
    int rc;
    if ((rc = make_map(&a.maps[0], x, (uint64_t)a.M, 128, 128, 128)) != CM_OK) return rc;
    if ((rc = make_map(&a.maps[1], w_rgproj, 256, 128, 128, 256)) != CM_OK) return rc;
    if ((rc = make_map(&a.maps[2], wcat, 256, 256, 256, 256)) != CM_OK) return rc;
    if ((rc = make_map(&a.maps[3], wv2, 256, 256, 256, 256)) != CM_OK) return rc;
    if ((rc = make_map(&a.maps[4], vot, 256, 128, 128, 256)) != CM_OK) return rc;
    if ((rc = make_map(&a.maps[5], ffn1, 512, 256, 256, 256)) != CM_OK) return rc;
    if ((rc = make_map(&a.maps[6], ffn2, 256, 512, 512, 256)) != CM_OK) return rc;
    fn_set_attr();
    {
        CM_TIMED("fusion_node", st);
        if (a.part_acc) fusion_node_kernel<2><<<ceil_div(a.M, 128), FN_THREADS, FN_SMEM, st>>>(a);
        else fusion_node_kernel<0><<<ceil_div(a.M, 128), FN_THREADS, FN_SMEM, st>>>(a);
    }
    return post_launch("fusion_node");
}

// KG-row chain: z = z0 + FFN(z0), z0 = LN(res_rows[row % res_mod] + x Wo^T + b_out1); x [M,256] fp16, z -> a.y
int launch_fusion_kgrows(FusionNodeArgs& a, const void* x, const void* wo, const void* ffn1, const void* ffn2, cudaStream_t st) {
    int rc;
    if ((rc = make_map(&a.maps[0], x, (uint64_t)a.M, 256, 256, 128)) != CM_OK) return rc;
    if ((rc = make_map(&a.maps[1], wo, 256, 256, 256, 256)) != CM_OK) return rc;
    if ((rc = make_map(&a.maps[2], ffn1, 512, 256, 256, 256)) != CM_OK) return rc;
    if ((rc = make_map(&a.maps[3], ffn2, 256, 512, 512, 256)) != CM_OK) return rc;
    fn_set_attr();
    {
        CM_TIMED("fusion_kgrows", st);
        fusion_node_kernel<1><<<ceil_div(a.M, 128), FN_THREADS, FN_SMEM, st>>>(a);
    }
    return post_launch("fusion_kgrows");
}

}  // namespace cm
