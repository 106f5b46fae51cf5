// Fused GCN layer on tcgen05 for hidden == 128 (see gcn_tc.cuh): neighbour gather -> split-fp16 A tile in shared
// memory -> tcgen05.mma against the BatchNorm-folded weight -> bias + ReLU.  The last layer chains the row-local
// tail (fc_shared -> three head hidden layers -> head outputs) in the same CTA, so the activations of a 128-node
// tile never leave the SM between the third GCNConv and the logits.
//
// One CTA = 128 nodes (CSR rows), 512 threads:
//   phase 0  thread 0 starts the TMA load of the layer weight ([W_hi | W_lo], 64 KB)
//   phase 1  eight lanes per row, four rows per warp at a time: sum_j s_ij x_j over fp32 rows [M,128] (L2 resident),
//            col / w / dis[j] read straight from the CSR arrays; the sum is split into fp16 hi + lo and both halves
//            are stored into the K-major SWIZZLE_128B A tile
//   phase 2  thread 0: 24 x tcgen05.mma (A_hi W_hi + A_lo W_hi + A_hi W_lo), fp32 accumulators in TMEM
//   phase 3  16 warps read TMEM (lane quarter = warp % 4, column chunk = warp / 4): bias + ReLU, then either the
//            fp32 row store (layers 1, 2) or hi/lo back into the A tile for the next chained product
#include <cuda.h>

#include "gcn_tc.cuh"
#include "tc_ptx.cuh"

namespace cm {

constexpr int GC_ROWS = 128, GC_H = 128, GC_THREADS = 512, GC_EDGE_CAP = 2048;
// Staging the tile's (j, dis_j * w_e) list in shared memory before the gather measured no faster than reading
// col / w / dis[j] in the gather itself (one dependent round trip fewer), so the gather reads them directly.
constexpr bool GC_STAGE_EDGES = false;
constexpr int GC_A_BLOCK = GC_ROWS * 128;                 // one K block (64 fp16) of the A tile: 16 KB
constexpr int GC_OFF_A = 0;                               // A_hi kb0, A_hi kb1, A_lo kb0, A_lo kb1
constexpr int GC_OFF_W = 4 * GC_A_BLOCK;                  // W_hi kb0, kb1, W_lo kb0, kb1 (N x 128 B each, N <= 192)
constexpr int GC_OFF_EDGE = GC_OFF_W + 4 * 192 * 128;     // int2 {j, float bits of dis[j] * w}
constexpr int GC_OFF_VEC = GC_OFF_EDGE + GC_EDGE_CAP * 8; // bias[128] fc_bias[128] h1_bias[192] w2[5*64] b2[8]
constexpr int GC_VEC_FLOATS = 128 + 128 + 192 + 320 + 8;
constexpr int GC_OFF_PART = GC_OFF_VEC + GC_VEC_FLOATS * 4;   // head partial sums [4 groups][128 rows][5]
constexpr int GC_OFF_BAR = GC_OFF_PART + 4 * 128 * 5 * 4;
constexpr int GC_SMEM = GC_OFF_BAR + 128 + 1024;          // + slack for 1 KB alignment
constexpr int GC_SMEM_GAT = 4 * GC_A_BLOCK + GC_VEC_FLOATS * 4 + 128 + 1024;

__device__ __forceinline__ void gc_store_split(uint8_t* a_tile, int row, int chunk_base, int kb, const float* x, int n8) {
    // x: n8 groups of 8 consecutive columns starting at 16-byte chunk `chunk_base` of K block kb
    for (int j = 0; j < n8; ++j) {
        float lo[8];
#pragma unroll
        for (int t = 0; t < 8; ++t) lo[t] = x[8 * j + t] - op2f(f2op(x[8 * j + t]));
        const uint32_t off = sw128_offset(row, chunk_base + j);
        *reinterpret_cast<uint4*>(a_tile + kb * GC_A_BLOCK + off) = pack8(x + 8 * j);
        *reinterpret_cast<uint4*>(a_tile + (2 + kb) * GC_A_BLOCK + off) = pack8(lo);
    }
}

// A_hi W_hi + A_lo W_hi + A_hi W_lo over K = 128 (two 64-wide K blocks, four UMMA_K = 16 steps each)
__device__ __forceinline__ void gc_issue(uint32_t tmem, uint32_t a_addr, uint32_t w_addr, int n, uint32_t idesc, int nkb = 2) {
    const uint32_t wb = (uint32_t)n * 128;
#pragma unroll
    for (int kb = 0; kb < 2; ++kb) {
        if (kb >= nkb) break;
        const uint64_t ahi = umma_desc_sw128(a_addr + kb * GC_A_BLOCK), alo = umma_desc_sw128(a_addr + (2 + kb) * GC_A_BLOCK);
        const uint64_t whi = umma_desc_sw128(w_addr + kb * wb), wlo = umma_desc_sw128(w_addr + (2 + kb) * wb);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            tc_mma_f16(tmem, ahi + 2 * k, whi + 2 * k, idesc, (kb | k) != 0);
            tc_mma_f16(tmem, alo + 2 * k, whi + 2 * k, idesc, 1);
            tc_mma_f16(tmem, ahi + 2 * k, wlo + 2 * k, idesc, 1);
        }
    }
}

// GAT layer weight [128, 3 x 64]: one K block of W_hi (columns 0..63) and one of W_lo (columns 128..191)
__device__ __forceinline__ void gc_load_weight_k64(uint8_t* w_dst, const CUtensorMap* map, uint64_t* bar, int n) {
    mbar_expect_tx(bar, (uint32_t)(2 * n * 128));
    tma_load_2d(w_dst, map, bar, 0, 0);
    tma_load_2d(w_dst + 2 * n * 128, map, bar, 128, 0);
}
__device__ __forceinline__ void gc_load_weight(uint8_t* w_dst, const CUtensorMap* map, uint64_t* bar, int n) {
    mbar_expect_tx(bar, (uint32_t)(4 * n * 128));
    tma_load_2d(w_dst, map, bar, 0, 0);                          // W_hi, K columns [0, 64)
    tma_load_2d(w_dst + n * 128, map, bar, 64, 0);               //        [64, 128)
    tma_load_2d(w_dst + 2 * n * 128, map, bar, 2 * GC_H, 0);     // W_lo (third K third of [W_hi | W_hi | W_lo])
    tma_load_2d(w_dst + 3 * n * 128, map, bar, 2 * GC_H + 64, 0);
}

#ifdef CM_GC_PROFILE
__device__ unsigned long long gc_prof[16];
#define GC_STAMP(i)                                                                   \
    do {                                                                              \
        if (blockIdx.x == 3 && threadIdx.x == 0) {                                    \
            unsigned long long t_;                                                    \
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));                    \
            gc_prof[i] = t_;                                                          \
        }                                                                             \
    } while (0)
extern "C" int cm_debug_gcn_profile(unsigned long long* out) {
    return cudaMemcpyFromSymbol(out, gc_prof, sizeof(gc_prof)) == cudaSuccess ? 0 : 1;
}
#else
#define GC_STAMP(i)
#endif

// MODE 0: GCN layer -> fp32 rows;  1: GCN layer + chained fc_shared / heads;  2: GAT layer (softmax-weighted 15-d
// neighbour sums per head as the A operand, K = 64) -> fp32 rows, also writes the GCN normalisation deg^-1/2
template <int MODE>
__global__ void __launch_bounds__(GC_THREADS, 1)
gcn_layer_tc_kernel(const __grid_constant__ CUtensorMap map_w, const __grid_constant__ CUtensorMap map_fc,
                    const __grid_constant__ CUtensorMap map_h1, GcnLayer p) {
    constexpr bool CHAIN = MODE == 1, GAT = MODE == 2;
    extern __shared__ uint8_t smem_raw[];
    const int n_live = p.n_dev ? min(*p.n_dev, p.M) : p.M;
    const int m0 = blockIdx.x * GC_ROWS;
    if (m0 >= n_live) return;
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    // GAT mode needs one K block per operand half: A_hi | W_hi | A_lo | W_lo interleaved in the first 64 KB, no edge list
    constexpr int OFF_W = GAT ? GC_A_BLOCK : GC_OFF_W, OFF_VEC = GAT ? 4 * GC_A_BLOCK : GC_OFF_VEC;
    constexpr int OFF_BAR = GAT ? OFF_VEC + GC_VEC_FLOATS * 4 : GC_OFF_BAR;
    constexpr uint32_t TMEM_COLS = CHAIN ? 256u : 128u;
    uint8_t* a_tile = smem + GC_OFF_A;
    uint8_t* w_tile = smem + OFF_W;
    int2* s_edge = reinterpret_cast<int2*>(smem + GC_OFF_EDGE);
    float* s_bias = reinterpret_cast<float*>(smem + OFF_VEC);
    float* s_fcb = s_bias + 128;
    float* s_h1b = s_fcb + 128;
    float* s_w2 = s_h1b + 192;
    float* s_b2 = s_w2 + 320;
    float* s_part = reinterpret_cast<float*>(smem + GC_OFF_PART);
    uint64_t* w_full = reinterpret_cast<uint64_t*>(smem + OFF_BAR);   // [3]
    uint64_t* acc = w_full + 3;                                          // [3]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc + 3);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    GC_STAMP(0);

    if (tid == 0) {
        for (int i = 0; i < 3; ++i) {
            mbar_init(w_full + i, 1);
            mbar_init(acc + i, 1);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        if (GAT) gc_load_weight_k64(w_tile, &map_w, w_full, GC_H); else gc_load_weight(w_tile, &map_w, w_full, GC_H);
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    // ---------------------------------------------------------------- phase 0: vectors + the tile's edge list
    const int rows_here = min(GC_ROWS, n_live - m0);
    const int e_lo = p.row_ptr[m0], e_hi = p.row_ptr[m0 + rows_here];
    if (!GAT && GC_STAGE_EDGES) {
        for (int e = e_lo + tid; e < min(e_hi, e_lo + GC_EDGE_CAP); e += GC_THREADS) {
            const int j = p.col[e];
            s_edge[e - e_lo] = make_int2(j, __float_as_int(p.dis[j] * (p.w ? p.w[e] : 1.0f)));
        }
    }
    for (int c = tid; c < 128; c += GC_THREADS) s_bias[c] = p.bias[c];
    if (CHAIN) {
        for (int c = tid; c < 128; c += GC_THREADS) s_fcb[c] = p.fc_bias ? p.fc_bias[c] : 0.f;
        if (p.do_heads) {
            for (int c = tid; c < 192; c += GC_THREADS) s_h1b[c] = p.h1_bias[c];
            for (int c = tid; c < 320; c += GC_THREADS) {
                const int o = c >> 6, k = c & 63;   // outputs 0,1: mask; 2,3: instance; 4: edge
                s_w2[c] = o < 2 ? p.w2[0][o * 64 + k] : (o < 4 ? p.w2[1][(o - 2) * 64 + k] : p.w2[2][k]);
            }
            if (tid < 5) s_b2[tid] = tid < 2 ? p.b2[0][tid] : (tid < 4 ? p.b2[1][tid - 2] : p.b2[2][0]);
        }
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    GC_STAMP(1);

    // ---------------------------------------------------------------- phase 1: gather, four rows per warp at a time
    // Eight lanes own a row: load t of a lane covers columns [32 t + 4 l8, +4), so every load instruction reads one
    // whole 128-byte line per row and a lane keeps 2 edges x 4 loads in flight.
    if (!GAT) {
        const int slot = lane >> 3, l8 = lane & 7;
        const float* xc = p.x + l8 * 4;
        for (int i = 0; i < 2; ++i) {
            const int r = warp * 8 + i * 4 + slot, g = m0 + r;
            const bool live = r < rows_here;
            float4 a[4];
#pragma unroll
            for (int t = 0; t < 4; ++t) a[t] = make_float4(0.f, 0.f, 0.f, 0.f);
            const int e0 = live ? p.row_ptr[g] : 0, e1 = live ? p.row_ptr[g + 1] : 0;
            const float di = live ? p.dis[g] : 0.f;
            int deg = e1 - e0;
            deg = max(deg, __shfl_xor_sync(0xffffffffu, deg, 8));
            deg = max(deg, __shfl_xor_sync(0xffffffffu, deg, 16));
            for (int k = 0; k < deg; k += 2) {
                float4 v[2][4];
                float s[2];
#pragma unroll
                for (int u = 0; u < 2; ++u) {
                    const int e = e0 + k + u;
                    int j = live ? g : 0;
                    float sc = 0.f;
                    if (e < e1) {
                        if (GC_STAGE_EDGES && e - e_lo < GC_EDGE_CAP) {
                            const int2 ed = s_edge[e - e_lo];
                            j = ed.x;
                            sc = __int_as_float(ed.y);
                        } else {
                            j = p.col[e];
                            sc = p.dis[j] * (p.w ? p.w[e] : 1.0f);
                        }
                    }
                    s[u] = j != g ? sc * di : 0.f;
                    const float* xr = xc + (size_t)j * GC_H;
#pragma unroll
                    for (int t = 0; t < 4; ++t) v[u][t] = *reinterpret_cast<const float4*>(xr + 32 * t);
                }
#pragma unroll
                for (int u = 0; u < 2; ++u)
#pragma unroll
                    for (int t = 0; t < 4; ++t) {
                        a[t].x = fmaf(s[u], v[u][t].x, a[t].x);
                        a[t].y = fmaf(s[u], v[u][t].y, a[t].y);
                        a[t].z = fmaf(s[u], v[u][t].z, a[t].z);
                        a[t].w = fmaf(s[u], v[u][t].w, a[t].w);
                    }
            }
            if (live) {
                const float ss = di * (p.self_w ? p.self_w[g] : 1.0f) * di;
                const float* xr = xc + (size_t)g * GC_H;
#pragma unroll
                for (int t = 0; t < 4; ++t) {
                    const float4 v = *reinterpret_cast<const float4*>(xr + 32 * t);
                    a[t].x = fmaf(ss, v.x, a[t].x);
                    a[t].y = fmaf(ss, v.y, a[t].y);
                    a[t].z = fmaf(ss, v.z, a[t].z);
                    a[t].w = fmaf(ss, v.w, a[t].w);
                }
            }
#pragma unroll
            for (int t = 0; t < 4; ++t) {
                const float av[4] = {a[t].x, a[t].y, a[t].z, a[t].w};
                __align__(8) op_t hi[4], lo[4];
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    hi[c] = f2op(av[c]);
                    lo[c] = f2op(av[c] - op2f(hi[c]));
                }
                const uint32_t off = sw128_offset(r, (t & 1) * 4 + (l8 >> 1)) + (l8 & 1) * 8;
                *reinterpret_cast<uint2*>(a_tile + (t >> 1) * GC_A_BLOCK + off) = *reinterpret_cast<const uint2*>(hi);
                *reinterpret_cast<uint2*>(a_tile + (2 + (t >> 1)) * GC_A_BLOCK + off) = *reinterpret_cast<const uint2*>(lo);
            }
        }
    }
    // ---------------------------------------------------------------- phase 1 (GAT): one row at a time per warp
    // Softmax weights with one lane per incoming edge (the added self loop is edge number deg), then lane c owns input
    // feature c and pulls (j, w_h) by shuffle -- gat_aggregate_kernel's scheme, the row written as fp16 hi / lo into
    // K block 0 of the A operand (columns h * C + c, zero padded to 64) instead of to HBM.
    if (GAT) {
        // sixteen lanes per row, two rows per warp at a time (C <= 16 features; edges beyond 16 take another trip);
        // loop bounds are the maximum over the two halves so every shuffle runs converged
        constexpr int HEADS = 4;
        const unsigned full = 0xffffffffu;
        const int C = p.gat_c, half = lane >> 4, l16 = lane & 15;
        for (int i = 0; i < GC_ROWS / 32; ++i) {
            const int r = warp * (GC_ROWS / 16) + 2 * i + half, g = m0 + r;
            const bool live = r < rows_here;
            const int gs = live ? g : m0;
            float acc[HEADS] = {0.f, 0.f, 0.f, 0.f}, den[HEADS] = {0.f, 0.f, 0.f, 0.f};
            const int e0 = p.row_ptr[gs], deg = live ? p.row_ptr[gs + 1] - e0 : -1;   // dead rows: no trips count
            const int deg_max = max(deg, __shfl_xor_sync(full, deg, 16));
            float ad[HEADS], mx[HEADS];
#pragma unroll
            for (int h = 0; h < HEADS; ++h) {
                ad[h] = p.a_dst[(size_t)gs * HEADS + h];
                mx[h] = -INFINITY;
            }
            float wsum = 0.f;
            for (int base = 0; base <= deg_max; base += 16) {
                const int idx = base + l16;
                const int j = idx < deg ? p.col[e0 + idx] : gs;
                const bool ok = idx <= deg && !(idx < deg && j == gs);
                const float4 as = *reinterpret_cast<const float4*>(p.a_src + (size_t)j * HEADS);
                const float t[HEADS] = {as.x + ad[0], as.y + ad[1], as.z + ad[2], as.w + ad[3]};
#pragma unroll
                for (int h = 0; h < HEADS; ++h) mx[h] = fmaxf(mx[h], ok ? (t[h] > 0.f ? t[h] : 0.2f * t[h]) : -INFINITY);
                if (idx < deg && j != gs) wsum += p.w ? p.w[e0 + idx] : 1.0f;
            }
#pragma unroll
            for (int o = 8; o > 0; o >>= 1) {
#pragma unroll
                for (int h = 0; h < HEADS; ++h) mx[h] = fmaxf(mx[h], __shfl_xor_sync(full, mx[h], o));
                wsum += __shfl_xor_sync(full, wsum, o);
            }
            if (l16 == 0 && live && p.dis_out) {   // gcn_norm: (sum of incoming weights + self weight)^-1/2, inf -> 0
                const float rs = 1.0f / sqrtf(wsum + (p.self_w ? p.self_w[g] : 1.0f));
                p.dis_out[g] = isinf(rs) ? 0.f : rs;
            }
            for (int base = 0; base <= deg_max; base += 16) {
                const int idx = base + l16;
                const int j = idx < deg ? p.col[e0 + idx] : gs;
                const bool ok = idx <= deg && !(idx < deg && j == gs);
                const float4 as = *reinterpret_cast<const float4*>(p.a_src + (size_t)j * HEADS);
                const float t[HEADS] = {as.x + ad[0], as.y + ad[1], as.z + ad[2], as.w + ad[3]};
                float wv[HEADS];
#pragma unroll
                for (int h = 0; h < HEADS; ++h) {
                    const float l = t[h] > 0.f ? t[h] : 0.2f * t[h];
                    wv[h] = ok ? expf(l - mx[h]) : 0.f;
                    den[h] += wv[h];
                }
                const int cnt = min(16, deg_max + 1 - base);
                for (int q = 0; q < cnt; ++q) {   // entries past this half's own edge count carry weight 0 and j = row itself
                    const int jt = __shfl_sync(full, j, q, 16);
                    const float xv = l16 < C ? p.x[(size_t)jt * C + l16] : 0.f;
#pragma unroll
                    for (int h = 0; h < HEADS; ++h) acc[h] = fmaf(__shfl_sync(full, wv[h], q, 16), xv, acc[h]);
                }
            }
#pragma unroll
            for (int o = 8; o > 0; o >>= 1)
#pragma unroll
                for (int h = 0; h < HEADS; ++h) den[h] += __shfl_xor_sync(full, den[h], o);
            if (l16 < C) {
#pragma unroll
                for (int h = 0; h < HEADS; ++h) {
                    const float v = live ? acc[h] / (den[h] + 1e-16f) : 0.f;
                    const op_t hi = f2op(v), lo = f2op(v - op2f(hi));
                    const int k = h * C + l16;
                    const uint32_t off = sw128_offset(r, k >> 3) + (k & 7) * 2;
                    *reinterpret_cast<op_t*>(a_tile + off) = hi;
                    *reinterpret_cast<op_t*>(a_tile + 2 * GC_A_BLOCK + off) = lo;
                }
            }
            for (int k = HEADS * C + l16; k < 64; k += 16) {   // zero padding of the K block
                const uint32_t off = sw128_offset(r, k >> 3) + (k & 7) * 2;
                *reinterpret_cast<op_t*>(a_tile + off) = f2op(0.f);
                *reinterpret_cast<op_t*>(a_tile + 2 * GC_A_BLOCK + off) = f2op(0.f);
            }
        }
    }
    fence_proxy_async_smem();
    tc_fence_before();
    __syncthreads();
    GC_STAMP(2);

    // ---------------------------------------------------------------- phases 2 + 3, up to three chained products
    const int q = warp & 3, grp = warp >> 2;
    const int row = q * 32 + lane, grow = m0 + row;
    const bool row_ok = row < rows_here;
    const uint32_t trow = tmem_base + ((uint32_t)(q * 32) << 16);
    const int n_stage = CHAIN ? (p.do_heads ? 3 : 2) : 1;
    float hp[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
    for (int stage = 0; stage < n_stage; ++stage) {
        const int n = stage == 2 ? 192 : 128;
        const uint32_t tcol = stage == 1 ? 128u : 0u;
        if (tid == 0) {
            mbar_wait(w_full + stage, 0);
            tc_fence_after();
            gc_issue(tmem_base + tcol, smem_u32(a_tile), smem_u32(w_tile), n,
                     stage == 2 ? umma_idesc_f16(GC_ROWS, 192) : umma_idesc_f16(GC_ROWS, 128), GAT ? 1 : 2);
            tc_commit(acc + stage);
            if (stage + 1 < n_stage) {   // the weight buffer is free once these products retire
                mbar_wait(acc + stage, 0);
                gc_load_weight(w_tile, stage == 0 ? &map_fc : &map_h1, w_full + stage + 1, stage == 0 ? 128 : 192);
            }
        }
        __syncwarp();
        mbar_wait(acc + stage, 0);
        __syncwarp();
        tc_fence_after();
        GC_STAMP(3 + 2 * stage);
        const float* sb = stage == 0 ? s_bias : (stage == 1 ? s_fcb : s_h1b);
        for (int c = grp; c < n / 32; c += 4) {
            uint32_t v[32];
            float x[32];
            tc_ld32(trow + tcol + 32 * c, v);
#pragma unroll
            for (int j = 0; j < 32; ++j) x[j] = fmaxf(__uint_as_float(v[j]) + sb[32 * c + j], 0.f);
            if (!CHAIN) {
                if (row_ok) {
                    float* o = p.out + (size_t)grow * GC_H + 32 * c;
#pragma unroll
                    for (int j = 0; j < 32; j += 4) *reinterpret_cast<float4*>(o + j) = make_float4(x[j], x[j + 1], x[j + 2], x[j + 3]);
                }
            } else if (stage < 2) {
                if (stage == 1 && p.emb && row_ok) {
                    float* o = p.emb + (size_t)grow * GC_H + 32 * c;
#pragma unroll
                    for (int j = 0; j < 32; j += 4) *reinterpret_cast<float4*>(o + j) = make_float4(x[j], x[j + 1], x[j + 2], x[j + 3]);
                }
                if (stage + 1 < n_stage) gc_store_split(a_tile, row, (c & 1) * 4, c >> 1, x, 4);
            } else {
                // head hidden columns [32c, 32c+32): head t = c / 2 (64 hidden units each), outputs 2t, 2t+1 (edge: 4)
                const int t = c >> 1, k0 = (c & 1) * 32;
                const float* w0 = s_w2 + (2 * t) * 64 + k0;
                float s0 = 0.f, s1 = 0.f;
#pragma unroll
                for (int j = 0; j < 32; ++j) s0 = fmaf(x[j], w0[j], s0);
                if (t < 2) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) s1 = fmaf(x[j], w0[64 + j], s1);
                }
                if (t == 0) { hp[0] += s0; hp[1] += s1; }
                else if (t == 1) { hp[2] += s0; hp[3] += s1; }
                else hp[4] += s0;
            }
        }
        if (stage + 1 < n_stage) {
            fence_proxy_async_smem();
            tc_fence_before();
            __syncthreads();
        }
        GC_STAMP(4 + 2 * stage);
    }
    if (CHAIN && p.do_heads) {
#pragma unroll
        for (int o = 0; o < 5; ++o) s_part[(grp * 128 + row) * 5 + o] = hp[o];
        __syncthreads();
        if (grp == 0 && row_ok) {
            float y[5];
#pragma unroll
            for (int o = 0; o < 5; ++o)
                y[o] = s_part[row * 5 + o] + s_part[(128 + row) * 5 + o] + s_part[(256 + row) * 5 + o] + s_part[(384 + row) * 5 + o] + s_b2[o];
            if (p.out_mask) *reinterpret_cast<float2*>(p.out_mask + (size_t)grow * 2) = make_float2(y[0], y[1]);
            if (p.out_inst) *reinterpret_cast<float2*>(p.out_inst + (size_t)grow * 2) = make_float2(y[2], y[3]);
            if (p.out_edge) p.out_edge[grow] = y[4];
            if (p.prob_out) {   // softmax over the two mask logits, class 1 (models/region_graph/test.py:238)
                const float m = fmaxf(y[0], y[1]);
                const float e0 = expf(y[0] - m), e1 = expf(y[1] - m);
                p.prob_out[grow] = e1 / (e0 + e1);
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    GC_STAMP(9);
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS) : "memory");
    }
}

int launch_gcn_layer_tc(const GcnLayer& p, cudaStream_t st) {
    if (p.M <= 0) return CM_OK;
    CM_REQUIRE(p.x && p.row_ptr && p.col && p.bias && p.w_split, "gcn layer: NULL operand");
    CM_REQUIRE(p.gat ? (p.a_src && p.a_dst && p.gat_c > 0 && 4 * p.gat_c <= 64 && p.out) : (p.dis != nullptr), "gcn layer: missing GAT / GCN operand");
    CM_REQUIRE(p.chain || p.out, "gcn layer: no output");
    CM_REQUIRE(!p.chain || (p.fc_split && (!p.do_heads || (p.h1_split && p.h1_bias))), "gcn layer: chained tail needs its weights");
    CUtensorMap mw, mfc, mh1;
    int rc = p.gat ? make_map(&mw, p.w_split, GC_H, 192, 192, GC_H) : make_map(&mw, p.w_split, GC_H, 3 * GC_H, 3 * GC_H, GC_H);
    if (rc != CM_OK) return rc;
    mfc = mw;
    mh1 = mw;
    if (p.chain) {
        if ((rc = make_map(&mfc, p.fc_split, GC_H, 3 * GC_H, 3 * GC_H, GC_H)) != CM_OK) return rc;
        if (p.do_heads && (rc = make_map(&mh1, p.h1_split, 192, 3 * GC_H, 3 * GC_H, 192)) != CM_OK) return rc;
    }
    static PerDeviceOnce attr;
    if (attr.first()) {
        cudaFuncSetAttribute(gcn_layer_tc_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, GC_SMEM);
        cudaFuncSetAttribute(gcn_layer_tc_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, GC_SMEM);
        cudaFuncSetAttribute(gcn_layer_tc_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, GC_SMEM_GAT);
    }
    const int grid = ceil_div(p.M, GC_ROWS);
    if (p.gat) {
        CM_TIMED("gat_layer_tc", st);
        gcn_layer_tc_kernel<2><<<grid, GC_THREADS, GC_SMEM_GAT, st>>>(mw, mfc, mh1, p);
    } else if (p.chain) {
        CM_TIMED("gcn_tail_tc", st);
        gcn_layer_tc_kernel<1><<<grid, GC_THREADS, GC_SMEM, st>>>(mw, mfc, mh1, p);
    } else {
        CM_TIMED("gcn_layer_tc", st);
        gcn_layer_tc_kernel<0><<<grid, GC_THREADS, GC_SMEM, st>>>(mw, mfc, mh1, p);
    }
    return post_launch("gcn_layer_tc");
}

}  // namespace cm
