// A3 on 8-bit images: the 15 region features from the RGB bytes a decoder produces (PIL `convert('RGB')`,
// extract_rg_embeddings.py:260-261), image value = byte / 255.0 -- what the reference's ToTensor -> Normalize ->
// de-normalise round trip (:263-276) reproduces to float32 rounding.  Also the int16 -> int32 label widening
// used by the host-facing pipeline (SLIC labels are < 32768; the reference holds them as int64).
//
// Why a second kernel: with byte inputs every moment the features need is an INTEGER
//   n, sum x, sum y, canny count, perimeter count, ring count, sum r|g|b, sum r^2|g^2|b^2, sum rg|rb|gb, ring sum r|g|b
// so the per-label table is 18 native 32-bit shared-memory atomics wide (the float kernel needs 39 rows of
// 16-bit limbs of 64-bit fixed point), no fp64 arithmetic happens per pixel, and the result is EXACT: variances come
// from the integer numerators n*sum(c^2) - sum(c)^2, the gray statistics (:151, 166-167, 193) from the same
// covariance integers, and only the final divisions round (once, in fp64, then to float32 like the reference's
// torch.FloatTensor(...), :213).  The host copies 196 KB per image instead of 786 KB.
//
// Layout: a cluster of CL CTAs owns one image (H/CL rows each).  A warp takes a band of 8 rows x 32 columns; a lane
// walks DOWN its column keeping the sums of at most two labels (member) and two foreign labels (perimeter / ring,
// emitted by the OUTSIDE pixel exactly as in features.cu) in registers, lanes holding the same label are combined
// with match.any + redux.sync, and one lane per label issues the shared-memory atomics.  After a cluster barrier one
// thread per node sums the CL partial tables over distributed shared memory and writes the 15 floats.
#include <cooperative_groups.h>
#include <limits.h>

#include <algorithm>

#include "common.cuh"

namespace cg = cooperative_groups;

namespace cm {

constexpr int U8_THREADS = 512, U8_NONE = INT_MIN, U8_ROWS = 18;
enum { UQ_N = 0, UQ_SX, UQ_SY, UQ_CAN, UQ_PER, UQ_RN, UQ_R, UQ_G, UQ_B, UQ_RR, UQ_GG, UQ_BB, UQ_RG, UQ_RB, UQ_GB, UQ_XR, UQ_XG, UQ_XB };

__global__ void __launch_bounds__(U8_THREADS)
node_features_u8_kernel(const uint8_t* __restrict__ image, const int* __restrict__ labels, const uint8_t* __restrict__ canny, int H, int W,
                        int cap, const int* __restrict__ node_ptr, const int* __restrict__ node_label, float norm_side,
                        float* __restrict__ xout, int node_cap) {
    extern __shared__ __align__(16) unsigned char u8_smem[];
    cg::cluster_group cluster = cg::this_cluster();
    const int ncl = (int)gridDim.x, rank = blockIdx.x, b = blockIdx.y, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int rpc = (H + ncl - 1) / ncl, y0 = rank * rpc, y1 = min(H, y0 + rpc), th = rpc + 4;
    unsigned int* tab = reinterpret_cast<unsigned int*>(u8_smem);                       // [U8_ROWS][cap]
    int* lab_s = reinterpret_cast<int*>(tab + (U8_ROWS * (size_t)cap + 3) / 4 * 4);     // [th][W], rows y0-2 .. y0+rpc+1
    const int* lab = labels + (size_t)b * H * W;
    const uint8_t* img = image + (size_t)b * H * W * 3;
    const uint8_t* can = canny + (size_t)b * H * W;

    for (int i = tid; i < U8_ROWS * cap; i += U8_THREADS) tab[i] = 0u;
    if ((W & 3) == 0) {
        for (int i = tid; i < th * (W / 4); i += U8_THREADS) {
            const int r = i / (W / 4), c4 = (i - r * (W / 4)) * 4, y = y0 - 2 + r;
            int4 v = make_int4(U8_NONE, U8_NONE, U8_NONE, U8_NONE);
            if (y >= 0 && y < H) v = *reinterpret_cast<const int4*>(lab + (size_t)y * W + c4);
            *reinterpret_cast<int4*>(lab_s + r * W + c4) = v;
        }
    } else {
        for (int i = tid; i < th * W; i += U8_THREADS) {
            const int r = i / W, y = y0 - 2 + r;
            lab_s[i] = (y >= 0 && y < H) ? lab[(size_t)y * W + (i - r * W)] : U8_NONE;
        }
    }
    __syncthreads();

    struct Member {
        int L;
        unsigned q[12];   // n | canny << 16, sx, sy, r, g, b, rr, gg, bb, rg, rb, gb
    };
    struct Ring {
        int L;
        unsigned q[4];    // ring count | perimeter << 16, r, g, b
    };
    auto row_of = [&](int q) { return tab + (size_t)q * cap; };
    auto flush_member = [&](Member& m) {
        if (__any_sync(0xffffffffu, m.L != U8_NONE)) {
            const unsigned grp = __match_any_sync(0xffffffffu, m.L);
            const bool lead = m.L != U8_NONE && lane == __ffs(grp) - 1;
            unsigned t[12];
#pragma unroll
            for (int k = 0; k < 12; ++k) t[k] = __reduce_add_sync(grp, m.q[k]);
            if (lead) {
                atomicAdd(row_of(UQ_N) + m.L, t[0] & 0xffffu);
                if (t[0] >> 16) atomicAdd(row_of(UQ_CAN) + m.L, t[0] >> 16);
                atomicAdd(row_of(UQ_SX) + m.L, t[1]);
                atomicAdd(row_of(UQ_SY) + m.L, t[2]);
#pragma unroll
                for (int k = 0; k < 9; ++k) atomicAdd(row_of(UQ_R + k) + m.L, t[3 + k]);
            }
        }
        m.L = U8_NONE;
#pragma unroll
        for (int k = 0; k < 12; ++k) m.q[k] = 0u;
    };
    auto flush_ring = [&](Ring& r) {
        if (__any_sync(0xffffffffu, r.L != U8_NONE)) {
            const unsigned grp = __match_any_sync(0xffffffffu, r.L);
            const bool lead = r.L != U8_NONE && lane == __ffs(grp) - 1;
            unsigned t[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) t[k] = __reduce_add_sync(grp, r.q[k]);
            if (lead) {
                atomicAdd(row_of(UQ_RN) + r.L, t[0] & 0xffffu);
                if (t[0] >> 16) atomicAdd(row_of(UQ_PER) + r.L, t[0] >> 16);
                atomicAdd(row_of(UQ_XR) + r.L, t[1]);
                atomicAdd(row_of(UQ_XG) + r.L, t[2]);
                atomicAdd(row_of(UQ_XB) + r.L, t[3]);
            }
        }
        r.L = U8_NONE;
        r.q[0] = r.q[1] = r.q[2] = r.q[3] = 0u;
    };

    const int n_bands = (y1 - y0 + 7) / 8, n_xg = (W + 31) / 32;
    for (int item = warp; item < n_bands * n_xg; item += U8_THREADS / 32) {
        const int yb = y0 + (item / n_xg) * 8;
        const int x = (item % n_xg) * 32 + lane;
        const bool in = x < W;
        Member ma, mb;
        Ring ra, rb;
        ma.L = mb.L = ra.L = rb.L = U8_NONE;
#pragma unroll
        for (int k = 0; k < 12; ++k) ma.q[k] = mb.q[k] = 0u;
#pragma unroll
        for (int k = 0; k < 4; ++k) ra.q[k] = rb.q[k] = 0u;
        for (int i = 0; i < 8; ++i) {
            const int y = yb + i;
            if (y >= y1) break;
            if (!in) continue;
            const int* row = lab_s + (y - y0 + 2) * W;
            const size_t p = (size_t)y * W + x;
            const int L = row[x];
            const unsigned r = img[3 * p], g = img[3 * p + 1], bl = img[3 * p + 2];
            if (L >= 0 && L < cap) {
                const unsigned t[12] = {1u | (can[p] != 0 ? 0x10000u : 0u), (unsigned)x, (unsigned)y, r, g, bl,
                                        r * r, g * g, bl * bl, r * g, r * bl, g * bl};
                if (ma.L == U8_NONE) ma.L = L;
                if (L == ma.L) {
#pragma unroll
                    for (int k = 0; k < 12; ++k) ma.q[k] += t[k];
                } else {
                    if (mb.L == U8_NONE) mb.L = L;
                    if (L == mb.L) {
#pragma unroll
                        for (int k = 0; k < 12; ++k) mb.q[k] += t[k];
                    } else {   // a third label inside 8 rows of one column: straight to the table
                        atomicAdd(row_of(UQ_N) + L, 1u);
                        if (t[0] >> 16) atomicAdd(row_of(UQ_CAN) + L, 1u);
                        atomicAdd(row_of(UQ_SX) + L, t[1]);
                        atomicAdd(row_of(UQ_SY) + L, t[2]);
#pragma unroll
                        for (int k = 0; k < 9; ++k) atomicAdd(row_of(UQ_R + k) + L, t[3 + k]);
                    }
                }
            }
            // ---- contributions to foreign neighbours: distinct labels among the 12 L1 <= 2 neighbours, the
            // 4-neighbours first (a label first seen there makes this a perimeter pixel of that region, :177-178;
            // any of them makes it a ring pixel, :185-186)
            int v[12];
            v[0] = x > 0 ? row[x - 1] : U8_NONE;
            v[1] = x + 1 < W ? row[x + 1] : U8_NONE;
            v[2] = row[x - W];
            v[3] = row[x + W];
            v[4] = x > 1 ? row[x - 2] : U8_NONE;
            v[5] = x + 2 < W ? row[x + 2] : U8_NONE;
            v[6] = row[x - 2 * W];
            v[7] = row[x + 2 * W];
            v[8] = x > 0 ? row[x - W - 1] : U8_NONE;
            v[9] = x + 1 < W ? row[x - W + 1] : U8_NONE;
            v[10] = x > 0 ? row[x + W - 1] : U8_NONE;
            v[11] = x + 1 < W ? row[x + W + 1] : U8_NONE;
            int u1 = U8_NONE, u2 = U8_NONE, u3 = U8_NONE;
            unsigned nf = 0, over = 0;
#pragma unroll
            for (int k = 0; k < 12; ++k) {
                const int u = v[k];
                if ((u != L) & (u >= 0) & (u < cap) & (u != u1) & (u != u2) & (u != u3)) {
                    if (u1 == U8_NONE) { u1 = u; nf |= k < 4 ? 1u : 0u; }
                    else if (u2 == U8_NONE) { u2 = u; nf |= k < 4 ? 2u : 0u; }
                    else if (u3 == U8_NONE) { u3 = u; nf |= k < 4 ? 4u : 0u; }
                    else over |= 1u << k;
                }
            }
            auto contribute = [&](int u, bool near) {
                const unsigned cp = 1u | (near ? 0x10000u : 0u);
                if (ra.L == U8_NONE) ra.L = u;
                if (u == ra.L) {
                    ra.q[0] += cp; ra.q[1] += r; ra.q[2] += g; ra.q[3] += bl;
                } else {
                    if (rb.L == U8_NONE) rb.L = u;
                    if (u == rb.L) {
                        rb.q[0] += cp; rb.q[1] += r; rb.q[2] += g; rb.q[3] += bl;
                    } else {
                        atomicAdd(row_of(UQ_RN) + u, 1u);
                        if (near) atomicAdd(row_of(UQ_PER) + u, 1u);
                        atomicAdd(row_of(UQ_XR) + u, r);
                        atomicAdd(row_of(UQ_XG) + u, g);
                        atomicAdd(row_of(UQ_XB) + u, bl);
                    }
                }
            };
            if (u1 != U8_NONE) contribute(u1, (nf & 1u) != 0);
            if (u2 != U8_NONE) contribute(u2, (nf & 2u) != 0);
            if (u3 != U8_NONE) contribute(u3, (nf & 4u) != 0);
            if (over) {
#pragma unroll
                for (int k = 0; k < 12; ++k) {
                    if (!((over >> k) & 1u)) continue;
                    bool first = true;
#pragma unroll
                    for (int j = 0; j < k; ++j) first &= v[j] != v[k];
                    if (first) contribute(v[k], k < 4);
                }
            }
        }
        __syncwarp();
        flush_member(ma);
        flush_member(mb);
        flush_ring(ra);
        flush_ring(rb);
    }
    cluster.sync();

    // ---- one thread per node of the image: sum the partial tables over DSMEM, then the 15 features (:154-211)
    const int n0 = node_ptr[b], nb = min(node_ptr[b + 1], node_cap) - n0;
    for (int i = rank * U8_THREADS + tid; i < nb; i += ncl * U8_THREADS) {
        const int gi = n0 + i, L = node_label[gi];
        unsigned long long q[U8_ROWS];
#pragma unroll
        for (int k = 0; k < U8_ROWS; ++k) q[k] = 0ull;
        if (L >= 0 && L < cap) {
            for (int r = 0; r < ncl; ++r) {
                const unsigned int* rt = cluster.map_shared_rank(tab, r);
#pragma unroll
                for (int k = 0; k < U8_ROWS; ++k) q[k] += rt[(size_t)k * cap + L];
            }
        }
        float* o = xout + (size_t)gi * CM_NODE_FEATURES;
        const double n = (double)q[UQ_N], s = 255.0 * n, s2 = s * s;
        const long long ni = (long long)q[UQ_N];
        const long long sr = (long long)q[UQ_R], sg = (long long)q[UQ_G], sb = (long long)q[UQ_B];
        // integer covariance numerators n*sum(ab) - sum(a)sum(b): exact
        const double crr = (double)(ni * (long long)q[UQ_RR] - sr * sr), cgg = (double)(ni * (long long)q[UQ_GG] - sg * sg),
                     cbb = (double)(ni * (long long)q[UQ_BB] - sb * sb), crg = (double)(ni * (long long)q[UQ_RG] - sr * sg),
                     crb = (double)(ni * (long long)q[UQ_RB] - sr * sb), cgb = (double)(ni * (long long)q[UQ_GB] - sg * sb);
        const double mean[3] = {(double)sr / s, (double)sg / s, (double)sb / s};
        const double var[3] = {crr / s2, cgg / s2, cbb / s2};
        const int rcnt = (int)q[UQ_RN];
        double contrast2 = 0.0;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            o[c] = (float)mean[c];
            o[3 + c] = (float)sqrt(var[c] > 0.0 ? var[c] : 0.0);   // population std, :165
            if (rcnt > 0) {
                const double d = mean[c] - (double)q[UQ_XR + c] / (255.0 * (double)rcnt);
                contrast2 += d * d;
            }
        }
        const double wr = 0.2989, wg = 0.5870, wb = 0.1140;   // :151
        const double gm = (wr * (double)sr + wg * (double)sg + wb * (double)sb) / s;
        double gvar = (wr * wr * crr + wg * wg * cgg + wb * wb * cbb + 2.0 * (wr * wg * crg + wr * wb * crb + wg * wb * cgb)) / s2;
        gvar = gvar > 0.0 ? gvar : 0.0;
        const double ns = (double)norm_side;
        o[6] = (float)gm;
        o[7] = (float)sqrt(gvar);
        o[8] = (float)(((double)q[UQ_SX] / n) / ns);    // centroid x, :173
        o[9] = (float)(((double)q[UQ_SY] / n) / ns);    // centroid y, :172
        o[10] = (float)(n / (ns * ns));                 // :174
        const double per = (double)q[UQ_PER];
        o[11] = (float)((per * per) / (4.0 * 3.14159265358979323846 * n + 1e-10));   // :179
        o[12] = (float)sqrt(contrast2);                 // :190 (0 when the ring is empty)
        o[13] = (float)((double)q[UQ_CAN] / n);         // :182
        o[14] = (float)gvar;                            // :193
    }
    cluster.sync();   // peers may still be reading this CTA's tables
}

// ---------------------------------------------------------------------------------------------------------------
// Region-parallel variant (the default when the bounding boxes from cm_rag_build are passed): one warp per region
// scans the region's bounding box grown by two pixels.  A member pixel adds its 12 integer moments; a non-member
// pixel asks "is one of my 4 / 12 neighbours in THIS region" -- no de-duplication of foreign labels, no atomics,
// no shared memory, ~1.7 pixel visits per pixel for compact superpixels.  Label and image windows are small and
// re-read through L1/L2; DRAM sees each byte once.  ~9x fewer instructions than the tile kernel on SLIC-shaped
// regions; the tile kernel remains the better choice for a few huge regions.
// the image b with node_ptr[b] <= g < node_ptr[b + 1].  Images hold about the same number of regions, so the proportional
// guess is right or one off: two independent loads instead of the log2(batch) dependent ones of a bisection.
__device__ __forceinline__ int u8_find_image(const int* __restrict__ node_ptr, int batch, int g, int total) {
    int b = min(__float2int_rz((float)g * (float)batch / (float)total), batch - 1);
    int lo = node_ptr[b], hi = node_ptr[b + 1];
    while (lo > g) {
        hi = lo;
        lo = node_ptr[--b];
    }
    while (hi <= g) {
        lo = hi;
        hi = node_ptr[++b + 1];
    }
    return b;
}
// warp sum of 32 values < 2^32 whose total may exceed 2^32
__device__ __forceinline__ unsigned long long warp_sum_wide(unsigned v) {
    const unsigned lo = __reduce_add_sync(0xffffffffu, v & 0xffffu), hi = __reduce_add_sync(0xffffffffu, v >> 16);
    return (unsigned long long)lo + ((unsigned long long)hi << 16);
}

struct U8Moments {
    unsigned long long n, sx, sy, can, per, rn, r, g, b, rr, gg, bb, rg, rb, gb, xr, xg, xb;
};
// the 15 features from the exact integer moments (extract_rg_embeddings.py:154-211)
__device__ __forceinline__ void u8_write_features(const U8Moments& m, float norm_side, float* __restrict__ o) {
    const double n = (double)m.n, s = 255.0 * n, s2 = s * s;
    const long long ni = (long long)m.n, sr = (long long)m.r, sg = (long long)m.g, sb = (long long)m.b;
    // integer covariance numerators n*sum(ab) - sum(a)sum(b): exact
    const double crr = (double)(ni * (long long)m.rr - sr * sr), cgg = (double)(ni * (long long)m.gg - sg * sg),
                 cbb = (double)(ni * (long long)m.bb - sb * sb), crg = (double)(ni * (long long)m.rg - sr * sg),
                 crb = (double)(ni * (long long)m.rb - sr * sb), cgb = (double)(ni * (long long)m.gb - sg * sb);
    const double mean[3] = {(double)sr / s, (double)sg / s, (double)sb / s};
    const double var[3] = {crr / s2, cgg / s2, cbb / s2};
    const unsigned long long ring[3] = {m.xr, m.xg, m.xb};
    double contrast2 = 0.0;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        o[c] = (float)mean[c];
        o[3 + c] = (float)sqrt(var[c] > 0.0 ? var[c] : 0.0);   // population std, :165
        if (m.rn > 0) {
            const double d = mean[c] - (double)ring[c] / (255.0 * (double)m.rn);
            contrast2 += d * d;
        }
    }
    const double wr = 0.2989, wg = 0.5870, wb = 0.1140;   // :151
    const double gm = (wr * (double)sr + wg * (double)sg + wb * (double)sb) / s;
    double gvar = (wr * wr * crr + wg * wg * cgg + wb * wb * cbb + 2.0 * (wr * wg * crg + wr * wb * crb + wg * wb * cgb)) / s2;
    gvar = gvar > 0.0 ? gvar : 0.0;
    const double ns = (double)norm_side;
    o[6] = (float)gm;
    o[7] = (float)sqrt(gvar);
    o[8] = (float)(((double)m.sx / n) / ns);    // centroid x, :173
    o[9] = (float)(((double)m.sy / n) / ns);    // centroid y, :172
    o[10] = (float)(n / (ns * ns));             // :174
    const double per = (double)m.per;
    o[11] = (float)((per * per) / (4.0 * 3.14159265358979323846 * n + 1e-10));   // :179
    o[12] = (float)sqrt(contrast2);             // :190 (0 when the ring is empty)
    o[13] = (float)((double)m.can / n);         // :182
    o[14] = (float)gvar;                        // :193
}

// Same arithmetic as u8_write_features with the moments warp-uniform and the 15 outputs spread over the lanes: every
// lane runs ONE fp64 division / square root on its own operands instead of lane 0 running sixteen of them in turn.
__device__ __forceinline__ void u8_write_features_warp(const U8Moments& m, float norm_side, float* __restrict__ o, int lane) {
    const unsigned full = 0xffffffffu;
    const double n = (double)m.n, s = 255.0 * n, s2 = s * s, ns = (double)norm_side;
    const long long ni = (long long)m.n, sr = (long long)m.r, sg = (long long)m.g, sb = (long long)m.b;
    const double crr = (double)(ni * (long long)m.rr - sr * sr), cgg = (double)(ni * (long long)m.gg - sg * sg),
                 cbb = (double)(ni * (long long)m.bb - sb * sb), crg = (double)(ni * (long long)m.rg - sr * sg),
                 crb = (double)(ni * (long long)m.rb - sr * sb), cgb = (double)(ni * (long long)m.gb - sg * sb);
    const double wr = 0.2989, wg = 0.5870, wb = 0.1140;   // :151
    const double gnum = wr * (double)sr + wg * (double)sg + wb * (double)sb;
    const double vnum = wr * wr * crr + wg * wg * cgg + wb * wb * cbb + 2.0 * (wr * wg * crg + wr * wb * crb + wg * wb * cgb);
    const double per = (double)m.per, rden = m.rn > 0 ? 255.0 * (double)m.rn : 1.0;
    double num = 0.0, den = 1.0;
    switch (lane) {
        case 0: num = (double)sr; den = s; break;
        case 1: num = (double)sg; den = s; break;
        case 2: num = (double)sb; den = s; break;
        case 3: num = crr; den = s2; break;
        case 4: num = cgg; den = s2; break;
        case 5: num = cbb; den = s2; break;
        case 6: num = gnum; den = s; break;
        case 7: case 14: num = vnum; den = s2; break;
        case 8: num = (double)m.sx; den = n; break;
        case 9: num = (double)m.sy; den = n; break;
        case 10: num = n; den = ns * ns; break;
        case 11: num = per * per; den = 4.0 * 3.14159265358979323846 * n + 1e-10; break;   // :179
        case 13: num = (double)m.can; den = n; break;
        case 16: num = (double)m.xr; den = rden; break;
        case 17: num = (double)m.xg; den = rden; break;
        case 18: num = (double)m.xb; den = rden; break;
        default: break;
    }
    double a = num / den;
    const double a_ns = a / ns;                       // centroids, :172-173
    if (lane == 8 || lane == 9) a = a_ns;
    if ((lane >= 3 && lane <= 5) || lane == 7 || lane == 14) a = a > 0.0 ? a : 0.0;
    const double d = a - __shfl_down_sync(full, a, 16);   // lanes 0..2: mean_c - ring mean_c
    const double d0 = __shfl_sync(full, d, 0), d1 = __shfl_sync(full, d, 1), d2 = __shfl_sync(full, d, 2);
    if (lane == 12) a = m.rn > 0 ? __fma_rn(d2, d2, __fma_rn(d1, d1, d0 * d0)) : 0.0;   // :190
    const double rt = sqrt(a);
    if ((lane >= 3 && lane <= 5) || lane == 7 || lane == 12) a = rt;
    if (lane < CM_NODE_FEATURES) o[lane] = (float)a;
}

template <typename LT>
__global__ void __launch_bounds__(256) node_features_u8_bbox_kernel(const uint8_t* __restrict__ image, const LT* __restrict__ labels,
                                                                    const uint8_t* __restrict__ canny, int batch, int H, int W,
                                                                    const int* __restrict__ node_ptr, const int* __restrict__ node_label,
                                                                    const int* __restrict__ node_bbox, float norm_side,
                                                                    float* __restrict__ xout, int node_cap) {
    const int lane = threadIdx.x & 31;
    const int gi = blockIdx.x * 8 + (threadIdx.x >> 5);
    const int total = min(node_ptr[batch], node_cap);
    if (gi >= total) return;
    const int b = u8_find_image(node_ptr, batch, gi, node_ptr[batch]);
    const int L = node_label[gi];
    const int4 bb = *reinterpret_cast<const int4*>(node_bbox + 4 * (size_t)gi);
    const int wx0 = max(bb.x - 2, 0), wy0 = max(bb.y - 2, 0), wx1 = min(bb.z + 2, W - 1), wy1 = min(bb.w + 2, H - 1);
    const int ww = wx1 - wx0 + 1, npx = ww * (wy1 - wy0 + 1);
    const float inv_ww = 1.0f / (float)ww;
    const LT* lab = labels + (size_t)b * H * W;
    const uint8_t* img = image + (size_t)b * H * W * 3;
    const uint8_t* can = canny + (size_t)b * H * W;

    unsigned n = 0, sx = 0, sy = 0, sc = 0, per = 0, rn = 0, r1 = 0, g1 = 0, b1 = 0, rr = 0, gg = 0, b2 = 0, rg = 0, rb = 0, gb = 0,
             xr = 0, xg = 0, xb = 0;
    const int wh = wy1 - wy0 + 1;
    __shared__ unsigned short visit[8][1024];
    if (ww <= 32 && wh <= 32) {
        // Window of at most 32 x 32 pixels (every SLIC-sized region).  Lane y keeps row y of the membership mask as
        // one 32-bit word (ballot over the row's labels); the 4-connected and L1 <= 2 dilations (:177, :185) are
        // shifts and ORs of the words above / below (shuffles) and the three counts are popc.  The member and ring
        // pixels are then written as a compact list (row, column, member bit) to shared memory so that the moment
        // pass keeps all 32 lanes on pixels that count: one label load per window pixel instead of thirteen, and
        // ~0.7 visits per window pixel instead of one.
        const unsigned full = 0xffffffffu;
        const bool colin = lane < ww;
        const int base = wy0 * W + wx0;
        const LT* lp0 = lab + base + (colin ? lane : 0);
        unsigned m = 0;
        for (int y8 = 0; y8 < wh; y8 += 16) {   // up to sixteen rows of label loads in flight; rows past the window are skipped
            int lw[16];
            const LT* lp = lp0 + y8 * W;
#pragma unroll
            for (int k4 = 0; k4 < 16; k4 += 4) {
                if (y8 + k4 >= wh) break;       // warp-uniform
#pragma unroll
                for (int k = k4; k < k4 + 4; ++k) lw[k] = __ldg(lp + min(k, wh - 1 - y8) * W);
            }
#pragma unroll
            for (int k4 = 0; k4 < 16; k4 += 4) {
                if (y8 + k4 >= wh) break;
#pragma unroll
                for (int k = k4; k < k4 + 4; ++k) {
                    const unsigned bal = __ballot_sync(full, colin && lw[k] == L);
                    if (lane == y8 + k) m = bal;
                }
            }
        }
        if (lane >= wh) m = 0;
        unsigned up1 = __shfl_up_sync(full, m, 1), up2 = __shfl_up_sync(full, m, 2);
        unsigned dn1 = __shfl_down_sync(full, m, 1), dn2 = __shfl_down_sync(full, m, 2);
        if (lane < 1) up1 = 0;
        if (lane < 2) up2 = 0;
        if (lane > 30) dn1 = 0;
        if (lane > 29) dn2 = 0;
        const unsigned keep = lane < wh ? (ww == 32 ? full : ((1u << ww) - 1u)) & ~m : 0u;
        const unsigned h1 = (m << 1) | (m >> 1), v1 = up1 | dn1;
        const unsigned m1 = keep & (h1 | v1);
        const unsigned m2 = keep & (h1 | v1 | (m << 2) | (m >> 2) | (v1 << 1) | (v1 >> 1) | up2 | dn2);
        n = __popc(m);
        per = __popc(m1);
        rn = __popc(m2);
        // compact list of the pixels to visit: exclusive scan of the per-row counts, each lane writes its row
        unsigned vv = m | m2;
        int incl = __popc(vv);
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int t = __shfl_up_sync(full, incl, d);
            if (lane >= d) incl += t;
        }
        const int total_px = __shfl_sync(full, incl, 31);
        unsigned short* lst = visit[threadIdx.x >> 5];
        int o = incl - __popc(vv);
        while (vv) {
            const int bit = __ffs(vv) - 1;
            vv &= vv - 1;
            lst[o++] = (unsigned short)((lane << 5) | bit | (((m >> bit) & 1u) ? 0x400u : 0x800u));
        }
        __syncwarp();
        const uint8_t* ip0 = img + 3 * (size_t)base;
        const uint8_t* cp0 = can + base;
        for (int k0 = 0; k0 < total_px; k0 += 64) {   // two list entries per lane per trip, their loads issued together
            unsigned e[2], r[2], g[2], bl[2], cb[2];
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                const int k = k0 + 32 * j + lane;
                e[j] = k < total_px ? lst[k] : 0u;   // 0: pixel (0, 0) of the window with both weights 0
                const int off = ((e[j] >> 5) & 31) * W + (e[j] & 31);
                const uint8_t* ip = ip0 + 3 * off;
                r[j] = ip[0];
                g[j] = ip[1];
                bl[j] = ip[2];
                cb[j] = cp0[off];
            }
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                const unsigned mm = (e[j] >> 10) & 1u, q2 = e[j] >> 11;
                const unsigned rm = r[j] * mm, gm = g[j] * mm, bm = bl[j] * mm;
                sx += (unsigned)(wx0 + (e[j] & 31)) * mm; sy += (unsigned)(wy0 + ((e[j] >> 5) & 31)) * mm; sc += cb[j] != 0 ? mm : 0u;
                r1 += rm; g1 += gm; b1 += bm;
                rr += r[j] * rm; gg += g[j] * gm; b2 += bl[j] * bm;
                rg += r[j] * gm; rb += r[j] * bm; gb += g[j] * bm;
                xr += r[j] * q2; xg += g[j] * q2; xb += bl[j] * q2;
            }
        }
    } else
    for (int i = lane; i < npx; i += 32) {
        int yy = __float2int_rz(((float)i + 0.5f) * inv_ww);   // i / ww, fixed up below (npx < 2^22 keeps it within +-1)
        int xx = i - yy * ww;
        if (xx < 0) { --yy; xx += ww; }
        if (xx >= ww) { ++yy; xx -= ww; }
        const int y = wy0 + yy, x = wx0 + xx;
        const LT* lp = lab + (size_t)y * W + x;
        const size_t p = (size_t)y * W + x;
        // every load of the iteration is issued up front (out-of-image neighbours are clamped onto the pixel itself,
        // whose label is != L whenever the neighbour tests matter): one memory round trip per pixel, not thirteen
        const int ox1 = x > 0 ? -1 : 0, ox2 = x + 1 < W ? 1 : 0, oy1 = y > 0 ? -W : 0, oy2 = y + 1 < H ? W : 0;
        const int l0 = __ldg(lp);
        const int a0 = __ldg(lp + ox1), a1 = __ldg(lp + ox2), a2 = __ldg(lp + oy1), a3 = __ldg(lp + oy2);
        const int c0 = __ldg(lp + (x > 1 ? -2 : 0)), c1 = __ldg(lp + (x + 2 < W ? 2 : 0)), c2 = __ldg(lp + (y > 1 ? -2 * W : 0)),
                  c3 = __ldg(lp + (y + 2 < H ? 2 * W : 0));
        const int d0 = __ldg(lp + ((ox1 != 0 && oy1 != 0) ? oy1 + ox1 : 0)), d1 = __ldg(lp + ((ox2 != 0 && oy1 != 0) ? oy1 + ox2 : 0)),
                  d2 = __ldg(lp + ((ox1 != 0 && oy2 != 0) ? oy2 + ox1 : 0)), d3 = __ldg(lp + ((ox2 != 0 && oy2 != 0) ? oy2 + ox2 : 0));
        const unsigned r = img[3 * p], g = img[3 * p + 1], bl = img[3 * p + 2];
        const unsigned cn = can[p] != 0;
        const bool member = l0 == L;
        // inside the 4-connected cross dilation of the region (:177)?  within L1 distance 2 (:185)?
        const bool near1 = !member & ((a0 == L) | (a1 == L) | (a2 == L) | (a3 == L));
        const bool near2 = !member & (near1 | (c0 == L) | (c1 == L) | (c2 == L) | (c3 == L) | (d0 == L) | (d1 == L) | (d2 == L) | (d3 == L));
        const unsigned mm = member ? 1u : 0u, m2 = near2 ? 1u : 0u;
        const unsigned rm = r * mm, gm = g * mm, bm = bl * mm;
        n += mm; sx += (unsigned)x * mm; sy += (unsigned)y * mm; sc += cn & mm;
        r1 += rm; g1 += gm; b1 += bm;
        rr += r * rm; gg += g * gm; b2 += bl * bm;
        rg += r * gm; rb += r * bm; gb += g * bm;
        per += near1 ? 1u : 0u;
        rn += m2; xr += r * m2; xg += g * m2; xb += bl * m2;
    }
    U8Moments m;
    if (ww <= 32 && wh <= 32) {
        // at most 1024 pixels of 8-bit values: every warp total is below 2^27 -- one redux each, no 16-bit halves
        m.n = __reduce_add_sync(0xffffffffu, n); m.can = __reduce_add_sync(0xffffffffu, sc);
        m.per = __reduce_add_sync(0xffffffffu, per); m.rn = __reduce_add_sync(0xffffffffu, rn);
        m.sx = __reduce_add_sync(0xffffffffu, sx); m.sy = __reduce_add_sync(0xffffffffu, sy);
        m.r = __reduce_add_sync(0xffffffffu, r1); m.g = __reduce_add_sync(0xffffffffu, g1); m.b = __reduce_add_sync(0xffffffffu, b1);
        m.rr = __reduce_add_sync(0xffffffffu, rr); m.gg = __reduce_add_sync(0xffffffffu, gg); m.bb = __reduce_add_sync(0xffffffffu, b2);
        m.rg = __reduce_add_sync(0xffffffffu, rg); m.rb = __reduce_add_sync(0xffffffffu, rb); m.gb = __reduce_add_sync(0xffffffffu, gb);
        m.xr = __reduce_add_sync(0xffffffffu, xr); m.xg = __reduce_add_sync(0xffffffffu, xg); m.xb = __reduce_add_sync(0xffffffffu, xb);
        u8_write_features_warp(m, norm_side, xout + (size_t)gi * CM_NODE_FEATURES, lane);
        return;
    }
    m.n = __reduce_add_sync(0xffffffffu, n);
    m.can = __reduce_add_sync(0xffffffffu, sc);
    m.per = __reduce_add_sync(0xffffffffu, per);
    m.rn = __reduce_add_sync(0xffffffffu, rn);
    m.sx = warp_sum_wide(sx); m.sy = warp_sum_wide(sy);
    m.r = warp_sum_wide(r1); m.g = warp_sum_wide(g1); m.b = warp_sum_wide(b1);
    m.rr = warp_sum_wide(rr); m.gg = warp_sum_wide(gg); m.bb = warp_sum_wide(b2);
    m.rg = warp_sum_wide(rg); m.rb = warp_sum_wide(rb); m.gb = warp_sum_wide(gb);
    m.xr = warp_sum_wide(xr); m.xg = warp_sum_wide(xg); m.xb = warp_sum_wide(xb);
    u8_write_features_warp(m, norm_side, xout + (size_t)gi * CM_NODE_FEATURES, lane);
}

__global__ void labels_widen_kernel(const short* __restrict__ in, long long n, int* __restrict__ out) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long n8 = n >> 3;   // 8 labels per thread per trip: one 16-byte load, two 16-byte stores
    for (; i < n8; i += stride) {
        const uint4 v = reinterpret_cast<const uint4*>(in)[i];
        const short* s = reinterpret_cast<const short*>(&v);
        reinterpret_cast<int4*>(out)[2 * i] = make_int4(s[0], s[1], s[2], s[3]);
        reinterpret_cast<int4*>(out)[2 * i + 1] = make_int4(s[4], s[5], s[6], s[7]);
    }
    i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < (n & 7)) out[(n8 << 3) + i] = in[(n8 << 3) + i];
}

}  // namespace cm

static int node_features_u8_impl(const uint8_t* image_u8, const void* labels_v, int label_bytes, const uint8_t* canny, int batch, int h,
                                   int w, int label_cap, const int32_t* node_ptr, const int32_t* node_label,
                                   const int32_t* node_bbox, float norm_side, float* x, int node_cap, void* stream) {
    const int32_t* labels = reinterpret_cast<const int32_t*>(labels_v);
    using namespace cm;
    CM_DEVICE_OR_RETURN();
    CM_REQUIRE(image_u8 && labels && canny && node_ptr && node_label && x, "cm_node_features_u8: NULL pointer");
    CM_REQUIRE(batch > 0 && batch <= 65535 && h > 0 && w > 0 && node_cap > 0 && norm_side > 0.f, "cm_node_features_u8: bad sizes");
    CM_REQUIRE((long long)h * w < (1 << 22), "cm_node_features_u8: images up to 4M pixels");
    if (node_bbox) {
        CM_TIMED("node_features_u8", as_stream(stream));
        if (label_bytes == 2)
            node_features_u8_bbox_kernel<short><<<ceil_div(node_cap, 8), 256, 0, as_stream(stream)>>>(
                image_u8, reinterpret_cast<const short*>(labels_v), canny, batch, h, w, node_ptr, node_label, node_bbox, norm_side, x, node_cap);
        else
            node_features_u8_bbox_kernel<int><<<ceil_div(node_cap, 8), 256, 0, as_stream(stream)>>>(
                image_u8, labels, canny, batch, h, w, node_ptr, node_label, node_bbox, norm_side, x, node_cap);
        return post_launch("cm_node_features_u8");
    }
    CM_REQUIRE(label_bytes == 4, "cm_node_features_u8_i16: int16 label maps go through the region kernel (pass node_bbox)");
    CM_REQUIRE(label_cap > 0, "cm_node_features_u8: the tile kernel (node_bbox == NULL) needs label_cap > 0");
    // a CTA's 32-bit sums of squares hold (pixels per CTA) * 255^2 < 2^32; its labels + table must fit in shared memory
    int ncl = 0;
    size_t smem = 0;
    for (int c = 4; c <= 8; c *= 2) {
        const long long rpc = (h + c - 1) / c;
        const size_t need = ((size_t)label_cap * U8_ROWS + 3) / 4 * 16 + (size_t)(rpc + 4) * w * 4;
        if (rpc * w <= 66051 && need <= 200 * 1024) {
            ncl = c;
            smem = need;
            break;
        }
    }
    CM_REQUIRE(ncl > 0, "cm_node_features_u8: image %dx%d with label_cap %d does not fit the integer tables; use cm_node_features "
               "on float images", h, w, label_cap);
    cudaStream_t st = as_stream(stream);
    static PerDeviceMax smem_set;
    if (smem_set.raise(smem)) {
        cudaFuncSetAttribute(node_features_u8_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(ncl, batch);
    cfg.blockDim = dim3(U8_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = ncl;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    {
        CM_TIMED("node_features_u8_tile", st);
        cudaLaunchKernelEx(&cfg, node_features_u8_kernel, image_u8, labels, canny, h, w, label_cap, node_ptr, node_label, norm_side, x,
                           node_cap);
    }
    return post_launch("cm_node_features_u8");
}

extern "C" int cm_labels_widen_i16(const int16_t* labels16, int64_t n, int32_t* labels32, void* stream) {
    using namespace cm;
    CM_DEVICE_OR_RETURN();
    CM_REQUIRE(labels16 && labels32 && n >= 0, "cm_labels_widen_i16: bad arguments");
    CM_REQUIRE(((uintptr_t)labels16 & 15) == 0 && ((uintptr_t)labels32 & 15) == 0, "cm_labels_widen_i16: buffers must be 16-byte aligned");
    if (n == 0) return CM_OK;
    cudaStream_t st = as_stream(stream);
    const int blocks = (int)std::min<long long>((n / 8 + 255) / 256 + 1, 4LL * num_sms());
    {
        CM_TIMED("labels_widen", st);
        labels_widen_kernel<<<blocks, 256, 0, st>>>(labels16, (long long)n, labels32);
    }
    return post_launch("cm_labels_widen_i16");
}

extern "C" int cm_node_features_u8(const uint8_t* image_u8, const int32_t* labels, const uint8_t* canny, int batch, int h,
                                   int w, int label_cap, const int32_t* node_ptr, const int32_t* node_label,
                                   const int32_t* node_bbox, float norm_side, float* x, int node_cap, void* stream) {
    return node_features_u8_impl(image_u8, labels, 4, canny, batch, h, w, label_cap, node_ptr, node_label, node_bbox, norm_side, x, node_cap, stream);
}

extern "C" int cm_node_features_u8_i16(const uint8_t* image_u8, const int16_t* labels, const uint8_t* canny, int batch, int h,
                                   int w, int label_cap, const int32_t* node_ptr, const int32_t* node_label,
                                   const int32_t* node_bbox, float norm_side, float* x, int node_cap, void* stream) {
    return node_features_u8_impl(image_u8, labels, 2, canny, batch, h, w, label_cap, node_ptr, node_label, node_bbox, norm_side, x, node_cap, stream);
}
