// A3 + A4: the 15 region features and the Gaussian-kernel edge weights; plus the general
// edge_index -> CSR-by-target conversion used when a graph does not come from cm_rag_build.
//
// Reference: models/region_graph/extract_rg_embeddings.py:154-211 (features, computed there in float64 and
// stored as float32) and :227-235 (weights, float32 arithmetic).
//
// Design: one warp per region scans the region's bounding box grown by two pixels (the reach of the
// reference's binary_dilation(iterations=2)).  Member pixels feed the moment sums; non-member pixels test
// their 4-neighbourhood (perimeter, :177) and L1<=2 neighbourhood (contrast ring, :185-186) for the region's
// label.  No float atomics: every sum is a fixed-order warp tree over float64 partials, so the result is
// deterministic and agrees with numpy's float64 arithmetic to ~1e-15 before the final float32 rounding.
// The batch (38 MB at B=32) is L2-resident, so HBM sees each input byte once.
#include <cooperative_groups.h>
#include <limits.h>

#include "common.cuh"

namespace cg = cooperative_groups;

namespace cm {

__device__ __forceinline__ int find_image(const int* __restrict__ node_ptr, int batch, int g) {
    int lo = 0, hi = batch;   // largest b with node_ptr[b] <= g
    while (hi - lo > 1) {
        int mid = (lo + hi) >> 1;
        if (node_ptr[mid] <= g) lo = mid; else hi = mid;
    }
    return lo;
}

// block 256 = 8 warps, one region per warp
__global__ void __launch_bounds__(256) node_features_kernel(const float* __restrict__ image, const int* __restrict__ labels,
                                                            const uint8_t* __restrict__ canny, int batch, int H, int W,
                                                            const int* __restrict__ node_ptr, const int* __restrict__ node_label,
                                                            const int* __restrict__ node_bbox, float norm_side,
                                                            float* __restrict__ xout, int node_cap) {
    const int lane = threadIdx.x & 31;
    const int g = blockIdx.x * 8 + (threadIdx.x >> 5);
    const int total = min(node_ptr[batch], node_cap);
    if (g >= total) return;
    const int b = find_image(node_ptr, batch, g);
    const int L = node_label[g];
    const int bx0 = node_bbox[4 * g + 0], by0 = node_bbox[4 * g + 1], bx1 = node_bbox[4 * g + 2], by1 = node_bbox[4 * g + 3];
    const int wx0 = max(bx0 - 2, 0), wy0 = max(by0 - 2, 0), wx1 = min(bx1 + 2, W - 1), wy1 = min(by1 + 2, H - 1);
    const int ww = wx1 - wx0 + 1, wh = wy1 - wy0 + 1;
    const int* lab = labels + (size_t)b * H * W;
    const float* img = image + (size_t)b * H * W * 3;
    const uint8_t* can = canny + (size_t)b * H * W;

    double s1[3] = {0, 0, 0}, s2[3] = {0, 0, 0}, sg = 0, sg2 = 0, rs[3] = {0, 0, 0};
    int cnt = 0, sx = 0, sy = 0, sc = 0, perim = 0, rcnt = 0;
    const int npx = ww * wh;
    for (int i = lane; i < npx; i += 32) {
        const int yy = i / ww;
        const int y = wy0 + yy, x = wx0 + (i - yy * ww);
        const size_t p = (size_t)y * W + x;
        if (lab[p] == L) {
            const float r = img[3 * p], gch = img[3 * p + 1], bl = img[3 * p + 2];
            const double dr = r, dg = gch, db = bl;
            s1[0] += dr; s1[1] += dg; s1[2] += db;
            s2[0] += dr * dr; s2[1] += dg * dg; s2[2] += db * db;
            const double gray = dr * 0.2989 + dg * 0.5870 + db * 0.1140;   // :151
            sg += gray;
            sg2 += gray * gray;
            ++cnt;
            sx += x;
            sy += y;
            sc += can[p] != 0;
        } else {
            // inside the 4-connected cross dilation of the region?
            bool near1 = (x > 0 && lab[p - 1] == L) || (x + 1 < W && lab[p + 1] == L) ||
                         (y > 0 && lab[p - W] == L) || (y + 1 < H && lab[p + W] == L);
            bool near2 = near1;
            if (!near2) {
                near2 = (x > 1 && lab[p - 2] == L) || (x + 2 < W && lab[p + 2] == L) ||
                        (y > 1 && lab[p - 2 * W] == L) || (y + 2 < H && lab[p + 2 * W] == L) ||
                        (x > 0 && y > 0 && lab[p - W - 1] == L) || (x + 1 < W && y > 0 && lab[p - W + 1] == L) ||
                        (x > 0 && y + 1 < H && lab[p + W - 1] == L) || (x + 1 < W && y + 1 < H && lab[p + W + 1] == L);
            }
            perim += near1;
            if (near2) {
                ++rcnt;
                rs[0] += (double)img[3 * p];
                rs[1] += (double)img[3 * p + 1];
                rs[2] += (double)img[3 * p + 2];
            }
        }
    }
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        s1[c] = warp_sum(s1[c]);
        s2[c] = warp_sum(s2[c]);
        rs[c] = warp_sum(rs[c]);
    }
    sg = warp_sum(sg);
    sg2 = warp_sum(sg2);
    cnt = warp_sum(cnt);
    sx = warp_sum(sx);
    sy = warp_sum(sy);
    sc = warp_sum(sc);
    perim = warp_sum(perim);
    rcnt = warp_sum(rcnt);
    if (lane != 0) return;
    float* o = xout + (size_t)g * CM_NODE_FEATURES;
    const double n = (double)cnt;
    double mean[3], contrast2 = 0.0;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        mean[c] = s1[c] / n;
        double var = s2[c] / n - mean[c] * mean[c];
        o[c] = (float)mean[c];
        o[3 + c] = (float)sqrt(var > 0.0 ? var : 0.0);   // population std, :165
        if (rcnt > 0) {
            double d = mean[c] - rs[c] / (double)rcnt;
            contrast2 += d * d;
        }
    }
    const double gm = sg / n;
    double gvar = sg2 / n - gm * gm;
    gvar = gvar > 0.0 ? gvar : 0.0;
    const double ns = (double)norm_side;
    o[6] = (float)gm;
    o[7] = (float)sqrt(gvar);
    o[8] = (float)(((double)sx / n) / ns);    // centroid x, :173
    o[9] = (float)(((double)sy / n) / ns);    // centroid y, :172
    o[10] = (float)(n / (ns * ns));           // :174
    const double per = (double)perim;
    o[11] = (float)((per * per) / (4.0 * 3.14159265358979323846 * n + 1e-10));   // :179
    o[12] = (float)sqrt(contrast2);            // :190 (0 when the ring is empty)
    o[13] = (float)((double)sc / n);           // :182
    o[14] = (float)gvar;                       // :193
}

// ---------------------------------------------------------------------------------------------------------------
// Pixel-parallel variant (the default): every pixel is read once.  A cluster of 4 CTAs owns one image (a quarter of
// the rows each).  The CTA's label rows (+2 halo rows) sit in shared memory next to per-label accumulators; a lane
// walks a strip of 8 consecutive pixels, keeps the moment sums of the current label run in fp64 registers and adds
// them to the label's accumulators as 64-bit FIXED-POINT integers (2^-40 for sums, 2^-44 for sums of squares) with
// shared-memory integer atomics -- integer addition is associative, so the result is independent of scheduling and
// within ~1e-11 of the fp64 sums numpy forms.  Non-member contributions (perimeter :177, contrast ring :185-186)
// are emitted by the OUTSIDE pixel: it looks at its 12 L1<=2 neighbours and contributes once to every distinct
// foreign label (the 4-neighbours come first, so a first occurrence among them is a perimeter pixel).  The four
// partial tables are then summed over distributed shared memory and one thread per node writes the 15 features.
constexpr int NF_THREADS = 256, NF_CLUSTER = 4, NF_STRIP = 8, NF_NONE = INT_MIN;
constexpr double NF_S1 = 1099511627776.0, NF_S2 = 17592186044416.0;   // 2^40, 2^44

struct NfTables {
    unsigned long long* a64;   // [11][cap]: s1 r,g,b | gray | s2 r,g,b | gray^2 | ring r,g,b
    unsigned int* a32;         // [6][cap]:  count | sum x | sum y | canny | perimeter | ring count
};

__device__ __forceinline__ unsigned long long nf_fix(double v, double scale) { return (unsigned long long)__double2ll_rn(v * scale); }

__global__ void __cluster_dims__(NF_CLUSTER, 1, 1) __launch_bounds__(NF_THREADS)
node_features_tile_kernel(const float* __restrict__ image, const int* __restrict__ labels, const uint8_t* __restrict__ canny, int H, int W,
                          int cap, const int* __restrict__ node_ptr, const int* __restrict__ node_label, float norm_side,
                          float* __restrict__ xout, int node_cap) {
    extern __shared__ __align__(16) unsigned char nf_smem[];
    cg::cluster_group cluster = cg::this_cluster();
    const int rank = blockIdx.x, b = blockIdx.y, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int rpc = (H + NF_CLUSTER - 1) / NF_CLUSTER, y0 = rank * rpc, y1 = min(H, y0 + rpc), th = rpc + 4;
    NfTables t;
    t.a64 = reinterpret_cast<unsigned long long*>(nf_smem);
    t.a32 = reinterpret_cast<unsigned int*>(t.a64 + 11 * (size_t)cap);
    int* lab_s = reinterpret_cast<int*>(t.a32 + 6 * (size_t)cap);
    const int* lab = labels + (size_t)b * H * W;
    const float* img = image + (size_t)b * H * W * 3;
    const uint8_t* can = canny + (size_t)b * H * W;

    for (int i = tid; i < 11 * cap; i += NF_THREADS) t.a64[i] = 0ull;
    for (int i = tid; i < 6 * cap; i += NF_THREADS) t.a32[i] = 0u;
    for (int i = tid; i < th * (W / 4); i += NF_THREADS) {   // label rows y0-2 .. y0+rpc+1; rows outside the image -> NF_NONE
        const int r = i / (W / 4), c4 = (i - r * (W / 4)) * 4, y = y0 - 2 + r;
        int4 v = make_int4(NF_NONE, NF_NONE, NF_NONE, NF_NONE);
        if (y >= 0 && y < H) v = *reinterpret_cast<const int4*>(lab + (size_t)y * W + c4);
        *reinterpret_cast<int4*>(lab_s + r * W + c4) = v;
    }
    __syncthreads();

    auto flush = [&](int L, const double* s1, const double* s2, double sg, double sg2, int cnt, int sx, int y, int sc) {
        if (cnt == 0 || L < 0 || L >= cap) return;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            atomicAdd(t.a64 + (size_t)c * cap + L, nf_fix(s1[c], NF_S1));
            atomicAdd(t.a64 + (size_t)(4 + c) * cap + L, nf_fix(s2[c], NF_S2));
        }
        atomicAdd(t.a64 + (size_t)3 * cap + L, nf_fix(sg, NF_S1));
        atomicAdd(t.a64 + (size_t)7 * cap + L, nf_fix(sg2, NF_S2));
        atomicAdd(t.a32 + L, (unsigned)cnt);
        atomicAdd(t.a32 + (size_t)cap + L, (unsigned)sx);
        atomicAdd(t.a32 + (size_t)2 * cap + L, (unsigned)(cnt * y));
        if (sc) atomicAdd(t.a32 + (size_t)3 * cap + L, (unsigned)sc);
    };

    for (int y = y0 + warp; y < y1; y += NF_THREADS / 32) {
        const int* row = lab_s + (y - y0 + 2) * W;
        for (int x0 = lane * NF_STRIP; x0 < W; x0 += 32 * NF_STRIP) {
            float px[3 * NF_STRIP];
            const float4* ip = reinterpret_cast<const float4*>(img + ((size_t)y * W + x0) * 3);
#pragma unroll
            for (int i = 0; i < 6; ++i) {
                const float4 v = __ldg(ip + i);
                px[4 * i] = v.x; px[4 * i + 1] = v.y; px[4 * i + 2] = v.z; px[4 * i + 3] = v.w;
            }
            const uint2 cw = __ldg(reinterpret_cast<const uint2*>(can + (size_t)y * W + x0));
            const unsigned long long cbits = ((unsigned long long)cw.y << 32) | cw.x;
            double s1[3] = {0, 0, 0}, s2[3] = {0, 0, 0}, sg = 0, sg2 = 0;
            int curL = NF_NONE, cnt = 0, sx = 0, sc = 0;
#pragma unroll
            for (int i = 0; i < NF_STRIP; ++i) {
                const int x = x0 + i, L = row[x];
                const double dr = px[3 * i], dg = px[3 * i + 1], db = px[3 * i + 2];
                if (L != curL) {
                    flush(curL, s1, s2, sg, sg2, cnt, sx, y, sc);
                    curL = L; cnt = 0; sx = 0; sc = 0; sg = 0; sg2 = 0;
                    s1[0] = s1[1] = s1[2] = 0; s2[0] = s2[1] = s2[2] = 0;
                }
                s1[0] += dr; s1[1] += dg; s1[2] += db;
                s2[0] += dr * dr; s2[1] += dg * dg; s2[2] += db * db;
                const double gray = dr * 0.2989 + dg * 0.5870 + db * 0.1140;   // :151
                sg += gray;
                sg2 += gray * gray;
                ++cnt;
                sx += x;
                sc += ((cbits >> (8 * i)) & 0xffull) != 0ull;
                // ---- what this pixel contributes to its foreign neighbours' regions
                int v[12];
                v[0] = x > 0 ? row[x - 1] : NF_NONE;
                v[1] = x + 1 < W ? row[x + 1] : NF_NONE;
                v[2] = row[x - W];
                v[3] = row[x + W];
                v[4] = x > 1 ? row[x - 2] : NF_NONE;
                v[5] = x + 2 < W ? row[x + 2] : NF_NONE;
                v[6] = row[x - 2 * W];
                v[7] = row[x + 2 * W];
                v[8] = x > 0 ? row[x - W - 1] : NF_NONE;
                v[9] = x + 1 < W ? row[x - W + 1] : NF_NONE;
                v[10] = x > 0 ? row[x + W - 1] : NF_NONE;
                v[11] = x + 1 < W ? row[x + W + 1] : NF_NONE;
                bool any = false;
#pragma unroll
                for (int k = 0; k < 12; ++k) any |= (v[k] != L) & (v[k] != NF_NONE);
                if (any) {
#pragma unroll
                    for (int k = 0; k < 12; ++k) {
                        const int u = v[k];
                        bool first = (u != L) & (u >= 0) & (u < cap);
#pragma unroll
                        for (int j = 0; j < k; ++j) first &= v[j] != u;
                        if (first) {
                            atomicAdd(t.a32 + (size_t)5 * cap + u, 1u);
                            atomicAdd(t.a64 + (size_t)8 * cap + u, nf_fix(dr, NF_S1));
                            atomicAdd(t.a64 + (size_t)9 * cap + u, nf_fix(dg, NF_S1));
                            atomicAdd(t.a64 + (size_t)10 * cap + u, nf_fix(db, NF_S1));
                            if (k < 4) atomicAdd(t.a32 + (size_t)4 * cap + u, 1u);
                        }
                    }
                }
            }
            flush(curL, s1, s2, sg, sg2, cnt, sx, y, sc);
        }
    }
    cluster.sync();

    // ---- one thread per node of the image: sum the four partial tables (DSMEM), then the 15 features
    const int n0 = node_ptr[b], nb = min(node_ptr[b + 1], node_cap) - n0;
    const unsigned long long* r64[NF_CLUSTER];
    const unsigned int* r32[NF_CLUSTER];
#pragma unroll
    for (int r = 0; r < NF_CLUSTER; ++r) {
        r64[r] = cluster.map_shared_rank(t.a64, r);
        r32[r] = cluster.map_shared_rank(t.a32, r);
    }
    for (int i = rank * NF_THREADS + tid; i < nb; i += NF_CLUSTER * NF_THREADS) {
        const int g = n0 + i, L = node_label[g];
        unsigned long long q64[11];
        unsigned int q32[6];
#pragma unroll
        for (int k = 0; k < 11; ++k) q64[k] = 0ull;
#pragma unroll
        for (int k = 0; k < 6; ++k) q32[k] = 0u;
        if (L >= 0 && L < cap) {
#pragma unroll
            for (int r = 0; r < NF_CLUSTER; ++r) {
#pragma unroll
                for (int k = 0; k < 11; ++k) q64[k] += r64[r][(size_t)k * cap + L];
#pragma unroll
                for (int k = 0; k < 6; ++k) q32[k] += r32[r][(size_t)k * cap + L];
            }
        }
        float* o = xout + (size_t)g * CM_NODE_FEATURES;
        const double n = (double)q32[0];
        const int rcnt = (int)q32[5];
        double mean[3], contrast2 = 0.0;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            mean[c] = ((double)(long long)q64[c] / NF_S1) / n;
            double var = ((double)(long long)q64[4 + c] / NF_S2) / n - mean[c] * mean[c];
            o[c] = (float)mean[c];
            o[3 + c] = (float)sqrt(var > 0.0 ? var : 0.0);   // population std, :165
            if (rcnt > 0) {
                const double d = mean[c] - ((double)(long long)q64[8 + c] / NF_S1) / (double)rcnt;
                contrast2 += d * d;
            }
        }
        const double gm = ((double)(long long)q64[3] / NF_S1) / n;
        double gvar = ((double)(long long)q64[7] / NF_S2) / n - gm * gm;
        gvar = gvar > 0.0 ? gvar : 0.0;
        const double ns = (double)norm_side;
        o[6] = (float)gm;
        o[7] = (float)sqrt(gvar);
        o[8] = (float)(((double)q32[1] / n) / ns);    // centroid x, :173
        o[9] = (float)(((double)q32[2] / n) / ns);    // centroid y, :172
        o[10] = (float)(n / (ns * ns));               // :174
        const double per = (double)q32[4];
        o[11] = (float)((per * per) / (4.0 * 3.14159265358979323846 * n + 1e-10));   // :179
        o[12] = (float)sqrt(contrast2);               // :190 (0 when the ring is empty)
        o[13] = (float)((double)q32[3] / n);          // :182
        o[14] = (float)gvar;                          // :193
    }
    cluster.sync();   // peers may still be reading this CTA's tables
}

static size_t nf_tile_smem(int H, int W, int cap) {
    const int rpc = (H + NF_CLUSTER - 1) / NF_CLUSTER;
    return (size_t)cap * (11 * 8 + 6 * 4) + (size_t)(rpc + 4) * W * 4;
}

// one thread per node row; weights for every directed CSR entry
__global__ void edge_weights_kernel(const float* __restrict__ x, const int* __restrict__ row_ptr, const int* __restrict__ col,
                                    const int* __restrict__ n_nodes_dev, int node_cap, float* __restrict__ w) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int n = min(*n_nodes_dev, node_cap);
    if (i >= n) return;
    const float* xi = x + (size_t)i * CM_NODE_FEATURES;
    const float r = xi[0], g = xi[1], b = xi[2], t = xi[6], c = xi[12];
    for (int e = row_ptr[i]; e < row_ptr[i + 1]; ++e) {
        const float* xj = x + (size_t)col[e] * CM_NODE_FEATURES;
        const float dr = __fsub_rn(r, xj[0]), dg = __fsub_rn(g, xj[1]), db = __fsub_rn(b, xj[2]);
        // float32 sum of squares without fma contraction, like numpy's elementwise ops (:229)
        const float ss = __fadd_rn(__fadd_rn(__fmul_rn(dr, dr), __fmul_rn(dg, dg)), __fmul_rn(db, db));
        const float color = __fsqrt_rn(ss);
        const float tex = fabsf(__fsub_rn(t, xj[6]));
        const float con = fabsf(__fsub_rn(c, xj[12]));
        const float e1 = expf(-__fdiv_rn(color, 0.15f));
        const float e2 = expf(-__fdiv_rn(tex, 0.08f));
        const float e3 = expf(-__fdiv_rn(con, 0.1f));
        w[e] = __fmul_rn(__fmul_rn(e1, e2), e3);
    }
}

// ---------------------------------------------------------------- edge_index -> CSR by target
__global__ void csr_count_kernel(const long long* __restrict__ src, const long long* __restrict__ dst, int n_nodes,
                                 int n_edges, int* deg, int* self_idx, uint32_t* status) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n_edges) return;
    const long long s = src[e], d = dst[e];
    if (s < 0 || d < 0 || s >= n_nodes || d >= n_nodes) {
        atomicOr(status, CM_STATUS_NODE_OVERFLOW);
        return;
    }
    if (s == d) atomicMax(self_idx + d, e + 1);
    else atomicAdd(deg + d, 1);
}

// single CTA exclusive scan of deg[0..n) -> row_ptr[0..n]; also resets deg to 0 for reuse as a cursor
__global__ void __launch_bounds__(1024) csr_scan_kernel(int* deg, int n, int* row_ptr) {
    __shared__ int swarp[33];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int carry = 0;
    for (int base = 0; base < n; base += 1024) {
        const int i = base + threadIdx.x;
        const int v = i < n ? deg[i] : 0;
        int inc = warp_scan_incl(v, lane);
        if (lane == 31) swarp[warp] = inc;
        __syncthreads();
        if (warp == 0) {
            int t = swarp[lane];
            int ti = warp_scan_incl(t, lane);
            swarp[lane] = ti - t;
            if (lane == 31) swarp[32] = ti;
        }
        __syncthreads();
        if (i < n) {
            row_ptr[i] = carry + swarp[warp] + inc - v;
            deg[i] = 0;
        }
        carry += swarp[32];
        __syncthreads();
    }
    if (threadIdx.x == 0) row_ptr[n] = carry;
}

__global__ void csr_fill_kernel(const long long* __restrict__ src, const long long* __restrict__ dst, int n_nodes, int n_edges,
                                const int* __restrict__ row_ptr, int* cursor, int* tmp_edge) {
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n_edges) return;
    const long long s = src[e], d = dst[e];
    if (s < 0 || d < 0 || s >= n_nodes || d >= n_nodes || s == d) return;
    const int p = atomicAdd(cursor + d, 1);
    tmp_edge[row_ptr[d] + p] = e;
}

// one thread per row: order the row's entries by input edge id (deterministic), then resolve
__global__ void csr_finish_kernel(const long long* __restrict__ src, const float* __restrict__ edge_w, int n_nodes,
                                  const int* __restrict__ row_ptr, int* tmp_edge, const int* __restrict__ self_idx,
                                  int* __restrict__ col, float* __restrict__ w, float* __restrict__ self_w) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_nodes) return;
    const int a = row_ptr[i], z = row_ptr[i + 1];
    for (int p = a + 1; p < z; ++p) {   // insertion sort; rows are short
        int v = tmp_edge[p], q = p - 1;
        while (q >= a && tmp_edge[q] > v) {
            tmp_edge[q + 1] = tmp_edge[q];
            --q;
        }
        tmp_edge[q + 1] = v;
    }
    for (int p = a; p < z; ++p) {
        const int e = tmp_edge[p];
        col[p] = (int)src[e];
        if (w) w[p] = edge_w ? edge_w[e] : 1.0f;
    }
    if (self_w) {
        const int si = self_idx[i];
        self_w[i] = (si > 0 && edge_w) ? edge_w[si - 1] : 1.0f;
    }
}

}  // namespace cm

extern "C" int cm_node_features(const float* image, const int32_t* labels, const uint8_t* canny, int batch, int h,
                                int w, int label_cap, const int32_t* node_ptr, const int32_t* node_label,
                                const int32_t* node_bbox, float norm_side, float* x, int node_cap, void* stream) {
    using namespace cm;
    CM_DEVICE_OR_RETURN();
    CM_REQUIRE(image && labels && canny && node_ptr && node_label && node_bbox && x, "cm_node_features: NULL pointer");
    CM_REQUIRE(batch > 0 && h > 0 && w > 0 && node_cap > 0 && norm_side > 0.f, "cm_node_features: bad sizes");
    cudaStream_t st = as_stream(stream);
    const size_t smem = label_cap > 0 ? nf_tile_smem(h, w, label_cap) : 0;
    if (label_cap > 0 && w % 8 == 0 && smem <= 200 * 1024 && batch <= 65535) {
        static size_t smem_set = 0;
        if (smem > smem_set) {
            cudaFuncSetAttribute(node_features_tile_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            smem_set = smem;
        }
        {
            CM_TIMED("node_features", st);
            node_features_tile_kernel<<<dim3(NF_CLUSTER, batch), NF_THREADS, smem, st>>>(image, labels, canny, h, w, label_cap, node_ptr,
                                                                                        node_label, norm_side, x, node_cap);
        }
        return post_launch("cm_node_features");
    }
    // label tables that do not fit in shared memory (or odd widths): one warp per region over its bounding box
    {
        CM_TIMED("node_features_bbox", st);
        node_features_kernel<<<ceil_div(node_cap, 8), 256, 0, st>>>(image, labels, canny, batch, h, w, node_ptr, node_label, node_bbox,
                                                                   norm_side, x, node_cap);
    }
    return post_launch("cm_node_features");
}

extern "C" int cm_edge_weights(const float* x, const int32_t* csr_row_ptr, const int32_t* csr_col,
                               const int32_t* n_nodes_dev, int node_cap, float* edge_w, void* stream) {
    using namespace cm;
    CM_DEVICE_OR_RETURN();
    CM_REQUIRE(x && csr_row_ptr && csr_col && n_nodes_dev && edge_w, "cm_edge_weights: NULL pointer");
    CM_REQUIRE(node_cap > 0, "cm_edge_weights: node_cap must be positive");
    {
        CM_TIMED("edge_weights", as_stream(stream));
        edge_weights_kernel<<<ceil_div(node_cap, 128), 128, 0, as_stream(stream)>>>(x, csr_row_ptr, csr_col, n_nodes_dev,
                                                                                   node_cap, edge_w);
    }
    return post_launch("cm_edge_weights");
}

namespace cm {
struct CsrWs {
    int* deg;
    int* self_idx;
    int* tmp_edge;
    size_t zero_bytes, total;
};
static CsrWs carve_csr(void* ws, int n, int e) {
    Carver c(ws);
    CsrWs r;
    r.deg = c.take<int>((size_t)n + 1);
    r.self_idx = c.take<int>((size_t)n + 1);
    r.zero_bytes = align_up(c.off, 256);
    r.tmp_edge = c.take<int>((size_t)e + 1);
    r.total = align_up(c.off, 256);
    return r;
}
}  // namespace cm

extern "C" size_t cm_csr_workspace_bytes(int n_nodes, int n_edges) {
    if (n_nodes < 0 || n_edges < 0) return 0;
    return cm::carve_csr(nullptr, n_nodes, n_edges).total;
}

extern "C" int cm_csr_from_edge_index(const int64_t* edge_index, int64_t row_stride, const float* edge_w, int n_nodes,
                                      int n_edges, void* workspace, size_t workspace_bytes, int32_t* csr_row_ptr,
                                      int32_t* csr_col, float* csr_w, float* self_w, uint32_t* status, void* stream) {
    using namespace cm;
    CM_DEVICE_OR_RETURN();
    CM_REQUIRE(workspace && csr_row_ptr && status, "cm_csr_from_edge_index: NULL pointer");
    CM_REQUIRE(n_nodes >= 0 && n_edges >= 0, "cm_csr_from_edge_index: negative size");
    CM_REQUIRE(n_edges == 0 || (edge_index && csr_col && row_stride >= n_edges), "cm_csr_from_edge_index: bad edge_index");
    CsrWs ws = carve_csr(workspace, n_nodes, n_edges);
    if (workspace_bytes < ws.total) {
        set_error("cm_csr_from_edge_index: workspace %zu < %zu bytes", workspace_bytes, ws.total);
        return CM_ERR_WORKSPACE;
    }
    cudaStream_t st = as_stream(stream);
    cudaMemsetAsync(workspace, 0, ws.zero_bytes, st);
    const long long* src = reinterpret_cast<const long long*>(edge_index);
    const long long* dst = src + row_stride;
    if (n_edges > 0)
        {
            CM_TIMED("csr_count", st);
            csr_count_kernel<<<ceil_div(n_edges, 256), 256, 0, st>>>(src, dst, n_nodes, n_edges, ws.deg, ws.self_idx, status);
        }
    {
        CM_TIMED("csr_scan", st);
        csr_scan_kernel<<<1, 1024, 0, st>>>(ws.deg, n_nodes, csr_row_ptr);
    }
    if (n_edges > 0)
        {
            CM_TIMED("csr_fill", st);
            csr_fill_kernel<<<ceil_div(n_edges, 256), 256, 0, st>>>(src, dst, n_nodes, n_edges, csr_row_ptr, ws.deg, ws.tmp_edge);
        }
    if (n_nodes > 0)
        {
            CM_TIMED("csr_finish", st);
            csr_finish_kernel<<<ceil_div(n_nodes, 128), 128, 0, st>>>(src, edge_w, n_nodes, csr_row_ptr, ws.tmp_edge, ws.self_idx,
                                                                     csr_col, csr_w, self_w);
        }
    return post_launch("cm_csr_from_edge_index");
}
