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
constexpr int NF_THREADS = 512, NF_CLUSTER = 4, NF_NONE = INT_MIN;
constexpr double NF_S1 = 1099511627776.0, NF_S2 = 17592186044416.0;   // 2^40, 2^44

// Per-label accumulators, all 32-bit so that every update is a native shared-memory ATOMS.ADD (64-bit shared atomics
// are CAS loops): the 11 fixed-point sums (s1 r,g,b | gray | s2 r,g,b | gray^2 | ring r,g,b) are kept as three 16-bit
// limbs each -- limb sums stay below 2^32 for up to 65535 pixels per CTA -- rows 3q .. 3q+2 of a32; rows 33..38 hold
// count | sum x | sum y | canny | perimeter | ring count.
constexpr int NF_ROWS = 39, NF_CNT = 33;

__device__ __forceinline__ unsigned long long nf_fix(double v, double scale) { return (unsigned long long)__double2ll_rn(v * scale); }

__global__ void __cluster_dims__(NF_CLUSTER, 1, 1) __launch_bounds__(NF_THREADS)
node_features_tile_kernel(const float* __restrict__ image, const int* __restrict__ labels, const uint8_t* __restrict__ canny, int H, int W,
                          int cap, const int* __restrict__ node_ptr, const int* __restrict__ node_label, float norm_side,
                          float* __restrict__ xout, int node_cap) {
    extern __shared__ __align__(16) unsigned char nf_smem[];
    cg::cluster_group cluster = cg::this_cluster();
    const int rank = blockIdx.x, b = blockIdx.y, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int rpc = (H + NF_CLUSTER - 1) / NF_CLUSTER, y0 = rank * rpc, y1 = min(H, y0 + rpc), th = rpc + 4;
    unsigned int* tab = reinterpret_cast<unsigned int*>(nf_smem);   // [NF_ROWS][cap]
    int* lab_s = reinterpret_cast<int*>(tab + (NF_ROWS * (size_t)cap + 3) / 4 * 4);   // 16-byte aligned for the int4 row copies
    const int* lab = labels + (size_t)b * H * W;
    const float* img = image + (size_t)b * H * W * 3;
    const uint8_t* can = canny + (size_t)b * H * W;

    for (int i = tid; i < NF_ROWS * cap; i += NF_THREADS) tab[i] = 0u;
    for (int i = tid; i < th * (W / 4); i += NF_THREADS) {   // label rows y0-2 .. y0+rpc+1; rows outside the image -> NF_NONE
        const int r = i / (W / 4), c4 = (i - r * (W / 4)) * 4, y = y0 - 2 + r;
        int4 v = make_int4(NF_NONE, NF_NONE, NF_NONE, NF_NONE);
        if (y >= 0 && y < H) v = *reinterpret_cast<const int4*>(lab + (size_t)y * W + c4);
        *reinterpret_cast<int4*>(lab_s + r * W + c4) = v;
    }
    __syncthreads();

    // ---- helpers -------------------------------------------------------------------------------------------
    auto add64 = [&](int q, int L, unsigned long long v) {   // quantity q of label L += v, as three 16-bit limbs
        atomicAdd(tab + (size_t)(3 * q) * cap + L, (unsigned)(v & 0xffffull));
        atomicAdd(tab + (size_t)(3 * q + 1) * cap + L, (unsigned)((v >> 16) & 0xffffull));
        atomicAdd(tab + (size_t)(3 * q + 2) * cap + L, (unsigned)(v >> 32));
    };
    // Lanes that hold partial sums for the same label in neighbouring columns are combined with a segmented
    // inclusive scan (shuffles, no divergence); only the last lane of every run touches the table.
    auto seg_begin = [&](int key, bool& head, bool& tail) {
        const int up = __shfl_up_sync(0xffffffffu, key, 1), dn = __shfl_down_sync(0xffffffffu, key, 1);
        head = lane == 0 || up != key;
        tail = lane == 31 || dn != key;
    };
    auto seg_scan64 = [&](unsigned long long v, bool head) -> unsigned long long {
        bool f = head;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const unsigned long long vu = __shfl_up_sync(0xffffffffu, v, d);
            const bool fu = __shfl_up_sync(0xffffffffu, (int)f, d) != 0;
            if (lane >= d && !f) {
                v += vu;
                f = fu;
            }
        }
        return v;
    };
    auto seg_scan32 = [&](unsigned v, bool head) -> unsigned {
        bool f = head;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const unsigned vu = __shfl_up_sync(0xffffffffu, v, d);
            const bool fu = __shfl_up_sync(0xffffffffu, (int)f, d) != 0;
            if (lane >= d && !f) {
                v += vu;
                f = fu;
            }
        }
        return v;
    };
    struct Member {   // moment sums of one label inside this lane's column segment
        int L;
        unsigned long long q[8];   // s1 r,g,b | gray | s2 r,g,b | gray^2 (fixed point)
        unsigned cs, sx, sy;       // count | canny << 16, sum x, sum y
    };
    struct Ring {     // what this lane's pixels contribute to one foreign label
        int L;
        unsigned long long q[3];   // rgb (fixed point)
        unsigned cp;               // ring count | perimeter << 16
    };
    auto flush_member = [&](Member& m) {
        if (!__any_sync(0xffffffffu, m.L != NF_NONE)) return;
        bool head, tail;
        seg_begin(m.L, head, tail);
        unsigned long long tq[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) tq[k] = seg_scan64(m.q[k], head);
        const unsigned cs = seg_scan32(m.cs, head), sx = seg_scan32(m.sx, head), sy = seg_scan32(m.sy, head);
        if (tail && m.L != NF_NONE) {
#pragma unroll
            for (int k = 0; k < 8; ++k) add64(k, m.L, tq[k]);
            atomicAdd(tab + (size_t)NF_CNT * cap + m.L, cs & 0xffffu);
            atomicAdd(tab + (size_t)(NF_CNT + 1) * cap + m.L, sx);
            atomicAdd(tab + (size_t)(NF_CNT + 2) * cap + m.L, sy);
            if (cs >> 16) atomicAdd(tab + (size_t)(NF_CNT + 3) * cap + m.L, cs >> 16);
        }
        m.L = NF_NONE;
#pragma unroll
        for (int k = 0; k < 8; ++k) m.q[k] = 0ull;
        m.cs = m.sx = m.sy = 0u;
    };
    auto flush_ring = [&](Ring& r) {
        if (!__any_sync(0xffffffffu, r.L != NF_NONE)) return;
        bool head, tail;
        seg_begin(r.L, head, tail);
        unsigned long long tq[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) tq[k] = seg_scan64(r.q[k], head);
        const unsigned cp = seg_scan32(r.cp, head);
        if (tail && r.L != NF_NONE) {
#pragma unroll
            for (int k = 0; k < 3; ++k) add64(8 + k, r.L, tq[k]);
            atomicAdd(tab + (size_t)(NF_CNT + 5) * cap + r.L, cp & 0xffffu);
            if (cp >> 16) atomicAdd(tab + (size_t)(NF_CNT + 4) * cap + r.L, cp >> 16);
        }
        r.L = NF_NONE;
        r.q[0] = r.q[1] = r.q[2] = 0ull;
        r.cp = 0u;
    };

    // ---- a warp owns bands of 8 adjacent rows; a lane walks DOWN its column, so the pixels it sees belong to one or
    // two labels (regions are taller than a few pixels) and their sums stay in registers until the band ends
    const float S1f = 1099511627776.0f;   // 2^40: float x 2^40 is exact, so is its conversion to a 64-bit integer
    const int n_bands = (y1 - y0 + 7) / 8;
    // work item = (band, group of 32 columns), dealt round-robin to the 16 warps
    const int n_xg = (W + 31) / 32;
    for (int item = warp; item < n_bands * n_xg; item += NF_THREADS / 32) {
        const int yb = y0 + (item / n_xg) * 8;
        {
            const int x = (item % n_xg) * 32 + lane;
            const bool in = x < W;
            Member ma, mb;
            Ring ra, rb;
            ma.L = mb.L = ra.L = rb.L = NF_NONE;
#pragma unroll
            for (int k = 0; k < 8; ++k) ma.q[k] = mb.q[k] = 0ull;
            ma.cs = ma.sx = ma.sy = mb.cs = mb.sx = mb.sy = 0u;
            ra.q[0] = ra.q[1] = ra.q[2] = rb.q[0] = rb.q[1] = rb.q[2] = 0ull;
            ra.cp = rb.cp = 0u;
            for (int i = 0; i < 8; ++i) {
                const int y = yb + i;
                if (y >= y1) break;
                const int* row = lab_s + (y - y0 + 2) * W;
                const size_t p = (size_t)y * W + (in ? x : 0);
                const int L = in ? row[x] : NF_NONE;
                const float r = __ldg(img + 3 * p), g = __ldg(img + 3 * p + 1), bl = __ldg(img + 3 * p + 2);
                const unsigned long long fr = (unsigned long long)__float2ll_rn(r * S1f), fg = (unsigned long long)__float2ll_rn(g * S1f),
                                         fb = (unsigned long long)__float2ll_rn(bl * S1f);
                if (in && L >= 0 && L < cap) {
                    const double dr = r, dg = g, db = bl;
                    const double gray = dr * 0.2989 + dg * 0.5870 + db * 0.1140;   // :151
                    unsigned long long t8[8] = {fr, fg, fb, nf_fix(gray, NF_S1), nf_fix(dr * dr, NF_S2), nf_fix(dg * dg, NF_S2),
                                                nf_fix(db * db, NF_S2), nf_fix(gray * gray, NF_S2)};
                    const unsigned cs = 1u | (can[p] != 0 ? 0x10000u : 0u);
                    if (ma.L == NF_NONE) ma.L = L;
                    if (L == ma.L) {
#pragma unroll
                        for (int k = 0; k < 8; ++k) ma.q[k] += t8[k];
                        ma.cs += cs; ma.sx += (unsigned)x; ma.sy += (unsigned)y;
                    } else {
                        if (mb.L == NF_NONE) mb.L = L;
                        if (L == mb.L) {
#pragma unroll
                            for (int k = 0; k < 8; ++k) mb.q[k] += t8[k];
                            mb.cs += cs; mb.sx += (unsigned)x; mb.sy += (unsigned)y;
                        } else {   // a third label inside 8 rows of one column: straight to the table
#pragma unroll
                            for (int k = 0; k < 8; ++k) add64(k, L, t8[k]);
                            atomicAdd(tab + (size_t)NF_CNT * cap + L, 1u);
                            atomicAdd(tab + (size_t)(NF_CNT + 1) * cap + L, (unsigned)x);
                            atomicAdd(tab + (size_t)(NF_CNT + 2) * cap + L, (unsigned)y);
                            if (cs >> 16) atomicAdd(tab + (size_t)(NF_CNT + 3) * cap + L, 1u);
                        }
                    }
                }
                // ---- what this pixel contributes to its foreign neighbours' regions: distinct labels among the 12
                // L1 <= 2 neighbours, the 4-neighbours first (a label first seen there makes this a perimeter pixel)
                if (in) {
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
                    // first three distinct foreign labels (near flags in nf); a 4th+ distinct label is rare -> `over`
                    int u1 = NF_NONE, u2 = NF_NONE, u3 = NF_NONE;
                    unsigned nf = 0, over = 0;
#pragma unroll
                    for (int k = 0; k < 12; ++k) {
                        const int u = v[k];
                        if ((u != L) & (u >= 0) & (u < cap) & (u != u1) & (u != u2) & (u != u3)) {
                            if (u1 == NF_NONE) { u1 = u; nf |= k < 4 ? 1u : 0u; }
                            else if (u2 == NF_NONE) { u2 = u; nf |= k < 4 ? 2u : 0u; }
                            else if (u3 == NF_NONE) { u3 = u; nf |= k < 4 ? 4u : 0u; }
                            else over |= 1u << k;
                        }
                    }
                    auto contribute = [&](int u, bool near) {
                        const unsigned cp = 1u | (near ? 0x10000u : 0u);
                        if (ra.L == NF_NONE) ra.L = u;
                        if (u == ra.L) {
                            ra.q[0] += fr; ra.q[1] += fg; ra.q[2] += fb; ra.cp += cp;
                        } else {
                            if (rb.L == NF_NONE) rb.L = u;
                            if (u == rb.L) {
                                rb.q[0] += fr; rb.q[1] += fg; rb.q[2] += fb; rb.cp += cp;
                            } else {
                                add64(8, u, fr); add64(9, u, fg); add64(10, u, fb);
                                atomicAdd(tab + (size_t)(NF_CNT + 5) * cap + u, 1u);
                                if (near) atomicAdd(tab + (size_t)(NF_CNT + 4) * cap + u, 1u);
                            }
                        }
                    };
                    if (u1 != NF_NONE) contribute(u1, (nf & 1u) != 0);
                    if (u2 != NF_NONE) contribute(u2, (nf & 2u) != 0);
                    if (u3 != NF_NONE) contribute(u3, (nf & 4u) != 0);
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
            }
            flush_member(ma);
            flush_member(mb);
            flush_ring(ra);
            flush_ring(rb);
        }
    }
    cluster.sync();

    // ---- one thread per node of the image: sum the four partial tables (DSMEM), then the 15 features
    const int n0 = node_ptr[b], nb = min(node_ptr[b + 1], node_cap) - n0;
    const unsigned int* rt[NF_CLUSTER];
#pragma unroll
    for (int r = 0; r < NF_CLUSTER; ++r) rt[r] = cluster.map_shared_rank(tab, r);
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
                for (int k = 0; k < 11; ++k)
                    q64[k] += (unsigned long long)rt[r][(size_t)(3 * k) * cap + L] + ((unsigned long long)rt[r][(size_t)(3 * k + 1) * cap + L] << 16) +
                              ((unsigned long long)rt[r][(size_t)(3 * k + 2) * cap + L] << 32);
#pragma unroll
                for (int k = 0; k < 6; ++k) q32[k] += rt[r][(size_t)(NF_CNT + k) * cap + L];
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
    return ((size_t)cap * NF_ROWS + 3) / 4 * 16 + (size_t)(rpc + 4) * W * 4;
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
    if (label_cap > 0 && w % 8 == 0 && smem <= 200 * 1024 && batch <= 65535 && (long long)((h + NF_CLUSTER - 1) / NF_CLUSTER) * w <= 65535) {
        static PerDeviceMax smem_set;
        if (smem_set.raise(smem)) {
            cudaFuncSetAttribute(node_features_tile_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
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
