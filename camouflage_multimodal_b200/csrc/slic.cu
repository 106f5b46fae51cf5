// SLIC superpixels on the device -- skimage.segmentation.slic(img_u8, n_segments, compactness, sigma)
// as the reference calls it (models/region_graph/extract_rg_embeddings.py:143-144), restated in oracle/slic.py.
// float64 throughout, every sum in the order of the sequential algorithm (no FMA contraction), so the result equals the
// oracle's up to libm differences in pow / cbrt:
//   slic_minmax_kernel    per-image min / max byte (skimage >= 0.19 rescales the image to [0,1] first)
//   slic_lab_kernel       byte * (1/255), min-max rescale, rgb2lab (sRGB, D65), then the singleton z pass of the 3-D
//                         Gaussian (sum of all taps on one value) -- Lab BEFORE the blur, as skimage >= 0.19 does
//   slic_blur_kernel      scipy correlate1d (symmetric form), mode='reflect', along y then x, per Lab channel; the x pass
//                         multiplies by 1 / compactness on the way out
//   slic_assign_kernel    per pixel: the cluster of least distance among those whose +-2*step window holds the pixel,
//                         lowest index on ties (the sequential loop's strict '<' in index order)
//   slic_update_kernel    one warp per cluster walks the OLD window in raster order: sums / count -> new centre
//   slic_connect_kernel   _enforce_label_connectivity_cython: raster-order BFS, one warp per image (the algorithm is
//                         sequential by definition: the label a small segment joins depends on the visiting order)
#include <math.h>
#include <stdlib.h>

#include <algorithm>
#include <vector>

#include "common.cuh"

namespace cm {

constexpr int SLIC_MAX_RADIUS = 16;
struct SlicGauss {
    double w[2 * SLIC_MAX_RADIUS + 1];   // index r + radius
    int radius;
};
struct SlicGrid {
    int start_y, start_x, seed_step_y, seed_step_x, ny, nx, K;   // seeds
    int step_y, step_x, step;                                    // window half-extent / spatial weight
};

__device__ __forceinline__ int slic_reflect(int i, int n) {   // scipy mode='reflect': (d c b a | a b c d | d c b a)
    while (i < 0 || i >= n) {
        if (i < 0) i = -i - 1;
        if (i >= n) i = 2 * n - 1 - i;
    }
    return i;
}

__global__ void slic_mm_init_kernel(int* __restrict__ mm, int batch) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < batch) {
        mm[2 * i] = 255;
        mm[2 * i + 1] = 0;
    }
}
// per-image min / max byte (skimage >= 0.19 rescales the whole image to [0,1] before anything else); mm [B,2] = (min, max),
// pre-set to (255, 0)
__global__ void slic_minmax_kernel(const uint8_t* __restrict__ img, int per_image, int* __restrict__ mm) {
    const int b = blockIdx.y;
    const uint8_t* p = img + (size_t)b * per_image;
    int lo = 255, hi = 0;
    for (int i = (blockIdx.x * blockDim.x + threadIdx.x) * 4; i < per_image; i += gridDim.x * blockDim.x * 4) {
        if (i + 3 < per_image && (reinterpret_cast<uintptr_t>(p + i) & 3) == 0) {
            const uchar4 v = *reinterpret_cast<const uchar4*>(p + i);
            lo = min(lo, min(min((int)v.x, (int)v.y), min((int)v.z, (int)v.w)));
            hi = max(hi, max(max((int)v.x, (int)v.y), max((int)v.z, (int)v.w)));
        } else {
            for (int j = i; j < min(i + 4, per_image); ++j) {
                lo = min(lo, (int)p[j]);
                hi = max(hi, (int)p[j]);
            }
        }
    }
    lo = __reduce_min_sync(~0u, lo);
    hi = __reduce_max_sync(~0u, hi);
    if ((threadIdx.x & 31) == 0) {
        atomicMin(mm + 2 * b, lo);
        atomicMax(mm + 2 * b + 1, hi);
    }
}
// axis 0: along y, axis 1: along x; in / out [B,H,W,3]
__global__ void slic_blur_kernel(const double* __restrict__ in, int H, int W, int axis, const SlicGauss g, double scale, double* __restrict__ out) {
    const int xc = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y, b = blockIdx.z;   // xc = x * 3 + channel
    if (xc >= 3 * W) return;
    const int x = xc / 3, c = xc - 3 * x;
    const double* p = in + (size_t)b * H * W * 3;
    auto at = [&](int yy, int xx) { return p[((size_t)yy * W + xx) * 3 + c]; };
    double t = __dmul_rn(at(y, x), g.w[g.radius]);
    for (int j = g.radius; j >= 1; --j) {
        double a, d;
        if (axis == 0) {
            a = at(slic_reflect(y - j, H), x);
            d = at(slic_reflect(y + j, H), x);
        } else {
            a = at(y, slic_reflect(x - j, W));
            d = at(y, slic_reflect(x + j, W));
        }
        t = __dadd_rn(t, __dmul_rn(__dadd_rn(a, d), g.w[g.radius - j]));
    }
    out[((size_t)b * H + y) * W * 3 + xc] = __dmul_rn(t, scale);   // 1.0 (exact) or 1 / compactness after the last pass
}
// bytes -> img_as_float (* 1/255) -> min-max rescale -> rgb2lab (sRGB, D65) -> the singleton z pass of the 3-D Gaussian
// (sum of all taps on one value) -> * ratio (1.0 while a blur follows).  out [B,H,W,3] f64
__global__ void slic_lab_kernel(const uint8_t* __restrict__ img, const int* __restrict__ mm, int hw, long long npx, const SlicGauss g, int z_pass,
                                double ratio, double* __restrict__ v) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= npx) return;
    const int b = (int)(i / hw);
    const double imin = __dmul_rn((double)mm[2 * b], 1.0 / 255.0), imax = __dmul_rn((double)mm[2 * b + 1], 1.0 / 255.0);
    const double range = __dadd_rn(imax, -imin);
    double c[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        double a = __dadd_rn(__dmul_rn((double)img[3 * i + k], 1.0 / 255.0), -imin);
        if (imax != imin) a = __ddiv_rn(a, range);
        c[k] = a > 0.04045 ? pow(__ddiv_rn(__dadd_rn(a, 0.055), 1.055), 2.4) : __ddiv_rn(a, 12.92);
    }
    const double M[3][3] = {{0.412453, 0.357580, 0.180423}, {0.212671, 0.715160, 0.072169}, {0.019334, 0.119193, 0.950227}};
    const double white[3] = {0.95047, 1.0, 1.08883};
    double f[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        double t = __dadd_rn(__dadd_rn(__dmul_rn(c[0], M[k][0]), __dmul_rn(c[1], M[k][1])), __dmul_rn(c[2], M[k][2]));
        t = __ddiv_rn(t, white[k]);
        f[k] = t > 0.008856 ? cbrt(t) : __dadd_rn(__dmul_rn(7.787, t), 16.0 / 116.0);
    }
    double lab[3] = {__dadd_rn(__dmul_rn(116.0, f[1]), -16.0), __dmul_rn(500.0, __dadd_rn(f[0], -f[1])), __dmul_rn(200.0, __dadd_rn(f[1], -f[2]))};
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        double x = lab[k];
        if (z_pass) {   // every tap of the z axis (length 1, reflect) reads the same value
            double t = __dmul_rn(x, g.w[g.radius]);
            for (int j = g.radius; j >= 1; --j) t = __dadd_rn(t, __dmul_rn(__dadd_rn(x, x), g.w[g.radius - j]));
            x = t;
        }
        v[3 * i + k] = __dmul_rn(x, ratio);
    }
}

// seg [B,K,5] = (y, x, L, a, b)
__global__ void slic_seed_kernel(SlicGrid g, int batch, double* __restrict__ seg) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= batch * g.K) return;
    const int k = i % g.K;
    double* s = seg + (size_t)i * 5;
    s[0] = (double)(g.start_y + (k / g.nx) * g.seed_step_y);
    s[1] = (double)(g.start_x + (k % g.nx) * g.seed_step_x);
    s[2] = s[3] = s[4] = 0.0;
}

// window of cluster centre c along an axis of length n: [max(c - 2 step, 0), min(c + 2 step + 1, n)) truncated to int
__device__ __forceinline__ void slic_window(double c, int step, int n, int* lo, int* hi) {
    if (!isfinite(c)) {   // an emptied cluster (0 / 0): no window
        *lo = 0;
        *hi = 0;
        return;
    }
    *lo = (int)fmax(c - 2.0 * step, 0.0);
    *hi = (int)fmin(c + 2.0 * step + 1.0, (double)n);
}

// one CTA = 16 x 16 pixel tile; the clusters whose window meets the tile are compacted into shared memory (any order:
// the pixel keeps the lexicographically smallest (distance, cluster index), which is what the sequential loop's strict
// '<' in index order leaves behind), their centres and integer windows staged 64 at a time
// (Measured and dropped: a float32 sweep over the candidates that leaves only those within the float32 error bound of the
// least distance to the float64 arithmetic -- exact, bit-equal labels, but slower: 2.57 against 2.2 ms for the ten rounds.
// The kernel is bound by the window tests and the candidate walk, not by the float64 pipe.)
constexpr int SLIC_CAND_CAP = 1024;
__global__ void __launch_bounds__(256) slic_assign_kernel(const double* __restrict__ feat, const double* __restrict__ seg, SlicGrid g, int H,
                                                          int W, int* __restrict__ nearest) {
    __shared__ int s_list[SLIC_CAND_CAP];
    __shared__ int s_count;
    __shared__ double s_seg[64][5];
    __shared__ int s_win[64][4];
    __shared__ int s_k[64];
    const int b = blockIdx.z, ty0 = blockIdx.y * 16, tx0 = blockIdx.x * 16;
    const int tid = threadIdx.x, y = ty0 + (tid >> 4), x = tx0 + (tid & 15);
    const double* sg = seg + (size_t)b * g.K * 5;
    if (tid == 0) s_count = 0;
    __syncthreads();
    for (int k = tid; k < g.K; k += 256) {
        int ylo, yhi, xlo, xhi;
        slic_window(sg[k * 5], g.step_y, H, &ylo, &yhi);
        slic_window(sg[k * 5 + 1], g.step_x, W, &xlo, &xhi);
        if (ylo < ty0 + 16 && yhi > ty0 && xlo < tx0 + 16 && xhi > tx0) {
            const int idx = atomicAdd(&s_count, 1);
            if (idx < SLIC_CAND_CAP) s_list[idx] = k;
        }
    }
    __syncthreads();
    const bool listed = s_count <= SLIC_CAND_CAP;   // otherwise (more than 1024 overlapping windows) every cluster is tested
    const int n_cand = listed ? s_count : g.K;
    const bool in = y < H && x < W;
    double f0 = 0, f1 = 0, f2 = 0;
    if (in) {
        const double* p = feat + (((size_t)b * H + y) * W + x) * 3;
        f0 = p[0];
        f1 = p[1];
        f2 = p[2];
    }
    const double sw = 1.0 / (double)(g.step * g.step);
    double best = 1.7976931348623157e308;
    int best_k = -1;
    for (int c0 = 0; c0 < n_cand; c0 += 64) {
        const int nc = min(64, n_cand - c0);
        __syncthreads();
        if (tid < nc) {
            const int k = listed ? s_list[c0 + tid] : c0 + tid;
            s_k[tid] = k;
#pragma unroll
            for (int j = 0; j < 5; ++j) s_seg[tid][j] = sg[(size_t)k * 5 + j];
            slic_window(s_seg[tid][0], g.step_y, H, &s_win[tid][0], &s_win[tid][1]);
            slic_window(s_seg[tid][1], g.step_x, W, &s_win[tid][2], &s_win[tid][3]);
        }
        __syncthreads();
        if (in) {
            for (int i = 0; i < nc; ++i) {
                if (y < s_win[i][0] || y >= s_win[i][1] || x < s_win[i][2] || x >= s_win[i][3]) continue;
                const double dy = __dadd_rn(s_seg[i][0], -(double)y), dx = __dadd_rn(s_seg[i][1], -(double)x);
                double d = __dmul_rn(__dadd_rn(__dmul_rn(dy, dy), __dmul_rn(dx, dx)), sw);
                const double e0 = __dadd_rn(f0, -s_seg[i][2]), e1 = __dadd_rn(f1, -s_seg[i][3]), e2 = __dadd_rn(f2, -s_seg[i][4]);
                const double dc = __dadd_rn(__dadd_rn(__dmul_rn(e0, e0), __dmul_rn(e1, e1)), __dmul_rn(e2, e2));
                d = __dadd_rn(d, dc);
                const int k = s_k[i];
                if (d < best || (d == best && k < best_k)) {   // the lowest cluster index keeps a tie
                    best = d;
                    best_k = k;
                }
            }
        }
    }
    if (in) nearest[((size_t)b * H + y) * W + x] = best_k;
}

// One warp per (image, cluster): raster-order sums over the cluster's OLD window (every member lies inside it).  The
// lanes load a window row 32 pixels at a time (coalesced: position + three colours per lane) and ballot the members;
// lane c then owns running sum c (y, x, three colours) and adds the members in raster order -- the values come from the
// member's lane by shuffle, so the chain of float64 additions never waits for memory -- the same sequence of additions
// as the scalar loop.  The count is an exact integer either way.
__global__ void __launch_bounds__(256) slic_update_kernel(const double* __restrict__ feat, const int* __restrict__ nearest, SlicGrid g, int batch,
                                                          int H, int W, const double* __restrict__ seg_old, double* __restrict__ seg_new) {
    const int i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (i >= batch * g.K) return;
    const int b = i / g.K, k = i - b * g.K, c = min(lane, 4);
    const double* so = seg_old + (size_t)i * 5;
    int ylo, yhi, xlo, xhi;
    slic_window(so[0], g.step_y, H, &ylo, &yhi);
    slic_window(so[1], g.step_x, W, &xlo, &xhi);
    __shared__ double s_val[8][32][3];   // per warp: the three colours of the chunk's member pixels
    double (*sv)[3] = s_val[threadIdx.x >> 5];
    double acc = 0;
    int cnt = 0;
    for (int y = ylo; y < yhi; ++y) {
        const size_t row = ((size_t)b * H + y) * W;
        for (int x0 = xlo; x0 < xhi; x0 += 32) {
            const int x = x0 + lane;
            const bool mine = x < xhi && nearest[row + x] == k;
            unsigned m = __ballot_sync(~0u, mine);
            if (!m) continue;
            cnt += __popc(m);
            if (mine) {
                const double* p = feat + 3 * (row + x);
                sv[lane][0] = p[0];
                sv[lane][1] = p[1];
                sv[lane][2] = p[2];
            }
            __syncwarp();
            const int cc = max(c - 2, 0);
            const double yv = (double)y;
            while (m) {   // one pass for all five sums: the colour comes from the staging rows, y / x are known without memory
                const int src = __ffs(m) - 1;
                m &= m - 1;
                const double col = sv[src][cc];
                acc = __dadd_rn(acc, c == 0 ? yv : (c == 1 ? (double)(x0 + src) : col));
            }
            __syncwarp();
        }
    }
    if (lane < 5) seg_new[(size_t)i * 5 + lane] = __ddiv_rn(acc, (double)cnt);   // 0 / 0 = NaN for an emptied cluster, as in the reference loop
}

// One CTA per image; the walk is sequential by definition (one warp replays it eight queue entries at a time in the
// shared-memory variant, thread 0 alone in the global-memory fallback; the other threads only initialise / copy).
// `lab` holds any labels (equality is all that matters); while the image fits (about H*W <= 88K) the new labels (int16), a 4-bit neighbour-equality mask per pixel
// and the queue all live in shared memory, else the walk runs on `out` / `labels_in` / `coords` in global memory.
// coords [B, max_size] packed y << 16 | x.
#ifndef CM_SLIC_WARP_WALK
#define CM_SLIC_WARP_WALK 1
#endif
constexpr bool SLIC_WARP_WALK = CM_SLIC_WARP_WALK != 0;

template <bool SMEM>
__global__ void __launch_bounds__(256) slic_connect_kernel(const int* __restrict__ labels_in, int H, int W, int min_size, int max_size,
                                                           int start_label, int* __restrict__ coords, const int* __restrict__ only_if,
                                                           int* __restrict__ out) {
    extern __shared__ short s_out[];
    const int b = blockIdx.x, n = H * W;
    if (only_if && !only_if[b]) return;   // fallback pass behind the parallel kernel: only the images it could not finish
    const int* lab = labels_in + (size_t)b * n;
    int* o = out + (size_t)b * n;
    int* q = coords + (size_t)b * max(max_size, 1);
    // SMEM variant: besides the new labels (int16), a 4-bit "same input label as my E / W / S / N neighbour" mask per
    // pixel (two pixels per byte), built by all threads, so the sequential walk below never leaves shared memory
    unsigned char* s_same = reinterpret_cast<unsigned char*>(s_out + n);
    if (SMEM) q = reinterpret_cast<int*>(s_same + ((size_t)n / 2 + 8) / 4 * 4);   // the walk's queue too
    if (SMEM) {
        for (int i = threadIdx.x; i < (n + 1) / 2; i += blockDim.x) s_same[i] = 0;
        __syncthreads();
    }
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        if (SMEM) {
            s_out[i] = -1;
            const int y = i / W, x = i - y * W, l = lab[i];
            unsigned m = 0;
            if (x + 1 < W && lab[i + 1] == l) m |= 1u;
            if (x > 0 && lab[i - 1] == l) m |= 2u;
            if (y + 1 < H && lab[i + W] == l) m |= 4u;
            if (y > 0 && lab[i - W] == l) m |= 8u;
            atomicOr(reinterpret_cast<unsigned*>(s_same) + (i >> 3), m << (4 * (i & 7)));
        } else {
            o[i] = -1;
        }
    }
    __syncthreads();
    if (SMEM && SLIC_WARP_WALK && threadIdx.x < 32) {
        // One warp replays the sequential walk: lane = (queue entry % 8) * 4 + direction, so eight queue entries are
        // expanded per step.  Lane order == the order the scalar loop visits (entry, direction) pairs, so "first lane
        // claims a pixel", "last lane sets adjacent" and "stop at the lane that fills the queue" reproduce it exactly:
        // a pixel claimed by an earlier lane reads as `cur` for the later ones in the scalar loop, i.e. no effect.
        const int lane = threadIdx.x, ent = lane >> 2, dir = lane & 3;
        const int dy = dir == 2 ? 1 : (dir == 3 ? -1 : 0), dx = dir == 0 ? 1 : (dir == 1 ? -1 : 0);
        const unsigned below = (1u << lane) - 1u;
        int cur = start_label;
        for (int base = 0; base < n; base += 32) {
            unsigned pending = __ballot_sync(~0u, base + lane < n && s_out[base + lane] < 0);
            while (pending) {
                const int p0 = base + __ffs(pending) - 1;
                if (lane == 0) {
                    s_out[p0] = (short)cur;
                    q[0] = p0;
                }
                __syncwarp();
                int size = 1, visited = 0, adjacent = start_label;   // no labelled neighbour yet: join the first label (oracle/slic.py)
                while (visited < size && size < max_size) {
                    const int nent = min(8, size - visited);
                    const bool act = ent < nent;
                    const int pi = act ? q[visited + ent] : 0;
                    const int py = pi / W, px = pi - py * W, qy = py + dy, qx = px + dx;
                    const bool inb = act && qy >= 0 && qy < H && qx >= 0 && qx < W;
                    const int qi = inb ? qy * W + qx : 0;
                    const int oq = inb ? (int)s_out[qi] : cur;
                    const bool same = inb && ((s_same[pi >> 1] >> (4 * (pi & 1) + dir)) & 1u);
                    const bool cand = oq == -1 && same;
                    const unsigned cm = __ballot_sync(~0u, cand);
                    bool winner = false;
                    if (cand) winner = (__ffs(__match_any_sync(cm, qi)) - 1) == lane;
                    const unsigned wm = __ballot_sync(~0u, winner);
                    const int room = max_size - size, rank = __popc(wm & below);
                    const unsigned fill = __ballot_sync(~0u, winner && rank == room - 1);   // the append that fills the queue
                    const int cut = fill ? __ffs(fill) - 1 : 31;
                    if (winner && lane <= cut) {
                        s_out[qi] = (short)cur;
                        q[size + rank] = qi;
                    }
                    const unsigned am = __ballot_sync(~0u, inb && !cand && oq >= 0 && oq != cur && lane <= cut);
                    if (am) adjacent = __shfl_sync(~0u, oq, 31 - __clz(am));
                    size += min(__popc(wm), room);
                    visited += nent;
                    __syncwarp();
                }
                if (size < min_size) {
                    for (int i = lane; i < size; i += 32) s_out[q[i]] = (short)adjacent;
                } else {
                    ++cur;
                }
                __syncwarp();
                pending = __ballot_sync(~0u, base + lane < n && s_out[base + lane] < 0);
            }
        }
    }
    if (!(SMEM && SLIC_WARP_WALK) && threadIdx.x == 0) {
        auto get = [&](int i) -> int { return SMEM ? (int)s_out[i] : o[i]; };
        auto put = [&](int i, int v) { if (SMEM) s_out[i] = (short)v; else o[i] = v; };
        int cur = start_label;
        const int ddy[4] = {0, 0, 1, -1}, ddx[4] = {1, -1, 0, 0};
        for (int y = 0; y < H; ++y)
            for (int x = 0; x < W; ++x) {
                if (get(y * W + x) >= 0) continue;
                int adjacent = start_label;
                const int label = SMEM ? 0 : lab[y * W + x];
                put(y * W + x, cur);
                int size = 1, visited = 0;
                q[0] = (y << 16) | x;
                while (visited < size && size < max_size) {
                    const int py = q[visited] >> 16, px = q[visited] & 0xffff, pi = py * W + px;
                    const unsigned same = SMEM ? (s_same[pi >> 1] >> (4 * (pi & 1))) & 15u : 0u;
                    for (int i = 0; i < 4; ++i) {
                        const int qy = py + ddy[i], qx = px + ddx[i];
                        if (qy < 0 || qy >= H || qx < 0 || qx >= W) continue;
                        const int oq = get(qy * W + qx);
                        const bool same_label = SMEM ? ((same >> i) & 1u) != 0 : lab[qy * W + qx] == label;
                        if (oq == -1 && same_label) {
                            put(qy * W + qx, cur);
                            q[size++] = (qy << 16) | qx;
                            if (size >= max_size) break;
                        } else if (oq >= 0 && oq != cur) {
                            adjacent = oq;
                        }
                    }
                    ++visited;
                }
                if (size < min_size) {
                    for (int i = 0; i < size; ++i) put((q[i] >> 16) * W + (q[i] & 0xffff), adjacent);
                } else {
                    ++cur;
                }
            }
    }
    __syncthreads();
    if (SMEM)
        for (int i = threadIdx.x; i < n; i += blockDim.x) o[i] = s_out[i];
}

// ------------------------------------------------------------------------------------------------ parallel exact connectivity
// The same result as the sequential walk (oracle/slic.py enforce_connectivity), computed from its structure instead of
// replaying it pixel by pixel.  A "unit" is what one BFS of the walk labels.  For a 4-connected component of equal input
// labels with at most max_size pixels the unit is the whole component, found by a union-find in shared memory; a larger
// component is carved into units by replaying the walk's BFS on that component alone (it only ever expands into
// unlabelled pixels of its own component, so the rest of the image does not matter) -- rare, and done by one warp.
// Units are processed by the walk in the order of their first pixel, so
//   * kept units (size >= min_size) are numbered start_label, start_label + 1, ... by a prefix count over first pixels;
//   * a small unit takes `adjacent`: the label of the last neighbour pixel its BFS meets (queue order, directions
//     E W S N) that belongs to a unit with an EARLIER first pixel (later units are still unlabelled at that time) --
//     found by replaying the BFS of that one small unit (< min_size pixels) -- or start_label if there is none; small
//     units that point at small units are resolved by following the (strictly earlier) pointers.
// One CTA per image, H*W <= 65536 (16-bit pixel indices): parent table 128 KB + 4-bit neighbour-equality masks 32 KB +
// carve bits 8 KB of shared memory; unit sizes / pointers / labels in a global scratch area.
constexpr int CC_THREADS = 1024;
struct CcScratch {   // per image, all indexed by a unit's first pixel
    int* usz;              // [n] unit size (0 = not a unit start)
    unsigned short* adj;   // [n] small unit -> first pixel of the unit it joins (CC_NONE: none)
    int* fin;              // [n] final label of the unit
    unsigned short* queue; // [n] BFS queue of the carving warp
};
constexpr unsigned short CC_NONE = 0xffff;

__device__ __forceinline__ unsigned cc_same(const unsigned char* s_same, int i) { return (s_same[i >> 1] >> (4 * (i & 1))) & 15u; }
__device__ __forceinline__ int cc_find(const unsigned short* par, int x) {
    int p = par[x];
    while (p != x) {
        x = p;
        p = par[x];
    }
    return x;
}
__device__ __forceinline__ void cc_unite(unsigned short* par, int a, int b) {
    while (true) {
        a = cc_find(par, a);
        b = cc_find(par, b);
        if (a == b) return;
        if (a > b) {
            const int t = a;
            a = b;
            b = t;
        }
        const unsigned short old = atomicCAS(par + b, (unsigned short)b, (unsigned short)a);   // the larger root goes under the smaller
        if (old == (unsigned short)b) return;
        b = old;
    }
}

__global__ void __launch_bounds__(CC_THREADS) slic_connect_ccl_kernel(const int* __restrict__ labels_in, int H, int W, int min_size, int max_size,
                                                                      int start_label, int* __restrict__ usz_all, unsigned short* __restrict__ adj_all,
                                                                      int* __restrict__ fin_all, unsigned short* __restrict__ queue_all,
                                                                      int* __restrict__ fallback, int* __restrict__ out) {
    extern __shared__ unsigned short cc_par[];
    const int b = blockIdx.x, n = H * W, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    unsigned short* par = cc_par;
    unsigned char* s_same = reinterpret_cast<unsigned char*>(par + ((n + 7) & ~7));
    unsigned* s_done = reinterpret_cast<unsigned*>(s_same + (((n + 1) / 2 + 15) & ~15));   // one bit per pixel (carving only)
    __shared__ int s_cut[64], s_ncut, s_scan[CC_THREADS / 32], s_over;
    const int* lab = labels_in + (size_t)b * n;
    int* o = out + (size_t)b * n;
    int* usz = usz_all + (size_t)b * n;
    unsigned short* adj = adj_all + (size_t)b * n;
    int* fin = fin_all + (size_t)b * n;
    unsigned short* queue = queue_all + (size_t)b * n;
    // ---- neighbour-equality masks (E W S N = bits 0..3), parents, cleared tables
    for (int i = tid; i < (n + 1) / 2; i += CC_THREADS) s_same[i] = 0;
    for (int i = tid; i < (n + 31) / 32; i += CC_THREADS) s_done[i] = 0;
    if (tid == 0) {
        s_ncut = 0;
        s_over = 0;
    }
    __syncthreads();
    for (int i = tid; i < n; i += CC_THREADS) {
        const int y = i / W, x = i - y * W, l = lab[i];
        unsigned m = 0;
        if (x + 1 < W && lab[i + 1] == l) m |= 1u;
        if (x > 0 && lab[i - 1] == l) m |= 2u;
        if (y + 1 < H && lab[i + W] == l) m |= 4u;
        if (y > 0 && lab[i - W] == l) m |= 8u;
        atomicOr(reinterpret_cast<unsigned*>(s_same) + (i >> 3), m << (4 * (i & 7)));
        par[i] = (unsigned short)i;
        usz[i] = 0;
    }
    __syncthreads();
    // ---- union-find over the E and S links; the root of a component is its first pixel in raster order
    for (int i = tid; i < n; i += CC_THREADS) {
        const unsigned m = cc_same(s_same, i);
        if (m & 1u) cc_unite(par, i, i + 1);
        if (m & 4u) cc_unite(par, i, i + W);
    }
    __syncthreads();
    for (int i = tid; i < n; i += CC_THREADS) par[i] = (unsigned short)cc_find(par, i);   // flatten (roots are fixed points)
    __syncthreads();
    // ---- component sizes: one atomic per run of equal roots inside a warp's 32 consecutive pixels
    for (int base = 0; base < n; base += CC_THREADS) {
        const int i = base + tid;
        const int r = i < n ? (int)par[i] : -1;
        const int prev = __shfl_up_sync(0xffffffffu, r, 1);
        const unsigned heads = __ballot_sync(0xffffffffu, lane == 0 || r != prev);
        if (r >= 0 && ((heads >> lane) & 1u)) {
            const unsigned later = lane == 31 ? 0u : heads >> (lane + 1);
            atomicAdd(usz + r, later ? __ffs(later) : 32 - lane);
        }
    }
    __syncthreads();
    // ---- components larger than max_size are carved by the walk itself
    for (int i = tid; i < n; i += CC_THREADS)
        if (par[i] == i && usz[i] > max_size) {
            const int k = atomicAdd(&s_ncut, 1);
            if (k < 64) s_cut[k] = i; else s_over = 1;
        }
    __syncthreads();
    if (tid == 0) fallback[b] = 0;
    if (s_over) {   // more oversized components than the list holds: this image goes through the sequential walk kernel
        if (tid == 0) fallback[b] = 1;
        return;
    }
    if (warp == 0 && s_ncut > 0) {
        // The walk on one component: BFS from its first unlabelled pixel (eight queue entries x four directions per step,
        // lane order = visiting order, cut at the append that fills the queue), repeated until the component is used up.
        const int ent = lane >> 2, dir = lane & 3;
        const int dy = dir == 2 ? 1 : (dir == 3 ? -1 : 0), dx = dir == 0 ? 1 : (dir == 1 ? -1 : 0);
        const unsigned below = (1u << lane) - 1u;
        for (int c = 0; c < s_ncut; ++c) {
            const int root = s_cut[c];
            int remaining = usz[root], scan_from = root;
            while (remaining > 0) {
                // first pixel of the component that is still unlabelled, in raster order
                int p0 = -1;
                for (int base = scan_from & ~31; base < n && p0 < 0; base += 32) {
                    const int i = base + lane;
                    const bool hit = i < n && i >= scan_from && par[i] == root && !((s_done[i >> 5] >> (i & 31)) & 1u);
                    const unsigned m = __ballot_sync(0xffffffffu, hit);
                    if (m) p0 = base + __ffs(m) - 1;
                }
                if (p0 < 0) break;
                scan_from = p0 + 1;
                if (lane == 0) {
                    atomicOr(&s_done[p0 >> 5], 1u << (p0 & 31));
                    queue[0] = (unsigned short)p0;
                }
                __syncwarp();
                int size = 1, visited = 0;
                while (visited < size && size < max_size) {
                    const int nent = min(8, size - visited);
                    const bool act = ent < nent;
                    const int pi = act ? (int)queue[visited + ent] : 0;
                    const int py = pi / W, px = pi - py * W, qy = py + dy, qx = px + dx;
                    const bool inb = act && qy >= 0 && qy < H && qx >= 0 && qx < W;
                    const int qi = inb ? qy * W + qx : 0;
                    const bool same = inb && ((cc_same(s_same, pi) >> dir) & 1u);
                    const bool cand = same && !((s_done[qi >> 5] >> (qi & 31)) & 1u);
                    const unsigned cm = __ballot_sync(0xffffffffu, cand);
                    bool winner = false;
                    if (cand) winner = (__ffs(__match_any_sync(cm, qi)) - 1) == lane;
                    const unsigned wm = __ballot_sync(0xffffffffu, winner);
                    const int room = max_size - size, rank = __popc(wm & below);
                    const unsigned fill = __ballot_sync(0xffffffffu, winner && rank == room - 1);
                    const int cut = fill ? __ffs(fill) - 1 : 31;
                    if (winner && lane <= cut) {
                        atomicOr(&s_done[qi >> 5], 1u << (qi & 31));
                        queue[size + rank] = (unsigned short)qi;
                    }
                    size += min(__popc(wm), room);
                    visited += nent;
                    __syncwarp();
                }
                for (int i = lane; i < size; i += 32) par[queue[i]] = (unsigned short)p0;   // the unit's first pixel names it
                if (lane == 0) usz[p0] = size;
                remaining -= size;
                __syncwarp();
            }
        }
    }
    __syncthreads();
    // ---- small units: replay their own BFS (< min_size pixels) for `adjacent`.  Their first pixels are compacted into a
    // list (one unit per thread, no idle lanes between them); the carve bits become the units' visited marks.
    for (int i = tid; i < (n + 31) / 32; i += CC_THREADS) s_done[i] = 0;
    if (tid == 0) s_ncut = 0;
    __syncthreads();
    unsigned short* small = queue;   // the carving queue is free again: list of small-unit first pixels
    for (int i = tid; i < n; i += CC_THREADS)
        if (par[i] == i && usz[i] < min_size) small[atomicAdd(&s_ncut, 1)] = (unsigned short)i;
    __syncthreads();
    const int n_small = s_ncut;
    for (int k = tid; k < n_small; k += CC_THREADS) {
        const int s = small[k], sz = usz[s];
        unsigned short q[64];
        int last = CC_NONE;
        if (sz <= 64) {
            int size = 1, visited = 0;
            q[0] = (unsigned short)s;
            atomicOr(&s_done[s >> 5], 1u << (s & 31));
            while (visited < size) {
                const int pi = q[visited], py = pi / W, px = pi - py * W;
                const unsigned m = cc_same(s_same, pi);
#pragma unroll
                for (int d = 0; d < 4; ++d) {
                    const int qy = py + (d == 2 ? 1 : (d == 3 ? -1 : 0)), qx = px + (d == 0 ? 1 : (d == 1 ? -1 : 0));
                    if (qy < 0 || qy >= H || qx < 0 || qx >= W) continue;
                    const int qi = qy * W + qx, u = par[qi];
                    if (((m >> d) & 1u) && u == s) {
                        if (!((s_done[qi >> 5] >> (qi & 31)) & 1u)) {   // only this thread ever marks pixels of unit s
                            atomicOr(&s_done[qi >> 5], 1u << (qi & 31));
                            q[size++] = (unsigned short)qi;
                        }
                    } else if (u != s && u < s) {
                        last = u;   // already labelled when this unit is walked: its unit starts earlier
                    }
                }
                ++visited;
            }
        } else {
            s_over = 2;   // min_size > 64: above the local queue
        }
        adj[s] = (unsigned short)last;
    }
    __syncthreads();
    if (s_over) {   // a small unit above the local queue (min_size > 64): sequential walk for this image
        if (tid == 0) fallback[b] = 1;
        return;
    }
    // ---- kept units numbered by first pixel: prefix count over raster order (bit per pixel, 64 pixels per thread)
    for (int i = tid; i < (n + 31) / 32; i += CC_THREADS) s_done[i] = 0;
    __syncthreads();
    for (int i = tid; i < n; i += CC_THREADS)
        if (par[i] == i && usz[i] >= min_size) atomicOr(&s_done[i >> 5], 1u << (i & 31));
    __syncthreads();
    {
        const int words = (n + 31) / 32, wper = (words + CC_THREADS - 1) / CC_THREADS, w0 = tid * wper, w1 = min(words, w0 + wper);
        int cnt = 0;
        for (int w = w0; w < w1; ++w) cnt += __popc(s_done[w]);
        const int incl = warp_scan_incl(cnt, lane);
        if (lane == 31) s_scan[warp] = incl;
        __syncthreads();
        if (warp == 0) {
            const int v = s_scan[lane], sc = warp_scan_incl(v, lane);
            s_scan[lane] = sc - v;
        }
        __syncthreads();
        int run = start_label + s_scan[warp] + incl - cnt;
        for (int w = w0; w < w1; ++w) {
            unsigned m = s_done[w];
            while (m) {
                fin[w * 32 + __ffs(m) - 1] = run++;
                m &= m - 1;
            }
        }
    }
    __syncthreads();
    // ---- small units follow their pointers to a kept unit (strictly earlier first pixels: the chain ends)
    for (int s = tid; s < n; s += CC_THREADS) {
        if (par[s] != s || usz[s] >= min_size) continue;
        int u = adj[s], l = start_label;
        while (u != CC_NONE) {
            if (usz[u] >= min_size) {
                l = fin[u];
                break;
            }
            u = adj[u];
        }
        fin[s] = l;   // chains read `adj` and the kept units' labels only, never another small unit's `fin`
    }
    __syncthreads();
    for (int i = tid; i < n; i += CC_THREADS) {
        o[i] = fin[par[i]];
    }
}

static int launch_slic_walk(const int* labels_in, int batch, int h, int w, int min_size, int max_size, int start_label, int* coords,
                            const int* only_if, int* out, cudaStream_t st) {
    // labels (int16) + 4-bit neighbour masks + the queue; at most n / min_size + 1 labels are ever handed out
    const size_t smem = (size_t)h * w * sizeof(short) + ((size_t)h * w / 2 + 8) / 4 * 4 + (size_t)std::max(max_size, 1) * sizeof(int);
    const long long max_label = (long long)start_label + (long long)h * w / std::max(min_size, 1) + 1;
    if (smem <= 224 * 1024 && max_label < 32767 && ((size_t)h * w * sizeof(short)) % 4 == 0) {
        static PerDeviceOnce attr;
        if (attr.first()) {
            cudaFuncSetAttribute(slic_connect_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 224 * 1024);
        }
        slic_connect_kernel<true><<<batch, 256, smem, st>>>(labels_in, h, w, min_size, max_size, start_label, coords, only_if, out);
    } else {
        slic_connect_kernel<false><<<batch, 256, 0, st>>>(labels_in, h, w, min_size, max_size, start_label, coords, only_if, out);
    }
    return CM_OK;
}

#ifndef CM_SLIC_PARALLEL_CONNECT
#define CM_SLIC_PARALLEL_CONNECT 1
#endif
static int launch_slic_connect(const int* labels_in, int batch, int h, int w, int min_size, int max_size, int start_label, int* coords,
                               int* out, cudaStream_t st) {
    CM_TIMED("slic_connectivity", st);
    const size_t n = (size_t)h * w;
    const bool parallel = CM_SLIC_PARALLEL_CONNECT && n <= 65536 && min_size >= 0 && max_size >= 1 && !getenv("CM_SLIC_SEQUENTIAL_WALK");
    if (!parallel) return launch_slic_walk(labels_in, batch, h, w, min_size, max_size, start_label, coords, nullptr, out, st);
    // scratch of the parallel kernel, stream-ordered: unit sizes / final labels (int32), pointers / carving queue (uint16), fallback flags
    const size_t per = n * (4 + 4 + 2 + 2), flags_off = align_up(per * batch, 256), bytes = flags_off + sizeof(int) * batch;
    char* buf = nullptr;
    if (cudaMallocAsync(reinterpret_cast<void**>(&buf), bytes, st) != cudaSuccess) {
        cudaGetLastError();
        return launch_slic_walk(labels_in, batch, h, w, min_size, max_size, start_label, coords, nullptr, out, st);
    }
    int* usz = reinterpret_cast<int*>(buf);
    int* fin = usz + n * batch;
    unsigned short* adj = reinterpret_cast<unsigned short*>(fin + n * batch);
    unsigned short* queue = adj + n * batch;
    int* fallback = reinterpret_cast<int*>(buf + flags_off);
    const size_t smem = (((n + 7) & ~(size_t)7) * 2) + ((((n + 1) / 2) + 15) & ~(size_t)15) + ((n + 31) / 32) * 4;
    static PerDeviceOnce attr;
    if (attr.first()) cudaFuncSetAttribute(slic_connect_ccl_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    slic_connect_ccl_kernel<<<batch, CC_THREADS, smem, st>>>(labels_in, h, w, min_size, max_size, start_label, usz, adj, fin, queue, fallback, out);
    launch_slic_walk(labels_in, batch, h, w, min_size, max_size, start_label, coords, fallback, out, st);   // exits at once unless flagged
    cudaFreeAsync(buf, st);
    return CM_OK;
}

// skimage.util.regular_grid for the volume (1, H, W): seed start and step per axis
static void regular_grid_hw(int H, int W, int n_points, int* start_y, int* start_x, int* step_y, int* step_x) {
    double dims[3] = {1.0, (double)std::min(H, W), (double)std::max(H, W)};   // ascending
    double space = dims[0] * dims[1] * dims[2];
    double st[3];
    if (space <= (double)n_points) {
        *start_y = *start_x = 0;
        *step_y = *step_x = 1;
        return;
    }
    for (int i = 0; i < 3; ++i) st[i] = pow(space / n_points, 1.0 / 3.0);
    if (dims[0] < st[0] || dims[1] < st[1] || dims[2] < st[2]) {
        for (int d = 0; d < 3; ++d) {
            st[d] = dims[d];
            double rest = 1.0;
            for (int e = d + 1; e < 3; ++e) rest *= dims[e];
            for (int e = d + 1; e < 3; ++e) st[e] = pow(rest / n_points, 1.0 / (double)(3 - d - 1));
            bool ok = true;
            for (int e = 0; e < 3; ++e) ok = ok && dims[e] >= st[e];
            if (ok) break;
        }
    }
    int starts[3], steps[3];
    for (int i = 0; i < 3; ++i) {
        starts[i] = (int)floor(st[i] / 2.0);
        steps[i] = (int)nearbyint(st[i]);   // np.round: half to even
    }
    const int iy = H <= W ? 1 : 2, ix = H <= W ? 2 : 1;   // position of each axis in the sorted order (ties: y first)
    *start_y = starts[iy];
    *step_y = std::max(steps[iy], 1);
    *start_x = starts[ix];
    *step_x = std::max(steps[ix], 1);
}

struct SlicWs {
    double *a, *b, *seg0, *seg1;
    int *nearest, *coords, *mm;
    size_t total;
};
static SlicWs carve_slic(void* ws, int batch, int h, int w, int K, int max_size) {
    Carver c(ws);
    SlicWs r;
    const size_t n = (size_t)batch * h * w;
    r.a = c.take<double>(n * 3);
    r.b = c.take<double>(n * 3);
    r.seg0 = c.take<double>((size_t)batch * K * 5);
    r.seg1 = c.take<double>((size_t)batch * K * 5);
    r.nearest = c.take<int>(n);
    r.coords = c.take<int>((size_t)batch * std::max(max_size, 1));
    r.mm = c.take<int>((size_t)batch * 2);
    r.total = align_up(c.off, 256);
    return r;
}
static SlicGrid slic_grid(int h, int w, int n_segments) {
    SlicGrid g;
    regular_grid_hw(h, w, n_segments, &g.start_y, &g.start_x, &g.seed_step_y, &g.seed_step_x);
    g.ny = (h - g.start_y + g.seed_step_y - 1) / g.seed_step_y;
    g.nx = (w - g.start_x + g.seed_step_x - 1) / g.seed_step_x;
    g.K = g.ny * g.nx;
    int sy, sx;
    regular_grid_hw(h, w, g.K, &sy, &sx, &g.step_y, &g.step_x);   // the clustering loop recomputes its steps from the seed count
    g.step = std::max(std::max(g.seed_step_y, g.seed_step_x), 1);
    return g;
}

}  // namespace cm

extern "C" size_t cm_slic_workspace_bytes(int batch, int h, int w, int n_segments) {
    if (batch <= 0 || h <= 0 || w <= 0 || n_segments <= 0) return 0;
    const cm::SlicGrid g = cm::slic_grid(h, w, n_segments);
    const int max_size = (int)(3.0 * ((double)h * w / g.K));
    return cm::carve_slic(nullptr, batch, h, w, g.K, max_size).total;
}

extern "C" int cm_slic(const uint8_t* image_u8, int batch, int h, int w, int n_segments, double compactness, double sigma,
                       int max_num_iter, void* workspace, size_t workspace_bytes, int32_t* labels, void* stream) {
    using namespace cm;
    CM_DEVICE_OR_RETURN();
    CM_REQUIRE(image_u8 && workspace && labels, "cm_slic: NULL pointer");
    CM_REQUIRE(batch > 0 && h > 0 && w > 0 && h < 65536 && w < 65536 && n_segments > 0 && compactness > 0.0 && sigma >= 0.0 && max_num_iter > 0,
               "cm_slic: bad arguments");
    const SlicGrid g = slic_grid(h, w, n_segments);
    const double segment_size = (double)h * w / g.K;
    const int min_size = (int)(0.5 * segment_size), max_size = (int)(3.0 * segment_size);
    SlicWs ws = carve_slic(workspace, batch, h, w, g.K, max_size);
    if (workspace_bytes < ws.total) {
        set_error("cm_slic: workspace %zu < %zu bytes", workspace_bytes, ws.total);
        return CM_ERR_WORKSPACE;
    }
    SlicGauss gw;
    gw.radius = 0;
    gw.w[0] = 1.0;
    if (sigma > 0.0) {   // scipy _gaussian_kernel1d, truncate = 4
        gw.radius = (int)(4.0 * sigma + 0.5);
        CM_REQUIRE(gw.radius <= SLIC_MAX_RADIUS, "cm_slic: sigma %.3f needs a kernel radius above %d", sigma, SLIC_MAX_RADIUS);
        const int nw = 2 * gw.radius + 1;
        for (int r = -gw.radius; r <= gw.radius; ++r) gw.w[r + gw.radius] = exp(-0.5 / (sigma * sigma) * (double)(r * r));
        // phi.sum() in numpy's order: eight running sums over blocks of 8, combined as a tree, then the remainder
        double sum = 0.0;
        if (nw < 8) {
            for (int i = 0; i < nw; ++i) sum += gw.w[i];
        } else {
            double r8[8];
            for (int j = 0; j < 8; ++j) r8[j] = gw.w[j];
            int i = 8;
            for (; i < nw - (nw % 8); i += 8)
                for (int j = 0; j < 8; ++j) r8[j] += gw.w[i + j];
            sum = ((r8[0] + r8[1]) + (r8[2] + r8[3])) + ((r8[4] + r8[5]) + (r8[6] + r8[7]));
            for (; i < nw; ++i) sum += gw.w[i];
        }
        for (int r = 0; r < nw; ++r) gw.w[r] /= sum;
    }
    cudaStream_t st = as_stream(stream);
    const long long npx = (long long)batch * h * w;
    {
        CM_TIMED("slic_preprocess", st);
        slic_mm_init_kernel<<<ceil_div(batch, 256), 256, 0, st>>>(ws.mm, batch);
        slic_minmax_kernel<<<dim3(std::min(ceil_div((long long)h * w * 3, 1024), 64), batch), 256, 0, st>>>(image_u8, h * w * 3, ws.mm);
        slic_lab_kernel<<<ceil_div(npx, 256), 256, 0, st>>>(image_u8, ws.mm, h * w, npx, gw, sigma > 0.0 ? 1 : 0, sigma > 0.0 ? 1.0 : 1.0 / compactness, ws.a);
        if (sigma > 0.0) {   // scipy gaussian_filter over the Lab volume: y pass, then x pass (* 1 / compactness on the way out)
            dim3 grid(ceil_div(3 * w, 128), h, batch);
            slic_blur_kernel<<<grid, 128, 0, st>>>(ws.a, h, w, 0, gw, 1.0, ws.b);
            slic_blur_kernel<<<grid, 128, 0, st>>>(ws.b, h, w, 1, gw, 1.0 / compactness, ws.a);
        }
        slic_seed_kernel<<<ceil_div(batch * g.K, 256), 256, 0, st>>>(g, batch, ws.seg0);
    }
    double *cur = ws.seg0, *nxt = ws.seg1;
    {
        CM_TIMED("slic_iterations", st);
        for (int it = 0; it < max_num_iter; ++it) {
            slic_assign_kernel<<<dim3(ceil_div(w, 16), ceil_div(h, 16), batch), 256, 0, st>>>(ws.a, cur, g, h, w, ws.nearest);
            slic_update_kernel<<<ceil_div(batch * g.K, 8), 256, 0, st>>>(ws.a, ws.nearest, g, batch, h, w, cur, nxt);
            std::swap(cur, nxt);
        }
    }
    // labels = nearest + 1 only shifts every value: equality tests are unchanged, so the BFS reads `nearest` itself
    launch_slic_connect(ws.nearest, batch, h, w, min_size, max_size, 1, ws.coords, labels, st);
    return post_launch("cm_slic");
}

extern "C" int cm_slic_enforce_connectivity(const int32_t* labels_in, int batch, int h, int w, int min_size, int max_size,
                                            int start_label, int32_t* coords_workspace, int32_t* labels_out, void* stream) {
    using namespace cm;
    CM_DEVICE_OR_RETURN();
    CM_REQUIRE(labels_in && coords_workspace && labels_out, "cm_slic_enforce_connectivity: NULL pointer");
    CM_REQUIRE(batch > 0 && h > 0 && w > 0 && h < 65536 && w < 65536 && max_size > 0 && start_label >= 0, "cm_slic_enforce_connectivity: bad arguments");
    launch_slic_connect(labels_in, batch, h, w, min_size, max_size, start_label, coords_workspace, labels_out, as_stream(stream));
    return post_launch("cm_slic_enforce_connectivity");
}
