// A9': S-measure, E-measure, F-measure (beta^2 = 0.3) and weighted F-measure as device reductions.
//
// NOT IN THE REFERENCE (utils/metrics.py:9-92 stops at IoU / Dice / P / R / F1 / MAE / accuracy); BASELINE.json's
// north_star names them, so the published definitions are restated in oracle/sod_metrics.py (parity unpinned) and
// computed here on the same conventions: pred [B,H,W] float32 in [0,1] used as it is, GT foreground = byte >= 128
// (gt / 255 > 0.5, utils/metrics.py:12), 8-bit quantisation q = floor(pred * 255) for the threshold curves.
//
//   sod_pass1   one pass: GT count / centroid sums, object-term moments of S, mean(pred), the two 256-bin histograms
//   sod_pass2   needs the centroid + adaptive threshold from pass 1: SSIM moments of the four quadrants (shifted by each
//               quadrant's corner pixel so a constant quadrant gives exactly 0 variance), adaptive-threshold counts
//   sod_edt_cols / sod_edt_rows   exact Euclidean feature transform of the GT (nearest foreground pixel, ties to the
//               lexicographically smallest (y, x)) for the weighted F-measure
//   sod_wf      pixel dependency (7x7 Gaussian, sigma 5, zero border) + pixel importance -> weighted error sums
//   sod_finalize / sod_partial   closed forms per image, curves, batch partial sums
// The prediction is either a dense float map or PAINTED on the fly from the label map and the per-node probabilities
// (models/region_graph/test.py:241-244, the same gather cm_paint_metrics does) -- the pipeline never writes the mask.
// The weighted F-measure (feature transform + 7x7 filter: the expensive part) is optional.
// (A single cluster-of-8 launch per batch -- pass 1, cluster barrier, pass 2, barrier, closed forms -- was built and measured
// slower in the pipeline: 151 us against 78 us for these three small launches, and the step fell from 151 k to 116 k img/s,
// because eight co-scheduled CTAs per image have to wait for SMs that fusion_node's full-shared-memory CTAs hold, whereas
// independent small CTAs slip into the gaps.)
// Every float64 sum is a per-thread running sum -> warp shuffle tree -> fixed-order block / chunk combination: the
// result does not depend on scheduling (no floating-point atomics).  Bound: HBM/L2 on pred (4 B) + gt (1 B) per pixel,
// read three times (the second and third from L2 at batch 32: 10 MB).
#include <math.h>

#include <algorithm>

#include "common.cuh"
#include "metrics_body.cuh"

namespace cm {

constexpr int SOD_CHUNKS = 16;   // blocks per image in the two pixel passes (256 x 256: sixteen pixels per thread, four in flight together)
constexpr int SOD_ILP = 4;
constexpr int SOD_P1 = 8;        // doubles per chunk, pass 1
constexpr int SOD_P2 = 18;       // doubles per chunk, pass 2
constexpr int SOD_KEYS = 8;
constexpr int SOD_PARTIAL = 2 * SOD_KEYS + 512;
constexpr double SOD_EPS = 2.220446049250313e-16;

struct SodWs {
    double *p1, *p2, *p4, *curves, *shift, *tot1, *tot2;
    int* hist;   // [B,512] histograms, then [2B + 1] completion counters (zeroed together)
    short *col_d, *col_y;
    int *d2, *near;
    size_t total;
    int wf_tiles;
};
static SodWs carve_sod(void* ws, int batch, int h, int w) {
    Carver c(ws);
    SodWs r;
    const size_t n = (size_t)batch * h * w;
    r.wf_tiles = ceil_div(w, 32) * ceil_div(h, 8);
    r.p1 = c.take<double>((size_t)batch * SOD_CHUNKS * SOD_P1);
    r.p2 = c.take<double>((size_t)batch * SOD_CHUNKS * SOD_P2);
    r.p4 = c.take<double>((size_t)batch * r.wf_tiles * 2);
    r.curves = c.take<double>((size_t)batch * 512);
    r.shift = c.take<double>((size_t)batch * 4);
    r.tot1 = c.take<double>((size_t)batch * SOD_P1);
    r.tot2 = c.take<double>((size_t)batch * SOD_P2);
    r.hist = c.take<int>((size_t)batch * 512 + 2 * (size_t)batch + 1);
    r.col_d = c.take<short>(n);
    r.col_y = c.take<short>(n);
    r.d2 = c.take<int>(n);
    r.near = c.take<int>(n);
    r.total = align_up(c.off, 256);
    return r;
}


// Where a pixel's prediction comes from: a dense float map, or the paint of test.py:241-244 through the label map.
struct SodSrc {
    const float* pred;        // dense [B,H,W] or NULL
    const void* labels;       // [B,H,W] int32 / int16
    int label_bytes;
    const float* mask_prob;   // [nodes]
    const int* node_ptr;      // [B+1]
    const int* id_of_label;   // [B,cap] or NULL (raw-label indexing, the reference's quirk)
    int cap;
};
struct SodPix {
    const float* pr;
    const int* l32;
    const short* l16;
    const float* prob;
    const int* idm;
    int nb, cap;
    __device__ __forceinline__ SodPix(const SodSrc& s, int b, size_t HW) {
        pr = s.pred ? s.pred + (size_t)b * HW : nullptr;
        l32 = (!s.pred && s.label_bytes == 4) ? reinterpret_cast<const int*>(s.labels) + (size_t)b * HW : nullptr;
        l16 = (!s.pred && s.label_bytes == 2) ? reinterpret_cast<const short*>(s.labels) + (size_t)b * HW : nullptr;
        nb = 0;
        prob = nullptr;
        idm = nullptr;
        cap = s.cap;
        if (!s.pred) {
            const int n0 = s.node_ptr[b];
            nb = s.node_ptr[b + 1] - n0;
            prob = s.mask_prob + n0;
            idm = s.id_of_label ? s.id_of_label + (size_t)b * s.cap : nullptr;
        }
    }
    __device__ __forceinline__ float at(int i) const {
        if (pr) return pr[i];
        int l = l32 ? l32[i] : (int)l16[i];
        if (idm) l = (l >= 0 && l < cap) ? idm[l] : -1;
        return (l >= 0 && l < nb) ? prob[l] : 0.f;
    }
};

// block-wide sum of NV doubles per thread, result valid in thread 0 (fixed order: lanes by shuffle tree, warps 0..7 serially)
template <int NV>
__device__ __forceinline__ void block_sum(double (&v)[NV], double* s_red /* [8 * NV] */) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < NV; ++k) {
        const double t = warp_sum(v[k]);
        if (lane == 0) s_red[wid * NV + k] = t;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
#pragma unroll
        for (int k = 0; k < NV; ++k) {
            double t = 0.0;
            for (int wi = 0; wi < (int)(blockDim.x >> 5); ++wi) t += s_red[wi * NV + k];
            v[k] = t;
        }
    }
}

// The block that finishes an image's last chunk adds the chunk partials up in chunk order (so the result does not
// depend on which block that is) and publishes the image totals.  `counter` is zeroed before the launch.
template <int NV>
__device__ __forceinline__ void sod_publish_totals(const double* __restrict__ part, int b, int* counter, double* __restrict__ tot) {
    __shared__ int s_last;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) s_last = atomicAdd(counter, 1) == SOD_CHUNKS - 1;
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    if (threadIdx.x < NV) {
        double t = 0.0;
#pragma unroll 4
        for (int c = 0; c < SOD_CHUNKS; ++c) t += __ldcg(part + ((size_t)b * SOD_CHUNKS + c) * NV + threadIdx.x);
        tot[(size_t)b * NV + threadIdx.x] = t;
    }
}

// grid (SOD_CHUNKS, B), block 256
__global__ void __launch_bounds__(256) sod_pass1_kernel(const SodSrc src, const uint8_t* __restrict__ gt, int H, int W, double* __restrict__ p1,
                                                        int* __restrict__ hist, int* __restrict__ counters, double* __restrict__ tot1) {
    __shared__ int s_hist[512];
    __shared__ double s_red[8 * SOD_P1];
    const int b = blockIdx.y, HW = H * W, lane = threadIdx.x & 31;
    const SodPix px(src, b, (size_t)HW);
    const uint8_t* g = gt + (size_t)b * HW;
    for (int i = threadIdx.x; i < 512; i += 256) s_hist[i] = 0;
    __syncthreads();
    // v: n_fg, sum y_fg, sum x_fg, sum p (fg), sum p^2 (fg), sum (1-p) (bg), sum (1-p)^2 (bg), sum p
    double v[SOD_P1] = {0, 0, 0, 0, 0, 0, 0, 0};
    int n_fg = 0, sy = 0, sx = 0;   // a chunk holds < 2^31 / max(H, W) pixels
    const int per = (HW + SOD_CHUNKS - 1) / SOD_CHUNKS, i0 = blockIdx.x * per, i1 = min(i0 + per, HW);
    for (int base = i0; base < i1; base += 256 * SOD_ILP) {   // whole warps stay in the loop: the histogram vote needs all lanes
        float pv[SOD_ILP];
        unsigned char gv[SOD_ILP];
#pragma unroll
        for (int u = 0; u < SOD_ILP; ++u) {   // the label -> node -> probability gathers of SOD_ILP pixels in flight together
            const int i = base + u * 256 + threadIdx.x;
            pv[u] = i < i1 ? px.at(i) : 0.f;
            gv[u] = i < i1 ? g[i] : 0;
        }
#pragma unroll
        for (int u = 0; u < SOD_ILP; ++u) {
            const int i = base + u * 256 + threadIdx.x;
            const bool in = i < i1;
            const double p = (double)pv[u];
            const bool fg = in && gv[u] >= 128;
            int q = (int)(p * 255.0);
            q = min(max(q, 0), 255);
            // a painted mask is constant over runs of pixels: one shared-memory atomic per run of equal (class, bin) keys
            const int key = in ? (fg ? 0 : 256) + q : -1;
            const int prev = __shfl_up_sync(0xffffffffu, key, 1);
            const unsigned heads = __ballot_sync(0xffffffffu, lane == 0 || key != prev);
            if (key >= 0 && ((heads >> lane) & 1u)) {
                const unsigned later = lane == 31 ? 0u : heads >> (lane + 1);
                atomicAdd(&s_hist[key], later ? __ffs(later) : 32 - lane);
            }
            if (!in) continue;
            const int y = i / W;
            v[7] += p;
            n_fg += fg;
            sy += fg ? y : 0;
            sx += fg ? i - y * W : 0;
            v[3] += fg ? p : 0.0;
            v[4] += fg ? p * p : 0.0;
            const double o = 1.0 - p;
            v[5] += fg ? 0.0 : o;
            v[6] += fg ? 0.0 : o * o;
        }
    }
    v[0] = (double)n_fg;
    v[1] = (double)sy;
    v[2] = (double)sx;
    block_sum<SOD_P1>(v, s_red);
    if (threadIdx.x == 0) {
        double* o = p1 + ((size_t)b * SOD_CHUNKS + blockIdx.x) * SOD_P1;
#pragma unroll
        for (int k = 0; k < SOD_P1; ++k) o[k] = v[k];
    }
    __syncthreads();
    for (int i = threadIdx.x; i < 512; i += 256)
        if (s_hist[i]) atomicAdd(&hist[(size_t)b * 512 + i], s_hist[i]);
    sod_publish_totals<SOD_P1>(p1, b, counters + b, tot1);
}

// GT centroid the way the published code rounds it: round-half-even of the mean, + 1; image centre when the GT is empty
__device__ __forceinline__ void sod_centroid(const double* t, int H, int W, int* cx, int* cy) {
    if (t[0] == 0.0) {
        *cx = (int)rint((double)W / 2.0) + 1;
        *cy = (int)rint((double)H / 2.0) + 1;
    } else {
        *cy = (int)rint(t[1] / t[0]) + 1;
        *cx = (int)rint(t[2] / t[0]) + 1;
    }
}

// grid (SOD_CHUNKS, B), block 256.  Per quadrant k (LT, RT, LB, RB): sum p', sum p'^2, sum p' g (p' = p - corner pixel),
// sum g; then adaptive-threshold tp, fp.
__global__ void __launch_bounds__(256) sod_pass2_kernel(const SodSrc src, const uint8_t* __restrict__ gt, int H, int W,
                                                        const double* __restrict__ tot1, double* __restrict__ p2, double* __restrict__ shift, int* __restrict__ counters,
                                                        double* __restrict__ tot2) {
    __shared__ double s_red[8 * SOD_P2];
    __shared__ double s_tot[SOD_P1];
    __shared__ double s_shift[4];
    const int b = blockIdx.y, HW = H * W;
    const SodPix px(src, b, (size_t)HW);
    const uint8_t* g = gt + (size_t)b * HW;
    if (threadIdx.x < SOD_P1) s_tot[threadIdx.x] = tot1[(size_t)b * SOD_P1 + threadIdx.x];
    __syncthreads();
    int cx, cy;
    sod_centroid(s_tot, H, W, &cx, &cy);
    const double thr = fmin(2.0 * (s_tot[7] / (double)HW), 1.0);
    if (threadIdx.x < 4) {   // the corner pixel of quadrant k
        const int k = threadIdx.x, y0 = (k & 2) ? cy : 0, x0 = (k & 1) ? cx : 0;
        const double sh = (y0 < H && x0 < W) ? (double)px.at(y0 * W + x0) : 0.0;
        s_shift[k] = sh;
        if (blockIdx.x == 0) shift[(size_t)b * 4 + k] = sh;
    }
    __syncthreads();
    double v[SOD_P2];
#pragma unroll
    for (int k = 0; k < SOD_P2; ++k) v[k] = 0.0;
    int gq[4] = {0, 0, 0, 0}, tpa = 0, fpa = 0;
    const int per = (HW + SOD_CHUNKS - 1) / SOD_CHUNKS, i0 = blockIdx.x * per, i1 = min(i0 + per, HW);
    for (int base = i0; base < i1; base += 256 * SOD_ILP) {
        float pv[SOD_ILP];
        unsigned char gv[SOD_ILP];
#pragma unroll
        for (int u = 0; u < SOD_ILP; ++u) {
            const int i = base + u * 256 + threadIdx.x;
            pv[u] = i < i1 ? px.at(i) : 0.f;
            gv[u] = i < i1 ? g[i] : 0;
        }
#pragma unroll
        for (int u = 0; u < SOD_ILP; ++u) {
            const int i = base + u * 256 + threadIdx.x;
            if (i >= i1) continue;
            const double p = (double)pv[u];
            const bool fg = gv[u] >= 128;
            const int y = i / W, x = i - y * W;
            const int k = (y >= cy ? 2 : 0) + (x >= cx ? 1 : 0);
            const double d = p - s_shift[k], dd = d * d, dg = fg ? d : 0.0;
            // a warp's 32 consecutive pixels share the quadrant except where they straddle cx: near-uniform branches
            switch (k) {
                case 0: v[0] += d; v[1] += dd; v[2] += dg; gq[0] += fg; break;
                case 1: v[3] += d; v[4] += dd; v[5] += dg; gq[1] += fg; break;
                case 2: v[6] += d; v[7] += dd; v[8] += dg; gq[2] += fg; break;
                default: v[9] += d; v[10] += dd; v[11] += dg; gq[3] += fg; break;
            }
            if (p >= thr) {
                tpa += fg;
                fpa += !fg;
            }
        }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) v[12 + j] = (double)gq[j];
    v[16] = (double)tpa;
    v[17] = (double)fpa;
    block_sum<SOD_P2>(v, s_red);
    if (threadIdx.x == 0) {
        double* o = p2 + ((size_t)b * SOD_CHUNKS + blockIdx.x) * SOD_P2;
#pragma unroll
        for (int k = 0; k < SOD_P2; ++k) o[k] = v[k];
    }
    sod_publish_totals<SOD_P2>(p2, b, counters + b, tot2);
}

// ---------------------------------------------------------------------------------------------- weighted F: feature transform
constexpr short SOD_NONE = 0x7fff;
// one thread per (image, column): nearest foreground row in the column (distance, row), the upper one on ties
__global__ void sod_edt_cols_kernel(const uint8_t* __restrict__ gt, int batch, int H, int W, short* __restrict__ col_d, short* __restrict__ col_y) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= batch * W) return;
    const int b = i / W, x = i - b * W;
    const uint8_t* g = gt + (size_t)b * H * W + x;
    short* cd = col_d + (size_t)b * H * W + x;
    short* cyy = col_y + (size_t)b * H * W + x;
    int last = -1;
    for (int y = 0; y < H; ++y) {   // nearest at or above
        if (g[(size_t)y * W] >= 128) last = y;
        cd[(size_t)y * W] = last >= 0 ? (short)(y - last) : SOD_NONE;
        cyy[(size_t)y * W] = (short)max(last, 0);
    }
    last = -1;
    for (int y = H - 1; y >= 0; --y) {   // nearest below wins only when strictly closer
        if (g[(size_t)y * W] >= 128) last = y;
        if (last >= 0 && (last - y) < (int)cd[(size_t)y * W]) {
            cd[(size_t)y * W] = (short)(last - y);
            cyy[(size_t)y * W] = (short)last;
        }
    }
}
// one block per (row, image): each thread owns pixels x, searches columns outward from x while they can still win
__global__ void __launch_bounds__(256) sod_edt_rows_kernel(const short* __restrict__ col_d, const short* __restrict__ col_y, int H, int W,
                                                           int* __restrict__ d2_out, int* __restrict__ near_out) {
    extern __shared__ short s_row[];   // [W] distances, [W] rows
    short* s_d = s_row;
    short* s_y = s_row + W;
    const int y = blockIdx.x, b = blockIdx.y;
    const size_t base = ((size_t)b * H + y) * W;
    for (int x = threadIdx.x; x < W; x += blockDim.x) {
        s_d[x] = col_d[base + x];
        s_y[x] = col_y[base + x];
    }
    __syncthreads();
    for (int x = threadIdx.x; x < W; x += blockDim.x) {
        long long best = (long long)1 << 40;
        int key = 0x7fffffff;
        auto consider = [&](int xx) {
            const int cd = s_d[xx];
            if (cd == SOD_NONE) return;
            const long long d = (long long)(x - xx) * (x - xx) + (long long)cd * cd;
            const int k = (int)s_y[xx] * W + xx;
            if (d < best || (d == best && k < key)) {
                best = d;
                key = k;
            }
        };
        consider(x);
        for (int k = 1; k < W && (long long)k * k <= best; ++k) {
            if (x - k >= 0) consider(x - k);
            if (x + k < W) consider(x + k);
        }
        d2_out[base + x] = best >= ((long long)1 << 40) ? -1 : (int)best;
        near_out[base + x] = key == 0x7fffffff ? 0 : key;
    }
}

struct Gauss7 {
    double k[49];
};
// block = 32 x 8 pixel tile; Et with a 3-pixel halo in shared memory (zero outside the image)
__global__ void __launch_bounds__(256) sod_wf_kernel(const SodSrc src, const uint8_t* __restrict__ gtb, const int* __restrict__ d2b,
                                                     const int* __restrict__ nearb, int H, int W, const Gauss7 gk, double* __restrict__ p4) {
    __shared__ double s_et[14][38];
    __shared__ double s_red[8 * 2];
    const int b = blockIdx.z, tx0 = blockIdx.x * 32, ty0 = blockIdx.y * 8;
    const size_t off0 = (size_t)b * H * W, off = 0;
    const SodPix pred(src, b, (size_t)H * W);
    const uint8_t* gt = gtb + off0;
    const int* d2 = d2b + off0;
    const int* near = nearb + off0;
    for (int i = threadIdx.x; i < 14 * 38; i += 256) {
        const int ly = i / 38, lx = i - ly * 38, y = ty0 + ly - 3, x = tx0 + lx - 3;
        double et = 0.0;
        if (y >= 0 && y < H && x >= 0 && x < W) {
            const size_t p = off + (size_t)y * W + x;
            if (gt[p] >= 128) {
                et = fabs((double)pred.at((int)p) - 1.0);
            } else {
                et = fabs((double)pred.at(near[p]) - 1.0);   // the error of the nearest foreground pixel
            }
        }
        s_et[ly][lx] = et;
    }
    __syncthreads();
    const int lx = threadIdx.x & 31, ly = threadIdx.x >> 5, x = tx0 + lx, y = ty0 + ly;
    double v[2] = {0.0, 0.0};   // sum Ew over GT, sum Ew over background
    if (x < W && y < H) {
        const size_t p = off + (size_t)y * W + x;
        const bool fg = gt[p] >= 128;
        const double e = fabs((double)pred.at((int)p) - (fg ? 1.0 : 0.0));
        double ea = 0.0;
#pragma unroll
        for (int dy = 0; dy < 7; ++dy)
#pragma unroll
            for (int dx = 0; dx < 7; ++dx) ea += gk.k[dy * 7 + dx] * s_et[ly + dy][lx + dx];
        if (fg) {
            v[0] = ea < e ? ea : e;
        } else {
            const double dst = sqrt((double)d2[p]);
            v[1] = e * (2.0 - exp(log(0.5) / 5.0 * dst));
        }
    }
    block_sum<2>(v, s_red);
    if (threadIdx.x == 0) {
        double* o = p4 + ((size_t)b * gridDim.x * gridDim.y + (size_t)blockIdx.y * gridDim.x + blockIdx.x) * 2;
        o[0] = v[0];
        o[1] = v[1];
    }
}

// ---------------------------------------------------------------------------------------------- closed forms
__device__ __forceinline__ double sod_e_from_counts(double tp, double fp, double n_fg, double n) {
    const double pf = tp + fp, pb = n - pf;
    double s;
    if (n_fg == 0.0) {
        s = pb;
    } else if (n_fg == n) {
        s = pf;
    } else {
        const double fn = n_fg - tp, tn = pb - fn, mp = pf / n, mg = n_fg / n;
        const double cnt[4] = {tp, fp, fn, tn};
        const double a[4] = {1.0 - mp, 1.0 - mp, 0.0 - mp, 0.0 - mp}, c[4] = {1.0 - mg, 0.0 - mg, 1.0 - mg, 0.0 - mg};
        s = 0.0;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const double align = 2.0 * (a[k] * c[k]) / (a[k] * a[k] + c[k] * c[k] + SOD_EPS);
            s = s + (align + 1.0) * (align + 1.0) / 4.0 * cnt[k];
        }
    }
    return s / (n - 1.0 + SOD_EPS);
}
__device__ __forceinline__ double sod_f_from_counts(double tp, double fp, double n_fg, double beta2) {
    double ps = tp + fp;
    if (ps == 0.0) ps = 1.0;
    const double prec = tp / ps, rec = tp / fmax(n_fg, 1.0);
    double den = beta2 * prec + rec;
    if (den == 0.0) den = 1.0;
    return (1.0 + beta2) * prec * rec / den;
}
__device__ __forceinline__ double sod_s_object(double sum, double sum2, double n) {
    const double x = sum / n;
    double sigma = 0.0;
    if (n > 1.0) sigma = sqrt(fmax((sum2 - n * x * x) / (n - 1.0), 0.0));
    return 2.0 * x / (x * x + 1.0 + sigma + SOD_EPS);
}

// one block of 256 threads per image: thread t owns curve entry t (threshold q >= 255 - t).  Returns true in the block that
// finished last (it has written the batch partial).  tot1 / tot2 / shift / hist may have been written by this very block
// (the fused region-table kernel): plain loads after a __syncthreads().
__device__ __forceinline__ bool sod_finalize_body(const double* tot1, const double* tot2, const double* shift,
                                                  const double* __restrict__ p4, int wf_tiles, const int* hist, int H, int W,
                                                  double beta2_f, double beta2_wf, double* __restrict__ per_image,
                                                  double* __restrict__ curves, int* __restrict__ done_counter, int batch,
                                                  double* __restrict__ partial) {
    __shared__ int s_tp[256], s_fp[256], s_wtp[8], s_wfp[8], s_last;
    __shared__ double s_f[256], s_e[256], s_fm[256], s_em[256];
    __shared__ double s_tot[SOD_P1], s_q[SOD_P2], s_wf[2];
    const int b = blockIdx.x, t = threadIdx.x;
    const double n = (double)H * (double)W;
    s_tp[t] = hist[(size_t)b * 512 + 255 - t];
    s_fp[t] = hist[(size_t)b * 512 + 256 + 255 - t];
    if (t < SOD_P1) s_tot[t] = tot1[(size_t)b * SOD_P1 + t];
    if (t >= 32 && t < 32 + SOD_P2) s_q[t - 32] = tot2[(size_t)b * SOD_P2 + t - 32];
    {   // inclusive scan of the two histograms (integers: order-free): warp scans + the warp totals
        const int lane = t & 31, wid = t >> 5;
        int a = warp_scan_incl(s_tp[t], lane), c = warp_scan_incl(s_fp[t], lane);
        if (lane == 31) {
            s_wtp[wid] = a;
            s_wfp[wid] = c;
        }
        __syncthreads();
        for (int wi = 0; wi < wid; ++wi) {
            a += s_wtp[wi];
            c += s_wfp[wi];
        }
        __syncthreads();
        s_tp[t] = a;
        s_fp[t] = c;
    }
    const double n_fg = s_tot[0];
    const double f = sod_f_from_counts((double)s_tp[t], (double)s_fp[t], n_fg, beta2_f);
    const double e = sod_e_from_counts((double)s_tp[t], (double)s_fp[t], n_fg, n);
    curves[(size_t)b * 512 + t] = f;
    curves[(size_t)b * 512 + 256 + t] = e;
    s_f[t] = f;
    s_e[t] = e;
    s_fm[t] = f;
    s_em[t] = e;
    // weighted-F tile sums: threads 0 / 1 own the two columns, strided partial sums combined in a fixed order below
    if (p4) {
        double a = 0.0, c = 0.0;
        for (int i = t; i < wf_tiles; i += 256) {
            a += p4[((size_t)b * wf_tiles + i) * 2];
            c += p4[((size_t)b * wf_tiles + i) * 2 + 1];
        }
        __syncthreads();
        // reuse the tree below: stash in registers until the curve reduction is done
        for (int o = 128; o > 0; o >>= 1) {   // curve sums / maxima first (fixed tree)
            __syncthreads();
            if (t < o) {
                s_f[t] += s_f[t + o];
                s_e[t] += s_e[t + o];
                s_fm[t] = fmax(s_fm[t], s_fm[t + o]);
                s_em[t] = fmax(s_em[t], s_em[t + o]);
            }
        }
        __syncthreads();
        const double fsum = s_f[0], esum = s_e[0], fmx = s_fm[0], emx = s_em[0];
        __syncthreads();
        s_f[t] = a;
        s_e[t] = c;
        for (int o = 128; o > 0; o >>= 1) {
            __syncthreads();
            if (t < o) {
                s_f[t] += s_f[t + o];
                s_e[t] += s_e[t + o];
            }
        }
        __syncthreads();
        if (t == 0) {
            s_wf[0] = s_f[0];
            s_wf[1] = s_e[0];
            s_f[0] = fsum;
            s_e[0] = esum;
            s_fm[0] = fmx;
            s_em[0] = emx;
        }
    } else {
        for (int o = 128; o > 0; o >>= 1) {
            __syncthreads();
            if (t < o) {
                s_f[t] += s_f[t + o];
                s_e[t] += s_e[t + o];
                s_fm[t] = fmax(s_fm[t], s_fm[t + o]);
                s_em[t] = fmax(s_em[t], s_em[t + o]);
            }
        }
    }
    __syncthreads();
    if (t < 32) {
    // The closed forms are ~40 dependent float64 divisions / square roots: spread over the lanes of warp 0 (one quadrant of
    // the region term per lane 0..3, the two object terms on lanes 4 / 5, the adaptive-threshold F / E on 6 / 7, weighted F
    // on 8) and gathered by lane 0 in the order a single thread would have added them.
    const unsigned full = 0xffffffffu;
    const double fsum = s_f[0], esum = s_e[0], fmx = s_fm[0], emx = s_em[0];
    const double* tot = s_tot;
    const double* q = s_q;
    const double mean_p = tot[7] / n;
    const bool mixed = n_fg != 0.0 && n_fg != n;
    double piece = 0.0;
    if (mixed && t < 4) {            // ---- S-measure, region term of quadrant t
        const int k = t;
        int cx, cy;
        sod_centroid(tot, H, W, &cx, &cy);
        const double w1 = (double)cx * cy / n, w2 = (double)cy * (W - cx) / n, w3 = (double)(H - cy) * cx / n, w4 = 1.0 - w1 - w2 - w3;
        const double wk = k == 0 ? w1 : k == 1 ? w2 : k == 2 ? w3 : w4;
        const double nn = k == 0 ? (double)cx * cy : k == 1 ? (double)(W - cx) * cy : k == 2 ? (double)cx * (H - cy) : (double)(W - cx) * (H - cy);
        if (nn > 0.0) {              // centroid on the last row / column: empty quadrant of weight 0
            const double nm1 = fmax(nn - 1.0, 1.0);
            const double md = q[3 * k] / nn, x = shift[(size_t)b * 4 + k] + md, y = q[12 + k] / nn;
            const double sx = fmax((q[3 * k + 1] - nn * md * md) / nm1, 0.0);
            const double sy = fmax(q[12 + k] * (1.0 - y) / nm1, 0.0);   // g is 0/1: sum (g - y)^2 = sum g - n y^2
            const double sxy = (q[3 * k + 2] - nn * md * y) / nm1;
            const double alpha = 4.0 * x * y * sxy, beta = (x * x + y * y) * (sx + sy);
            double score;
            if (alpha != 0.0) {
                score = alpha / (beta + SOD_EPS);
            } else {
                score = beta == 0.0 ? 1.0 : 0.0;
            }
            piece = wk * score;
        }
    } else if (mixed && t == 4) {    // ---- S-measure, object terms
        piece = sod_s_object(tot[3], tot[4], n_fg);
    } else if (mixed && t == 5) {
        piece = sod_s_object(tot[5], tot[6], n - n_fg);
    } else if (t == 6) {             // ---- adaptive threshold
        piece = q[16] == 0.0 ? 0.0 : sod_f_from_counts(q[16], q[17], n_fg, beta2_f);
    } else if (t == 7) {
        piece = sod_e_from_counts(q[16], q[17], n_fg, n);
    } else if (t == 8) {             // ---- weighted F (NaN when it was not asked for)
        piece = p4 ? 0.0 : nan("");
        if (p4 && n_fg > 0.0) {
            const double ew_fg = s_wf[0], ew_bg = s_wf[1];
            const double tpw = n_fg - ew_fg, fpw = ew_bg, r = 1.0 - ew_fg / n_fg, pw = tpw / (tpw + fpw + SOD_EPS);
            piece = (1.0 + beta2_wf) * r * pw / (r + beta2_wf * pw + SOD_EPS);
        }
    }
    double pc[9];
#pragma unroll
    for (int k = 0; k < 9; ++k) pc[k] = __shfl_sync(full, piece, k);
    if (t == 0) {
        double sm;
        if (n_fg == 0.0) {
            sm = 1.0 - mean_p;
        } else if (n_fg == n) {
            sm = mean_p;
        } else {
            const double yv = n_fg / n;
            const double so = yv * pc[4] + (1.0 - yv) * pc[5];
            double sr = 0.0;
#pragma unroll
            for (int k = 0; k < 4; ++k) sr += pc[k];
            sm = fmax(0.0, 0.5 * so + 0.5 * sr);
        }
        double* o = per_image + (size_t)b * SOD_KEYS;
        o[0] = sm;
        o[1] = pc[7];
        o[2] = esum / 256.0;
        o[3] = emx;
        o[4] = pc[6];
        o[5] = fsum / 256.0;
        o[6] = fmx;
        o[7] = pc[8];
    }
    }
    // ---- the last image's block adds the batch up in image order: partial [16 + 512]
    __threadfence();
    __syncthreads();
    if (t == 0) s_last = atomicAdd(done_counter, 1) == batch - 1;
    __syncthreads();
    if (!s_last) return false;
    __threadfence();
    // image order, one accumulator per entry (the sum does not depend on the batch's split into blocks); sixteen loads are in
    // flight before the first add -- with one load per add this loop alone was ~10 us at batch 32 and 0.5 ms at batch 512
    for (int i = t; i < SOD_PARTIAL; i += 256) {
        const bool key = i < 2 * SOD_KEYS;
        const double* src = key ? per_image + (i % SOD_KEYS) : curves + (i - 2 * SOD_KEYS);
        const size_t stride = key ? SOD_KEYS : 512;
        const bool sq = key && i >= SOD_KEYS;
        double sum = 0.0;
        for (int bb0 = 0; bb0 < batch; bb0 += 16) {
            double v[16];
#pragma unroll
            for (int u = 0; u < 16; ++u) v[u] = bb0 + u < batch ? __ldcg(src + (size_t)(bb0 + u) * stride) : 0.0;
#pragma unroll
            for (int u = 0; u < 16; ++u)
                if (bb0 + u < batch) sum += sq ? v[u] * v[u] : v[u];
        }
        partial[i] = sum;
    }
    return true;
}
__global__ void __launch_bounds__(256) sod_finalize_kernel(const double* __restrict__ tot1, const double* __restrict__ tot2, const double* __restrict__ shift,
                                                           const double* __restrict__ p4, int wf_tiles, const int* __restrict__ hist, int H, int W,
                                                           double beta2_f, double beta2_wf, double* __restrict__ per_image,
                                                           double* __restrict__ curves, int* __restrict__ done_counter, int batch,
                                                           double* __restrict__ partial) {
    sod_finalize_body(tot1, tot2, shift, p4, wf_tiles, hist, H, W, beta2_f, beta2_wf, per_image, curves, done_counter, batch, partial);
}


// ---------------------------------------------------------------------------------------------- region x ground-truth table
// A painted prediction is constant over a region, so every pixel sum of A9 / A9' factors through the contingency table
//   c[label][fg][quadrant] = pixels of that label, ground-truth class and S-measure quadrant
// (quadrants are cut at the GROUND TRUTH's centroid, so the table needs nothing from the model): confusion counts,
// |pred - gt|, the object / quadrant moments of S, both 256-bin histograms and the adaptive-threshold counts are sums over
// <= label_cap + 1 table rows of count x f(probability).  The table is built from labels + gt only, i.e. NEXT TO the graph
// build instead of after the model (rt_table_kernel), and what remains behind the GNN is one block per image
// (rt_metrics_kernel: table rows x probabilities, the closed forms, the batch sums) -- the three pixel passes
// (paint_metrics, sod_pass1, sod_pass2: 7 B/pixel each) and two finalize launches leave the step.
// Exactness: counts are integers; the float64 sums differ from the pixel-order sums by rounding only (tests: <= 1e-12
// relative).  |pred - gt| needs gt in {0, 255}: the table kernel counts the other bytes per image, and for an image that has
// any (a resized mask with soft borders) its block of rt_metrics_kernel walks that image's pixels for the A9 counts / MAE
// exactly as paint_metrics does.  Labels outside [0, label_cap) share the last row (probability 0, like SodPix::at /
// paint_metrics).
constexpr int RT_MAX_SPLITS = 16, RT_SPLIT_PIXELS = 1 << 18, RT_THREADS = 1024;

struct RtWs {
    long long* gtstat;   // [B,4] n_fg, sum y, sum x, bytes outside {0, 255}
    int* part;           // [B, splits, (cap + 1) * 8]
    double *curves, *shift, *tot1, *tot2;
    int* hist;           // [B,512] + [1] finalize counter
    int splits;
    size_t total;
};
static int rt_splits(int h, int w) {
    const long long hw = (long long)h * w;
    return (int)std::min<long long>(RT_MAX_SPLITS, (hw + RT_SPLIT_PIXELS - 1) / RT_SPLIT_PIXELS);
}
static RtWs carve_rt(void* ws, int batch, int h, int w, int cap) {
    Carver c(ws);
    RtWs r;
    r.splits = rt_splits(h, w);
    r.gtstat = c.take<long long>((size_t)batch * 4);
    r.part = c.take<int>((size_t)batch * r.splits * (cap + 1) * 8);
    r.curves = c.take<double>((size_t)batch * 512);
    r.shift = c.take<double>((size_t)batch * 4);
    r.tot1 = c.take<double>((size_t)batch * SOD_P1);
    r.tot2 = c.take<double>((size_t)batch * SOD_P2);
    r.hist = c.take<int>((size_t)batch * 512 + 1);
    r.total = align_up(c.off, 256);
    return r;
}

// foreground bits of eight consecutive pixels starting at pixel 8 * q (gt: one byte or one bit per pixel), and how many
// of the eight bytes are neither 0 nor 255
template <bool PACKED>
__device__ __forceinline__ unsigned rt_fg8(const uint8_t* __restrict__ g, int q, int* gray) {
    if (PACKED) return g[q];
    const uint2 w = reinterpret_cast<const uint2*>(g)[q];
    const unsigned hx = (w.x >> 7) & 0x01010101u, hy = (w.y >> 7) & 0x01010101u;
    *gray += (__popc(__vcmpne4(w.x, hx * 255u)) + __popc(__vcmpne4(w.y, hy * 255u))) >> 3;
    return ((hx * 0x01020408u) >> 24) | (((hy * 0x01020408u) >> 24) << 4);
}
template <bool PACKED>
__device__ __forceinline__ unsigned rt_fg1(const uint8_t* __restrict__ g, int i, int* gray) {
    if (PACKED) return (g[i >> 3] >> (i & 7)) & 1u;
    const unsigned v = g[i];
    *gray += v != 0u && v != 255u;
    return v >> 7;
}
__device__ __forceinline__ void rt_load8(const int* __restrict__ lab, int q, int* l) {
    const int4 a = reinterpret_cast<const int4*>(lab)[2 * q], c = reinterpret_cast<const int4*>(lab)[2 * q + 1];
    l[0] = a.x; l[1] = a.y; l[2] = a.z; l[3] = a.w; l[4] = c.x; l[5] = c.y; l[6] = c.z; l[7] = c.w;
}
__device__ __forceinline__ void rt_load8(const short* __restrict__ lab, int q, int* l) {
    const uint4 a = reinterpret_cast<const uint4*>(lab)[q];
    const unsigned w[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        l[2 * j] = (int)(short)(w[j] & 0xffffu);
        l[2 * j + 1] = (int)(short)(w[j] >> 16);
    }
}

// grid (splits, B), block 1024, dynamic shared memory (cap + 1) * 8 ints.  Phase 1: the image's ground-truth count /
// centroid sums / non-binary bytes (every split block computes them for the whole image -- the ground truth is 1 B or
// 1 bit per pixel -- so no block waits for another).  Phase 2: this block's slice of the pixels -> its own table (plain
// stores: no zeroing, no global atomics).  A thread takes 16 consecutive pixels (one 32-byte sector of int16 labels);
// runs of equal (label, class, quadrant) inside them cost one shared-memory atomic, and neighbouring lanes are 16 pixels
// apart -- wider than most superpixels, so same-address conflicts inside a warp stay rare.
template <typename LT, bool PACKED>
__global__ void __launch_bounds__(RT_THREADS) rt_table_kernel(const LT* __restrict__ labels, const uint8_t* __restrict__ gt, int H, int W, int cap,
                                                              long long* __restrict__ gtstat, int* __restrict__ part, int* __restrict__ done_counter) {
    extern __shared__ int s_tab[];
    __shared__ long long s_red[32][4];
    __shared__ long long s_stat[4];
    const int b = blockIdx.y, HW = H * W, nt = (cap + 1) * 8, lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const LT* lab = labels + (size_t)b * HW;
    const uint8_t* g = gt + (PACKED ? (size_t)b * (HW >> 3) : (size_t)b * HW);
    for (int i = threadIdx.x; i < nt; i += RT_THREADS) s_tab[i] = 0;
    // ---- phase 1
    long long n = 0, sy = 0, sx = 0;
    int gray = 0;
    const bool vec8 = (W & 7) == 0 && (PACKED || (reinterpret_cast<uintptr_t>(g) & 7) == 0);
    if (vec8) {
#pragma unroll 4
        for (int q = threadIdx.x; q < (HW >> 3); q += RT_THREADS) {
            const unsigned m = rt_fg8<PACKED>(g, q, &gray);
            const int i = q << 3, y = i / W, x0 = i - y * W, c = __popc(m);
            n += c;
            sy += (long long)y * c;
            sx += (long long)x0 * c + __popc(m & 0xaau) + 2 * __popc(m & 0xccu) + 4 * __popc(m & 0xf0u);
        }
    } else {
        for (int i = threadIdx.x; i < HW; i += RT_THREADS) {
            const unsigned f = rt_fg1<PACKED>(g, i, &gray);
            const int y = i / W;
            n += f;
            sy += f ? y : 0;
            sx += f ? i - y * W : 0;
        }
    }
    {
        long long v[4] = {n, sy, sx, (long long)gray};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            for (int o = 16; o > 0; o >>= 1) v[k] += __shfl_xor_sync(0xffffffffu, v[k], o);
            if (lane == 0) s_red[wid][k] = v[k];
        }
        __syncthreads();
        if (threadIdx.x < 4) {
            long long t = 0;
            for (int wi = 0; wi < RT_THREADS / 32; ++wi) t += s_red[wi][threadIdx.x];
            s_stat[threadIdx.x] = t;
            if (blockIdx.x == 0) {
                gtstat[(size_t)b * 4 + threadIdx.x] = t;
                if (threadIdx.x == 1 && b == 0) *done_counter = 0;
            }
        }
        __syncthreads();
    }
    int cx, cy;
    {
        const double t[3] = {(double)s_stat[0], (double)s_stat[1], (double)s_stat[2]};
        sod_centroid(t, H, W, &cx, &cy);
    }
    // ---- phase 2
    int dummy = 0;
    if ((W & 15) == 0 && ((reinterpret_cast<uintptr_t>(lab) | (PACKED ? (uintptr_t)0 : reinterpret_cast<uintptr_t>(g))) & 31) == 0) {
        const int nq = HW >> 4, per = (nq + gridDim.x - 1) / gridDim.x, q0 = blockIdx.x * per, q1 = min(q0 + per, nq);
        for (int q = q0 + threadIdx.x; q < q1; q += RT_THREADS) {
            int l[16];
            rt_load8(lab, 2 * q, l);
            rt_load8(lab, 2 * q + 1, l + 8);
            const unsigned m = rt_fg8<PACKED>(g, 2 * q, &dummy) | (rt_fg8<PACKED>(g, 2 * q + 1, &dummy) << 8);
            const int i = q << 4, y = i / W, x0 = i - y * W, ky = y >= cy ? 2 : 0;
            int prev = -1, cnt = 0;
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                const int lb = (l[j] >= 0 && l[j] < cap) ? l[j] : cap;
                const int key = (lb << 3) | (int)(((m >> j) & 1u) << 2) | ky | (x0 + j >= cx ? 1 : 0);
                if (key != prev) {
                    if (cnt) atomicAdd(&s_tab[prev], cnt);
                    prev = key;
                    cnt = 0;
                }
                ++cnt;
            }
            atomicAdd(&s_tab[prev], cnt);
        }
    } else {
        const int per = (HW + gridDim.x - 1) / gridDim.x, i0 = blockIdx.x * per, i1 = min(i0 + per, HW);
        for (int i = i0 + threadIdx.x; i < i1; i += RT_THREADS) {
            const int l = (int)lab[i], y = i / W, x = i - y * W;
            const int lb = (l >= 0 && l < cap) ? l : cap;
            atomicAdd(&s_tab[(lb << 3) | (int)(rt_fg1<PACKED>(g, i, &dummy) << 2) | (y >= cy ? 2 : 0) | (x >= cx ? 1 : 0)], 1);
        }
    }
    __syncthreads();
    int* o = part + ((size_t)b * gridDim.x + blockIdx.x) * nt;
    for (int i = threadIdx.x; i < nt; i += RT_THREADS) o[i] = s_tab[i];
}

// one block of 256 threads per image, after the model: table rows x node probabilities -> everything sod_pass1 /
// sod_pass2 / paint_metrics would have summed over the pixels, then the closed forms of sod_finalize and (last block)
// the batch sums of both metric families.  Dynamic shared memory: (cap + 1) * 8 ints + (cap + 1) floats.
__global__ void __launch_bounds__(256) rt_metrics_kernel(const int* __restrict__ part, int splits, int cap, const long long* __restrict__ gtstat,
                                                         const void* __restrict__ labels, int label_bytes, const float* __restrict__ mask_prob,
                                                         const int* __restrict__ node_ptr, const int* __restrict__ id_of_label, float threshold,
                                                         int H, int W, const uint8_t* __restrict__ gt_bytes,
                                                         unsigned long long* __restrict__ counts, double* __restrict__ abs_err,
                                                         double* tot1, double* tot2, double* shift, int* hist, double* __restrict__ curves,
                                                         int* __restrict__ done_counter, int batch, double* __restrict__ per_image,
                                                         double* __restrict__ partial, double* __restrict__ per_image_sod,
                                                         double* __restrict__ partial_sod) {
    extern __shared__ int s_dyn[];
    __shared__ int s_hist[512];
    __shared__ double s_red[8 * SOD_P2];
    __shared__ double s_shift[4], s_thr;
    const int b = blockIdx.x, t = threadIdx.x, HW = H * W, rows = cap + 1, nt = rows * 8;
    int* s_tab = s_dyn;
    float* s_p = reinterpret_cast<float*>(s_dyn + nt);
    const int n0 = node_ptr[b], nb = node_ptr[b + 1] - n0;
    const float* prob = mask_prob + n0;
    const int* idm = id_of_label ? id_of_label + (size_t)b * cap : nullptr;
    auto prob_of = [&](int l) {
        int idx = (l >= 0 && l < cap) ? l : -1;
        if (idm && idx >= 0) idx = idm[idx];
        return (idx >= 0 && idx < nb) ? prob[idx] : 0.f;
    };
    for (int i = t; i < 512; i += 256) s_hist[i] = 0;
    int cx, cy;
    const long long gray = gtstat[(size_t)b * 4 + 3];
    const double gt3[3] = {(double)gtstat[(size_t)b * 4], (double)gtstat[(size_t)b * 4 + 1], (double)gtstat[(size_t)b * 4 + 2]};
    sod_centroid(gt3, H, W, &cx, &cy);
    if (t < 4) {   // the corner pixel of quadrant k (sod_pass2)
        const int y0 = (t & 2) ? cy : 0, x0 = (t & 1) ? cx : 0;
        double sh = 0.0;
        if (y0 < H && x0 < W) {
            const size_t p = (size_t)b * HW + (size_t)y0 * W + x0;
            sh = (double)prob_of(label_bytes == 2 ? (int)reinterpret_cast<const short*>(labels)[p] : reinterpret_cast<const int*>(labels)[p]);
        }
        s_shift[t] = sh;
        shift[(size_t)b * 4 + t] = sh;
    }
    // the slices' tables added up, each row's probability next to it
    const int* pb = part + (size_t)b * splits * nt;
    for (int i = t; i < nt / 4; i += 256) {
        int4 a = reinterpret_cast<const int4*>(pb)[i];
        for (int s = 1; s < splits; ++s) {
            const int4 c = reinterpret_cast<const int4*>(pb + (size_t)s * nt)[i];
            a.x += c.x; a.y += c.y; a.z += c.z; a.w += c.w;
        }
        reinterpret_cast<int4*>(s_tab)[i] = a;
    }
    for (int l = t; l < rows; l += 256) s_p[l] = prob_of(l);
    __syncthreads();
    // ---- pass-1 sums, the two histograms, the A9 counts
    const double thr = (double)threshold;
    double v[SOD_P2];
#pragma unroll
    for (int k = 0; k < SOD_P2; ++k) v[k] = 0.0;
    long long tp = 0, sp = 0, eq = 0;
    double ae = 0.0;
    for (int l = t; l < rows; l += 256) {
        const int* c = s_tab + l * 8;
        const int cb = c[0] + c[1] + c[2] + c[3], cf = c[4] + c[5] + c[6] + c[7];
        if ((cb | cf) == 0) continue;
        const double p = (double)s_p[l], o = 1.0 - p;
        v[3] += p * cf;
        v[4] += p * p * cf;
        v[5] += o * cb;
        v[6] += o * o * cb;
        v[7] += p * (cb + cf);
        int q = (int)(p * 255.0);
        q = min(max(q, 0), 255);
        if (cf) atomicAdd(&s_hist[q], cf);
        if (cb) atomicAdd(&s_hist[256 + q], cb);
        const bool pos = p > thr;
        tp += pos ? cf : 0;
        sp += pos ? cb + cf : 0;
        eq += pos ? cf : cb;
        ae += o * cf + p * cb;   // |p - 1| on the foreground, |p - 0| on the background
    }
    double sg = gt3[0];
    if (gray != 0 && gt_bytes) {
        // non-binary ground truth: |p - g / 255| does not factor through the table -- this image's pixels once more, the
        // way paint_metrics accounts them (utils/metrics.py:11-12, 47, 52: both masks thresholded at `threshold`)
        tp = sp = eq = 0;
        ae = 0.0;
        long long sgi = 0;
        const uint8_t* g = gt_bytes + (size_t)b * HW;
        for (int i = t; i < HW; i += 256) {
            const size_t p = (size_t)b * HW + i;
            const double pr = (double)prob_of(label_bytes == 2 ? (int)reinterpret_cast<const short*>(labels)[p] : reinterpret_cast<const int*>(labels)[p]);
            const double gv = (double)g[i] / 255.0;
            const bool pb = pr > thr, gb = gv > thr;
            tp += pb && gb;
            sp += pb;
            sgi += gb;
            eq += pb == gb;
            ae += fabs(pr - gv);
        }
        v[9] = (double)sgi;
    }
    v[0] = (double)tp;
    v[1] = (double)sp;
    v[2] = (double)eq;
    v[8] = ae;
    block_sum<SOD_P2>(v, s_red);
    if (gray != 0 && gt_bytes) sg = v[9];
    if (t == 0) {
        double* o1 = tot1 + (size_t)b * SOD_P1;
        o1[0] = gt3[0];
        o1[1] = gt3[1];
        o1[2] = gt3[2];
#pragma unroll
        for (int k = 3; k < SOD_P1; ++k) o1[k] = v[k];
        s_thr = fmin(2.0 * (v[7] / (double)HW), 1.0);
        counts[4 * (size_t)b] = (unsigned long long)v[0];
        counts[4 * (size_t)b + 1] = (unsigned long long)v[1];
        counts[4 * (size_t)b + 2] = (unsigned long long)sg;
        counts[4 * (size_t)b + 3] = (unsigned long long)v[2];
        abs_err[b] = v[8];
    }
    __syncthreads();
    for (int i = t; i < 512; i += 256) hist[(size_t)b * 512 + i] = s_hist[i];
    // ---- pass-2 sums: quadrant moments about the corner pixel's value, adaptive-threshold counts
    const double athr = s_thr;
#pragma unroll
    for (int k = 0; k < SOD_P2; ++k) v[k] = 0.0;
    long long gq[4] = {0, 0, 0, 0}, tpa = 0, fpa = 0;
    for (int l = t; l < rows; l += 256) {
        const int* c = s_tab + l * 8;
        const double p = (double)s_p[l];
        int cb = 0, cf = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int b0 = c[k], f0 = c[4 + k];
            cb += b0;
            cf += f0;
            if ((b0 | f0) == 0) continue;
            const double d = p - s_shift[k];
            v[3 * k] += d * (b0 + f0);
            v[3 * k + 1] += d * d * (b0 + f0);
            v[3 * k + 2] += d * f0;
            gq[k] += f0;
        }
        if (p >= athr) {
            tpa += cf;
            fpa += cb;
        }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) v[12 + j] = (double)gq[j];
    v[16] = (double)tpa;
    v[17] = (double)fpa;
    __syncthreads();
    block_sum<SOD_P2>(v, s_red);
    if (t == 0) {
#pragma unroll
        for (int k = 0; k < SOD_P2; ++k) tot2[(size_t)b * SOD_P2 + k] = v[k];
    }
    __syncthreads();
    // ---- closed forms (sod_finalize) and, in the block that finishes last, the batch sums of both families
    const bool last = sod_finalize_body(tot1, tot2, shift, nullptr, 0, hist, H, W, 0.3, 1.0, per_image_sod, curves, done_counter, batch, partial_sod);
    if (last) {
        __syncthreads();
        metrics_finalize_block(counts, abs_err, batch, (long long)HW, per_image, partial, reinterpret_cast<double(*)[14]>(s_red));
    }
}

}  // namespace cm

extern "C" size_t cm_sod_workspace_bytes(int batch, int h, int w) {
    if (batch <= 0 || h <= 0 || w <= 0) return 0;
    return cm::carve_sod(nullptr, batch, h, w).total;
}

static int sod_metrics_impl(const cm::SodSrc& src, const uint8_t* gt, int batch, int h, int w, int with_weighted_f, void* workspace,
                            size_t workspace_bytes, double* per_image, double* partial, void* stream, const char* who) {
    using namespace cm;
    CM_DEVICE_OR_RETURN();
    CM_REQUIRE(gt && workspace && per_image && partial, "%s: NULL pointer", who);
    CM_REQUIRE(batch > 0 && batch <= 65535 && h > 0 && w > 0 && h < 32767 && w < 32767 && (long long)h * w < (1ll << 30), "%s: bad shape", who);
    CM_REQUIRE(((long long)h * w / SOD_CHUNKS + 1) * (h > w ? h : w) < (1ll << 31), "%s: image too large for the 32-bit centroid sums of a chunk", who);
    SodWs ws = carve_sod(workspace, batch, h, w);
    if (workspace_bytes < ws.total) {
        set_error("%s: workspace %zu < %zu bytes", who, workspace_bytes, ws.total);
        return CM_ERR_WORKSPACE;
    }
    cudaStream_t st = as_stream(stream);
    cudaMemsetAsync(ws.hist, 0, ((size_t)batch * 512 + 2 * (size_t)batch + 1) * sizeof(int), st);
    int* counters = ws.hist + (size_t)batch * 512;   // [B] pass 1, [B] pass 2, [1] finalize
    {
        CM_TIMED("sod_pass1", st);
        sod_pass1_kernel<<<dim3(SOD_CHUNKS, batch), 256, 0, st>>>(src, gt, h, w, ws.p1, ws.hist, counters, ws.tot1);
    }
    {
        CM_TIMED("sod_pass2", st);
        sod_pass2_kernel<<<dim3(SOD_CHUNKS, batch), 256, 0, st>>>(src, gt, h, w, ws.tot1, ws.p2, ws.shift, counters + batch, ws.tot2);
    }
    if (with_weighted_f) {
        Gauss7 gk;   // Matlab fspecial('gaussian', 7, 5)
        double sum = 0.0;
        for (int dy = -3; dy <= 3; ++dy)
            for (int dx = -3; dx <= 3; ++dx) {
                gk.k[(dy + 3) * 7 + dx + 3] = exp(-(double)(dy * dy + dx * dx) / (2.0 * 5.0 * 5.0));
                sum += gk.k[(dy + 3) * 7 + dx + 3];
            }
        for (int i = 0; i < 49; ++i) gk.k[i] /= sum;
        {
            CM_TIMED("sod_edt", st);
            sod_edt_cols_kernel<<<ceil_div((long long)batch * w, 128), 128, 0, st>>>(gt, batch, h, w, ws.col_d, ws.col_y);
            sod_edt_rows_kernel<<<dim3(h, batch), 256, 2 * (size_t)w * sizeof(short), st>>>(ws.col_d, ws.col_y, h, w, ws.d2, ws.near);
        }
        {
            CM_TIMED("sod_wf", st);
            sod_wf_kernel<<<dim3(ceil_div(w, 32), ceil_div(h, 8), batch), 256, 0, st>>>(src, gt, ws.d2, ws.near, h, w, gk, ws.p4);
        }
    }
    {
        CM_TIMED("sod_finalize", st);
        sod_finalize_kernel<<<batch, 256, 0, st>>>(ws.tot1, ws.tot2, ws.shift, with_weighted_f ? ws.p4 : nullptr, ws.wf_tiles, ws.hist, h, w, 0.3, 1.0,
                                                   per_image, ws.curves, counters + 2 * batch, batch, partial);
    }
    return post_launch(who);
}

extern "C" int cm_sod_metrics(const float* pred, const uint8_t* gt, int batch, int h, int w, int with_weighted_f, void* workspace,
                              size_t workspace_bytes, double* per_image, double* partial, void* stream) {
    if (!pred) {
        cm::set_error("cm_sod_metrics: NULL pointer");
        return CM_ERR_ARG;
    }
    cm::SodSrc src = {};
    src.pred = pred;
    return sod_metrics_impl(src, gt, batch, h, w, with_weighted_f, workspace, workspace_bytes, per_image, partial, stream, "cm_sod_metrics");
}

// ---------------------------------------------------------------------------------------------- no-allocation variant
// cm_sod_metrics (without the weighted F-measure) for callers that cannot allocate: the workspace is a file-scope device
// array that every GPU loading the library holds until exit.  All calls on a device share it (and find the previous call's
// contents there: the histogram / counter block is cleared by sod_metrics_impl as for any workspace) -- calls that overlap on
// the GPU race.
constexpr size_t SOD_STATIC_BYTES = 2u << 20;   // 2 MiB: ~170 images of 256 x 256 per call
static __device__ __align__(256) unsigned char g_sod_static_scratch[SOD_STATIC_BYTES];

// workspace bytes of the S / E / F passes alone: carve_sod lays the weighted-F arrays (col_d onwards) out last
static size_t sod_bytes_without_wf(int batch, int h, int w) {
    const cm::SodWs ws = cm::carve_sod(nullptr, batch, h, w);
    return cm::align_up((size_t)reinterpret_cast<uintptr_t>(ws.col_d), 256);
}

extern "C" size_t cm_sod_metrics_static_max_bytes(void) { return SOD_STATIC_BYTES; }

extern "C" int cm_sod_metrics_static(const float* pred, const uint8_t* gt, int batch, int h, int w, double* per_image, double* partial,
                                     void* stream) {
    using namespace cm;
    CM_DEVICE_OR_RETURN();
    if (!pred) {
        set_error("cm_sod_metrics_static: NULL pointer");
        return CM_ERR_ARG;
    }
    CM_REQUIRE(batch > 0 && h > 0 && w > 0 && batch <= 65535 && (long long)h * w < (1ll << 30), "cm_sod_metrics_static: bad shape");
    const size_t need = sod_bytes_without_wf(batch, h, w);
    if (need > SOD_STATIC_BYTES) {
        set_error("cm_sod_metrics_static: batch %d of %d x %d needs %zu bytes of scratch, the built-in array holds %zu -- use cm_sod_metrics "
                  "with a caller workspace", batch, h, w, need, (size_t)SOD_STATIC_BYTES);
        return CM_ERR_WORKSPACE;
    }
    // the array's name is a host-side stand-in here: its device address, looked up once per device with that device current
    static void* dev_ptr[64] = {};
    const int slot = current_device_slot();
    if (!dev_ptr[slot]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_sod_static_scratch);
        if (e != cudaSuccess || !p) {
            set_error("cm_sod_metrics_static: cudaGetSymbolAddress failed: %s", cudaGetErrorString(e));
            return CM_ERR_LAUNCH;
        }
        dev_ptr[slot] = p;
    }
    SodSrc src = {};
    src.pred = pred;
    // passing `need` as the size would fail carve_sod's full-size check: the weighted-F slices (unused with with_weighted_f = 0)
    // lie beyond the array and are never touched
    return sod_metrics_impl(src, gt, batch, h, w, 0, dev_ptr[slot], carve_sod(nullptr, batch, h, w).total, per_image, partial, stream,
                            "cm_sod_metrics_static");
}

extern "C" int cm_sod_metrics_painted(const void* labels, int label_bytes, const uint8_t* gt, int batch, int h, int w, const float* mask_prob,
                                      const int32_t* node_ptr, const int32_t* id_of_label, int label_cap, int with_weighted_f, void* workspace,
                                      size_t workspace_bytes, double* per_image, double* partial, void* stream) {
    if (!labels || !mask_prob || !node_ptr || (label_bytes != 2 && label_bytes != 4) || (id_of_label && label_cap <= 0)) {
        cm::set_error("cm_sod_metrics_painted: NULL pointer or label_bytes not 2 / 4");
        return CM_ERR_ARG;
    }
    cm::SodSrc src = {};
    src.labels = labels;
    src.label_bytes = label_bytes;
    src.mask_prob = mask_prob;
    src.node_ptr = node_ptr;
    src.id_of_label = id_of_label;
    src.cap = label_cap;
    return sod_metrics_impl(src, gt, batch, h, w, with_weighted_f, workspace, workspace_bytes, per_image, partial, stream, "cm_sod_metrics_painted");
}

extern "C" size_t cm_region_table_workspace_bytes(int batch, int h, int w, int label_cap) {
    if (batch <= 0 || h <= 0 || w <= 0 || label_cap <= 0) return 0;
    return cm::carve_rt(nullptr, batch, h, w, label_cap).total;
}

static const int RT_MAX_CAP = 6000;   // (cap + 1) * 36 bytes of shared memory per block

extern "C" int cm_region_table_build(const void* labels, int label_bytes, const uint8_t* gt, int gt_packed, int batch, int h, int w, int label_cap,
                                     void* workspace, size_t workspace_bytes, void* stream) {
    using namespace cm;
    CM_DEVICE_OR_RETURN();
    CM_REQUIRE(labels && gt && workspace && (label_bytes == 2 || label_bytes == 4), "cm_region_table_build: NULL pointer or label_bytes not 2 / 4");
    CM_REQUIRE(batch > 0 && batch <= 65535 && h > 0 && w > 0 && h < 32767 && w < 32767 && (long long)h * w < (1ll << 30), "cm_region_table_build: bad shape");
    CM_REQUIRE(label_cap > 0 && label_cap <= RT_MAX_CAP, "cm_region_table_build: label_cap must be in 1..%d", RT_MAX_CAP);
    CM_REQUIRE(!gt_packed || ((long long)h * w) % 8 == 0, "cm_region_table_build: a bit-packed ground truth needs H*W divisible by 8");
    RtWs ws = carve_rt(workspace, batch, h, w, label_cap);
    if (workspace_bytes < ws.total) {
        set_error("cm_region_table_build: workspace %zu < %zu bytes", workspace_bytes, ws.total);
        return CM_ERR_WORKSPACE;
    }
    cudaStream_t st = as_stream(stream);
    const size_t smem = (size_t)(label_cap + 1) * 8 * sizeof(int);
    const int v = (label_bytes == 2 ? 0 : 2) + (gt_packed ? 1 : 0);
    {
        static PerDeviceMax smem_set[4];
        if (smem > 40 * 1024 && smem_set[v].raise(smem)) {
            const void* f = v == 0 ? (const void*)rt_table_kernel<short, false> : v == 1 ? (const void*)rt_table_kernel<short, true>
                          : v == 2 ? (const void*)rt_table_kernel<int, false> : (const void*)rt_table_kernel<int, true>;
            cudaFuncSetAttribute(f, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        }
    }
    {
        CM_TIMED("region_gt_table", st);
        const dim3 grid(ws.splits, batch);
        int* done = ws.hist + (size_t)batch * 512;
        if (v == 0) rt_table_kernel<short, false><<<grid, RT_THREADS, smem, st>>>(reinterpret_cast<const short*>(labels), gt, h, w, label_cap, ws.gtstat, ws.part, done);
        else if (v == 1) rt_table_kernel<short, true><<<grid, RT_THREADS, smem, st>>>(reinterpret_cast<const short*>(labels), gt, h, w, label_cap, ws.gtstat, ws.part, done);
        else if (v == 2) rt_table_kernel<int, false><<<grid, RT_THREADS, smem, st>>>(reinterpret_cast<const int*>(labels), gt, h, w, label_cap, ws.gtstat, ws.part, done);
        else rt_table_kernel<int, true><<<grid, RT_THREADS, smem, st>>>(reinterpret_cast<const int*>(labels), gt, h, w, label_cap, ws.gtstat, ws.part, done);
    }
    return post_launch("cm_region_table_build");
}

extern "C" int cm_region_table_metrics(const void* labels, int label_bytes, const uint8_t* gt_bytes, int batch, int h, int w, const float* mask_prob,
                                       const int32_t* node_ptr, const int32_t* id_of_label, int label_cap, float threshold, void* workspace,
                                       size_t workspace_bytes, int64_t* counts, double* abs_err, double* per_image, double* partial,
                                       double* per_image_sod, double* partial_sod, void* stream) {
    using namespace cm;
    CM_DEVICE_OR_RETURN();
    CM_REQUIRE(labels && workspace && mask_prob && node_ptr && counts && abs_err && per_image_sod && partial_sod && (per_image || partial),
               "cm_region_table_metrics: NULL pointer");
    CM_REQUIRE(label_bytes == 2 || label_bytes == 4, "cm_region_table_metrics: label_bytes not 2 / 4");
    CM_REQUIRE(batch > 0 && batch <= 65535 && h > 0 && w > 0 && (long long)h * w < (1ll << 30), "cm_region_table_metrics: bad shape");
    CM_REQUIRE(label_cap > 0 && label_cap <= RT_MAX_CAP, "cm_region_table_metrics: label_cap must be in 1..%d", RT_MAX_CAP);
    CM_REQUIRE(threshold >= 0.f && threshold < 1.f, "cm_region_table_metrics: threshold must be in [0, 1)");
    RtWs ws = carve_rt(workspace, batch, h, w, label_cap);
    if (workspace_bytes < ws.total) {
        set_error("cm_region_table_metrics: workspace %zu < %zu bytes", workspace_bytes, ws.total);
        return CM_ERR_WORKSPACE;
    }
    cudaStream_t st = as_stream(stream);
    const size_t smem = (size_t)(label_cap + 1) * (8 * sizeof(int) + sizeof(float));
    {
        static PerDeviceMax smem_set;
        if (smem > 32 * 1024 && smem_set.raise(smem)) cudaFuncSetAttribute(rt_metrics_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    }
    {
        CM_TIMED("region_table_metrics", st);
        rt_metrics_kernel<<<batch, 256, smem, st>>>(ws.part, ws.splits, label_cap, ws.gtstat, labels, label_bytes, mask_prob, node_ptr, id_of_label, threshold,
                                                    h, w, gt_bytes, reinterpret_cast<unsigned long long*>(counts), abs_err, ws.tot1,
                                                    ws.tot2, ws.shift, ws.hist, ws.curves, ws.hist + (size_t)batch * 512, batch, per_image, partial,
                                                    per_image_sod, partial_sod);
    }
    return post_launch("cm_region_table_metrics");
}
