// §8(f) N2: the Canny edge map that feeds region feature 13 -- skimage.feature.canny(gray, sigma=2) at
// models/region_graph/extract_rg_embeddings.py:151-152 (and its copies in train.py / test.py).
//
// scikit-image is not vendored by the reference (>=0.21.0, unpinned): this restates its published algorithm --
//   gray = image . [0.2989, 0.5870, 0.1140]                                            (:151)
//   smoothed = gaussian(gray, sigma, mode='constant') / (gaussian(ones, ...) + eps)    (bleed-over correction)
//   jsobel / isobel = scipy.ndimage.sobel(smoothed, axis=1 / 0)  (mode='reflect'), magnitude = sqrt(i^2 + j^2)
//   non-maximum suppression with linear interpolation along the gradient, 1-pixel border excluded,
//   hysteresis: pixels >= low (0.1) connected (8-neighbourhood) to a pixel >= high (0.2)
// in float64 with scipy's own operation order (symmetric / antisymmetric correlate1d: centre tap first, then the
// outermost pair inwards, pairs added before the multiply; no FMA contraction), so the boolean map agrees with the
// numpy/scipy restatement in oracle/canny.py except on exact floating-point ties.  PARITY UNPINNED (no skimage here).
//
// Four small kernels per batch (everything is L2-resident: 64 K pixels x 8 B per plane): vertical blur of the gray
// image and of the all-ones mask, horizontal blur + bleed-over division, Sobel + magnitude, NMS + thresholds, then
// one CTA per image grows the strong pixels through the weak ones in shared memory until nothing changes.
#include <math.h>
#include <stdlib.h>

#include "common.cuh"

namespace cm {

constexpr int CANNY_MAX_RADIUS = 32;
struct GaussW {   // weights, index r + radius; passed by value (a kernel argument is capture-safe, a symbol copy is not)
    double w[2 * CANNY_MAX_RADIUS + 1];
};

struct CannyWs {
    double *gray, *tmp_i, *tmp_m, *smooth, *isob, *jsob, *mag;
    uint8_t* cls;
    size_t total;
};
static CannyWs carve_canny(void* ws, int batch, int h, int w) {
    Carver c(ws);
    CannyWs r;
    const size_t n = (size_t)batch * h * w;
    r.gray = c.take<double>(n);
    r.tmp_i = c.take<double>(n);
    r.tmp_m = c.take<double>(n);
    r.smooth = c.take<double>(n);
    r.isob = c.take<double>(n);
    r.jsob = c.take<double>(n);
    r.mag = c.take<double>(n);
    r.cls = c.take<uint8_t>(n);
    r.total = align_up(c.off, 256);
    return r;
}

// gray (:151) and the vertical (axis 0) Gaussian pass of the image and of the all-ones mask, zero outside the image
template <bool U8>
__global__ void canny_gray_kernel(const void* __restrict__ image, long n_px, double* __restrict__ gray) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_px) return;
    double r, g, b;
    if (U8) {
        const uint8_t* p = reinterpret_cast<const uint8_t*>(image) + 3 * i;
        r = __ddiv_rn((double)p[0], 255.0); g = __ddiv_rn((double)p[1], 255.0); b = __ddiv_rn((double)p[2], 255.0);
    } else {
        const float* p = reinterpret_cast<const float*>(image) + 3 * i;
        r = p[0]; g = p[1]; b = p[2];
    }
    gray[i] = __dadd_rn(__dadd_rn(__dmul_rn(r, 0.2989), __dmul_rn(g, 0.5870)), __dmul_rn(b, 0.1140));
}

// scipy correlate1d, symmetric weights, mode='constant' cval 0: tmp = in[0]*w0; for j = radius..1: tmp += (in[-j] + in[+j]) * w_j
__global__ void canny_blur_v_kernel(const double* __restrict__ gray, int H, int W, int radius, const GaussW gw, double* __restrict__ out_i,
                                    double* __restrict__ out_m) {
    const double* c_gauss = gw.w;
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y, b = blockIdx.z;
    if (x >= W) return;
    const double* g = gray + (size_t)b * H * W;
    const size_t o = ((size_t)b * H + y) * W + x;
    double ti = __dmul_rn(g[(size_t)y * W + x], c_gauss[radius]), tm = c_gauss[radius];
    for (int j = radius; j >= 1; --j) {
        const int ya = y - j, yb = y + j;
        const double a = ya >= 0 ? g[(size_t)ya * W + x] : 0.0, c = yb < H ? g[(size_t)yb * W + x] : 0.0;
        const double ma = ya >= 0 ? 1.0 : 0.0, mc = yb < H ? 1.0 : 0.0;
        ti = __dadd_rn(ti, __dmul_rn(__dadd_rn(a, c), c_gauss[radius - j]));
        tm = __dadd_rn(tm, __dmul_rn(__dadd_rn(ma, mc), c_gauss[radius - j]));
    }
    out_i[o] = ti;
    out_m[o] = tm;
}
__global__ void canny_blur_h_kernel(const double* __restrict__ in_i, const double* __restrict__ in_m, int H, int W, int radius,
                                    const GaussW gw, double* __restrict__ smooth) {
    const double* c_gauss = gw.w;
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y, b = blockIdx.z;
    if (x >= W) return;
    const size_t row = ((size_t)b * H + y) * W;
    double ti = __dmul_rn(in_i[row + x], c_gauss[radius]), tm = __dmul_rn(in_m[row + x], c_gauss[radius]);
    for (int j = radius; j >= 1; --j) {
        const int xa = x - j, xb = x + j;
        const double a = xa >= 0 ? in_i[row + xa] : 0.0, c = xb < W ? in_i[row + xb] : 0.0;
        const double ma = xa >= 0 ? in_m[row + xa] : 0.0, mc = xb < W ? in_m[row + xb] : 0.0;
        ti = __dadd_rn(ti, __dmul_rn(__dadd_rn(a, c), c_gauss[radius - j]));
        tm = __dadd_rn(tm, __dmul_rn(__dadd_rn(ma, mc), c_gauss[radius - j]));
    }
    smooth[row + x] = __ddiv_rn(ti, __dadd_rn(tm, 2.220446049250313e-16));   // + np.finfo(float64).eps
}

__device__ __forceinline__ int reflect_idx(int i, int n) {   // scipy mode='reflect': (d c b a | a b c d | d c b a)
    if (i < 0) i = -i - 1;
    if (i >= n) i = 2 * n - 1 - i;
    return min(max(i, 0), n - 1);
}
// ndi.sobel(s, axis=1): d = s[.,x+1] - s[.,x-1], then 2*d[y] + (d[y-1] + d[y+1]); axis=0 alike with the roles swapped
__global__ void canny_sobel_kernel(const double* __restrict__ smooth, int H, int W, double* __restrict__ isob, double* __restrict__ jsob,
                                   double* __restrict__ mag) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y, b = blockIdx.z;
    if (x >= W) return;
    const double* s = smooth + (size_t)b * H * W;
    auto S = [&](int yy, int xx) { return s[(size_t)reflect_idx(yy, H) * W + reflect_idx(xx, W)]; };
    auto dx = [&](int yy) { return __dsub_rn(S(yy, x + 1), S(yy, x - 1)); };
    auto dy = [&](int xx) { return __dsub_rn(S(y + 1, xx), S(y - 1, xx)); };
    // the derivative pass runs over the whole image first, the smoothing pass reflects ITS input rows / columns
    const int ym = reflect_idx(y - 1, H), yp = reflect_idx(y + 1, H), xm = reflect_idx(x - 1, W), xp = reflect_idx(x + 1, W);
    const double js = __dadd_rn(__dmul_rn(dx(y), 2.0), __dadd_rn(dx(ym), dx(yp)));
    const double is = __dadd_rn(__dmul_rn(dy(x), 2.0), __dadd_rn(dy(xm), dy(xp)));
    const size_t o = ((size_t)b * H + y) * W + x;
    isob[o] = is;
    jsob[o] = js;
    mag[o] = __dsqrt_rn(__dadd_rn(__dmul_rn(is, is), __dmul_rn(js, js)));
}

// non-maximum suppression with bilinear interpolation + double threshold: 0 = no edge, 1 = weak (>= low), 2 = strong (>= high)
__global__ void canny_nms_kernel(const double* __restrict__ isob, const double* __restrict__ jsob, const double* __restrict__ mag, int H, int W,
                                 double low, double high, uint8_t* __restrict__ cls) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y, b = blockIdx.z;
    if (x >= W) return;
    const size_t base = (size_t)b * H * W, o = base + (size_t)y * W + x;
    uint8_t out = 0;
    const double m = mag[o];
    if (y > 0 && y < H - 1 && x > 0 && x < W - 1 && m >= low) {   // eroded mask: the 1-pixel border never carries an edge
        const double is = isob[o], js = jsob[o];
        const bool down = is <= 0, up = is >= 0, left = js <= 0, right = js >= 0;
        const bool c1 = (up && right) || (down && left), c2 = (down && right) || (up && left);
        if (c1 || c2) {
            const double ai = fabs(is), aj = fabs(js);
            auto M = [&](int dyy, int dxx) { return mag[base + (size_t)(y + dyy) * W + (x + dxx)]; };
            double w, n11, n12, n21, n22;
            if (c1) {
                if (ai > aj) { w = __ddiv_rn(aj, ai); n11 = M(1, 0); n12 = M(1, 1); n21 = M(-1, 0); n22 = M(-1, -1); }
                else { w = __ddiv_rn(ai, aj); n11 = M(0, 1); n12 = M(1, 1); n21 = M(0, -1); n22 = M(-1, -1); }
            } else {
                if (ai < aj) { w = __ddiv_rn(ai, aj); n11 = M(0, 1); n12 = M(-1, 1); n21 = M(0, -1); n22 = M(1, -1); }
                else { w = __ddiv_rn(aj, ai); n11 = M(-1, 0); n12 = M(-1, 1); n21 = M(1, 0); n22 = M(1, -1); }
            }
            const double omw = __dsub_rn(1.0, w);
            const bool plus = __dadd_rn(__dmul_rn(n12, w), __dmul_rn(n11, omw)) <= m;
            const bool minus = __dadd_rn(__dmul_rn(n22, w), __dmul_rn(n21, omw)) <= m;
            if (plus && minus) out = m >= high ? 2 : 1;
        }
    }
    cls[o] = out;
}

// hysteresis: one CTA per image; strong pixels grow through 8-connected weak pixels until a sweep changes nothing.
// A thread owns a run of consecutive pixels and walks it forwards then backwards, so growth travels along rows
// within a sweep.  State in shared memory when the image fits (<= 200 KB), in place in global memory otherwise.
__global__ void __launch_bounds__(1024) canny_hysteresis_kernel(uint8_t* __restrict__ cls, int H, int W, int use_smem, uint8_t* __restrict__ edges) {
    extern __shared__ uint8_t hy_s[];
    const int b = blockIdx.x, n = H * W, tid = threadIdx.x;
    uint8_t* g = cls + (size_t)b * n;
    uint8_t* s = use_smem ? hy_s : g;
    if (use_smem) {
        for (int i = tid; i < n; i += blockDim.x) s[i] = g[i];
        __syncthreads();
    }
    const int per = (n + blockDim.x - 1) / blockDim.x, i0 = tid * per, i1 = min(n, i0 + per);
    auto grow = [&](int i) -> bool {
        if (s[i] != 1) return false;
        const int y = i / W, x = i - y * W;
        bool hit = false;
        for (int dy = -1; dy <= 1 && !hit; ++dy) {
            const int yy = y + dy;
            if (yy < 0 || yy >= H) continue;
            for (int dx = -1; dx <= 1; ++dx) {
                const int xx = x + dx;
                if (xx < 0 || xx >= W) continue;
                if (s[yy * W + xx] == 2) { hit = true; break; }
            }
        }
        if (hit) s[i] = 2;   // benign race: a neighbour may read 1 or 2; either way it converges to the same fixed point
        return hit;
    };
    for (;;) {
        bool changed = false;
        for (int i = i0; i < i1; ++i) changed |= grow(i);
        for (int i = i1 - 1; i >= i0; --i) changed |= grow(i);
        if (!__syncthreads_or(changed)) break;
    }
    for (int i = tid; i < n; i += blockDim.x) edges[(size_t)b * n + i] = s[i] == 2 ? 1 : 0;
}

// hysteresis as connected components: union-find (16-bit pixel indices in shared memory, H*W <= 65536) over the pixels
// >= low under 8-connectivity, one bit per component root that holds a pixel >= high, then every weak pixel looks its
// root up.  Same fixed point as growing the strong pixels sweep by sweep, without the sweeps.
__device__ __forceinline__ int hy_find(const unsigned short* par, int x) {
    int p = par[x];
    while (p != x) {
        x = p;
        p = par[x];
    }
    return x;
}
__device__ __forceinline__ void hy_unite(unsigned short* par, int a, int b) {
    while (true) {
        a = hy_find(par, a);
        b = hy_find(par, b);
        if (a == b) return;
        if (a > b) {
            const int t = a;
            a = b;
            b = t;
        }
        const unsigned short old = atomicCAS(par + b, (unsigned short)b, (unsigned short)a);
        if (old == (unsigned short)b) return;
        b = old;
    }
}
__global__ void __launch_bounds__(1024) canny_hysteresis_uf_kernel(const uint8_t* __restrict__ cls, int H, int W, uint8_t* __restrict__ edges) {
    extern __shared__ unsigned short hy_par[];
    const int b = blockIdx.x, n = H * W, tid = threadIdx.x;
    const uint8_t* g = cls + (size_t)b * n;
    unsigned* strong = reinterpret_cast<unsigned*>(hy_par + ((n + 7) & ~7));
    for (int i = tid; i < n; i += blockDim.x) hy_par[i] = (unsigned short)i;
    for (int i = tid; i < (n + 31) / 32; i += blockDim.x) strong[i] = 0;
    __syncthreads();
    for (int i = tid; i < n; i += blockDim.x) {
        if (!g[i]) continue;
        const int y = i / W, x = i - y * W;
        if (x + 1 < W && g[i + 1]) hy_unite(hy_par, i, i + 1);
        if (y + 1 < H) {
            if (g[i + W]) hy_unite(hy_par, i, i + W);
            if (x > 0 && g[i + W - 1]) hy_unite(hy_par, i, i + W - 1);
            if (x + 1 < W && g[i + W + 1]) hy_unite(hy_par, i, i + W + 1);
        }
    }
    __syncthreads();
    for (int i = tid; i < n; i += blockDim.x)
        if (g[i] == 2) {
            const int r = hy_find(hy_par, i);
            atomicOr(&strong[r >> 5], 1u << (r & 31));
        }
    __syncthreads();
    for (int i = tid; i < n; i += blockDim.x) {
        uint8_t e = 0;
        if (g[i]) {
            const int r = hy_find(hy_par, i);
            e = (strong[r >> 5] >> (r & 31)) & 1u;
        }
        edges[(size_t)b * n + i] = e;
    }
}

}  // namespace cm

extern "C" size_t cm_canny_workspace_bytes(int batch, int h, int w) {
    if (batch <= 0 || h <= 0 || w <= 0) return 0;
    return cm::carve_canny(nullptr, batch, h, w).total;
}

extern "C" int cm_canny(const void* image, int image_is_u8, int batch, int h, int w, double sigma, double low_threshold, double high_threshold,
                        void* workspace, size_t workspace_bytes, uint8_t* edges, void* stream) {
    using namespace cm;
    CM_DEVICE_OR_RETURN();
    CM_REQUIRE(image && workspace && edges && batch > 0 && batch <= 65535 && h > 2 && w > 2 && h <= 65535, "cm_canny: bad arguments");
    CM_REQUIRE(sigma > 0.0 && low_threshold <= high_threshold, "cm_canny: sigma > 0 and low <= high required");
    const int radius = (int)(4.0 * sigma + 0.5);   // scipy gaussian_filter truncate=4.0
    CM_REQUIRE(radius >= 1 && radius <= CANNY_MAX_RADIUS, "cm_canny: sigma %.3f needs radius %d > %d", sigma, radius, CANNY_MAX_RADIUS);
    CannyWs ws = carve_canny(workspace, batch, h, w);
    if (workspace_bytes < ws.total) {
        set_error("cm_canny: workspace %zu < %zu bytes", workspace_bytes, ws.total);
        return CM_ERR_WORKSPACE;
    }
    cudaStream_t st = as_stream(stream);
    // scipy _gaussian_kernel1d: phi = exp(-0.5 / sigma^2 * x^2), normalised by its sum
    GaussW gw;
    double* wts = gw.w;
    double sum = 0.0;
    const double s2 = sigma * sigma;
    for (int i = -radius; i <= radius; ++i) {
        wts[i + radius] = exp(-0.5 / s2 * (double)(i * i));
        sum += wts[i + radius];
    }
    for (int i = 0; i <= 2 * radius; ++i) wts[i] /= sum;
    const long n_px = (long)batch * h * w;
    dim3 grid(ceil_div(w, 128), h, batch);
    {
        CM_TIMED("canny_blur", st);
        if (image_is_u8) canny_gray_kernel<true><<<ceil_div(n_px, 256), 256, 0, st>>>(image, n_px, ws.gray);
        else canny_gray_kernel<false><<<ceil_div(n_px, 256), 256, 0, st>>>(image, n_px, ws.gray);
        canny_blur_v_kernel<<<grid, 128, 0, st>>>(ws.gray, h, w, radius, gw, ws.tmp_i, ws.tmp_m);
        canny_blur_h_kernel<<<grid, 128, 0, st>>>(ws.tmp_i, ws.tmp_m, h, w, radius, gw, ws.smooth);
    }
    {
        CM_TIMED("canny_sobel_nms", st);
        canny_sobel_kernel<<<grid, 128, 0, st>>>(ws.smooth, h, w, ws.isob, ws.jsob, ws.mag);
        canny_nms_kernel<<<grid, 128, 0, st>>>(ws.isob, ws.jsob, ws.mag, h, w, low_threshold, high_threshold, ws.cls);
    }
    if ((size_t)h * w <= 65536 && !getenv("CM_CANNY_SWEEP_HYSTERESIS")) {
        const size_t n = (size_t)h * w, smem = ((n + 7) & ~(size_t)7) * 2 + ((n + 31) / 32) * 4;
        static PerDeviceOnce attr2;
        if (attr2.first()) cudaFuncSetAttribute(canny_hysteresis_uf_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 144 * 1024);
        CM_TIMED("canny_hysteresis", st);
        canny_hysteresis_uf_kernel<<<batch, 1024, smem, st>>>(ws.cls, h, w, edges);
    } else {
        const size_t smem = (size_t)h * w;
        const int use_smem = smem <= 200 * 1024;
        static PerDeviceOnce attr;
        if (attr.first()) {
            cudaFuncSetAttribute(canny_hysteresis_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        }
        CM_TIMED("canny_hysteresis", st);
        canny_hysteresis_kernel<<<batch, 1024, use_smem ? smem : 0, st>>>(ws.cls, h, w, use_smem, edges);
    }
    return post_launch("cm_canny");
}
