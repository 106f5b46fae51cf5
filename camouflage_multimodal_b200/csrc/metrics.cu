// A8 + A9 + A13: pixel-mask paint through the label map fused with the segmentation-metric reductions.
//
// Reference: models/region_graph/test.py:241-244 (paint), utils/metrics.py:9-92 (IoU, Dice, P/R/F1, MAE,
// accuracy, batch mean/std), models/multimodal/train_multimodal.py:197-220 (image-level F1 counts).
// One pass over the pixels; HBM-bound on labels (i32) + gt (u8).  Confusion counts are
// integers (bit-exact); the |pred-gt| sum is float64.  Per-thread partials -> warp-shuffle tree -> one
// atomic per warp-leader into the image's accumulators.
#include "common.cuh"
#include "metrics_body.cuh"

namespace cm {

__device__ __forceinline__ void flush_partials(int tp, int sp, int sg, int eq, double ae, unsigned long long* counts, double* abs_err) {
    tp = warp_sum(tp);
    sp = warp_sum(sp);
    sg = warp_sum(sg);
    eq = warp_sum(eq);
    ae = warp_sum(ae);
    if ((threadIdx.x & 31) == 0) {
        if (tp) atomicAdd(counts + 0, (unsigned long long)tp);
        if (sp) atomicAdd(counts + 1, (unsigned long long)sp);
        if (sg) atomicAdd(counts + 2, (unsigned long long)sg);
        if (eq) atomicAdd(counts + 3, (unsigned long long)eq);
        atomicAdd(abs_err, ae);
    }
}

// grid (chunks, B), block 256; each thread walks pixels with stride gridDim.x*256
__device__ __forceinline__ int4 load4_labels(const int* p, int q) { return reinterpret_cast<const int4*>(p)[q]; }
__device__ __forceinline__ int4 load4_labels(const short* p, int q) {
    const uint2 a = reinterpret_cast<const uint2*>(p)[q];
    return make_int4((int)(short)(a.x & 0xffffu), (int)(short)(a.x >> 16), (int)(short)(a.y & 0xffffu), (int)(short)(a.y >> 16));
}
template <typename LT>
__global__ void __launch_bounds__(256) paint_metrics_kernel(const LT* __restrict__ labels, const uint8_t* __restrict__ gt, int HW,
                                                            const float* __restrict__ mask_prob, const int* __restrict__ node_ptr,
                                                            const int* __restrict__ id_of_label, int cap, float threshold,
                                                            float* __restrict__ pred_mask, unsigned long long* __restrict__ counts,
                                                            double* __restrict__ abs_err) {
    const int b = blockIdx.y;
    const int n0 = node_ptr[b], nb = node_ptr[b + 1] - n0;
    const LT* lab = labels + (size_t)b * HW;
    const uint8_t* g = gt + (size_t)b * HW;
    const int* idm = id_of_label ? id_of_label + (size_t)b * cap : nullptr;
    const double thr = (double)threshold;
    int tp = 0, sp = 0, sg = 0, eq = 0;
    double ae = 0.0;
    __shared__ double gt_lut[256];   // byte -> byte / 255.0 (test.py:275), the same fp64 quotient without a division per pixel
    gt_lut[threadIdx.x] = (double)threadIdx.x / 255.0;
    __syncthreads();
    auto account = [&](float pr, unsigned gbyte) {
        const double gv = gt_lut[gbyte];
        const bool pb = (double)pr > thr, gb = gv > thr;
        tp += pb && gb;
        sp += pb;
        sg += gb;
        eq += pb == gb;
        ae += fabs((double)pr - gv);
    };
    auto prob_of = [&](int l) {
        int idx = l;
        if (idm) idx = (l >= 0 && l < cap) ? idm[l] : -1;
        return (idx >= 0 && idx < nb) ? mask_prob[n0 + idx] : 0.f;
    };
    if ((HW & 3) == 0 && ((reinterpret_cast<uintptr_t>(lab) | reinterpret_cast<uintptr_t>(g)) & 15) == 0) {
        // four pixels per thread per trip: one 16-byte label load, one 4-byte gt load, the four dependent
        // label -> node id -> probability gathers in flight together
        for (int q = blockIdx.x * 256 + threadIdx.x; q < HW / 4; q += gridDim.x * 256) {
            const int4 l4 = load4_labels(lab, q);
            const unsigned g4 = reinterpret_cast<const unsigned*>(g)[q];
            const float p0 = prob_of(l4.x), p1 = prob_of(l4.y), p2 = prob_of(l4.z), p3 = prob_of(l4.w);
            if (pred_mask) reinterpret_cast<float4*>(pred_mask + (size_t)b * HW)[q] = make_float4(p0, p1, p2, p3);
            account(p0, g4 & 255u);
            account(p1, (g4 >> 8) & 255u);
            account(p2, (g4 >> 16) & 255u);
            account(p3, g4 >> 24);
        }
    } else {
        for (int p = blockIdx.x * 256 + threadIdx.x; p < HW; p += gridDim.x * 256) {
            const float pr = prob_of((int)lab[p]);
            if (pred_mask) pred_mask[(size_t)b * HW + p] = pr;
            account(pr, g[p]);
        }
    }
    flush_partials(tp, sp, sg, eq, ae, counts + 4 * (size_t)b, abs_err + b);
}

__global__ void __launch_bounds__(256) metrics_dense_kernel(const float* __restrict__ pred, const float* __restrict__ gt, long long pixels,
                                                            float threshold, unsigned long long* __restrict__ counts,
                                                            double* __restrict__ abs_err) {
    const int b = blockIdx.y;
    const float* pr = pred + (size_t)b * pixels;
    const float* g = gt + (size_t)b * pixels;
    int tp = 0, sp = 0, sg = 0, eq = 0;
    double ae = 0.0;
    for (long long p = (long long)blockIdx.x * 256 + threadIdx.x; p < pixels; p += (long long)gridDim.x * 256) {
        const float a = pr[p], c = g[p];
        const bool pb = a > threshold, gb = c > threshold;
        tp += pb && gb;
        sp += pb;
        sg += gb;
        eq += pb == gb;
        ae += fabs((double)a - (double)c);
    }
    flush_partials(tp, sp, sg, eq, ae, counts + 4 * (size_t)b, abs_err + b);
}

// single CTA of 256 threads
__global__ void __launch_bounds__(256) metrics_finalize_kernel(const unsigned long long* __restrict__ counts, const double* __restrict__ abs_err,
                                                               int batch, long long pixels, double* __restrict__ per_image,
                                                               double* __restrict__ partial) {
    __shared__ double red[8][14];
    metrics_finalize_block(counts, abs_err, batch, pixels, per_image, partial, red);
}

__global__ void class_confusion_kernel(const float* __restrict__ mask_out, const long long* __restrict__ true_cls, int n, int c,
                                       unsigned long long* counts4) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    int tp = 0, fp = 0, fn = 0, tn = 0;
    if (i < n) {
        const float* r = mask_out + (size_t)i * c;
        int arg = 0;
        float best = r[0];
        for (int k = 1; k < c; ++k)
            if (r[k] > best) {   // first maximum wins, like torch.argmax
                best = r[k];
                arg = k;
            }
        const long long t = true_cls[i];
        tp = arg == 1 && t == 1;
        fp = arg == 1 && t == 0;
        fn = arg == 0 && t == 1;
        tn = arg == 0 && t == 0;
    }
    tp = warp_sum(tp);
    fp = warp_sum(fp);
    fn = warp_sum(fn);
    tn = warp_sum(tn);
    if ((threadIdx.x & 31) == 0) {
        if (tp) atomicAdd(counts4 + 0, (unsigned long long)tp);
        if (fp) atomicAdd(counts4 + 1, (unsigned long long)fp);
        if (fn) atomicAdd(counts4 + 2, (unsigned long long)fn);
        if (tn) atomicAdd(counts4 + 3, (unsigned long long)tn);
    }
}

// one thread per packed byte: bit i (little-endian bit order) -> byte 8 * idx + i = one_value or 0, one 8-byte store
__global__ void unpack_bits_kernel(const uint8_t* __restrict__ bits, long long n_bytes, unsigned one_value, uint8_t* __restrict__ out) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_bytes) return;
    const unsigned b = bits[i];
    uint2 v;
    v.x = ((b & 1u) ? one_value : 0u) | ((b & 2u) ? one_value << 8 : 0u) | ((b & 4u) ? one_value << 16 : 0u) | ((b & 8u) ? one_value << 24 : 0u);
    v.y = ((b & 16u) ? one_value : 0u) | ((b & 32u) ? one_value << 8 : 0u) | ((b & 64u) ? one_value << 16 : 0u) | ((b & 128u) ? one_value << 24 : 0u);
    reinterpret_cast<uint2*>(out)[i] = v;
}

// Host -> device staging by the SMs: `src` is pinned host memory the device can address (cudaHostAlloc under unified
// addressing), read with 16-byte loads, eight per thread in flight, by a fixed small grid -- the copy is then a node of
// the captured step like any other kernel: no copy-engine channel to schedule next to the compute channels, no extra call.
__global__ void __launch_bounds__(256) ingest_copy_kernel(const uint4* __restrict__ src, uint4* __restrict__ dst, long long n16) {
    const long long stride = (long long)gridDim.x * 256;
    long long i = (long long)blockIdx.x * 256 + threadIdx.x;
    for (; i + 7 * stride < n16; i += 8 * stride) {
        uint4 v[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) v[u] = __ldcv(src + i + u * stride);
#pragma unroll
        for (int u = 0; u < 8; ++u) dst[i + u * stride] = v[u];
    }
    for (; i < n16; i += stride) dst[i] = __ldcv(src + i);
}

// 10-bit labels, little-endian bit stream (label i = bits [10 i, 10 i + 10)): one thread per 16 labels = five 32-bit words
// in, two 16-byte stores out
__global__ void unpack_labels10_kernel(const uint32_t* __restrict__ packed, long long n_groups, short* __restrict__ out) {
    const long long gidx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (gidx >= n_groups) return;
    const uint32_t* p = packed + gidx * 5;
    const uint32_t w[6] = {p[0], p[1], p[2], p[3], p[4], 0u};
    uint32_t o[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        uint32_t v[2];
#pragma unroll
        for (int u = 0; u < 2; ++u) {
            const int bit = 10 * (2 * k + u), wi = bit >> 5, sh = bit & 31;
            const unsigned long long two = (unsigned long long)w[wi] | ((unsigned long long)w[wi + 1] << 32);
            v[u] = (uint32_t)(two >> sh) & 1023u;
        }
        o[k] = v[0] | (v[1] << 16);
    }
    uint4* dst = reinterpret_cast<uint4*>(out + gidx * 16);
    dst[0] = make_uint4(o[0], o[1], o[2], o[3]);
    dst[1] = make_uint4(o[4], o[5], o[6], o[7]);
}

static int metric_chunks(long long pixels, int batch) {
    // about four CTAs per SM over the whole batch, at least 1 per image, at most one per 256 pixels
    long long want = (4LL * num_sms() + batch - 1) / batch;
    long long most = (pixels + 255) / 256;
    if (want > most) want = most;
    if (want < 1) want = 1;
    return (int)want;
}

}  // namespace cm

static int paint_metrics_impl(const void* labels_v, int label_bytes, const uint8_t* gt, int batch, int h, int w, const float* mask_prob,
                                const int32_t* node_ptr, const int32_t* id_of_label, int label_cap, float threshold,
                                float* pred_mask, int64_t* counts, double* abs_err, void* stream) {
    const int32_t* labels = reinterpret_cast<const int32_t*>(labels_v);
    using namespace cm;
    CM_DEVICE_OR_RETURN();
    CM_REQUIRE(labels && gt && mask_prob && node_ptr && counts && abs_err, "cm_paint_metrics: NULL pointer");
    CM_REQUIRE(batch > 0 && h > 0 && w > 0 && batch <= 65535, "cm_paint_metrics: bad shape");
    CM_REQUIRE(!id_of_label || label_cap > 0, "cm_paint_metrics: id_of_label needs label_cap");
    cudaStream_t st = as_stream(stream);
    cudaMemsetAsync(counts, 0, sizeof(int64_t) * 4 * batch, st);
    cudaMemsetAsync(abs_err, 0, sizeof(double) * batch, st);
    dim3 grid(metric_chunks((long long)h * w, batch), batch);
    {
        CM_TIMED("paint_metrics", st);
        if (label_bytes == 2)
            paint_metrics_kernel<short><<<grid, 256, 0, st>>>(reinterpret_cast<const short*>(labels_v), gt, h * w, mask_prob, node_ptr, id_of_label,
                                                              label_cap, threshold, pred_mask, reinterpret_cast<unsigned long long*>(counts), abs_err);
        else
            paint_metrics_kernel<int><<<grid, 256, 0, st>>>(labels, gt, h * w, mask_prob, node_ptr, id_of_label, label_cap, threshold,
                                                            pred_mask, reinterpret_cast<unsigned long long*>(counts), abs_err);
    }
    return post_launch("cm_paint_metrics");
}

extern "C" int cm_unpack_bits(const uint8_t* bits, int64_t n_bytes, int one_value, uint8_t* out, void* stream) {
    using namespace cm;
    CM_DEVICE_OR_RETURN();
    CM_REQUIRE(bits && out && n_bytes > 0 && one_value >= 0 && one_value <= 255, "cm_unpack_bits: bad arguments");
    CM_REQUIRE((reinterpret_cast<uintptr_t>(out) & 7) == 0, "cm_unpack_bits: the output must be 8-byte aligned");
    {
        CM_TIMED("unpack_bits", as_stream(stream));
        unpack_bits_kernel<<<ceil_div(n_bytes, 256), 256, 0, as_stream(stream)>>>(bits, n_bytes, (unsigned)one_value, out);
    }
    return post_launch("cm_unpack_bits");
}

extern "C" int cm_unpack_labels10(const uint8_t* packed, int64_t n_labels, int16_t* out, void* stream) {
    using namespace cm;
    CM_DEVICE_OR_RETURN();
    CM_REQUIRE(packed && out && n_labels > 0 && n_labels % 16 == 0, "cm_unpack_labels10: NULL pointer or a label count that is not a multiple of 16");
    CM_REQUIRE((reinterpret_cast<uintptr_t>(packed) & 3) == 0 && (reinterpret_cast<uintptr_t>(out) & 15) == 0,
               "cm_unpack_labels10: packed must be 4-byte aligned, out 16-byte aligned");
    {
        CM_TIMED("unpack_labels10", as_stream(stream));
        unpack_labels10_kernel<<<ceil_div(n_labels / 16, 256), 256, 0, as_stream(stream)>>>(reinterpret_cast<const uint32_t*>(packed), n_labels / 16,
                                                                                             reinterpret_cast<short*>(out));
    }
    return post_launch("cm_unpack_labels10");
}

extern "C" int cm_ingest_copy(const void* host_src, void* dst, int64_t n_bytes, int ctas, void* stream) {
    using namespace cm;
    CM_DEVICE_OR_RETURN();
    CM_REQUIRE(host_src && dst && n_bytes > 0 && n_bytes % 16 == 0, "cm_ingest_copy: NULL pointer or a size that is not a multiple of 16");
    CM_REQUIRE(((reinterpret_cast<uintptr_t>(host_src) | reinterpret_cast<uintptr_t>(dst)) & 15) == 0, "cm_ingest_copy: pointers must be 16-byte aligned");
    if (ctas <= 0) ctas = 32;
    {
        CM_TIMED("ingest_copy", as_stream(stream));
        ingest_copy_kernel<<<ctas, 256, 0, as_stream(stream)>>>(reinterpret_cast<const uint4*>(host_src), reinterpret_cast<uint4*>(dst), n_bytes / 16);
    }
    return post_launch("cm_ingest_copy");
}

extern "C" int cm_metrics_dense(const float* pred, const float* gt, int batch, int64_t pixels, float threshold,
                                int64_t* counts, double* abs_err, void* stream) {
    using namespace cm;
    CM_DEVICE_OR_RETURN();
    CM_REQUIRE(pred && gt && counts && abs_err, "cm_metrics_dense: NULL pointer");
    CM_REQUIRE(batch > 0 && batch <= 65535 && pixels > 0, "cm_metrics_dense: bad shape");
    cudaStream_t st = as_stream(stream);
    cudaMemsetAsync(counts, 0, sizeof(int64_t) * 4 * batch, st);
    cudaMemsetAsync(abs_err, 0, sizeof(double) * batch, st);
    dim3 grid(metric_chunks(pixels, batch), batch);
    {
        CM_TIMED("metrics_dense", st);
        metrics_dense_kernel<<<grid, 256, 0, st>>>(pred, gt, pixels, threshold, reinterpret_cast<unsigned long long*>(counts), abs_err);
    }
    return post_launch("cm_metrics_dense");
}

extern "C" int cm_metrics_finalize(const int64_t* counts, const double* abs_err, int batch, int64_t pixels,
                                   double* per_image, double* partial, void* stream) {
    using namespace cm;
    CM_DEVICE_OR_RETURN();
    CM_REQUIRE(counts && abs_err && (per_image || partial), "cm_metrics_finalize: NULL pointer");
    CM_REQUIRE(batch > 0 && pixels > 0, "cm_metrics_finalize: bad shape");
    {
        CM_TIMED("metrics_finalize", as_stream(stream));
        metrics_finalize_kernel<<<1, 256, 0, as_stream(stream)>>>(reinterpret_cast<const unsigned long long*>(counts), abs_err, batch,
                                                                  pixels, per_image, partial);
    }
    return post_launch("cm_metrics_finalize");
}

extern "C" int cm_class_confusion(const float* mask_out, const int64_t* true_cls, int n, int num_classes,
                                  int64_t* counts4, void* stream) {
    using namespace cm;
    CM_DEVICE_OR_RETURN();
    CM_REQUIRE(mask_out && true_cls && counts4, "cm_class_confusion: NULL pointer");
    CM_REQUIRE(n > 0 && num_classes >= 2, "cm_class_confusion: bad shape");
    {
        CM_TIMED("class_confusion", as_stream(stream));
        class_confusion_kernel<<<ceil_div(n, 256), 256, 0, as_stream(stream)>>>(
            mask_out, reinterpret_cast<const long long*>(true_cls), n, num_classes, reinterpret_cast<unsigned long long*>(counts4));
    }
    return post_launch("cm_class_confusion");
}

extern "C" int cm_paint_metrics(const int32_t* labels, const uint8_t* gt, int batch, int h, int w, const float* mask_prob,
                                const int32_t* node_ptr, const int32_t* id_of_label, int label_cap, float threshold,
                                float* pred_mask, int64_t* counts, double* abs_err, void* stream) {
    return paint_metrics_impl(labels, 4, gt, batch, h, w, mask_prob, node_ptr, id_of_label, label_cap, threshold, pred_mask, counts, abs_err, stream);
}

extern "C" int cm_paint_metrics_i16(const int16_t* labels, const uint8_t* gt, int batch, int h, int w, const float* mask_prob,
                                const int32_t* node_ptr, const int32_t* id_of_label, int label_cap, float threshold,
                                float* pred_mask, int64_t* counts, double* abs_err, void* stream) {
    return paint_metrics_impl(labels, 2, gt, batch, h, w, mask_prob, node_ptr, id_of_label, label_cap, threshold, pred_mask, counts, abs_err, stream);
}
