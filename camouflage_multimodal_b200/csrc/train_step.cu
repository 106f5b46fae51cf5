// SURVEY 8(f) N4 -- the pieces of the reference's training loops around the model passes:
//   cm_region_targets   per-region ground-truth labels from the three byte masks (models/region_graph/train.py:169-176)
//   cm_rg_node_loss     weighted CE (mask, instance) + BCE-with-logits(pos_weight) on the edge logit, task weights,
//                       value and gradients in one pass pair (train.py:325-353)
//   cm_image_loss       AggressiveFocalLoss * 3 + CE + 0.5 BCE-with-logits + 0.3 MSE per image, summed over the batch
//                       (models/multimodal/train_multimodal.py:29-57, 253-268: one backward per sample)
//   cm_grad_sq_norm, cm_adamw_step   clip_grad_norm_(max_norm) and torch.optim.AdamW over a table of tensors, the clip
//                       coefficient taken from the device-side norm (train.py:356-357, train_multimodal.py:276-277)
#include <math.h>

#include <algorithm>

#include "common.cuh"

namespace cm {

// ------------------------------------------------------------------------------------------ region targets
// One warp per region over its bounding box: byte sums of the three masks over the member pixels.  The reference
// thresholds the float32 mean of byte/255 (> 0.5, > 0.5, > 0.3); on integers that is 2 S > 255 n and 10 S > 765 n.
__global__ void __launch_bounds__(256) region_targets_kernel(const int* __restrict__ labels, const uint8_t* __restrict__ mask,
                                                             const uint8_t* __restrict__ inst, const uint8_t* __restrict__ edge,
                                                             int batch, int H, int W, const int* __restrict__ node_ptr,
                                                             const int* __restrict__ node_label, const int* __restrict__ node_bbox,
                                                             int node_cap, long long* __restrict__ y, long long* __restrict__ inst_y,
                                                             float* __restrict__ edge_y) {
    const int lane = threadIdx.x & 31, gi = blockIdx.x * 8 + (threadIdx.x >> 5);
    const int total = min(node_ptr[batch], node_cap);
    if (gi >= total) return;
    int lo = 0, hi = batch;   // image of the region: largest b with node_ptr[b] <= gi
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (node_ptr[mid] <= gi) lo = mid; else hi = mid;
    }
    const int L = node_label[gi];
    const int4 bb = *reinterpret_cast<const int4*>(node_bbox + 4 * (size_t)gi);
    const int ww = bb.z - bb.x + 1, npx = ww * (bb.w - bb.y + 1);
    const size_t base = (size_t)lo * H * W;
    unsigned n = 0, sm = 0, si = 0, se = 0;
    for (int i = lane; i < npx; i += 32) {
        const int yy = i / ww, xx = i - yy * ww;
        const size_t p = base + (size_t)(bb.y + yy) * W + bb.x + xx;
        if (labels[p] == L) {
            ++n;
            sm += mask[p];
            si += inst[p];
            se += edge[p];
        }
    }
    n = __reduce_add_sync(0xffffffffu, n);
    sm = __reduce_add_sync(0xffffffffu, sm);
    si = __reduce_add_sync(0xffffffffu, si);
    se = __reduce_add_sync(0xffffffffu, se);
    if (lane == 0) {
        y[gi] = 2ull * sm > 255ull * n ? 1 : 0;
        inst_y[gi] = 2ull * si > 255ull * n ? 1 : 0;
        edge_y[gi] = 10ull * se > 765ull * n ? 1.f : 0.f;
    }
}

// ------------------------------------------------------------------------------------------ losses
__device__ __forceinline__ double block_sum(double v, double* sh) {
    v = warp_sum(v);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    __syncthreads();
    if (lane == 0) sh[warp] = v;
    __syncthreads();
    double t = 0.0;
    if (warp == 0) {
        t = lane < (int)(blockDim.x >> 5) ? sh[lane] : 0.0;
        t = warp_sum(t);
    }
    return t;   // valid in warp 0
}
// two-class log-softmax: returns -log p[y] and p[1]
__device__ __forceinline__ float ce2(float z0, float z1, int y, float* p1) {
    const float m = fmaxf(z0, z1), e0 = expf(z0 - m), e1 = expf(z1 - m), s = e0 + e1;
    *p1 = e1 / s;
    return logf(s) + m - (y ? z1 : z0);
}
// BCEWithLogits(pos_weight): loss and d loss / d x  (PyTorch's stable form: (1 - y) x + lw (log1p(exp(-|x|)) + max(-x, 0)))
__device__ __forceinline__ float bce_logits(float x, float yv, float pw, float* dx) {
    const float lw = 1.f + (pw - 1.f) * yv;
    const float sp = log1pf(expf(-fabsf(x))) + fmaxf(-x, 0.f);
    const float sg = 1.f / (1.f + expf(-x));
    *dx = (1.f - yv) - lw * (1.f - sg);
    return (1.f - yv) * x + lw * sp;
}

// sums[0..4] = sum w_m ce_m, sum w_m, sum w_i ce_i, sum w_i, sum bce   (double atomics; zeroed by the caller)
__global__ void __launch_bounds__(256) rg_loss_sums_kernel(const float* __restrict__ mask, const float* __restrict__ inst,
                                                           const float* __restrict__ edge, const long long* __restrict__ y,
                                                           const long long* __restrict__ iy, const float* __restrict__ ey,
                                                           const int* n_dev, int cap, float wm1, float wi1, float pw,
                                                           double* __restrict__ sums) {
    __shared__ double sh[8];
    const int n = n_dev ? min(*n_dev, cap) : cap;
    double a[5] = {0, 0, 0, 0, 0};
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        float p1, dx;
        const int ym = (int)y[i], yi = (int)iy[i];
        const float w0 = ym ? wm1 : 1.f, w1 = yi ? wi1 : 1.f;
        a[0] += (double)(w0 * ce2(mask[2 * i], mask[2 * i + 1], ym, &p1));
        a[1] += (double)w0;
        a[2] += (double)(w1 * ce2(inst[2 * i], inst[2 * i + 1], yi, &p1));
        a[3] += (double)w1;
        a[4] += (double)bce_logits(edge[i], ey[i], pw, &dx);
    }
    for (int k = 0; k < 5; ++k) {
        const double t = block_sum(a[k], sh);
        if (threadIdx.x == 0) atomicAdd(sums + k, t);
    }
}
// loss_out[0..3] = total, mask, instance, edge terms (task weights applied); gradients w.r.t. the logits
__global__ void __launch_bounds__(256) rg_loss_grad_kernel(const float* __restrict__ mask, const float* __restrict__ inst,
                                                           const float* __restrict__ edge, const long long* __restrict__ y,
                                                           const long long* __restrict__ iy, const float* __restrict__ ey,
                                                           const int* n_dev, int cap, float wm1, float wi1, float pw, float tm,
                                                           float ti, float te, const double* __restrict__ sums,
                                                           float* __restrict__ loss_out, float* __restrict__ d_mask,
                                                           float* __restrict__ d_inst, float* __restrict__ d_edge) {
    const int n = n_dev ? min(*n_dev, cap) : cap;
    const float inv_m = (float)(1.0 / sums[1]), inv_i = (float)(1.0 / sums[3]), inv_n = 1.f / (float)max(n, 1);
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        const float lm = tm * (float)(sums[0] / sums[1]), li = ti * (float)(sums[2] / sums[3]), le = te * (float)(sums[4] / (double)max(n, 1));
        loss_out[0] = lm + li + le;
        loss_out[1] = lm;
        loss_out[2] = li;
        loss_out[3] = le;
    }
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < cap; i += gridDim.x * blockDim.x) {
        if (i >= n) {   // capacity rows past the live count carry no gradient
            if (d_mask) d_mask[2 * i] = d_mask[2 * i + 1] = 0.f;
            if (d_inst) d_inst[2 * i] = d_inst[2 * i + 1] = 0.f;
            if (d_edge) d_edge[i] = 0.f;
            continue;
        }
        float p1, dx;
        const int ym = (int)y[i], yi = (int)iy[i];
        if (d_mask) {
            ce2(mask[2 * i], mask[2 * i + 1], ym, &p1);
            const float s = tm * (ym ? wm1 : 1.f) * inv_m;
            d_mask[2 * i] = s * ((1.f - p1) - (ym ? 0.f : 1.f));
            d_mask[2 * i + 1] = s * (p1 - (ym ? 1.f : 0.f));
        }
        if (d_inst) {
            ce2(inst[2 * i], inst[2 * i + 1], yi, &p1);
            const float s = ti * (yi ? wi1 : 1.f) * inv_i;
            d_inst[2 * i] = s * ((1.f - p1) - (yi ? 0.f : 1.f));
            d_inst[2 * i + 1] = s * (p1 - (yi ? 1.f : 0.f));
        }
        if (d_edge) {
            bce_logits(edge[i], ey[i], pw, &dx);
            d_edge[i] = te * inv_n * dx;
        }
    }
}

// one thread per image; loss_out[0] = sum over images (written by a single CTA: G is a batch size)
__global__ void __launch_bounds__(256) image_loss_kernel(const float* __restrict__ mask, const float* __restrict__ inst,
                                                         const float* __restrict__ edge, const float* __restrict__ score,
                                                         const long long* __restrict__ mask_label, const float* __restrict__ edge_label,
                                                         const float* __restrict__ score_label, int G, float alpha, float gamma,
                                                         float w_focal, float w_inst, float w_edge, float w_score,
                                                         float* __restrict__ loss_out, float* __restrict__ d_mask,
                                                         float* __restrict__ d_inst, float* __restrict__ d_edge,
                                                         float* __restrict__ d_score) {
    __shared__ double sh[8];
    double acc = 0.0;
    for (int g = threadIdx.x; g < G; g += blockDim.x) {
        const int t = (int)mask_label[g];
        float p1, dx;
        // focal: alpha_t (1 - pt)^gamma ce,  ce = -log pt
        const float ce = ce2(mask[2 * g], mask[2 * g + 1], t, &p1);
        const float pt = t ? p1 : 1.f - p1, at = t ? alpha : 1.f - alpha;
        const float om = fmaxf(1.f - pt, 0.f), fw = powf(om, gamma);
        const float focal = at * fw * ce;
        // d focal / d pt = at [ -gamma (1-pt)^(gamma-1) ce - (1-pt)^gamma / pt ];  d pt / d z_k = pt (delta_kt - p_k)
        const float dpt = at * (-gamma * powf(om, gamma - 1.f) * ce - fw / pt);
        const float pk[2] = {1.f - p1, p1};
        float q1;
        const float cei = ce2(inst[2 * g], inst[2 * g + 1], t, &q1);
        const float bce = bce_logits(edge[g], edge_label[g], 1.f, &dx);
        const float ds = score[g] - score_label[g];
        acc += (double)(w_focal * focal + w_inst * cei + w_edge * bce + w_score * ds * ds);
        if (d_mask) {
            d_mask[2 * g] = w_focal * dpt * pt * ((t == 0 ? 1.f : 0.f) - pk[0]);
            d_mask[2 * g + 1] = w_focal * dpt * pt * ((t == 1 ? 1.f : 0.f) - pk[1]);
        }
        if (d_inst) {
            d_inst[2 * g] = w_inst * ((1.f - q1) - (t ? 0.f : 1.f));
            d_inst[2 * g + 1] = w_inst * (q1 - (t ? 1.f : 0.f));
        }
        if (d_edge) d_edge[g] = w_edge * dx;
        if (d_score) d_score[g] = w_score * 2.f * ds;
    }
    const double t = block_sum(acc, sh);
    if (threadIdx.x == 0) loss_out[0] = (float)t;
}

// ------------------------------------------------------------------------------------------ clip + AdamW
struct TensorTable {   // device arrays, one entry per parameter tensor
    float* const* param;
    const float* const* grad;
    float* const* exp_avg;
    float* const* exp_avg_sq;
    const long long* numel;
};

// norms[t] = sum of squares of tensor t's gradient (double); grid (chunks, tensors)
__global__ void __launch_bounds__(256) grad_sq_norm_kernel(const float* const* __restrict__ grad, const long long* __restrict__ numel,
                                                           double* __restrict__ norms) {
    __shared__ double sh[8];
    const int t = blockIdx.y;
    const float* g = grad[t];
    const long long n = numel[t];
    double a = 0.0;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const float v = g[i];
        a += (double)v * (double)v;
    }
    const double s = block_sum(a, sh);
    if (threadIdx.x == 0 && s != 0.0) atomicAdd(norms + t, s);
}
// total_norm = || (||g_t||_2)_t ||_2;  clip_coef = min(1, max_norm / (total_norm + 1e-6))  (torch.nn.utils.clip_grad_norm_)
__global__ void clip_coef_kernel(const double* __restrict__ norms, int n_tensors, float max_norm, float* __restrict__ out) {
    __shared__ double sh[8];
    double a = 0.0;
    for (int t = threadIdx.x; t < n_tensors; t += blockDim.x) a += norms[t];
    const double s = block_sum(a, sh);
    if (threadIdx.x == 0) {
        const float total = (float)sqrt(s);
        out[0] = total;
        out[1] = max_norm > 0.f ? fminf(max_norm / (total + 1e-6f), 1.f) : 1.f;
    }
}
__global__ void __launch_bounds__(256) adamw_kernel(TensorTable tt, const float* __restrict__ clip, float lr, float beta1, float beta2,
                                                    float eps, float weight_decay, float bc1, float bc2_sqrt, int write_grad) {
    const int t = blockIdx.y;
    float* p = tt.param[t];
    float* g = const_cast<float*>(tt.grad[t]);
    float* m = tt.exp_avg[t];
    float* v = tt.exp_avg_sq[t];
    const long long n = tt.numel[t];
    const float coef = clip ? clip[1] : 1.f;
    const float step_size = lr / bc1;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const float gi = g[i] * coef;
        if (write_grad) g[i] = gi;   // clip_grad_norm_ scales .grad in place
        float pi = p[i] * (1.f - lr * weight_decay);
        const float mi = beta1 * m[i] + (1.f - beta1) * gi;   // lerp(m, g, 1 - beta1) as torch's foreach path rounds it
        const float vi = beta2 * v[i] + (1.f - beta2) * gi * gi;
        m[i] = mi;
        v[i] = vi;
        const float denom = sqrtf(vi) / bc2_sqrt + eps;
        pi -= step_size * (mi / denom);
        p[i] = pi;
    }
}

}  // namespace cm

extern "C" int cm_region_targets(const int32_t* labels, const uint8_t* mask, const uint8_t* instance_mask, const uint8_t* edge_mask,
                                 int batch, int h, int w, const int32_t* node_ptr, const int32_t* node_label,
                                 const int32_t* node_bbox, int node_cap, int64_t* y, int64_t* instance_y, float* edge_y, void* stream) {
    using namespace cm;
    CM_DEVICE_OR_RETURN();
    CM_REQUIRE(labels && mask && instance_mask && edge_mask && node_ptr && node_label && node_bbox && y && instance_y && edge_y,
               "cm_region_targets: NULL pointer");
    CM_REQUIRE(batch > 0 && h > 0 && w > 0 && node_cap > 0 && (long long)h * w < (1 << 24), "cm_region_targets: bad sizes");
    {
        CM_TIMED("region_targets", as_stream(stream));
        region_targets_kernel<<<ceil_div(node_cap, 8), 256, 0, as_stream(stream)>>>(
            labels, mask, instance_mask, edge_mask, batch, h, w, node_ptr, node_label, node_bbox, node_cap,
            reinterpret_cast<long long*>(y), reinterpret_cast<long long*>(instance_y), edge_y);
    }
    return post_launch("cm_region_targets");
}

extern "C" int cm_rg_node_loss(const float* mask_logits, const float* instance_logits, const float* edge_logits, const int64_t* y,
                               const int64_t* instance_y, const float* edge_y, const int32_t* n_nodes_dev, int node_cap,
                               const cm_rg_loss_config* cfg, void* workspace, size_t workspace_bytes, float* loss_out,
                               float* d_mask, float* d_instance, float* d_edge, void* stream) {
    using namespace cm;
    CM_DEVICE_OR_RETURN();
    CM_REQUIRE(mask_logits && instance_logits && edge_logits && y && instance_y && edge_y && cfg && workspace && loss_out,
               "cm_rg_node_loss: NULL pointer");
    CM_REQUIRE(node_cap > 0 && workspace_bytes >= 5 * sizeof(double), "cm_rg_node_loss: needs node_cap > 0 and a 40-byte workspace");
    cudaStream_t st = as_stream(stream);
    double* sums = reinterpret_cast<double*>(workspace);
    cudaMemsetAsync(sums, 0, 5 * sizeof(double), st);
    const int blocks = std::min(ceil_div(node_cap, 256), 4 * num_sms());
    {
        CM_TIMED("rg_loss", st);
        rg_loss_sums_kernel<<<blocks, 256, 0, st>>>(mask_logits, instance_logits, edge_logits, reinterpret_cast<const long long*>(y),
                                                    reinterpret_cast<const long long*>(instance_y), edge_y, n_nodes_dev, node_cap,
                                                    cfg->mask_pos_weight, cfg->instance_pos_weight, cfg->edge_pos_weight, sums);
        rg_loss_grad_kernel<<<blocks, 256, 0, st>>>(mask_logits, instance_logits, edge_logits, reinterpret_cast<const long long*>(y),
                                                    reinterpret_cast<const long long*>(instance_y), edge_y, n_nodes_dev, node_cap,
                                                    cfg->mask_pos_weight, cfg->instance_pos_weight, cfg->edge_pos_weight,
                                                    cfg->task_mask, cfg->task_instance, cfg->task_edge, sums, loss_out, d_mask,
                                                    d_instance, d_edge);
    }
    return post_launch("cm_rg_node_loss");
}

extern "C" int cm_image_loss(const float* mask_logits, const float* instance_logits, const float* edge_logits, const float* score,
                             const int64_t* mask_label, const float* edge_label, const float* score_label, int n_graphs,
                             const cm_image_loss_config* cfg, float* loss_out, float* d_mask, float* d_instance, float* d_edge,
                             float* d_score, void* stream) {
    using namespace cm;
    CM_DEVICE_OR_RETURN();
    CM_REQUIRE(mask_logits && instance_logits && edge_logits && score && mask_label && edge_label && score_label && cfg && loss_out,
               "cm_image_loss: NULL pointer");
    CM_REQUIRE(n_graphs > 0, "cm_image_loss: n_graphs must be positive");
    {
        CM_TIMED("image_loss", as_stream(stream));
        image_loss_kernel<<<1, 256, 0, as_stream(stream)>>>(mask_logits, instance_logits, edge_logits, score,
                                                            reinterpret_cast<const long long*>(mask_label), edge_label, score_label,
                                                            n_graphs, cfg->alpha, cfg->gamma, cfg->w_focal, cfg->w_instance, cfg->w_edge,
                                                            cfg->w_score, loss_out, d_mask, d_instance, d_edge, d_score);
    }
    return post_launch("cm_image_loss");
}

extern "C" int cm_adamw_step(float* const* params, const float* const* grads, float* const* exp_avg, float* const* exp_avg_sq,
                             const int64_t* numel, int n_tensors, int64_t max_numel, const cm_adamw_config* cfg, int step,
                             double* norm_workspace, float* norm_out, void* stream) {
    using namespace cm;
    CM_DEVICE_OR_RETURN();
    CM_REQUIRE(params && grads && exp_avg && exp_avg_sq && numel && cfg && norm_workspace && norm_out, "cm_adamw_step: NULL pointer");
    CM_REQUIRE(n_tensors > 0 && n_tensors <= 65535 && max_numel > 0 && step >= 1, "cm_adamw_step: bad sizes");
    cudaStream_t st = as_stream(stream);
    const int chunks = (int)std::min<long long>((max_numel + 1023) / 1024, 64);
    dim3 grid(chunks, n_tensors);
    cudaMemsetAsync(norm_workspace, 0, (size_t)n_tensors * sizeof(double), st);
    {
        CM_TIMED("grad_sq_norm", st);
        grad_sq_norm_kernel<<<grid, 256, 0, st>>>(grads, reinterpret_cast<const long long*>(numel), norm_workspace);
        clip_coef_kernel<<<1, 256, 0, st>>>(norm_workspace, n_tensors, cfg->max_norm, norm_out);
    }
    TensorTable tt{params, grads, exp_avg, exp_avg_sq, reinterpret_cast<const long long*>(numel)};
    const float bc1 = 1.f - powf(cfg->beta1, (float)step);
    const float bc2_sqrt = sqrtf(1.f - powf(cfg->beta2, (float)step));
    {
        CM_TIMED("adamw", st);
        adamw_kernel<<<grid, 256, 0, st>>>(tt, norm_out, cfg->lr, cfg->beta1, cfg->beta2, cfg->eps, cfg->weight_decay, bc1, bc2_sqrt, 1);
    }
    return post_launch("cm_adamw_step");
}
