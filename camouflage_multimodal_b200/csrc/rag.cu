// A1 + A2: region adjacency graph of a superpixel label map and label -> node-id compaction.
//
// Reference: skimage.graph.rag_mean_color(...).edges() as used at
// models/region_graph/extract_rg_embeddings.py:216-225 plus the region_id_map bookkeeping (:148-161,
// :210-211).  Integer work, bit-exact; HBM-bound on one read of the label map.
//
// Data layout (per image b, all in the caller's workspace, L2-resident at batch 32):
//   box  [B,cap,4] int32   max(x)+1, max(y)+1, W-min(x), H-min(y)   (0 = label absent; all atomicMax)
//   adj  [B,cap,cap_words] uint32 symmetric adjacency bitmap; row l = neighbours of label l
//   rowstart [B,cap] int32 exclusive scan of the row degrees inside the image
// Kernels: rag_scan (pixel-parallel, one read of the labels) -> rag_compact (one CTA per image; the last
// CTA to finish scans the per-image totals) -> rag_emit (one warp per label row).
#include <limits.h>

#include "common.cuh"

namespace cm {

struct RagWs {
    int* box;
    uint32_t* adj;
    int* rowstart;
    int* n_nodes;   // [B]
    int* n_edges;   // [B]
    unsigned* ticket;
    size_t zero_bytes;   // prefix of the workspace that must be zeroed (box, adj, ticket)
    size_t total;
};

static RagWs carve_rag(void* ws, int batch, int cap) {
    int cw = (cap + 31) / 32;
    Carver c(ws);
    RagWs r;
    r.box = c.take<int>((size_t)batch * cap * 4);
    r.adj = c.take<uint32_t>((size_t)batch * cap * cw);
    r.ticket = c.take<unsigned>(64);
    r.zero_bytes = align_up(c.off, 256);
    r.rowstart = c.take<int>((size_t)batch * cap);
    r.n_nodes = c.take<int>(batch);
    r.n_edges = c.take<int>(batch);
    r.total = align_up(c.off, 256);
    return r;
}

__device__ __forceinline__ void set_pair(uint32_t* adj, int cw, int a, int b) {
    uint32_t* wa = adj + (size_t)a * cw + (b >> 5);
    uint32_t bit = 1u << (b & 31);
    if ((*wa & bit) == 0u) {   // a stale L1 line only costs a redundant atomic
        atomicOr(wa, bit);
        atomicOr(adj + (size_t)b * cw + (a >> 5), 1u << (a & 31));
    }
}

// grid (strips, B), block 256; dynamic smem 4*cap ints
__global__ void __launch_bounds__(256) rag_scan_kernel(const int* __restrict__ labels, int H, int W, int cap,
                                                       int cw, int rows_per_cta, int* __restrict__ box_g,
                                                       uint32_t* __restrict__ adj_g, uint32_t* status) {
    extern __shared__ int sbox[];
    const int b = blockIdx.y;
    const int* lab = labels + (size_t)b * H * W;
    const int r0 = blockIdx.x * rows_per_cta;
    const int r1 = min(H, r0 + rows_per_cta);
    for (int i = threadIdx.x; i < 4 * cap; i += blockDim.x) sbox[i] = 0;
    __syncthreads();
    uint32_t* adj = adj_g + (size_t)b * cap * cw;
    const int npx = (r1 - r0) * W;
    const int lane = threadIdx.x & 31;
    bool overflow = false;
    for (int base = 0; base < npx; base += blockDim.x) {
        const int i = base + threadIdx.x;
        const bool act = i < npx;
        int y = 0, x = 0, a = -1;
        if (act) {
            y = r0 + i / W;
            x = i - (i / W) * W;
            a = lab[(size_t)y * W + x];
        }
        if (a >= cap) {
            overflow = true;
            a = -1;
        }
        const bool valid = a >= 0;
        const unsigned grp = __match_any_sync(0xffffffffu, valid ? a : -1);
        if (valid) {
            const int mx = __reduce_max_sync(grp, x + 1);
            const int my = __reduce_max_sync(grp, y + 1);
            const int nx = __reduce_max_sync(grp, W - x);
            const int ny = __reduce_max_sync(grp, H - y);
            if (lane == __ffs(grp) - 1) {
                int* e = sbox + 4 * a;
                atomicMax(e + 0, mx);
                atomicMax(e + 1, my);
                atomicMax(e + 2, nx);
                atomicMax(e + 3, ny);
            }
            // 8-neighbourhood adjacency: each unordered pixel pair is visited once (E, SW, S, SE)
            int prev = a;
            if (x + 1 < W) {
                int c = lab[(size_t)y * W + x + 1];
                if (c != a && c >= 0 && c < cap) {
                    set_pair(adj, cw, a, c);
                    prev = c;
                }
            }
            if (y + 1 < H) {
                const int* row = lab + (size_t)(y + 1) * W;
#pragma unroll
                for (int d = -1; d <= 1; ++d) {
                    int xx = x + d;
                    if (xx < 0 || xx >= W) continue;
                    int c = row[xx];
                    if (c != a && c != prev && c >= 0 && c < cap) {
                        set_pair(adj, cw, a, c);
                        prev = c;
                    }
                }
            }
        }
    }
    if (overflow) atomicOr(status, CM_STATUS_LABEL_OVERFLOW);
    __syncthreads();
    int* box = box_g + (size_t)b * cap * 4;
    for (int l = threadIdx.x; l < cap; l += blockDim.x) {
        if (sbox[4 * l] != 0) {
#pragma unroll
            for (int k = 0; k < 4; ++k) atomicMax(box + 4 * l + k, sbox[4 * l + k]);
        }
    }
}

// Run-based scan for W % 8 == 0: a thread takes 8 consecutive pixels of a row (two 16-byte loads) plus the 10 pixels
// below them.  Superpixel labels are piecewise constant, so the events that matter are sparse: segment starts / ends
// (bounding boxes) and label changes towards E / S / SE / SW (adjacency).  Both are collected as bit masks with
// branch-free compares and then drained bit by bit, so a warp spends its instructions on ~4 events per thread
// instead of walking 32 pixel x neighbour sites that almost always diverge.  Same outputs as rag_scan_kernel (the
// box maxima and the adjacency bit set are order independent).
// eight consecutive labels (16-byte aligned run start for int16, two 16-byte loads for int32)
__device__ __forceinline__ void load8_labels(const int* p, int* v) {
    const int4 a0 = *reinterpret_cast<const int4*>(p), a1 = *reinterpret_cast<const int4*>(p + 4);
    v[0] = a0.x; v[1] = a0.y; v[2] = a0.z; v[3] = a0.w; v[4] = a1.x; v[5] = a1.y; v[6] = a1.z; v[7] = a1.w;
}
__device__ __forceinline__ void load8_labels(const short* p, int* v) {
    const uint4 a = *reinterpret_cast<const uint4*>(p);
    const unsigned w[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        v[2 * k] = (int)(short)(w[k] & 0xffffu);
        v[2 * k + 1] = (int)(short)(w[k] >> 16);
    }
}

template <typename LT>
__global__ void __launch_bounds__(256) rag_scan8_kernel(const LT* __restrict__ labels, int H, int W, int cap, int cw,
                                                        int rows_per_cta, int* __restrict__ box_g,
                                                        uint32_t* __restrict__ adj_g, uint32_t* status) {
    extern __shared__ int sbox[];
    const int b = blockIdx.y;
    const LT* lab = labels + (size_t)b * H * W;
    const int r0 = blockIdx.x * rows_per_cta;
    const int r1 = min(H, r0 + rows_per_cta);
    for (int i = threadIdx.x; i < 4 * cap; i += blockDim.x) sbox[i] = 0;
    __syncthreads();
    uint32_t* adj = adj_g + (size_t)b * cap * cw;
    const int upr = W >> 3, units = (r1 - r0) * upr;
    bool overflow = false;
    for (int u = threadIdx.x; u < units; u += blockDim.x) {
        const int r = u / upr, x0 = (u - r * upr) << 3, y = r0 + r;
        // v[0] = west of the run, v[1..8] = the run, v[9] = east of it; v[10..19] the same columns one row down
        int v[20];
        const LT* pa = lab + (size_t)y * W + x0;
        load8_labels(pa, v + 1);
        v[0] = x0 > 0 ? (int)pa[-1] : -1;
        v[9] = x0 + 8 < W ? (int)pa[8] : -1;
        if (y + 1 < H) {
            const LT* pb = pa + W;
            load8_labels(pb, v + 11);
            v[10] = x0 > 0 ? (int)pb[-1] : -1;
            v[19] = x0 + 8 < W ? (int)pb[8] : -1;
        } else {
#pragma unroll
            for (int k = 10; k < 20; ++k) v[k] = -1;
        }
#pragma unroll
        for (int k = 1; k <= 8; ++k) overflow |= v[k] >= cap;
#pragma unroll
        for (int k = 0; k < 20; ++k) v[k] = (unsigned)v[k] < (unsigned)cap ? v[k] : -1;   // out-of-range labels take no part
        unsigned seg = 0, pm = 0;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int a = v[1 + i], w_ = v[i], e = v[2 + i], sw = v[10 + i], sd = v[11 + i], se = v[12 + i];
            const bool ok = a >= 0;
            seg |= (ok && a != w_ ? 1u : 0u) << i;            // a segment starts here
            seg |= (ok && a != e ? 1u : 0u) << (8 + i);       //           ends here
            // each unordered pixel pair once (E, S, SE, SW); a pair that the S test of this or a neighbouring pixel
            // reports anyway is skipped (same two labels): S when the column pair repeats its western neighbour,
            // SE / SW when the lower label repeats the one straight below or the upper label repeats sideways
            pm |= (ok && e >= 0 && a != e ? 1u : 0u) << i;
            pm |= (ok && sd >= 0 && a != sd && (a != w_ || sd != sw) ? 1u : 0u) << (8 + i);
            pm |= (ok && se >= 0 && a != se && se != sd && e != a ? 1u : 0u) << (16 + i);
            pm |= (ok && sw >= 0 && a != sw && sw != sd && w_ != a ? 1u : 0u) << (24 + i);
        }
        while (seg) {
            const int k = __ffs(seg) - 1;
            seg &= seg - 1;
            const int i = k & 7, x = x0 + i;
            int* e = sbox + 4 * v[1 + i];
            if (k < 8) {
                atomicMax(e + 1, y + 1);
                atomicMax(e + 2, W - x);
                atomicMax(e + 3, H - y);
            } else {
                atomicMax(e + 0, x + 1);
            }
        }
        while (pm) {
            const int k = __ffs(pm) - 1;
            pm &= pm - 1;
            const int i = k & 7, dir = k >> 3;
            const int c = v[(dir == 0 ? 2 : (dir == 1 ? 11 : (dir == 2 ? 12 : 10))) + i];
            set_pair(adj, cw, v[1 + i], c);
        }
    }
    if (overflow) atomicOr(status, CM_STATUS_LABEL_OVERFLOW);
    __syncthreads();
    int* box = box_g + (size_t)b * cap * 4;
    for (int l = threadIdx.x; l < cap; l += blockDim.x) {
        if (sbox[4 * l] != 0) {
#pragma unroll
            for (int k = 0; k < 4; ++k) atomicMax(box + 4 * l + k, sbox[4 * l + k]);
        }
    }
}

// Exclusive scan of one int per thread across the CTA; returns the exclusive prefix, total in *total.
template <int THREADS>
__device__ __forceinline__ int block_scan_excl(int v, int* total, int* swarp) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int inc = warp_scan_incl(v, lane);
    if (lane == 31) swarp[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        int t = lane < THREADS / 32 ? swarp[lane] : 0;
        int ti = warp_scan_incl(t, lane);
        swarp[lane] = ti - t;   // exclusive warp offsets
        if (lane == 31) swarp[32] = ti;
    }
    __syncthreads();
    int r = inc - v + swarp[warp];
    *total = swarp[32];
    __syncthreads();
    return r;
}

// grid B, block 256
__global__ void __launch_bounds__(256) rag_compact_kernel(int cap, int cw, int batch, const int* __restrict__ box_g,
                                                          const uint32_t* __restrict__ adj_g, int* __restrict__ id_of_label,
                                                          int* __restrict__ rowstart_g, int* n_nodes, int* n_edges,
                                                          unsigned* ticket, int* __restrict__ node_ptr,
                                                          int* __restrict__ edge_ptr, int node_cap, int edge_cap,
                                                          uint32_t* status) {
    __shared__ int swarp[33];
    __shared__ bool is_last;
    const int b = blockIdx.x;
    const int* box = box_g + (size_t)b * cap * 4;
    const uint32_t* adj = adj_g + (size_t)b * cap * cw;
    int node_carry = 0, edge_carry = 0;
    for (int base = 0; base < cap; base += 256) {
        const int l = base + threadIdx.x;
        int present = 0, deg = 0;
        if (l < cap) {
            present = box[4 * l] != 0;
            if (present) {
                const uint32_t* row = adj + (size_t)l * cw;
                for (int k = 0; k < cw; ++k) deg += __popc(row[k]);
            }
        }
        int tot_n, tot_e;
        int pn = block_scan_excl<256>(present, &tot_n, swarp);
        int pe = block_scan_excl<256>(deg, &tot_e, swarp);
        if (l < cap) {
            id_of_label[(size_t)b * cap + l] = present ? node_carry + pn : -1;
            rowstart_g[(size_t)b * cap + l] = edge_carry + pe;
        }
        node_carry += tot_n;
        edge_carry += tot_e;
    }
    if (threadIdx.x == 0) {
        n_nodes[b] = node_carry;
        n_edges[b] = edge_carry;
        __threadfence();
        unsigned t = atomicAdd(ticket, 1u);
        is_last = (t == (unsigned)batch - 1u);
    }
    __syncthreads();
    if (!is_last) return;
    __threadfence();
    // cross-image exclusive scan (batch is small: thousands at most)
    int nc = 0, ec = 0;
    for (int base = 0; base < batch; base += 256) {
        const int i = base + threadIdx.x;
        int vn = i < batch ? ((volatile int*)n_nodes)[i] : 0;
        int ve = i < batch ? ((volatile int*)n_edges)[i] : 0;
        int tn, te;
        int pn = block_scan_excl<256>(vn, &tn, swarp);
        int pe = block_scan_excl<256>(ve, &te, swarp);
        if (i < batch) {
            node_ptr[i] = nc + pn;
            edge_ptr[i] = ec + pe;
        }
        nc += tn;
        ec += te;
    }
    if (threadIdx.x == 0) {
        node_ptr[batch] = nc;
        edge_ptr[batch] = ec;
        uint32_t s = 0;
        if (nc > node_cap) s |= CM_STATUS_NODE_OVERFLOW;
        if (ec > edge_cap) s |= CM_STATUS_EDGE_OVERFLOW;
        if (s) atomicOr(status, s);
        *ticket = 0u;
    }
}

// one warp per (image, label); grid = ceil(B*cap / 8), block 256
__global__ void __launch_bounds__(256) rag_emit_kernel(int cap, int cw, int batch, int H, int W, const int* __restrict__ box_g,
                                                       const uint32_t* __restrict__ adj_g, const int* __restrict__ id_of_label,
                                                       const int* __restrict__ rowstart_g, const int* __restrict__ node_ptr,
                                                       const int* __restrict__ edge_ptr, int* __restrict__ node_label,
                                                       int* __restrict__ node_bbox, long long* __restrict__ batch_vec,
                                                       int* __restrict__ csr_row_ptr, int* __restrict__ csr_col,
                                                       long long* __restrict__ edge_index, int local_ids, int node_cap,
                                                       int edge_cap) {
    const int lane = threadIdx.x & 31;
    const long long wid = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (wid >= (long long)batch * cap) return;
    const int b = (int)(wid / cap), l = (int)(wid - (long long)b * cap);
    const int lid = id_of_label[(size_t)b * cap + l];
    const int n0 = node_ptr[b], e0 = edge_ptr[b];
    if (wid == 0 && lane == 0) {
        int tn = node_ptr[batch];
        if (tn <= node_cap) csr_row_ptr[tn] = edge_ptr[batch];
    }
    if (lid < 0) return;
    const int g = n0 + lid;
    if (g >= node_cap) return;
    const int estart = e0 + rowstart_g[(size_t)b * cap + l];
    if (lane == 0) {
        const int* bx = box_g + ((size_t)b * cap + l) * 4;
        node_label[g] = l;
        node_bbox[4 * g + 0] = W - bx[2];
        node_bbox[4 * g + 1] = H - bx[3];
        node_bbox[4 * g + 2] = bx[0] - 1;
        node_bbox[4 * g + 3] = bx[1] - 1;
        if (batch_vec) batch_vec[g] = b;
        csr_row_ptr[g] = estart;
    }
    const uint32_t* row = adj_g + ((size_t)b * cap + l) * cw;
    const int* idmap = id_of_label + (size_t)b * cap;
    int pos = estart;
    const long long src = local_ids ? lid : g;
    for (int base = 0; base < cw; base += 32) {
        const int k = base + lane;
        uint32_t word = k < cw ? row[k] : 0u;
        const int cnt = __popc(word);
        const int inc = warp_scan_incl(cnt, lane);
        int p = pos + inc - cnt;
        while (word) {
            const int bit = __ffs(word) - 1;
            word &= word - 1;
            const int nl = idmap[k * 32 + bit];   // neighbour labels are present by construction
            if (p < edge_cap) {
                csr_col[p] = n0 + nl;
                if (edge_index) {
                    edge_index[p] = src;
                    edge_index[(size_t)edge_cap + p] = local_ids ? nl : n0 + nl;
                }
            }
            ++p;
        }
        pos += __shfl_sync(0xffffffffu, inc, 31);
    }
}

}  // namespace cm

extern "C" size_t cm_rag_workspace_bytes(int batch, int label_cap) {
    if (batch <= 0 || label_cap <= 0) return 0;
    return cm::carve_rag(nullptr, batch, label_cap).total;
}

static int rag_build_impl(const void* labels, int label_bytes, int batch, int h, int w, int label_cap, void* workspace,
                            size_t workspace_bytes, int32_t* id_of_label, int32_t* node_ptr, int32_t* edge_ptr,
                            int32_t* node_label, int32_t* node_bbox, int64_t* batch_vec, int32_t* csr_row_ptr,
                            int32_t* csr_col, int64_t* edge_index, int local_ids, int node_cap, int edge_cap,
                            uint32_t* status, void* stream) {
    using namespace cm;
    CM_DEVICE_OR_RETURN();
    CM_REQUIRE(labels && workspace && id_of_label && node_ptr && edge_ptr && node_label && node_bbox && csr_row_ptr &&
                   csr_col && status,
               "cm_rag_build: NULL pointer argument");
    CM_REQUIRE(batch > 0 && h > 0 && w > 0 && (long long)h * w < INT_MAX, "cm_rag_build: bad shape B=%d H=%d W=%d",
               batch, h, w);
    CM_REQUIRE(label_cap > 0 && label_cap <= 12288, "cm_rag_build: label_cap=%d outside 1..12288", label_cap);
    CM_REQUIRE(node_cap > 0 && edge_cap > 0, "cm_rag_build: node_cap / edge_cap must be positive");
    RagWs ws = carve_rag(workspace, batch, label_cap);
    if (workspace_bytes < ws.total) {
        set_error("cm_rag_build: workspace %zu < %zu bytes", workspace_bytes, ws.total);
        return CM_ERR_WORKSPACE;
    }
    cudaStream_t st = as_stream(stream);
    const int cw = (label_cap + 31) / 32;
    cudaMemsetAsync(workspace, 0, ws.zero_bytes, st);
    // strips: enough CTAs to fill the machine twice over at small batches
    int rows = 8;
    while (rows > 1 && (long long)ceil_div(h, rows) * batch < 2LL * num_sms()) rows >>= 1;
    const size_t smem = (size_t)label_cap * 4 * sizeof(int);
    static PerDeviceOnce attr_done;
    if (attr_done.first()) {
        cudaFuncSetAttribute(rag_scan_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        cudaFuncSetAttribute(rag_scan8_kernel<int>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        cudaFuncSetAttribute(rag_scan8_kernel<short>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    }
    dim3 g1(ceil_div(h, rows), batch);
    {
        CM_TIMED("rag_scan", st);
        if (label_bytes == 2)
            rag_scan8_kernel<short><<<g1, 256, smem, st>>>(reinterpret_cast<const short*>(labels), h, w, label_cap, cw, rows, ws.box, ws.adj, status);
        else if ((w & 7) == 0 && (reinterpret_cast<uintptr_t>(labels) & 15) == 0)
            rag_scan8_kernel<int><<<g1, 256, smem, st>>>(reinterpret_cast<const int*>(labels), h, w, label_cap, cw, rows, ws.box, ws.adj, status);
        else
            rag_scan_kernel<<<g1, 256, smem, st>>>(reinterpret_cast<const int*>(labels), h, w, label_cap, cw, rows, ws.box, ws.adj, status);
    }
    {
        CM_TIMED("rag_compact", st);
        rag_compact_kernel<<<batch, 256, 0, st>>>(label_cap, cw, batch, ws.box, ws.adj, id_of_label, ws.rowstart,
                                                  ws.n_nodes, ws.n_edges, ws.ticket, node_ptr, edge_ptr, node_cap,
                                                  edge_cap, status);
    }
    const long long warps = (long long)batch * label_cap;
    {
        CM_TIMED("rag_emit", st);
        rag_emit_kernel<<<ceil_div(warps, 8), 256, 0, st>>>(label_cap, cw, batch, h, w, ws.box, ws.adj, id_of_label,
                                                            ws.rowstart, node_ptr, edge_ptr, node_label, node_bbox,
                                                            reinterpret_cast<long long*>(batch_vec), csr_row_ptr, csr_col,
                                                            reinterpret_cast<long long*>(edge_index), local_ids, node_cap,
                                                            edge_cap);
    }
    return post_launch("cm_rag_build");
}

extern "C" int cm_rag_build(const int32_t* labels, int batch, int h, int w, int label_cap, void* workspace,
                            size_t workspace_bytes, int32_t* id_of_label, int32_t* node_ptr, int32_t* edge_ptr,
                            int32_t* node_label, int32_t* node_bbox, int64_t* batch_vec, int32_t* csr_row_ptr,
                            int32_t* csr_col, int64_t* edge_index, int local_ids, int node_cap, int edge_cap,
                            uint32_t* status, void* stream) {
    return rag_build_impl(labels, 4, batch, h, w, label_cap, workspace, workspace_bytes, id_of_label, node_ptr, edge_ptr, node_label, node_bbox, batch_vec, csr_row_ptr, csr_col, edge_index, local_ids, node_cap, edge_cap, status, stream);
}

extern "C" int cm_rag_build_i16(const int16_t* labels, int batch, int h, int w, int label_cap, void* workspace,
                            size_t workspace_bytes, int32_t* id_of_label, int32_t* node_ptr, int32_t* edge_ptr,
                            int32_t* node_label, int32_t* node_bbox, int64_t* batch_vec, int32_t* csr_row_ptr,
                            int32_t* csr_col, int64_t* edge_index, int local_ids, int node_cap, int edge_cap,
                            uint32_t* status, void* stream) {
    using namespace cm;
    CM_REQUIRE(labels && (w & 7) == 0 && (reinterpret_cast<uintptr_t>(labels) & 15) == 0,
               "cm_rag_build_i16: int16 label maps need W % 8 == 0 and a 16-byte aligned buffer (else cm_labels_widen_i16 + cm_rag_build)");
    return rag_build_impl(labels, 2, batch, h, w, label_cap, workspace, workspace_bytes, id_of_label, node_ptr, edge_ptr, node_label, node_bbox, batch_vec, csr_row_ptr, csr_col, edge_index, local_ids, node_cap, edge_cap, status, stream);
}
