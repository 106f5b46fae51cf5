/* ORACLE (test infrastructure only): plain-C restatement of the two sequential loops of scikit-image's SLIC
 * (segmentation/_slic.pyx: _slic_cython and _enforce_label_connectivity_cython, scikit-image >= 0.19) as the reference
 * reaches them through skimage.segmentation.slic(img_u8, n_segments=500, compactness=10, sigma=1)
 * (models/region_graph/extract_rg_embeddings.py:143-144).  Same loop nests, same float64 operation order as
 * oracle/slic.py's numpy/pure-Python statements of them (tests check C == Python); this file only exists so that the
 * known-answer check against the reference's recorded per-image region counts
 * (models/region_graph/rg_embeddings/embedding_summary.json) runs over hundreds of images in seconds.
 * Built by oracle/build_c.py (gcc -O2 -ffp-contract=off) into oracle/_build/; never linked into the product.
 */
#include <float.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

/* feat [H,W,C] f64; seg [K, 2 + C] in/out: (y, x, colour...) seeded by the caller; nearest [H,W] out (0-based, -1 if
 * never reached).  step_y / step_x: window half extents / 2 (the cython routine recomputes them from the seed count),
 * step: max seed step (spatial weight). */
void slic_cluster(const double* feat, int H, int W, int C, int K, int step_y, int step_x, int step, int max_num_iter,
                  double* seg, int32_t* nearest) {
    const int F = 2 + C;
    const double sw = 1.0 / (double)(step * step);
    double* dist = (double*)malloc(sizeof(double) * (size_t)H * W);
    int64_t* cnt = (int64_t*)malloc(sizeof(int64_t) * (size_t)K);
    for (size_t i = 0; i < (size_t)H * W; ++i) nearest[i] = -1;
    for (int it = 0; it < max_num_iter; ++it) {
        for (size_t i = 0; i < (size_t)H * W; ++i) dist[i] = DBL_MAX;
        for (int k = 0; k < K; ++k) {
            const double cy = seg[k * F], cx = seg[k * F + 1];
            if (!isfinite(cy) || !isfinite(cx)) continue; /* emptied cluster: the casted NaN window is empty */
            const double ylo = cy - 2 * step_y > 0 ? cy - 2 * step_y : 0, yhi = cy + 2 * step_y + 1 < H ? cy + 2 * step_y + 1 : H;
            const double xlo = cx - 2 * step_x > 0 ? cx - 2 * step_x : 0, xhi = cx + 2 * step_x + 1 < W ? cx + 2 * step_x + 1 : W;
            const long y0 = (long)ylo, y1 = (long)yhi, x0 = (long)xlo, x1 = (long)xhi;
            for (long y = y0; y < y1; ++y) {
                const double dyv = cy - (double)y, dy = dyv * dyv;
                for (long x = x0; x < x1; ++x) {
                    const double dxv = cx - (double)x;
                    double d = (dy + dxv * dxv) * sw;
                    double dc = 0.0;
                    const double* p = feat + ((size_t)y * W + x) * C;
                    for (int c = 0; c < C; ++c) {
                        const double e = p[c] - seg[k * F + 2 + c];
                        dc += e * e;
                    }
                    d += dc;
                    if (dist[(size_t)y * W + x] > d) {
                        nearest[(size_t)y * W + x] = k;
                        dist[(size_t)y * W + x] = d;
                    }
                }
            }
        }
        memset(cnt, 0, sizeof(int64_t) * (size_t)K);
        for (int i = 0; i < K * F; ++i) seg[i] = 0.0;
        for (int y = 0; y < H; ++y)
            for (int x = 0; x < W; ++x) {
                const int k = nearest[(size_t)y * W + x];
                if (k < 0) continue;
                cnt[k] += 1;
                seg[k * F] += (double)y;
                seg[k * F + 1] += (double)x;
                const double* p = feat + ((size_t)y * W + x) * C;
                for (int c = 0; c < C; ++c) seg[k * F + 2 + c] += p[c];
            }
        for (int k = 0; k < K; ++k)
            for (int c = 0; c < F; ++c) seg[k * F + c] /= (double)cnt[k]; /* 0 / 0 = NaN, as in the cython loop */
    }
    free(dist);
    free(cnt);
}

/* labels [H,W] any ints -> out [H,W]: raster-order 4-connected BFS with the min / max size rule. */
void slic_enforce_connectivity(const int32_t* labels, int H, int W, int min_size, int max_size, int start_label, int32_t* out) {
    const int ddy[4] = {0, 0, 1, -1}, ddx[4] = {1, -1, 0, 0};
    int32_t* q = (int32_t*)malloc(sizeof(int32_t) * 2 * (size_t)(max_size > 1 ? max_size : 1));
    for (size_t i = 0; i < (size_t)H * W; ++i) out[i] = -1;
    int cur = start_label;
    for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x) {
            if (out[(size_t)y * W + x] >= 0) continue;
            int adjacent = start_label; /* no labelled neighbour yet: join the first label (oracle/slic.py) */
            const int label = labels[(size_t)y * W + x];
            out[(size_t)y * W + x] = cur;
            int size = 1, visited = 0;
            q[0] = y;
            q[1] = x;
            while (visited < size && size < max_size) {
                const int py = q[2 * visited], px = q[2 * visited + 1];
                for (int i = 0; i < 4; ++i) {
                    const int qy = py + ddy[i], qx = px + ddx[i];
                    if (qy < 0 || qy >= H || qx < 0 || qx >= W) continue;
                    const size_t qi = (size_t)qy * W + qx;
                    if (labels[qi] == label && out[qi] == -1) {
                        out[qi] = cur;
                        q[2 * size] = qy;
                        q[2 * size + 1] = qx;
                        ++size;
                        if (size >= max_size) break;
                    } else if (out[qi] >= 0 && out[qi] != cur) {
                        adjacent = out[qi];
                    }
                }
                ++visited;
            }
            if (size < min_size) {
                for (int i = 0; i < size; ++i) out[(size_t)q[2 * i] * W + q[2 * i + 1]] = adjacent;
            } else {
                ++cur;
            }
        }
    free(q);
}
