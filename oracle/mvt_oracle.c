/*
 * mvt_oracle.c -- CPU ORACLE (test infrastructure, NOT product code).
 *
 * A plain-C restatement of the MVTracker hot path, used ONLY by tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg as
 * the checker / timed CPU baseline.  Nothing under mvtracker_b200/ may link,
 * import or call this file.
 *
 * Reference lines restated (paths under /root/reference):
 *   unproject + view fusion : mvtracker/models/core/model_utils.py:420-482
 *   homogeneous helpers     : mvtracker/utils/basic.py:437-444
 *   kNN (semantic def.)     : mvtracker/models/core/mvtracker/mvtracker.py:75-79 (_knn_torch)
 *                             and :26-72 (_knn_pointops: direct differences, fp32)
 *   corr_sample             : mvtracker/models/core/mvtracker/mvtracker.py:800-846
 *   feature init (k = 1)    : mvtracker/models/core/mvtracker/mvtracker.py:606-645
 *
 * Parity pinning: tests/test_oracle_vs_golden.py checks every function here
 * against the .npz files in tests/golden, which were produced by importing and running the
 * reference's own Python in the build container (tests/golden/make_golden.py).
 *
 * Canonical kNN arithmetic (the bit-exact contract the CUDA path is held to):
 *   dx = p.x - q.x (fp32), dy, dz likewise
 *   d2 = fmaf(dz, dz, fmaf(dy, dy, dx * dx))           (fp32, explicit fma chain)
 *   neighbours ordered ascending by the pair (d2, index); ties -> lower index.
 *   returned distance = sqrtf(d2).
 * Points whose d2 is NaN never become neighbours.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define ORC_EXPORT __attribute__((visibility("default")))

ORC_EXPORT int orc_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

ORC_EXPORT void orc_set_num_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

/* ---- small dense inverse in fp32 (Gauss-Jordan, partial pivoting) ----------
 * The reference calls torch.inverse on fp32 matrices (model_utils.py:454-457);
 * LAPACK's LU and this elimination agree to a few ulp on these well-conditioned
 * camera matrices, far inside the 1e-4 tolerance.                            */
static int inv_n(const float *a, float *out, int n) {
    float m[4][8];
    for (int i = 0; i < n; ++i) {
        for (int j = 0; j < n; ++j) { m[i][j] = a[i * n + j]; m[i][n + j] = (i == j) ? 1.f : 0.f; }
    }
    for (int c = 0; c < n; ++c) {
        int p = c; float best = fabsf(m[c][c]);
        for (int r = c + 1; r < n; ++r) if (fabsf(m[r][c]) > best) { best = fabsf(m[r][c]); p = r; }
        if (best == 0.f) return -1;
        if (p != c) for (int j = 0; j < 2 * n; ++j) { float t = m[c][j]; m[c][j] = m[p][j]; m[p][j] = t; }
        float piv = m[c][c];
        for (int j = 0; j < 2 * n; ++j) m[c][j] /= piv;
        for (int r = 0; r < n; ++r) if (r != c) {
            float f = m[r][c];
            if (f != 0.f) for (int j = 0; j < 2 * n; ++j) m[r][j] -= f * m[c][j];
        }
    }
    for (int i = 0; i < n; ++i) for (int j = 0; j < n; ++j) out[i * n + j] = m[i][n + j];
    return 0;
}

/* ---- unprojection + multi-view fusion (model_utils.py:420-482) -------------
 * fmaps  [V,S,C,H,W]   depths [V,S,H,W]   intrs [V,S,3,3]   extrs [V,S,3,4]
 * out xyz [S, V*Hl*Wl, 3], fvec [S, V*Hl*Wl, C], valid [S, V*Hl*Wl] (may be NULL)
 * point index inside a frame = v*Hl*Wl + y*Wl + x           (:475-476)
 * depth_avg: 0 = 'nearest' (default, :444-445), 1 = 'avg' (:442-443).        */
ORC_EXPORT int orc_unproject(const float *fmaps, const float *depths, const float *intrs, const float *extrs,
                             int V, int S, int C, int H, int W, float stride, int level, int depth_avg,
                             float *xyz, float *fvec, uint8_t *valid) {
    int Hl = H, Wl = W;
    size_t nimg = (size_t)V * S;
    /* pyramid by chained 2x2 average pooling / nearest x0.5 (:439-446) */
    float *f = (float *)fmaps, *d = (float *)depths;
    float *fown = NULL, *down = NULL;
    for (int l = 0; l < level; ++l) {
        int Hn = Hl / 2, Wn = Wl / 2;
        float *fn = (float *)malloc(sizeof(float) * nimg * C * (size_t)Hn * Wn + 16);
        float *dn = (float *)malloc(sizeof(float) * nimg * (size_t)Hn * Wn + 16);
        if (!fn || !dn) return -2;
#pragma omp parallel for schedule(static)
        for (long ic = 0; ic < (long)(nimg * C); ++ic) {
            const float *src = f + (size_t)ic * Hl * Wl;
            float *dst = fn + (size_t)ic * Hn * Wn;
            for (int y = 0; y < Hn; ++y) for (int x = 0; x < Wn; ++x) {
                float s = src[(2 * y) * Wl + 2 * x];
                s += src[(2 * y) * Wl + 2 * x + 1];
                s += src[(2 * y + 1) * Wl + 2 * x];
                s += src[(2 * y + 1) * Wl + 2 * x + 1];
                dst[y * Wn + x] = s / 4.0f;
            }
        }
        for (size_t i = 0; i < nimg; ++i) {
            const float *src = d + i * Hl * Wl;
            float *dst = dn + i * Hn * Wn;
            for (int y = 0; y < Hn; ++y) for (int x = 0; x < Wn; ++x) {
                if (depth_avg) {
                    float s = src[(2 * y) * Wl + 2 * x];
                    s += src[(2 * y) * Wl + 2 * x + 1];
                    s += src[(2 * y + 1) * Wl + 2 * x];
                    s += src[(2 * y + 1) * Wl + 2 * x + 1];
                    dst[y * Wn + x] = s / 4.0f;
                } else {
                    /* F.interpolate(scale_factor=0.5, 'nearest'): src = min(floor(dst*2), in-1) */
                    int sy = 2 * y < Hl - 1 ? 2 * y : Hl - 1, sx = 2 * x < Wl - 1 ? 2 * x : Wl - 1;
                    dst[y * Wn + x] = src[sy * Wl + sx];
                }
            }
        }
        free(fown); free(down);
        f = fown = fn; d = down = dn;
        Hl = Hn; Wl = Wn;
        stride *= 2.0f;
    }
    size_t HW = (size_t)Hl * Wl, P = (size_t)V * HW;
    int rc = 0;
    for (int v = 0; v < V; ++v) for (int s = 0; s < S; ++s) {
        const float *K = intrs + ((size_t)v * S + s) * 9;
        const float *E = extrs + ((size_t)v * S + s) * 12;
        float Kinv[9], E4[16], Einv[16];
        memcpy(E4, E, sizeof(float) * 12);
        E4[12] = 0.f; E4[13] = 0.f; E4[14] = 0.f; E4[15] = 1.f;
        if (inv_n(K, Kinv, 3) || inv_n(E4, Einv, 4)) { rc = -1; continue; }
        const float *dimg = d + ((size_t)v * S + s) * HW;
        const float *fimg = f + ((size_t)v * S + s) * C * HW;
#pragma omp parallel for schedule(static)
        for (int y = 0; y < Hl; ++y) for (int x = 0; x < Wl; ++x) {
            float u = ((float)x + 0.5f) * stride - 0.5f;         /* :462-466 */
            float w = ((float)y + 0.5f) * stride - 0.5f;
            float dep = dimg[(size_t)y * Wl + x];
            float cam[3];
            for (int i = 0; i < 3; ++i)                          /* :469-470 */
                cam[i] = (Kinv[i * 3 + 0] * u + Kinv[i * 3 + 1] * w + Kinv[i * 3 + 2] * 1.0f) * dep;
            float wh[4];
            for (int i = 0; i < 4; ++i)                          /* :471-472 */
                wh[i] = Einv[i * 4 + 0] * cam[0] + Einv[i * 4 + 1] * cam[1] + Einv[i * 4 + 2] * cam[2] + Einv[i * 4 + 3] * 1.0f;
            size_t p = (size_t)s * P + (size_t)v * HW + (size_t)y * Wl + x;
            xyz[p * 3 + 0] = wh[0] / wh[3];                       /* :473 */
            xyz[p * 3 + 1] = wh[1] / wh[3];
            xyz[p * 3 + 2] = wh[2] / wh[3];
            if (valid) valid[p] = dep > 0.f;                      /* :478-480 */
            float *fo = fvec + p * C;
            for (int c = 0; c < C; ++c) fo[c] = fimg[(size_t)c * HW + (size_t)y * Wl + x];
        }
    }
    free(fown); free(down);
    return rc;
}

/* ---- canonical brute-force kNN (semantics of mvtracker.py:75-79) -----------
 * ref [B,N,3], qry [B,M,3] -> idx [B,M,k] (int32, local to batch element),
 * dist [B,M,k] (fp32, Euclidean, ascending).  mask [B,N] (NULL = all valid):
 * points with mask==0 are skipped.  When fewer than k candidates exist the
 * tail is filled with idx=-1, dist=+inf.                                     */
static inline int better(float d, int i, float bd, int bi) { return d < bd || (d == bd && i < bi); }

ORC_EXPORT int orc_knn(int k, const float *ref, const uint8_t *mask, const float *qry, int B, int N, int M,
                       int32_t *idx, float *dist) {
    if (k <= 0 || k > 64) return -1;
    long total = (long)B * M;
#pragma omp parallel for schedule(dynamic, 16)
    for (long t = 0; t < total; ++t) {
        int b = (int)(t / M);
        const float *R = ref + (size_t)b * N * 3;
        const uint8_t *Mk = mask ? mask + (size_t)b * N : NULL;
        const float qx = qry[t * 3 + 0], qy = qry[t * 3 + 1], qz = qry[t * 3 + 2];
        float bd[64]; int bi[64]; int n = 0;
        for (int i = 0; i < N; ++i) {
            if (Mk && !Mk[i]) continue;
            float dx = R[i * 3 + 0] - qx, dy = R[i * 3 + 1] - qy, dz = R[i * 3 + 2] - qz;
            float d2 = fmaf(dz, dz, fmaf(dy, dy, dx * dx));
            if (d2 != d2) continue;
            if (n == k && !better(d2, i, bd[k - 1], bi[k - 1])) continue;
            int j = (n < k) ? n++ : k - 1;
            while (j > 0 && better(d2, i, bd[j - 1], bi[j - 1])) { bd[j] = bd[j - 1]; bi[j] = bi[j - 1]; --j; }
            bd[j] = d2; bi[j] = i;
        }
        for (int j = 0; j < k; ++j) {
            idx[t * k + j] = j < n ? bi[j] : -1;
            dist[t * k + j] = j < n ? sqrtf(bd[j]) : INFINITY;
        }
    }
    return 0;
}

/* ---- corr_sample (mvtracker.py:800-846) -------------------------------------
 * xyz [B,N,3] fvec [B,N,C] targets [B,M,C] coords [B,M,3]  idx [B,M,k]
 * out [B,M,k,G + 3*add_offset + 3*add_xyz]:  [corr_0..corr_{G-1}, dx,dy,dz, x,y,z]
 * corr = <target_g, nbr_g> / sqrt(C/G)  (:832-835);  off = nbr_xyz - coord (:840).
 * The dot product is accumulated in double then rounded (order-neutral).     */
ORC_EXPORT int orc_corr_from_idx(const float *xyz, const float *fvec, const float *targets, const float *coords,
                                 const int32_t *idx, int B, int N, int M, int C, int k, int G,
                                 int add_offset, int add_xyz, float *out) {
    if (G <= 0 || C % G) return -1;
    int Cg = C / G, D = G + 3 * (add_offset != 0) + 3 * (add_xyz != 0);
    float scale = (float)sqrt((double)C / (double)G);
    long total = (long)B * M;
#pragma omp parallel for schedule(static)
    for (long t = 0; t < total; ++t) {
        int b = (int)(t / M);
        const float *tg = targets + t * C;
        for (int j = 0; j < k; ++j) {
            int i = idx[t * k + j];
            float *o = out + (t * k + j) * D;
            if (i < 0) { for (int c = 0; c < D; ++c) o[c] = 0.f; continue; }
            const float *fr = fvec + ((size_t)b * N + i) * C;
            const float *px = xyz + ((size_t)b * N + i) * 3;
            for (int g = 0; g < G; ++g) {
                double acc = 0.0;
                for (int c = 0; c < Cg; ++c) acc += (double)tg[g * Cg + c] * (double)fr[g * Cg + c];
                o[g] = (float)acc / scale;
            }
            int w = G;
            if (add_offset) { for (int a = 0; a < 3; ++a) o[w + a] = px[a] - coords[t * 3 + a]; w += 3; }
            if (add_xyz) { for (int a = 0; a < 3; ++a) o[w + a] = px[a]; }
        }
    }
    return 0;
}

/* knn + corr in one call (the timed CPU baseline unit: one corr_sample). */
ORC_EXPORT int orc_corr_sample(int k, const float *xyz, const uint8_t *mask, const float *fvec, const float *targets,
                               const float *coords, int B, int N, int M, int C, int G, int add_offset, int add_xyz,
                               int32_t *idx_scratch, float *dist_scratch, float *out) {
    int rc = orc_knn(k, xyz, mask, coords, B, N, M, idx_scratch, dist_scratch);
    if (rc) return rc;
    return orc_corr_from_idx(xyz, fvec, targets, coords, idx_scratch, B, N, M, C, k, G, add_offset, add_xyz, out);
}

/* float64 ground truth distances for adjudicating near-ties against the
 * reference's cdist (which uses the |x|^2+|y|^2-2xy expansion).              */
ORC_EXPORT void orc_dist2_f64(const float *ref, const float *q, int N, double *out) {
    for (int i = 0; i < N; ++i) {
        double dx = (double)ref[i * 3] - q[0], dy = (double)ref[i * 3 + 1] - q[1], dz = (double)ref[i * 3 + 2] - q[2];
        out[i] = dx * dx + dy * dy + dz * dz;
    }
}
