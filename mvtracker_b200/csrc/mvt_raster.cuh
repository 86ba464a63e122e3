// Raster kNN: exact k nearest neighbours in a cloud that was unprojected from V depth rasters (cloud_build's xyz4
// records), searched directly in the rasters -- no spatial index and no build step.
//
// Every cloud point lies on the camera ray of its pixel, so a point within distance r of the query q must have its
// pixel inside the projection of the ball B(q, r) into that view.  One warp per query:
//   scan     a (2m+1)^2 pixel window around q's projection in every view, m = 2 first (lane v owns view v's
//            rectangle; the rectangles of all views are flattened into one candidate list, 32 candidates per step);
//   bound    r = current k-th best distance;  lane v computes the pixel bounding box of B(q, r) in view v
//            (tangent directions of the ball seen from the camera centre, clipped to z > 0);
//   done     when every view's box lies inside what has been scanned -- the result is then exact;
//   grow     otherwise m -> 2m+1, windows clipped to the boxes; only the new pixels are scanned.  r shrinks as nearer
//            points turn up, so the scanned area stays within a small factor of the final ball's projection.
// Density adapts by construction: pixels are uniformly spaced in the raster wherever the surface is.
// Points with negative depth are not searchable (they are not on the forward ray); depth == 0 points collapse onto
// the camera centre and are handled exactly (a ball containing the centre maps to the whole raster).
#pragma once
#include "mvt_knn.cuh"

namespace mvt {

struct RasterView {
    const float4 *xyz4;   // [S, P] {x,y,z,depth}, P = V*H*W, point = v*H*W + y*W + x
    const float *cam;     // [V*S, kCamStride], record v*S + s
    int V, S, H, W;
    float stride;         // input-image pixels per raster pixel (stride * 2^level)
    int skip_invalid;     // 1: points with depth <= 0 are not searchable
};

// tan-range [lo, hi] of the directions (in the half plane z > 0) from the origin to the disk of radius r centred at
// (z = cz, a = ca).  false = the disk is not visible at all.
__device__ __forceinline__ bool tan_range(float ca, float cz, float r, float &lo, float &hi) {
    float rho2 = fmaf(ca, ca, cz * cz), r2 = r * r;
    lo = -INFINITY;
    hi = INFINITY;
    if (!(rho2 > r2)) return true;     // origin inside the disk (or NaN): every direction
    float s = sqrtf(rho2 - r2);
    float cl = fmaf(cz, s, ca * r), sl = fmaf(ca, s, -cz * r);     // cos, sin of (theta0 - alpha), scaled by rho^2
    float ch = fmaf(cz, s, -ca * r), sh = fmaf(ca, s, cz * r);     // cos, sin of (theta0 + alpha)
    if (cl > 0.f) lo = sl / cl; else if (sl > 0.f) return false;
    if (ch > 0.f) hi = sh / ch; else if (sh < 0.f) return false;
    return true;
}

struct Rect { int x0, y0, x1, y1; };   // [x0,x1) x [y0,y1); empty when x1 <= x0 or y1 <= y0
__device__ __forceinline__ bool rect_empty(const Rect &r) { return r.x1 <= r.x0 || r.y1 <= r.y0; }

// float interval [lo, hi] of raster coordinates -> clipped integer pixel range [i0, i1)
__device__ __forceinline__ void pix_range(float ulo, float uhi, float st, int n, int &i0, int &i1) {
    const float pm = 0.05f;
    float flo = ceilf((ulo + 0.5f) / st - 0.5f - pm), fhi = floorf((uhi + 0.5f) / st - 0.5f + pm);
    flo = fmaxf(flo, 0.f);                 // NaN-safe: fmaxf/fminf return the non-NaN operand
    fhi = fminf(fhi, (float)(n - 1));
    if (!(flo <= fhi)) { i0 = 0; i1 = 0; return; }
    i0 = (int)flo;
    i1 = (int)fhi + 1;
}

__device__ __forceinline__ void raster_search(const RasterView &rv, int s, float qx, float qy, float qz, int k, int lane,
                                              float &bd, int &bi, unsigned long long *stats, int qtag = 0) {
    bd = INFINITY;
    bi = kNoIdx;
    float td = INFINITY;
    int ti = kNoIdx;
    const int W = rv.W, H = rv.H, HW = W * H;
    const float4 *frame = rv.xyz4 + (size_t)s * rv.V * HW;
    const bool mine = lane < rv.V;

    // per-lane (= per-view) camera and projection of q
    float X = 0.f, Y = 0.f, Z = -1.f, fx = 1.f, sk = 0.f, cx0 = 0.f, fy = 1.f, cy0 = 0.f, tnorm = 0.f;
    bool kok = false;
    if (mine) {
        const float *cm = rv.cam + ((size_t)lane * rv.S + s) * kCamStride;
        X = fmaf(cm[24], qx, fmaf(cm[25], qy, fmaf(cm[26], qz, cm[33])));
        Y = fmaf(cm[27], qx, fmaf(cm[28], qy, fmaf(cm[29], qz, cm[34])));
        Z = fmaf(cm[30], qx, fmaf(cm[31], qy, fmaf(cm[32], qz, cm[35])));
        fx = cm[36]; sk = cm[37]; cx0 = cm[38]; fy = cm[39]; cy0 = cm[40];
        kok = cm[41] != 0.f;
        tnorm = fabsf(cm[33]) + fabsf(cm[34]) + fabsf(cm[35]);
    }
    const bool front = mine && kok && Z > 1e-12f;
    // integer window centre (clamped so that centre +- m cannot overflow)
    int ci = 0, cj = 0;
    if (front) {
        float a = X / Z, b = Y / Z;
        float fi = (fmaf(fx, a, fmaf(sk, b, cx0)) + 0.5f) / rv.stride - 0.5f;
        float fj = (fmaf(fy, b, cy0) + 0.5f) / rv.stride - 0.5f;
        ci = (int)fminf(fmaxf(rintf(fi), -1.0e8f), 1.0e8f);
        cj = (int)fminf(fmaxf(rintf(fj), -1.0e8f), 1.0e8f);
    }
    const Rect whole = {0, 0, W, H};
    Rect A = {0, 0, 0, 0}, B = {0, 0, 0, 0};
    int m = 2;
    auto window = [&](int mm) {
        Rect r = {0, 0, 0, 0};
        if (!mine) return r;
        if (!front) return r;       // behind the camera / odd intrinsics: no window centre (handled in the growth step)
        r.x0 = max(ci - mm, 0); r.x1 = min(ci + mm + 1, W);
        r.y0 = max(cj - mm, 0); r.y1 = min(cj + mm + 1, H);
        if (rect_empty(r)) r = Rect{0, 0, 0, 0};
        return r;
    };
    B = window(m);
    unsigned long long st_cand = 0, st_ins = 0;
    int st_pass = 0;

    while (true) {
        // ---- scan B \ A over all views ----
        const bool same = (B.x0 == A.x0 && B.y0 == A.y0 && B.x1 == A.x1 && B.y1 == A.y1);
        const int bw = B.x1 - B.x0;
        unsigned cnt = (rect_empty(B) || same) ? 0u : (unsigned)bw * (unsigned)(B.y1 - B.y0);
        unsigned incl = cnt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            unsigned t = __shfl_up_sync(FULL, incl, o);
            if (lane >= o) incl += t;
        }
        const unsigned excl = incl - cnt;
        const unsigned total = __shfl_sync(FULL, incl, 31);
        st_cand += total;
        ++st_pass;
        for (unsigned i0 = 0; i0 < total; i0 += 32) {
            unsigned t = i0 + lane;
            bool valid = t < total;
            int j = 0;
#pragma unroll
            for (int step = 16; step; step >>= 1) {
                int pr = j + step;
                unsigned e = __shfl_sync(FULL, excl, pr);
                if (e <= t) j = pr;
            }
            unsigned e0 = __shfl_sync(FULL, excl, j);
            int wj = __shfl_sync(FULL, bw, j), x0 = __shfl_sync(FULL, B.x0, j), y0 = __shfl_sync(FULL, B.y0, j);
            int ax0 = __shfl_sync(FULL, A.x0, j), ay0 = __shfl_sync(FULL, A.y0, j);
            int ax1 = __shfl_sync(FULL, A.x1, j), ay1 = __shfl_sync(FULL, A.y1, j);
            unsigned o = valid ? t - e0 : 0u;
            unsigned row = o / (unsigned)max(wj, 1);
            int px = x0 + (int)(o - row * (unsigned)max(wj, 1)), py = y0 + (int)row;
            valid = valid && !(px >= ax0 && px < ax1 && py >= ay0 && py < ay1);
            int idx = j * HW + py * W + px;
            float4 p = valid ? __ldg(frame + idx) : make_float4(0.f, 0.f, 0.f, 0.f);
            valid = valid && (rv.skip_invalid ? (p.w > 0.f) : !(p.w < 0.f));
            float d2 = dist2(p.x, p.y, p.z, qx, qy, qz);
            st_ins += topk_offer(bd, bi, td, ti, d2, idx, valid, k, lane);
        }
        if (!rect_empty(B)) A = B;

        const bool have_k = ti != kNoIdx;
        Rect R = whole;      // pixels of my view that can still hold a neighbour
        bool covered;
        if (have_k) {
            // pixel bounding box of the ball B(q, r), r = current k-th best distance (an upper bound of the final one)
            float r = sqrtf(td);
            float reff = fmaf(r, 1.0001f, 4e-6f * (fabsf(X) + fabsf(Y) + fabsf(Z) + tnorm + r));
            if (mine && kok) {
                R = Rect{0, 0, 0, 0};
                float alo, ahi, blo, bhi;
                bool vis = tan_range(X, Z, reff, alo, ahi) && tan_range(Y, Z, reff, blo, bhi);
                if (fmaf(X, X, fmaf(Y, Y, Z * Z)) <= reff * reff) { vis = true; alo = blo = -INFINITY; ahi = bhi = INFINITY; }
                if (vis) {
                    float ulo = fmaf(fx, alo, cx0), uhi = fmaf(fx, ahi, cx0);
                    if (sk != 0.f) {
                        float s0 = sk * blo, s1 = sk * bhi;
                        ulo += fminf(s0, s1); uhi += fmaxf(s0, s1);
                    }
                    float vlo = fmaf(fy, blo, cy0), vhi = fmaf(fy, bhi, cy0);
                    pix_range(ulo, uhi, rv.stride, W, R.x0, R.x1);
                    pix_range(vlo, vhi, rv.stride, H, R.y0, R.y1);
                    if (rect_empty(R)) R = Rect{0, 0, 0, 0};
                }
            }
            covered = !mine || rect_empty(R) ||
                      (!rect_empty(A) && R.x0 >= A.x0 && R.y0 >= A.y0 && R.x1 <= A.x1 && R.y1 <= A.y1);
        } else {
            // fewer than k searchable points so far; stop only when all rasters are covered
            covered = !mine || (A.x0 == 0 && A.y0 == 0 && A.x1 == W && A.y1 == H);
        }
        if (__all_sync(FULL, covered)) break;     // every view's ball projection has been scanned: exact
        // views without a window centre (q behind the camera, odd intrinsics) wait until the views that have one are
        // done, so that they are scanned once, with the final (tight) radius
        const bool front_pending = __any_sync(FULL, front && !covered);
        // grow the windows geometrically, but never beyond what the current ball can reach; the bound tightens as
        // nearer points are found, so the scanned area stays within a small factor of the final ball's projection
        m = (m < (1 << 24)) ? 2 * m + 1 : m;
        Rect g = {0, 0, 0, 0};
        if (front) g = (m >= (1 << 24)) ? whole : window(m);
        else if (mine && !front_pending) g = whole;
        if (!rect_empty(g)) {
            g.x0 = max(g.x0, R.x0); g.y0 = max(g.y0, R.y0); g.x1 = min(g.x1, R.x1); g.y1 = min(g.y1, R.y1);
        }
        if (rect_empty(g)) {
            B = A;
        } else if (rect_empty(A)) {
            B = g;
        } else {
            B.x0 = min(g.x0, A.x0); B.y0 = min(g.y0, A.y0); B.x1 = max(g.x1, A.x1); B.y1 = max(g.y1, A.y1);
        }
    }
    if (stats && lane == 0) {
        atomicAdd(stats + 0, 1ull);
        atomicAdd(stats + min(st_pass, 4), 1ull);     // [1] one pass, [2] two passes, [3] three, [4] four or more
        atomicAdd(stats + 5, st_cand);
        atomicAdd(stats + 6, st_ins);
        // [9] worst query: candidates << 32 | (frame << 24 ^ low bits of qx as a tag is not needed: host recomputes)
        atomicMax(stats + 9, (st_cand << 24) | (unsigned long long)(qtag & 0xFFFFFF));
        int bin = 0;
        unsigned long long c = st_cand;
        while (c > 1 && bin < 5) { c >>= 3; ++bin; }       // bins: <2, <16, <128, <1024, <8192, >=8192 candidates
        atomicAdd(stats + 10 + bin, 1ull);
    }
}

}  // namespace mvt
