// Raster kNN: exact k nearest neighbours in a cloud that was unprojected from V depth rasters (cloud_build's xyz4
// records), searched directly in the rasters -- no spatial index and no build step.
//
// Every cloud point lies on the camera ray of its pixel, so a point within distance r of the query q must have its
// pixel inside the projection of the ball B(q, r) into that view.  One group of GW lanes per query (a warp, or a half
// warp when k <= 16 and V <= 16 -- two queries per warp halve the per-query cost of everything that is not per candidate):
//   pass 1   a fixed kPW x kPH pixel window around q's projection in every view: one 16-byte record per lane, the
//            loads of several chunks in flight together.  Pixels are taken nearest-to-the-projection first across
//            all views, so the first batch (sorted as a whole, bitonic) already holds most of the answer and few of
//            the later candidates beat its k-th best (those are inserted one by one);
//   bound    r = current k-th best distance;  group lane v computes the pixel bounding box R_v of B(q, r) in view v
//            (tangent directions of the ball seen from the camera centre, clipped to z > 0);
//   done     when every R_v lies inside what has been scanned -- the result is then exact;
//   grow     otherwise the windows grow (m -> 2m+1) clipped to R_v; only the NEW pixels are scanned: the up to four
//            strips of (new window \ old window) of every view are flattened into one candidate list, GW per step.
//            r only shrinks, so once every R_v has been scanned the search is over without another bound.
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
    const float2 *ztab;   // [S, V, TH, TW] {min, max} view-space depth per 8x8 tile, or null (tile_depth_kernel)
    int TW, TH;
    const int *neg;       // [S] frame holds points with depth < 0 (behind the camera on their pixel's line), or null
};
#ifndef MVT_TOPK_BUF
#define MVT_TOPK_BUF 0
#endif
constexpr int kTile = 8;          // == MVT_TILE
#ifndef MVT_VCONST
#define MVT_VCONST 0        // 1 (A/B): lane-constant view in the first pass when V divides 32.  Measured SLOWER on B200 (panoptic
                            // 0.757 vs 0.715 ms, DROID -- which does not even take the path -- 0.301 vs 0.283): same 56 registers,
                            // the second copy of the loop body costs more than the ~80 instructions per query it saves
#endif
constexpr int kTileMinPixels = 384;   // a pass that would read this many new pixels of one view goes to tile_search instead

#ifndef MVT_M0
#define MVT_M0 2
#endif
constexpr int kM0 = MVT_M0;                            // first-pass half window
constexpr int kPW = 2 * kM0 + 2, kPH = 2 * kM0 + 1;    // 6 x 5 pixels: x in [floor(u)-2, floor(u)+3], y in [rint(v)-2, rint(v)+2]

// The first-pass window pixels ordered nearest-first from the query's projection (which lies in [0,1) x [-0.5,0.5]
// relative to pixel (kM0, kM0) of the window): tab[rank] = wx | wy << 4.  Candidates are then enumerated rank-major,
// view-minor, so the first batch -- the one that is sorted and sets the first threshold -- holds the pixels nearest
// to the projection in EVERY view and few later candidates survive it.
// The table is a compile-time constant (eight 32-bit immediates); every WARP keeps its own 32-byte copy in shared memory
// (lane-indexed lookups), so no block-wide barrier is needed before the first search.
constexpr int kNW = kPW * kPH;
static_assert(kNW <= 32 && kPW <= 16 && kPH <= 16, "window table: 32 one-byte entries");
__host__ __device__ constexpr int win_key(int x, int y) {
    return (2 * x - (2 * kM0 + 1)) * (2 * x - (2 * kM0 + 1)) + (2 * (y - kM0)) * (2 * (y - kM0));
}
// 32-bit word i of the table: entries 4i .. 4i+3 (entry = wx | wy << 4 of the pixel with that nearest-first rank)
__host__ __device__ constexpr unsigned win_tab_word(int i) {
    unsigned word = 0;
    for (int l = 0; l < kNW; ++l) {
        const int wy = l / kPW, wx = l - wy * kPW, mykey = win_key(wx, wy);
        int rank = 0;
        for (int o = 0; o < kNW; ++o) {
            const int oy = o / kPW, ox = o - oy * kPW, ok = win_key(ox, oy);
            rank += (ok < mykey || (ok == mykey && o < l)) ? 1 : 0;
        }
        if ((rank >> 2) == i) word |= (unsigned)(wx | (wy << 4)) << (8 * (rank & 3));
    }
    return word;
}
template <int I> struct WinWord { static constexpr unsigned value = win_tab_word(I); };
// Per-warp shared memory of the search: the window table and, right behind it, the pending buffer of the top-k list.
#ifndef MVT_CAM_SMEM
#define MVT_CAM_SMEM 1
#endif
#ifndef MVT_TOURNAMENT
#define MVT_TOURNAMENT 1
#endif
constexpr int kTQ = 8;            // first-pass candidates per lane that go through the tournament (kTQ * 32 in all)
#if MVT_CAM_SMEM
// cam[i][lane]: the lane's (= its view's) projection constants, read back once per pass for the ball's pixel box --
// ten registers per thread that need not stay live across the scans
// queue[i - 1][lane]: the lane's i-th best first-pass candidate (tournament selection, see tournament_select), which
// hands out its winners through pend[]
#if MVT_TOURNAMENT
struct __align__(16) WarpScratch { unsigned char tab[32]; u64 pend[32]; float cam[11][32]; u64 queue[kTQ - 1][32]; };
#else
struct __align__(16) WarpScratch { unsigned char tab[32]; u64 pend[32]; float cam[11][32]; };
#endif
#else
struct __align__(16) WarpScratch { unsigned char tab[32]; u64 pend[32]; };
#endif
// declares the block's scratch (one per warp), hands this warp its table pointer `name` and fills the table
#define MVT_WARP_SCRATCH(name, nwarps)                                   \
    __shared__ mvt::WarpScratch name##_all[nwarps];                      \
    unsigned char *name = name##_all[threadIdx.x >> 5].tab;              \
    mvt::init_window_table(name)
__device__ __forceinline__ u64 *scratch_pend(const unsigned char *win_tab) {
    return reinterpret_cast<u64 *>(const_cast<unsigned char *>(win_tab) + 32);
}
// tab: this warp's 32 bytes of shared memory (4-byte aligned).  Call with all lanes of the warp.
__device__ __forceinline__ void init_window_table(unsigned char *tab) {
    const int l = threadIdx.x & 31;
    const unsigned w = l == 0 ? WinWord<0>::value : l == 1 ? WinWord<1>::value : l == 2 ? WinWord<2>::value
                     : l == 3 ? WinWord<3>::value : l == 4 ? WinWord<4>::value : l == 5 ? WinWord<5>::value
                     : l == 6 ? WinWord<6>::value : WinWord<7>::value;
    if (l < 8) reinterpret_cast<unsigned *>(tab)[l] = w;
    __syncwarp();
}

// tan-range [lo, hi] of the directions (in the half plane z > 0) from the origin to the disk of radius r centred at
// (z = cz, a = ca).  false = the disk is not visible at all.  Approximate divisions: callers inflate r and the pixel
// range by margins that are orders of magnitude above their error.
__device__ __forceinline__ bool tan_range(float ca, float cz, float r, float &lo, float &hi) {
    float rho2 = fmaf(ca, ca, cz * cz), r2 = r * r;
    lo = -INFINITY;
    hi = INFINITY;
    if (!(rho2 > r2)) return true;     // origin inside the disk (or NaN): every direction
    float s = sqrtf(rho2 - r2);
    float cl = fmaf(cz, s, ca * r), sl = fmaf(ca, s, -cz * r);     // cos, sin of (theta0 - alpha), scaled by rho^2
    float ch = fmaf(cz, s, -ca * r), sh = fmaf(ca, s, cz * r);     // cos, sin of (theta0 + alpha)
    // The directions seen are the arc [theta0 - alpha, theta0 + alpha] (shorter than 180 degrees); keep its part with z > 0.
    if (cl > 0.f) {                        // the arc starts in front: it ends in front too (hi finite) or leaves over +90 deg
        lo = __fdividef(sl, cl);
        if (ch > 0.f) hi = __fdividef(sh, ch);
        return true;
    }
    if (ch > 0.f) {                        // the arc starts behind the camera and enters the front half over -90 deg
        hi = __fdividef(sh, ch);
        return true;
    }
    return false;                          // both ends behind: nothing of the disk has z > 0
}

// The same range when the disk lies strictly in front (cz > r, checked by the caller): tan(theta0 -+ alpha) by the
// addition formula, tan(alpha) = r / sqrt(rho^2 - r^2).  lo is always finite; hi is +inf when theta0 + alpha >= 90 deg.
__device__ __forceinline__ void tan_range_front(float ca, float cz, float icz, float r, float &lo, float &hi) {
    float t0 = ca * icz;
    float ta = r * rsqrtf(fmaf(ca, ca, fmaf(cz, cz, -r * r)));
    float dl = fmaf(t0, ta, 1.f), dh = fmaf(-t0, ta, 1.f);
    lo = (dl > 0.f) ? __fdividef(t0 - ta, dl) : -INFINITY;
    hi = (dh > 0.f) ? __fdividef(t0 + ta, dh) : INFINITY;
}

#if MVT_TOURNAMENT && MVT_CAM_SMEM
__device__ __forceinline__ void cswap(u64 &a, u64 &b) {
    const bool sw = b < a;
    const u64 lo = sw ? b : a, hi = sw ? a : b;
    a = lo; b = hi;
}
// The k smallest of the warp's 32 * kTQ first-pass keys, in canonical (d2, index) order, WITHOUT a warp-wide sort and
// without one-by-one insertion: every lane sorts its own kTQ keys in registers (a 19-exchange network, no shuffles) and
// parks them in its shared-memory queue; then k rounds of "the smallest head wins": one REDUX.MIN over the heads' distance
// bits, a vote, and -- only when two heads tie on the distance -- a second REDUX.MIN over their indices.  The winner writes
// its key to sel[round] and advances its queue.  ~11 warp instructions per round, against ~20 per insertion (and ~28
// insertions after a 200-instruction bitonic sort) of the list update it replaces.  Keys >= kEmptyKey64 (no candidate, NaN
// distance) never win before a real point.  Returns with lane j < k holding the j-th best key (kEmptyKey64 if fewer).
__device__ __forceinline__ u64 tournament_select(u64 (&key)[kTQ], int n_used, unsigned char *win_tab, int k) {
    const int lane = threadIdx.x & 31;
    u64 *sel = reinterpret_cast<u64 *>(win_tab + 32);
    u64 (*queue)[32] = reinterpret_cast<u64 (*)[32]>(win_tab + 32 + 32 * sizeof(u64) + 11 * 32 * sizeof(float));
    if (n_used > 4) {       // (uniform) optimal 8-input network, 19 exchanges
        cswap(key[0], key[1]); cswap(key[2], key[3]); cswap(key[4], key[5]); cswap(key[6], key[7]);
        cswap(key[0], key[2]); cswap(key[1], key[3]); cswap(key[4], key[6]); cswap(key[5], key[7]);
        cswap(key[1], key[2]); cswap(key[5], key[6]); cswap(key[0], key[4]); cswap(key[3], key[7]);
        cswap(key[1], key[5]); cswap(key[2], key[6]);
        cswap(key[1], key[4]); cswap(key[3], key[6]);
        cswap(key[2], key[4]); cswap(key[3], key[5]);
        cswap(key[3], key[4]);
    } else {                // keys 4..7 are empty: 4-input network, 5 exchanges
        cswap(key[0], key[1]); cswap(key[2], key[3]); cswap(key[0], key[2]); cswap(key[1], key[3]); cswap(key[1], key[2]);
    }
#pragma unroll
    for (int i = 1; i < kTQ; ++i) queue[i - 1][lane] = key[i];
    u64 cur = key[0];
    int head = 0, nsel = 0;
    __syncwarp();
    for (int j = 0; j < k; ++j, ++nsel) {
        const unsigned hi = (unsigned)(cur >> 32);
        const unsigned m = __reduce_min_sync(FULL, hi);
        if (m > 0x7f800000u) break;                     // (uniform) only NaN distances are left
        bool win = hi == m;
        unsigned bal = __ballot_sync(FULL, win);
        if (bal & (bal - 1u)) {                         // (uniform) several heads at this distance: the smallest index wins
            const unsigned lo = win ? (unsigned)cur : 0xffffffffu;
            const unsigned ml = __reduce_min_sync(FULL, lo);
            win = win && lo == ml;
            bal = __ballot_sync(FULL, win);
            if (bal & (bal - 1u)) win = lane == __ffs(bal) - 1;     // the same key twice (all queues empty): any one
        }
        if (win) {
            sel[j] = cur;
            cur = head < kTQ - 1 ? queue[head][lane] : kEmptyKey64;
            ++head;
        }
    }
    __syncwarp();
    const u64 mine = lane < nsel ? sel[lane] : kEmptyKey64;
    __syncwarp();
    return mine;
}
#endif

struct Rect { int x0, y0, x1, y1; };   // [x0,x1) x [y0,y1); empty when x1 <= x0 or y1 <= y0
__device__ __forceinline__ bool rect_empty(const Rect &r) { return r.x1 <= r.x0 || r.y1 <= r.y0; }
__device__ __forceinline__ bool rect_inside(const Rect &r, const Rect &a) {   // r (non-empty) inside a
    return !rect_empty(a) && r.x0 >= a.x0 && r.y0 >= a.y0 && r.x1 <= a.x1 && r.y1 <= a.y1;
}

// float interval [lo, hi] of input-image coordinates -> clipped integer raster pixel range [i0, i1)
__device__ __forceinline__ void pix_range(float ulo, float uhi, float inv_st, int n, int &i0, int &i1) {
    const float pm = 0.05f;
    float flo = ceilf(fmaf(ulo + 0.5f, inv_st, -0.5f - pm)), fhi = floorf(fmaf(uhi + 0.5f, inv_st, -0.5f + pm));
    flo = fmaxf(flo, 0.f);                 // NaN-safe: fmaxf/fminf return the non-NaN operand
    fhi = fminf(fhi, (float)(n - 1));
    if (!(flo <= fhi)) { i0 = 0; i1 = 0; return; }
    i0 = (int)flo;
    i1 = (int)fhi + 1;
}

// Pixel bounding box, in one view, of the ball of radius reff around the view-space point (X, Y, Z): every pixel whose
// ray can hold a point of the ball.  front: Z > 0 and iz = 1/Z is valid.  Empty ({0,0,0,0}) when the view cannot see it.
__device__ __forceinline__ Rect ball_box(bool front, float X, float Y, float Z, float iz, float reff, float fx, float sk,
                                         float cx0, float fy, float cy0, float inv_st, int W, int H) {
    Rect R = {0, 0, 0, 0};
    float alo, ahi, blo, bhi;
    bool vis = true;
    if (front && Z > 1.01f * reff) {   // the usual case: the ball is strictly in front of the camera
        tan_range_front(X, Z, iz, reff, alo, ahi);
        tan_range_front(Y, Z, iz, reff, blo, bhi);
    } else {
        vis = tan_range(X, Z, reff, alo, ahi) && tan_range(Y, Z, reff, blo, bhi);
        if (fmaf(X, X, fmaf(Y, Y, Z * Z)) <= reff * reff) { vis = true; alo = blo = -INFINITY; ahi = bhi = INFINITY; }
    }
    if (vis) {
        float ulo = fmaf(fx, alo, cx0), uhi = fmaf(fx, ahi, cx0);
        if (sk != 0.f) {
            float s0 = sk * blo, s1 = sk * bhi;
            ulo += fminf(s0, s1); uhi += fmaxf(s0, s1);
        }
        float vlo = fmaf(fy, blo, cy0), vhi = fmaf(fy, bhi, cy0);
        pix_range(ulo, uhi, inv_st, W, R.x0, R.x1);
        pix_range(vlo, vhi, inv_st, H, R.y0, R.y1);
        if (rect_empty(R)) R = Rect{0, 0, 0, 0};
    }
    return R;
}

// Finishes a query that the window search (raster_windows) gave up on because a pass would have read a large pixel box
// of some view (a query next to a camera, or one far from every surface).  Same idea one level up: the unit is the 8x8
// tile, a tile is read only if the depth interval the ball spans in that view, |Z_p - Z_q| <= r * |depth axis|, meets
// the tile's {min, max} depth (tile_depth_kernel), and tiles are visited in rings around the query's projection, ring
// rho of every view before ring rho + 1, so that the k-th distance r -- and with it every view's box and depth interval
// -- tightens after the first few tiles.  Restarts with an empty list capped at the k-th distance found so far: every
// true neighbour is at most that far and no point can enter twice (each tile of each view is visited at most once).
// Lane v holds view v.  Exact.  Full warps only.
static __device__ __noinline__ TopK tile_search(const RasterView rv, const int s, const float qx, const float qy, const float qz,
                                                const int k, TopK tk, unsigned long long *stats) {
    const int lane = threadIdx.x & 31;
    const Group<32> g(lane);
    const int W = rv.W, H = rv.H, HW = W * H, V = rv.V, TW = rv.TW, TH = rv.TH;
    const float4 *frame = rv.xyz4 + (size_t)s * V * HW;
    const float2 *ztab = rv.ztab + (size_t)s * V * TW * TH;
    const float inv_st = __frcp_rn(rv.stride);
    const bool skip_invalid = rv.skip_invalid != 0;
    u64 cap = tk.th | 0xffffffffull;          // tk holds k real points: th is the k-th key, finite
    if (tk.cap < cap) cap = tk.cap;
    tk.init(cap);
    float X = 0.f, Y = 0.f, Z = -1.f, iz = 0.f, fx = 1.f, sk = 0.f, cx0 = 0.f, fy = 1.f, cy0 = 0.f, zpad = 0.f, nz = 1.f, smax = 1.f;
    bool kok = false;
    const bool mine = lane < V;
    if (mine) {
        const float *cm = rv.cam + ((size_t)lane * rv.S + s) * kCamStride;
        smax = cm[42];
        X = fmaf(cm[24], qx, fmaf(cm[25], qy, fmaf(cm[26], qz, cm[33])));
        Y = fmaf(cm[27], qx, fmaf(cm[28], qy, fmaf(cm[29], qz, cm[34])));
        Z = fmaf(cm[30], qx, fmaf(cm[31], qy, fmaf(cm[32], qz, cm[35])));
        fx = cm[36]; sk = cm[37]; cx0 = cm[38]; fy = cm[39]; cy0 = cm[40];
        kok = cm[41] != 0.f;
        zpad = 4e-6f * (fabsf(X) + fabsf(Y) + fabsf(Z) + fabsf(cm[33]) + fabsf(cm[34]) + fabsf(cm[35]));
        nz = sqrtf(fmaf(cm[30], cm[30], fmaf(cm[31], cm[31], cm[32] * cm[32])));   // 1 for a rigid view transform
    }
    const bool front = kok && Z > 1e-12f;
    if (front) iz = __frcp_rn(Z);
    // this view's pixel box for the current threshold (odd intrinsics: the whole raster; depth culling still applies)
    auto box = [&]() {
        const float r = sqrtf(key_d2(tk.th));
        Rect R = {0, 0, 0, 0};
        if (mine) R = kok ? ball_box(front, X, Y, Z, iz, fmaf(r, 1.0001f, fmaf(4e-6f, r, zpad)) * smax, fx, sk, cx0, fy, cy0, inv_st, W, H)
                          : Rect{0, 0, W, H};
        return R;
    };
    // ring centre: the tile of q's projection, clamped into the first box (its middle when q is behind the camera)
    int tcx = 0, tcy = 0;
    {
        const Rect R = box();
        int px = (R.x0 + R.x1) >> 1, py = (R.y0 + R.y1) >> 1;
        if (front && !rect_empty(R)) {
            const float a = X * iz, b = Y * iz;
            const float fi = fmaf(fmaf(fx, a, fmaf(sk, b, cx0)) + 0.5f, inv_st, -0.5f);
            const float fj = fmaf(fmaf(fy, b, cy0) + 0.5f, inv_st, -0.5f);
            px = (int)fminf(fmaxf(fi, (float)R.x0), (float)(R.x1 - 1));
            py = (int)fminf(fmaxf(fj, (float)R.y0), (float)(R.y1 - 1));
        }
        tcx = px / kTile; tcy = py / kTile;
    }
    unsigned tested = 0, kept = 0, st_ins = 0;
    for (int rho = 0; rho <= max(TW, TH); ++rho) {
        topk_flush(tk, k);
        // views whose current box still reaches ring rho
        const Rect R = box();
        const int tx0 = R.x0 / kTile, tx1 = (R.x1 + kTile - 1) / kTile, ty0 = R.y0 / kTile, ty1 = (R.y1 + kTile - 1) / kTile;   // [ , )
        const bool active = !rect_empty(R) && rho <= max(max(tcx - tx0, tx1 - 1 - tcx), max(tcy - ty0, ty1 - 1 - tcy));
        const unsigned amask = g.ballot(active);
        if (amask == 0u) break;
        const float r = sqrtf(key_d2(tk.th));
        const float zhalf = fmaf(r, 1.0001f, fmaf(4e-6f, r, zpad)) * nz;
        const int cnt = rho == 0 ? 1 : 8 * rho;            // tiles of one ring
        const int total = cnt * __popc(amask);
        for (int t0 = 0; t0 < total; t0 += 32) {
            const int t = t0 + lane;
            const int vi = t / cnt, i = t - vi * cnt;      // vi-th active view, i-th tile of its ring
            const int v = (t < total) ? (int)__fns(amask, 0, vi + 1) : 0;
            const int side = rho ? i / (2 * rho) : 0, j = rho ? i - side * 2 * rho : 0;
            const int dx = side == 0 ? -rho + j : side == 1 ? rho : side == 2 ? rho - j : -rho;
            const int dy = side == 0 ? -rho : side == 1 ? -rho + j : side == 2 ? rho : rho - j;
            const int tx = g.shfl(tcx, v) + dx, ty = g.shfl(tcy, v) + dy;
            const int bx0 = g.shfl(tx0, v), bx1 = g.shfl(tx1, v), by0 = g.shfl(ty0, v), by1 = g.shfl(ty1, v);
            const float zq = g.shfl(Z, v), zh = g.shfl(zhalf, v);
            bool keep = t < total && tx >= bx0 && tx < bx1 && ty >= by0 && ty < by1;     // inside the view's box (and the raster)
            if (keep) {
                const float2 zr = __ldg(ztab + ((size_t)v * TH + ty) * TW + tx);
                keep = !(zr.x > zq + zh) && !(zr.y < zq - zh);                           // an empty tile holds {+inf, -inf}
            }
            unsigned mask = g.ballot(keep);
            tested += (unsigned)min(32, total - t0);
            kept += __popc(mask);
            // the kept tiles, kTB at a time: all 2 * kTB record loads of a round are in flight before the first list update
            // (a query that ends up here is latency-bound: hundreds of dependent load rounds otherwise)
            constexpr int kTB = 4;
            while (mask) {
                float4 p0[kTB], p1[kTB];
                int i0[kTB], i1[kTB];
                bool v0[kTB], v1[kTB];
#pragma unroll
                for (int u = 0; u < kTB; ++u) {
                    const bool on = mask != 0u;
                    const int src = on ? __ffs(mask) - 1 : 0;
                    mask &= mask - 1;
                    const int vv = g.shfl(v, src);
                    const int px = g.shfl(tx, src) * kTile + (lane & 7), py0 = g.shfl(ty, src) * kTile + (lane >> 3), py1 = py0 + 4;
                    const bool in0 = on && px < W && py0 < H, in1 = on && px < W && py1 < H;
                    i0[u] = vv * HW + py0 * W + px; i1[u] = vv * HW + py1 * W + px;
                    p0[u] = in0 ? __ldg(frame + i0[u]) : make_float4(0.f, 0.f, 0.f, 0.f);
                    p1[u] = in1 ? __ldg(frame + i1[u]) : make_float4(0.f, 0.f, 0.f, 0.f);
                    v0[u] = in0; v1[u] = in1;
                }
#pragma unroll
                for (int u = 0; u < kTB; ++u) {
                    const bool a0 = v0[u] && (skip_invalid ? (p0[u].w > 0.f) : !(p0[u].w < 0.f));
                    const bool a1 = v1[u] && (skip_invalid ? (p1[u].w > 0.f) : !(p1[u].w < 0.f));
                    st_ins += topk_offer<32, MVT_TOPK_BUF != 0>(g, tk, a0 ? pack_key(dist2(p0[u].x, p0[u].y, p0[u].z, qx, qy, qz), i0[u]) : kEmptyKey64, k);
                    st_ins += topk_offer<32, MVT_TOPK_BUF != 0>(g, tk, a1 ? pack_key(dist2(p1[u].x, p1[u].y, p1[u].z, qx, qy, qz), i1[u]) : kEmptyKey64, k);
                }
            }
        }
    }
    topk_flush(tk, k);
    if (stats && lane == 0) {
        atomicAdd(stats + 4, 1ull);        // [4]: queries finished here (raster_windows counts only those it finishes)
        atomicAdd(stats + 5, (unsigned long long)kept * 64ull);
        atomicAdd(stats + 6, (unsigned long long)st_ins);
        atomicAdd(stats + 7, (unsigned long long)tested);
        atomicAdd(stats + 8, (unsigned long long)kept);
    }
    return tk;
}

// win_tab: shared-memory table of init_window_table.
// cap: only points whose key is below it are wanted (see TopK).  budget != 0: bounded effort, see below.
// Returns true when the list is the exact answer (always, unless a budget made the search give up).
// live = false: the group has no query (padding at the end of a frame's track list); it only takes part in the
// collectives.  Must be called by all 32 lanes.
// soft_count (TILES only): instead of giving a large-box query up, count it there (once) and carry on with the windows.
template <int GW, bool TILES>
__device__ __forceinline__ bool raster_windows(const Group<GW> &g, const RasterView &rv, const unsigned char *win_tab, int s,
                                               bool live, float qx, float qy, float qz, int k, TopK &tk,
                                               unsigned long long *stats, int qtag, u64 cap, unsigned budget,
                                               unsigned *soft_count = nullptr) {
    tk.init(cap);
    constexpr bool BUF = GW == 32 && MVT_TOPK_BUF != 0;
    tk.pend = scratch_pend(win_tab);
    const int gl = g.gl;
    const int W = rv.W, H = rv.H, HW = W * H, V = rv.V;
    const float4 *frame = rv.xyz4 + (size_t)s * V * HW;
    const bool mine = live && gl < V;
    const float inv_st = __frcp_rn(rv.stride);
    const bool skip_invalid = rv.skip_invalid != 0;
    // points with depth < 0 are not on the forward ray of their pixel: a frame that holds some (and must return them) has
    // no usable pixel box -- every view is scanned whole, once, like a view with odd intrinsics (uniform in the warp)
    const bool negf = !skip_invalid && rv.neg != nullptr && __ldg(rv.neg + s) != 0;
    const float wmin = negf ? -INFINITY : 0.f;          // searchable: !(depth < wmin)

    // per-lane (= per-view) camera and projection of q
    float X = 0.f, Y = 0.f, Z = -1.f, iz = 0.f, fx = 1.f, sk = 0.f, cx0 = 0.f, fy = 1.f, cy0 = 0.f, tnorm = 0.f, smax = 1.f;
    bool kok = false;
    if (mine) {
        const float *cm = rv.cam + ((size_t)gl * rv.S + s) * kCamStride;
        smax = cm[42];      // view-space radius of a unit world ball (1 for a rigid view transform)
        X = fmaf(cm[24], qx, fmaf(cm[25], qy, fmaf(cm[26], qz, cm[33])));
        Y = fmaf(cm[27], qx, fmaf(cm[28], qy, fmaf(cm[29], qz, cm[34])));
        Z = fmaf(cm[30], qx, fmaf(cm[31], qy, fmaf(cm[32], qz, cm[35])));
        fx = cm[36]; sk = cm[37]; cx0 = cm[38]; fy = cm[39]; cy0 = cm[40];
        kok = cm[41] != 0.f && !negf;
        tnorm = fabsf(cm[33]) + fabsf(cm[34]) + fabsf(cm[35]);
    }
    const bool front = mine && kok && Z > 1e-12f;
    // integer window centre (clamped so that centre +- m cannot overflow): x windows are [ci-m, ci+m+1], y [cj-m, cj+m]
    int ci = 0, cj = 0;
    if (front) {
        iz = __frcp_rn(Z);
        float a = X * iz, b = Y * iz;
        float fi = fmaf(fmaf(fx, a, fmaf(sk, b, cx0)) + 0.5f, inv_st, -0.5f);
        float fj = fmaf(fmaf(fy, b, cy0) + 0.5f, inv_st, -0.5f);
        ci = (int)fminf(fmaxf(floorf(fi), -1.0e8f), 1.0e8f);
        cj = (int)fminf(fmaxf(rintf(fj), -1.0e8f), 1.0e8f);
    }
#if MVT_CAM_SMEM
    {
        float (*cs)[32] = reinterpret_cast<float (*)[32]>(const_cast<unsigned char *>(win_tab) + 32 + 32 * sizeof(u64));
        const int ln = threadIdx.x & 31;
        cs[0][ln] = X; cs[1][ln] = Y; cs[2][ln] = Z; cs[3][ln] = iz; cs[4][ln] = fx; cs[5][ln] = sk; cs[6][ln] = cx0;
        cs[7][ln] = fy; cs[8][ln] = cy0; cs[9][ln] = tnorm; cs[10][ln] = smax;
    }
#endif
    const Rect whole = {0, 0, W, H};
    auto window = [&](int mm) {
        Rect r = {0, 0, 0, 0};
        if (!front) return r;       // behind the camera / odd intrinsics: no window centre (handled in the growth step)
        r.x0 = max(ci - mm, 0); r.x1 = min(ci + mm + 2, W);
        r.y0 = max(cj - mm, 0); r.y1 = min(cj + mm + 1, H);
        if (rect_empty(r)) r = Rect{0, 0, 0, 0};
        return r;
    };
    unsigned st_cand = 0, st_ins = 0;
    int st_pass = 1;
    bool soft_counted = false;

    // ---- pass 1: the fixed window of every view that sees q in front of it; candidate t = rank * V + view ----
    {
        constexpr int kCB = 4;                              // chunks whose loads are in flight together
        const unsigned front_mask = g.ballot(front);
        const int total = kNW * V;
        const unsigned magic = (65536u + (unsigned)V - 1u) / (unsigned)V;     // t / V == (t * magic) >> 16 for t < 960
        const int x00 = ci - kM0, y00 = cj - kM0;
        int t_first = 0;
#if MVT_TOURNAMENT && MVT_CAM_SMEM
        if (GW == 32) {
            // the first 32 * kTQ candidates (all of them for V <= 8): tournament selection instead of sort + insertions
            u64 key[kTQ];
#if MVT_VCONST
            // V divides 32 (1, 2, 4, 8, 16, 32 views): candidate t = j * 32 + lane belongs to view lane % V whatever j is, so the
            // view's window origin, front bit and point offset are lane constants -- two shuffles per query, not per candidate
            const bool vconst = (32 % V) == 0;                         // (uniform)
            const int vl = gl & (V - 1);
            const int x0v = g.shfl(x00, vconst ? vl : 0), y0v = g.shfl(y00, vconst ? vl : 0);
            const bool frv = (front_mask >> vl) & 1u;
            const unsigned basev = (unsigned)vl * (unsigned)HW;
            const int per = vconst ? 32 / V : 0, rk0 = vconst ? gl / V : 0;
#endif
#pragma unroll
            for (int h = 0; h < kTQ / kCB; ++h) {
                float4 p[kCB];
                int id[kCB];
                bool ok[kCB];
#pragma unroll
                for (int c = 0; c < kCB; ++c) {
#if MVT_VCONST
                    if (vconst) {
                        const int rk = (h * kCB + c) * per + rk0;
                        const unsigned wo = win_tab[min(rk, kNW - 1)];
                        const int px = x0v + (int)(wo & 15u), py = y0v + (int)(wo >> 4);
                        ok[c] = rk < kNW && frv && (unsigned)px < (unsigned)W && (unsigned)py < (unsigned)H;
                        id[c] = (int)(basev + (unsigned)py * (unsigned)W + (unsigned)px);
                        p[c] = ok[c] ? __ldg(frame + id[c]) : make_float4(0.f, 0.f, 0.f, 0.f);
                        continue;
                    }
#endif
                    const int t = (h * kCB + c) * 32 + gl;
                    const int rk = (int)(((unsigned)t * magic) >> 16), v = t - rk * V;
                    const unsigned wo = win_tab[min(rk, kNW - 1)];
                    const int px = g.shfl(x00, v) + (int)(wo & 15u), py = g.shfl(y00, v) + (int)(wo >> 4);
                    ok[c] = t < total && ((front_mask >> v) & 1u) && (unsigned)px < (unsigned)W && (unsigned)py < (unsigned)H;
                    id[c] = (int)((unsigned)v * (unsigned)HW + (unsigned)py * (unsigned)W + (unsigned)px);
                    p[c] = ok[c] ? __ldg(frame + id[c]) : make_float4(0.f, 0.f, 0.f, 0.f);
                }
#pragma unroll
                for (int c = 0; c < kCB; ++c) {
                    const float d2 = dist2(p[c].x, p[c].y, p[c].z, qx, qy, qz);
                    const bool valid = ok[c] && (skip_invalid ? (p[c].w > 0.f) : !(p[c].w < wmin)) && d2 == d2;
                    u64 kk = valid ? pack_key(d2, id[c]) : kEmptyKey64;
                    if (cap != kEmptyKey64 && !(kk < cap)) kk = kEmptyKey64;
                    key[h * kCB + c] = kk;
                }
                if (h == 0 && total <= kCB * 32) {          // (uniform) few views: the second half is empty
#pragma unroll
                    for (int c = kCB; c < kTQ; ++c) key[c] = kEmptyKey64;
                    break;
                }
            }
            tk.t = tournament_select(key, (total + 31) >> 5, const_cast<unsigned char *>(win_tab), k);
            tk.empty = false;
            const u64 kth = g.shfl(tk.t, k - 1);
            tk.th = kth < tk.cap ? kth : tk.cap;
            t_first = kTQ * 32;
        }
#endif
        for (int t0 = t_first; t0 < total; t0 += kCB * GW) {
            float4 p[kCB];
            int id[kCB];
            bool ok[kCB];
#pragma unroll
            for (int c = 0; c < kCB; ++c) {
                const int t = t0 + c * GW + gl;
                const int rk = (int)(((unsigned)t * magic) >> 16), v = t - rk * V;
                const unsigned wo = win_tab[min(rk, kNW - 1)];
                const int px = g.shfl(x00, v) + (int)(wo & 15u), py = g.shfl(y00, v) + (int)(wo >> 4);
                ok[c] = t < total && ((front_mask >> v) & 1u) && (unsigned)px < (unsigned)W && (unsigned)py < (unsigned)H;
                id[c] = (int)((unsigned)v * (unsigned)HW + (unsigned)py * (unsigned)W + (unsigned)px);
                p[c] = ok[c] ? __ldg(frame + id[c]) : make_float4(0.f, 0.f, 0.f, 0.f);
            }
            if (TILES) {
                // search-only kernels: ONE call site of the list update for the whole window (a rolled loop over the chunks,
                // the keys rotate through key[0]) -- the update is the largest piece of code in the kernel; 4-5 % faster
                // there, 1-3 % slower in the fused kernel, which keeps the unrolled form below
                u64 key[kCB];
#pragma unroll
                for (int c = 0; c < kCB; ++c) {
                    const bool valid = ok[c] && (skip_invalid ? (p[c].w > 0.f) : !(p[c].w < wmin));
                    key[c] = valid ? pack_key(dist2(p[c].x, p[c].y, p[c].z, qx, qy, qz), id[c]) : kEmptyKey64;
                }
                const int nc = min(kCB, (total - t0 + GW - 1) / GW);
#pragma unroll 1
                for (int c = 0; c < nc; ++c) {
                    st_ins += topk_offer<GW, BUF>(g, tk, key[0], k);
#pragma unroll
                    for (int r = 0; r + 1 < kCB; ++r) key[r] = key[r + 1];
                }
            } else {
#pragma unroll
                for (int c = 0; c < kCB; ++c) {
                    if (t0 + c * GW >= total) break;
                    bool valid = ok[c] && (skip_invalid ? (p[c].w > 0.f) : !(p[c].w < wmin));
                    float d2 = dist2(p[c].x, p[c].y, p[c].z, qx, qy, qz);
                    st_ins += topk_offer<GW, BUF>(g, tk, valid ? pack_key(d2, id[c]) : kEmptyKey64, k);
                }
            }
        }
        st_cand += kNW * __popc(front_mask);
        if (BUF) topk_flush(tk, k);
    }
    Rect A = window(kM0);     // scanned so far (per view)
    int m = kM0;
    bool done = false;        // this group's result is final (uniform in the group)

    while (true) {
        const bool have_k = tk.have_k();
        // budget (view-sharded search, phase A): give up once k points are known and `budget` candidates have been
        // looked at -- the list is then k real points (an upper bound of the k-th best), not the exact answer
        if (budget != 0u && Group<GW>::wall(have_k && st_cand >= budget)) return false;
        Rect R = whole;      // pixels of my view that can still hold a neighbour
        bool covered;
        if (have_k) {
            // pixel bounding box of the ball B(q, r), r = current k-th best distance (an upper bound of the final one)
            float r = sqrtf(key_d2(tk.th));
#if MVT_CAM_SMEM
            {
                float (*cs)[32] = reinterpret_cast<float (*)[32]>(const_cast<unsigned char *>(win_tab) + 32 + 32 * sizeof(u64));
                const int ln = threadIdx.x & 31;
                const float X_ = cs[0][ln], Y_ = cs[1][ln], Z_ = cs[2][ln];
                float reff = fmaf(r, 1.0001f, 4e-6f * (fabsf(X_) + fabsf(Y_) + fabsf(Z_) + cs[9][ln] + r)) * cs[10][ln];
                if (mine && kok) R = ball_box(front, X_, Y_, Z_, cs[3][ln], reff, cs[4][ln], cs[5][ln], cs[6][ln], cs[7][ln], cs[8][ln], inv_st, W, H);
            }
#else
            float reff = fmaf(r, 1.0001f, 4e-6f * (fabsf(X) + fabsf(Y) + fabsf(Z) + tnorm + r)) * smax;
            if (mine && kok) R = ball_box(front, X, Y, Z, iz, reff, fx, sk, cx0, fy, cy0, inv_st, W, H);
#endif
            covered = !mine || rect_empty(R) || rect_inside(R, A);
        } else {
            // fewer than k searchable points so far; stop only when all rasters are covered
            covered = !mine || (A.x0 == 0 && A.y0 == 0 && A.x1 == W && A.y1 == H);
        }
        const bool all_covered = g.all(covered);  // every view's ball projection has been scanned: exact
        done = done || all_covered;
        if (Group<GW>::wall(done)) break;
        // views without a window centre (q behind the camera, odd intrinsics) wait until the views that have one are
        // done, so that they are scanned once, with the final (tight) radius
        const bool front_pending = g.any(front && !covered);
        // grow the windows geometrically, but never beyond what the current ball can reach; the bound tightens as
        // nearer points are found, so the scanned area stays within a small factor of the final ball's projection
        m = (m < (1 << 24)) ? 2 * m + 1 : m;
        Rect gr = {0, 0, 0, 0};
        if (front) gr = (m >= (1 << 24)) ? whole : window(m);
        else if (mine && !front_pending) gr = whole;
        if (!rect_empty(gr)) {
            gr.x0 = max(gr.x0, R.x0); gr.y0 = max(gr.y0, R.y0); gr.x1 = min(gr.x1, R.x1); gr.y1 = min(gr.y1, R.y1);
        }
        Rect B;
        if (rect_empty(gr) || done) {
            B = A;
        } else if (rect_empty(A)) {
            B = gr;
        } else {
            B.x0 = min(gr.x0, A.x0); B.y0 = min(gr.y0, A.y0); B.x1 = max(gr.x1, A.x1); B.y1 = max(gr.y1, A.y1);
        }
        // r only shrinks: if this pass reaches every R, nothing outside can matter afterwards
        const bool reaches_all = g.all(!mine || rect_empty(R) || rect_inside(R, B));
        const bool last = have_k && reaches_all;
        // a large box of new pixels in some view: give the query to tile_search (depth-culled tiles), see raster_search
        if (TILES && GW == 32 && rv.ztab != nullptr && budget == 0u && !negf) {
            const int newpx = rect_empty(B) ? 0 : (B.x1 - B.x0) * (B.y1 - B.y0) - (rect_empty(A) ? 0 : (A.x1 - A.x0) * (A.y1 - A.y0));
            if (g.any(have_k && mine && !done && newpx >= kTileMinPixels)) {
                if (soft_count == nullptr) return false;
                if (!soft_counted) {
                    soft_counted = true;
                    if (gl == 0) atomicAdd(soft_count, 1u);
                }
            }
        }

        // ---- scan B \ A of every view: four strips per view (top, bottom, left, right), GW/4 views per round ----
        for (int v0 = 0; v0 < V; v0 += GW / 4) {
            const int ev = v0 + (gl >> 2), es = gl & 3, el = ev & (GW - 1);
            int ax0 = g.shfl(A.x0, el), ay0 = g.shfl(A.y0, el), ax1 = g.shfl(A.x1, el), ay1 = g.shfl(A.y1, el);
            const int bx0 = g.shfl(B.x0, el), by0 = g.shfl(B.y0, el), bx1 = g.shfl(B.x1, el), by1 = g.shfl(B.y1, el);
            if (ax1 <= ax0 || ay1 <= ay0) { ax0 = ax1 = bx0; ay0 = ay1 = by1; }   // nothing scanned yet: all of B is "top"
            const int sx0 = (es == 3) ? ax1 : bx0, sx1 = (es == 2) ? ax0 : bx1;
            const int sy0 = (es == 0) ? by0 : (es == 1) ? ay1 : ay0, sy1 = (es == 0) ? ay0 : (es == 1) ? by1 : ay1;
            const int sw = sx1 - sx0, sh = sy1 - sy0;
            const unsigned cnt = (!done && ev < V && sw > 0 && sh > 0) ? (unsigned)sw * (unsigned)sh : 0u;
            const float inv_w = __frcp_rn((float)max(sw, 1));
            unsigned incl = cnt;
#pragma unroll
            for (int o = 1; o < GW; o <<= 1) {
                unsigned t = g.shfl_up(incl, o);
                if (gl >= o) incl += t;
            }
            const unsigned excl = incl - cnt;
            const unsigned total = g.shfl(incl, GW - 1);
            st_cand += total;
            for (unsigned i0 = 0; Group<GW>::wany(i0 < total); i0 += GW) {
                // bounded effort: a single pass can be large (a view behind the query gets its whole box at once)
                if (budget != 0u && i0 >= budget && Group<GW>::wall(tk.have_k())) {
                    if (BUF) topk_flush(tk, k);
                    return false;
                }
                const unsigned t = i0 + gl;
                bool valid = t < total;
                int j = 0;
#pragma unroll
                for (int step = GW / 2; step; step >>= 1) {
                    int pr = j + step;
                    unsigned e = g.shfl(excl, pr);
                    if (e <= t) j = pr;
                }
                const unsigned e0 = g.shfl(excl, j);
                const int wj = g.shfl(sw, j), xj = g.shfl(sx0, j), yj = g.shfl(sy0, j);
                const float iw = g.shfl(inv_w, j);
                const unsigned o = valid ? t - e0 : 0u;
                unsigned row;
                if (o < (1u << 22)) {
                    row = (unsigned)(((float)o + 0.5f) * iw);      // within one of o / wj; fixed below
                } else {
                    row = o / (unsigned)max(wj, 1);
                }
                int col = (int)o - (int)row * wj;
                if (col < 0) { --row; col += wj; } else if (col >= wj) { ++row; col -= wj; }
                const int idx = (v0 + (j >> 2)) * HW + (yj + (int)row) * W + xj + col;
                float4 p = valid ? __ldg(frame + idx) : make_float4(0.f, 0.f, 0.f, 0.f);
                valid = valid && (skip_invalid ? (p.w > 0.f) : !(p.w < wmin));
                float d2 = dist2(p.x, p.y, p.z, qx, qy, qz);
                st_ins += topk_offer<GW, BUF>(g, tk, valid ? pack_key(d2, idx) : kEmptyKey64, k);
            }
        }
        if (BUF) topk_flush(tk, k);
        if (!rect_empty(B)) A = B;
        if (!done) ++st_pass;
        done = done || last;
        if (Group<GW>::wall(done)) break;
    }
    if (stats && live && gl == 0) {
        atomicAdd(stats + 0, 1ull);
        atomicAdd(stats + min(st_pass, 4), 1ull);     // [1] one pass, [2] two passes, [3] three, [4] four or more
        atomicAdd(stats + 5, (unsigned long long)st_cand);
        atomicAdd(stats + 6, (unsigned long long)st_ins);
        // [9] worst query: candidates << 24 | query tag (the host recomputes the rest)
        atomicMax(stats + 9, ((unsigned long long)st_cand << 24) | (unsigned long long)(qtag & 0xFFFFFF));
        int bin = 0;
        unsigned c = st_cand;
        while (c > 1 && bin < 5) { c >>= 3; ++bin; }       // bins: <2, <16, <128, <1024, <8192, >=8192 candidates
        atomicAdd(stats + 10 + bin, 1ull);
    }
    return true;
}

// The search: growing windows for the (vast majority of) queries whose neighbourhood is a few hundred pixels and, with
// TILES (and depth tables in rv), the depth-culled tile scan for those whose ball projects onto a large box in some view
// (next to a camera, far from every surface): that bounds the time of the slowest query, which is what a kernel with
// few queries per SM (a track shard of a multi-GPU run) waits for.  The fused single-GPU kernel is throughput-bound
// and keeps TILES off: the rare large scans cost it less than the registers of the call.  Same result either way.
// Arguments and return value as raster_windows.
template <int GW, bool TILES = false>
__device__ __forceinline__ bool raster_search(const Group<GW> &g, const RasterView &rv, const unsigned char *win_tab, int s,
                                              bool live, float qx, float qy, float qz, int k, TopK &tk,
                                              unsigned long long *stats, int qtag = 0, u64 cap = kEmptyKey64,
                                              unsigned budget = 0u) {
    bool exact = raster_windows<GW, TILES>(g, rv, win_tab, s, live, qx, qy, qz, k, tk, stats, qtag, cap, budget);
    if (TILES && GW == 32 && budget == 0u && !exact) {       // (uniform in the warp)
        tk = tile_search(rv, s, qx, qy, qz, k, tk, stats);
        exact = true;
    }
    return exact;
}

}  // namespace mvt
