// Warp-per-query exact kNN over the hashed grid (device side).  One warp owns one query; lane i holds the i-th best
// (d2, index) pair of a sorted top-32 list in registers; candidates are offered 32 at a time and inserted with warp
// shuffles.  Candidate order never matters: the list is ordered by the total order (d2, index).
//
// Canonical arithmetic (bit-exact contract with oracle/mvt_oracle.c):
//     d2 = fma(dz, dz, fma(dy, dy, dx*dx)),  dx = p.x - q.x  (fp32, round-to-nearest)
#pragma once
#include "mvt_common.cuh"

namespace mvt {

constexpr unsigned FULL = 0xffffffffu;
constexpr int kRMax = 3;                 // rings (Chebyshev radius in cells) before the exhaustive fallback
constexpr float kMarginCells = 2e-3f;    // slack on the ring bound: covers fp32 rounding of cell assignment (<= 3e-4 cells)
constexpr unsigned kBruteN = 512;        // segments this small are scanned exhaustively
constexpr int kNoIdx = 0x7fffffff;

__device__ __forceinline__ bool better(float d, int i, float bd, int bi) { return d < bd || (d == bd && i < bi); }

__device__ __forceinline__ float dist2(float px, float py, float pz, float qx, float qy, float qz) {
    float dx = __fsub_rn(px, qx), dy = __fsub_rn(py, qy), dz = __fsub_rn(pz, qz);
    return __fmaf_rn(dz, dz, __fmaf_rn(dy, dy, __fmul_rn(dx, dx)));
}

// Insert (cd, ci) -- known to every lane and strictly better than the current k-th -- into the sorted list.
__device__ __forceinline__ void topk_insert(float &bd, int &bi, float cd, int ci, int lane) {
    bool displaced = better(cd, ci, bd, bi);
    float ud = __shfl_up_sync(FULL, bd, 1);
    int ui = __shfl_up_sync(FULL, bi, 1);
    unsigned dm = __ballot_sync(FULL, displaced);
    if (displaced) {
        bool first = (lane == 0) || !((dm >> (lane - 1)) & 1u);
        bd = first ? cd : ud;
        bi = first ? ci : ui;
    }
}

// Offer one candidate per lane.  (td, ti) = current k-th best, uniform across the warp.
__device__ __forceinline__ int topk_offer(float &bd, int &bi, float &td, int &ti, float d2, int idx, bool valid, int k,
                                          int lane) {
    unsigned m = __ballot_sync(FULL, valid && better(d2, idx, td, ti));
    int ins = 0;
    while (m) {
        int src = __ffs(m) - 1;
        m &= m - 1;
        float cd = __shfl_sync(FULL, d2, src);
        int ci = __shfl_sync(FULL, idx, src);
        if (better(cd, ci, td, ti)) {
            topk_insert(bd, bi, cd, ci, lane);
            td = __shfl_sync(FULL, bd, k - 1);
            ti = __shfl_sync(FULL, bi, k - 1);
            ++ins;
        }
    }
    return ins;
}

__device__ __forceinline__ uint2 cell_lookup(const unsigned *__restrict__ keys, const uint2 *__restrict__ vals,
                                             unsigned tmask, int tshift, unsigned key) {
    unsigned slot = cell_hash(key, tshift);
    while (true) {
        unsigned cur = __ldg(keys + slot);
        if (cur == key) return __ldg(vals + slot);
        if (cur == kEmptyKey) return make_uint2(0u, 0u);
        slot = (slot + 1) & tmask;
    }
}

// Scan `n` consecutive records starting at pts.
__device__ __forceinline__ int scan_range(const float4 *__restrict__ pts, unsigned n, float qx, float qy, float qz,
                                          float &bd, int &bi, float &td, int &ti, int k, int lane) {
    int ins = 0;
    for (unsigned i0 = 0; i0 < n; i0 += 32) {
        unsigned t = i0 + lane;
        bool valid = t < n;
        float4 p = valid ? __ldg(pts + t) : make_float4(0.f, 0.f, 0.f, 0.f);
        float d2 = dist2(p.x, p.y, p.z, qx, qy, qz);
        ins += topk_offer(bd, bi, td, ti, d2, __float_as_int(p.w), valid, k, lane);
    }
    return ins;
}

// Exact k nearest neighbours of (qx,qy,qz) in segment b.  On return lane i < k holds the i-th neighbour
// (bi == kNoIdx when fewer than k searchable points exist).
__device__ __forceinline__ void knn_search(const GridView &g, int b, float qx, float qy, float qz, int k, int lane,
                                           float &bd, int &bi) {
    bd = INFINITY;
    bi = kNoIdx;
    float td = INFINITY;
    int ti = kNoIdx;
    const GridSeg *sp = g.segs + b;
    const unsigned n_valid = sp->n_valid;
    if (n_valid == 0) return;
    const float4 *pts = g.sorted + (size_t)b * g.N;
    bool done = false;
    unsigned long long st_cand = 0, st_ins = 0, st_look = 0;
    int st_ring = 0;
    if (n_valid > kBruteN) {
        const float ox = sp->ox, oy = sp->oy, oz = sp->oz, h = sp->h, inv_h = sp->inv_h;
        const int nx = sp->nx, ny = sp->ny, nz = sp->nz;
        const unsigned *keys = g.keys + (size_t)b * g.T;
        const uint2 *vals = g.vals + (size_t)b * g.T;
        const unsigned tmask = g.T - 1;
        const int cx = cell_coord(qx, ox, inv_h, nx), cy = cell_coord(qy, oy, inv_h, ny), cz = cell_coord(qz, oz, inv_h, nz);
        for (int r = 1; r <= kRMax && !done; ++r) {
            const int side = 2 * r + 1, ncell = side * side * side;
            for (int base = 0; base < ncell; base += 32) {
                int c = base + lane;
                unsigned start = 0, cnt = 0;
                if (c < ncell) {
                    int dz = c / (side * side);
                    int rem = c - dz * side * side;
                    int dy = rem / side;
                    int dx = rem - dy * side;
                    dx -= r; dy -= r; dz -= r;
                    bool shell = (r == 1) || max(abs(dx), max(abs(dy), abs(dz))) == r;
                    int x = cx + dx, y = cy + dy, z = cz + dz;
                    if (shell && x >= 0 && x < nx && y >= 0 && y < ny && z >= 0 && z < nz) {
                        uint2 v = cell_lookup(keys, vals, tmask, g.tshift, cell_key(x, y, z));
                        start = v.x;
                        cnt = v.y;
                    }
                }
                // flatten the (up to 32) cell ranges into one candidate list
                unsigned incl = cnt;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    unsigned t = __shfl_up_sync(FULL, incl, o);
                    if (lane >= o) incl += t;
                }
                const unsigned excl = incl - cnt;
                const unsigned total = __shfl_sync(FULL, incl, 31);
                st_cand += total;
                st_look += min(32, ncell - base);
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
                    unsigned s0 = __shfl_sync(FULL, start, j), e0 = __shfl_sync(FULL, excl, j);
                    float4 p = valid ? __ldg(pts + s0 + (t - e0)) : make_float4(0.f, 0.f, 0.f, 0.f);
                    float d2 = dist2(p.x, p.y, p.z, qx, qy, qz);
                    st_ins += topk_offer(bd, bi, td, ti, d2, __float_as_int(p.w), valid, k, lane);
                }
            }
            st_ring = r;
            // every unvisited point lies beyond a face of the searched cube [c-r, c+r]^3 that is inside the grid
            float bound = INFINITY;
            if (cx - r > 0) bound = fminf(bound, qx - (ox + (float)(cx - r) * h));
            if (cx + r < nx - 1) bound = fminf(bound, (ox + (float)(cx + r + 1) * h) - qx);
            if (cy - r > 0) bound = fminf(bound, qy - (oy + (float)(cy - r) * h));
            if (cy + r < ny - 1) bound = fminf(bound, (oy + (float)(cy + r + 1) * h) - qy);
            if (cz - r > 0) bound = fminf(bound, qz - (oz + (float)(cz - r) * h));
            if (cz + r < nz - 1) bound = fminf(bound, (oz + (float)(cz + r + 1) * h) - qz);
            if (bound == INFINITY) {
                done = true;   // the cube covers the whole grid
            } else {
                float bm = bound - kMarginCells * h;
                if (ti != kNoIdx && bm > 0.f && td < bm * bm) done = true;
            }
        }
    }
    if (!done) {   // exhaustive fallback (tiny segment, or a query far from the surface)
        bd = INFINITY; bi = kNoIdx; td = INFINITY; ti = kNoIdx;
        st_ins += scan_range(pts, n_valid, qx, qy, qz, bd, bi, td, ti, k, lane);
        st_cand += n_valid;
        st_ring = (n_valid > kBruteN) ? 4 : 8;
    }
    if (g.stats && lane == 0) {
        atomicAdd(g.stats + 0, 1ull);
        atomicAdd(g.stats + st_ring, 1ull);
        atomicAdd(g.stats + 5, st_cand);
        atomicAdd(g.stats + 6, st_ins);
        atomicAdd(g.stats + 7, st_look);
    }
}

}  // namespace mvt
