// Exact kNN, device side: the register top-k list shared by both searches, and the warp-per-query search over the
// hashed grid.  A group of lanes (a warp, or a half warp in the raster search) owns one query; group lane i holds the
// i-th best (d2, index) pair of a sorted list in registers; candidates are offered one per lane and merged with
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

__device__ __forceinline__ float dist2(float px, float py, float pz, float qx, float qy, float qz) {
    float dx = __fsub_rn(px, qx), dy = __fsub_rn(py, qy), dz = __fsub_rn(pz, qz);
    return __fmaf_rn(dz, dz, __fmaf_rn(dy, dy, __fmul_rn(dx, dx)));
}

// A (d2, index) pair as one 64-bit key: d2 >= 0 (or NaN), so its bit pattern orders like the float and one unsigned
// compare is the total order (d2, index).  NaN distances (0x7fc00000 or above) order after every real entry and
// after the empty key, i.e. they are never inserted -- the same as `d < bd || (d == bd && i < bi)` in the oracle.
typedef unsigned long long u64;
constexpr u64 kEmptyKey64 = 0x7f8000007fffffffull;   // (+inf, kNoIdx)
__device__ __forceinline__ u64 pack_key(float d2, int idx) { return ((u64)__float_as_uint(d2) << 32) | (u64)(unsigned)idx; }
__device__ __forceinline__ float key_d2(u64 key) { return __uint_as_float((unsigned)(key >> 32)); }
__device__ __forceinline__ int key_idx(u64 key) { return (int)(unsigned)(key & 0xffffffffull); }

// A query is owned by a GROUP of GW lanes: a whole warp (GW = 32: any k <= 32, up to 32 views) or a half warp
// (GW = 16: k <= 16, up to 16 views).  With half-warp groups the two queries of a warp run in lock step: every
// collective is executed by all 32 lanes (full mask, width GW), control flow is decided by warp-level votes
// (wany / wall) and each group's work is predicated on its own state -- so one instruction stream serves two queries.
template <int GW>
struct Group {
    int gl;          // lane index inside the group
    int shift;       // first lane of the group
    __device__ __forceinline__ explicit Group(int lane) : gl(lane & (GW - 1)), shift(GW == 32 ? 0 : (lane & 16)) {}
    template <typename T> __device__ __forceinline__ T shfl(T v, int src) const { return __shfl_sync(FULL, v, src, GW); }
    template <typename T> __device__ __forceinline__ T shfl_up(T v, int d) const { return __shfl_up_sync(FULL, v, d, GW); }
    template <typename T> __device__ __forceinline__ T shfl_xor(T v, int j) const { return __shfl_xor_sync(FULL, v, j, GW); }
    __device__ __forceinline__ unsigned ballot(bool p) const {
        unsigned b = __ballot_sync(FULL, p);
        return GW == 32 ? b : ((b >> shift) & 0xffffu);
    }
    __device__ __forceinline__ bool all(bool p) const { return GW == 32 ? __all_sync(FULL, p) : (ballot(!p) == 0u); }
    __device__ __forceinline__ bool any(bool p) const { return GW == 32 ? __any_sync(FULL, p) : (ballot(p) != 0u); }
    // warp-level votes on group-uniform predicates
    static __device__ __forceinline__ bool wany(bool p) { return GW == 32 ? p : __any_sync(FULL, p); }
    static __device__ __forceinline__ bool wall(bool p) { return GW == 32 ? p : __all_sync(FULL, p); }
};

// Ascending bitonic sort of one key per lane of the group (15 compare-exchange stages for 32 lanes, 10 for 16).
template <int GW>
__device__ __forceinline__ u64 group_sort(const Group<GW> &g, u64 v) {
#pragma unroll
    for (int k2 = 2; k2 <= GW; k2 <<= 1) {
#pragma unroll
        for (int j = k2 >> 1; j > 0; j >>= 1) {
            u64 o = g.shfl_xor(v, j);
            bool take_min = ((g.gl & k2) == 0) == ((g.gl & j) == 0);
            v = ((o < v) == take_min) ? o : v;
        }
    }
    return v;
}

// The running result of one query: group lane i holds the i-th best key of a sorted top-GW list (kEmptyKey64 = no
// entry); th = the k-th best, uniform across the group; empty = nothing has been offered yet (uniform).
// cap: an optional a-priori bound -- only keys below it are ever accepted (th = min(k-th best, cap)); a search that
// is given the cap of a wider search (other views, an earlier iteration) stops as early as that wider search would.
// pend / np (buffered update, see topk_offer<GW, true>): candidates that beat th wait in this warp's 32-entry shared
// memory buffer and are merged into the list as ONE sorted batch (topk_flush); t and th lag behind until then.
struct TopK {
    u64 t, th, cap;
    u64 *pend;
    int np;
    bool empty;
    __device__ __forceinline__ void init(u64 cap_key = kEmptyKey64) { t = kEmptyKey64; th = cap_key; cap = cap_key; empty = true; np = 0; }
    __device__ __forceinline__ bool have_k() const { return th != kEmptyKey64; }
};

#ifndef MVT_FLUSH_AT
#define MVT_FLUSH_AT 32
#endif
#ifndef MVT_FLUSH_INLINE
#define MVT_FLUSH_INLINE __noinline__
#endif
// Sorted merge of the np pending keys (this warp's buffer) into the sorted list t: sort the batch (bitonic, only the
// stages np needs: lanes >= np hold the empty key), min with the reversed batch = the 32 smallest of both as a bitonic
// sequence, 5 merge stages.  Full warps only.
static __device__ MVT_FLUSH_INLINE u64 topk_merge_pending(u64 t, const u64 *pend, int np) {
    const int lane = threadIdx.x & 31;
    __syncwarp();
    u64 b = lane < np ? pend[lane] : kEmptyKey64;
    __syncwarp();
#pragma unroll
    for (int k2 = 2; k2 <= 32; k2 <<= 1) {
        if ((k2 >> 1) >= np) break;        // (uniform) blocks of k2/2 lanes are sorted and block 0 holds every key
#pragma unroll
        for (int j = k2 >> 1; j > 0; j >>= 1) {
            u64 o = __shfl_xor_sync(FULL, b, j);
            bool take_min = ((lane & k2) == 0) == ((lane & j) == 0);
            b = ((o < b) == take_min) ? o : b;
        }
    }
    u64 r = __shfl_sync(FULL, b, 31 - lane);
    u64 v = r < t ? r : t;
#pragma unroll
    for (int j = 16; j > 0; j >>= 1) {
        u64 o = __shfl_xor_sync(FULL, v, j);
        bool take_min = (lane & j) == 0;
        v = ((o < v) == take_min) ? o : v;
    }
    return v;
}
// Brings t and th up to date (no-op without pending keys).  Call before reading tk.t / tk.th / have_k() for a decision
// that needs the current k-th best; a stale th is still a valid upper bound.
__device__ __forceinline__ void topk_flush(TopK &tk, int k) {
    if (tk.np == 0) return;
    tk.t = topk_merge_pending(tk.t, tk.pend, tk.np);
    tk.np = 0;
    const u64 kth = __shfl_sync(FULL, tk.t, k - 1);
    tk.th = kth < tk.cap ? kth : tk.cap;
}

// Offer one candidate key per lane (kEmptyKey64 for "no candidate").  Returns the number of candidates that beat the
// k-th best at entry.  The first non-empty batch is sorted as a whole; later ones are inserted one by one (few survive
// the threshold).  Within a batch the threshold is not refreshed: a candidate that no longer makes the top k after an
// earlier insertion of the same batch lands at a position >= k, which is harmless.
// Must be called by all 32 lanes (see Group).
// BUF (GW == 32, tk.pend set): after the first batch the survivors are appended to the pending buffer instead and merged
// as a sorted batch when it fills up or when the caller flushes (topk_flush).
template <int GW, bool BUF = false>
__device__ __forceinline__ int topk_offer(const Group<GW> &g, TopK &tk, u64 key, int k) {
    unsigned m = g.ballot(key < tk.th);
    if (!Group<GW>::wany(m != 0u)) return 0;
    if (BUF && GW == 32 && !tk.empty) {
        if (tk.np + __popc(m) > 32) {
            topk_flush(tk, k);
            m = __ballot_sync(FULL, key < tk.th);
            if (m == 0u) return 0;
        }
        const int nb = __popc(m);
        const unsigned below = m & ((1u << (threadIdx.x & 31)) - 1u);
        if (key < tk.th) tk.pend[tk.np + __popc(below)] = key;
        tk.np += nb;
        if (MVT_FLUSH_AT < 32 && tk.np >= MVT_FLUSH_AT) topk_flush(tk, k);
        return nb;
    }
    const int n = __popc(m);
    const bool do_sort = tk.empty && m != 0u;
    if (Group<GW>::wany(do_sort)) {
        u64 sorted = group_sort(g, key < tk.th ? key : kEmptyKey64);
        if (do_sort) {
            tk.t = sorted;
            tk.empty = false;
            m = 0u;
        }
    }
    while (Group<GW>::wany(m != 0u)) {
        const bool act = m != 0u;
        const int src = act ? __ffs(m) - 1 : 0;
        m &= m - 1;
        u64 c = g.shfl(key, src);
        bool disp = act && c < tk.t;                // entries after c move up one lane
        u64 up = g.shfl_up(tk.t, 1);
        unsigned dm = g.ballot(disp);               // a suffix of the group's lanes (the list is sorted)
        if (disp) tk.t = (g.gl == __ffs(dm) - 1) ? c : up;
    }
    const u64 kth = g.shfl(tk.t, k - 1);
    tk.th = kth < tk.cap ? kth : tk.cap;
    return n;
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
                                          TopK &tk, int k, int lane) {
    const Group<32> g(lane);
    int ins = 0;
    for (unsigned i0 = 0; i0 < n; i0 += 32) {
        unsigned t = i0 + lane;
        bool valid = t < n;
        float4 p = valid ? __ldg(pts + t) : make_float4(0.f, 0.f, 0.f, 0.f);
        float d2 = dist2(p.x, p.y, p.z, qx, qy, qz);
        ins += topk_offer(g, tk, valid ? pack_key(d2, __float_as_int(p.w)) : kEmptyKey64, k);
    }
    return ins;
}

// Exact k nearest neighbours of (qx,qy,qz) in segment b.  On return lane i < k holds the i-th neighbour
// (key_idx == kNoIdx when fewer than k searchable points exist).
__device__ __forceinline__ void knn_search(const GridView &g, int b, float qx, float qy, float qz, int k, int lane,
                                           TopK &tk) {
    tk.init();
    const Group<32> g32(lane);
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
                    st_ins += topk_offer(g32, tk, valid ? pack_key(d2, __float_as_int(p.w)) : kEmptyKey64, k);
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
                if (tk.have_k() && bm > 0.f && key_d2(tk.th) < bm * bm) done = true;
            }
        }
    }
    if (!done) {   // exhaustive fallback (tiny segment, or a query far from the surface)
        tk.init();
        st_ins += scan_range(pts, n_valid, qx, qy, qz, tk, k, lane);
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
