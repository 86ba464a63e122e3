// View-sharded kNN correlation across GPUs: every rank keeps ONLY the clouds of the views it unprojected (features
// never cross NVLink); what crosses are 8-byte (d2, index) keys and 16-byte result records.  Per refinement iteration
//
//   K0  mvt_knn_keys(budget)  every rank: bounded-effort search of its views: exact where it converges within the budget
//                          (a rank that sees the query's surface), k real but far points where it does not
//   C0  all-to-all + K2 merge + all-gather: the k-th best of the union is an upper bound ("cap") of the global k-th
//                          distance, known to every rank.  Without it a rank whose views do not see a query's surface
//                          would search out to ITS k-th distance (metres): measured 16x the whole single-GPU kernel.
//   K1  mvt_knn_keys(caps, skip)  every rank: the queries that did not converge in K0, exactly, within the cap -> sorted keys,
//                          written in the [track shard][level][frame][track][k] layout that all_to_all_single sends as is
//   C1  all-to-all         rank j receives, for its track shard, the n_ranks candidate lists of every query
//   K2  mvt_topk_merge     exact global top-k of each query of the shard (global top-k is a subset of the union of the
//                          local top-k's; (d2, index) is a total order, so the merge is deterministic) -> int32 indices
//   C2  all-gather         every rank gets the merged index lists of all queries
//   K3  mvt_corr_from_idx  every rank: correlation + offsets of the neighbours that live in ITS views, zeros elsewhere
//   C3  all-reduce (sum)   each record has exactly one non-zero contributor, so the sum is exact; every rank ends with
//                          the full [S, M, levels*k*4] block the (replicated) transformer consumes
// (mvtracker_b200/dist.py drives this with torch.distributed / NCCL.)  Replaces, for N > 1, the single-GPU fused kernel
// of mvt_knn.cu; same arithmetic, bit-identical neighbour sets and offsets, identical correlations.
#include "mvt_search.cuh"
#include "mvt_corr.cuh"
#include <stdlib.h>

namespace mvt {

struct KeysParams {
    SearchDev sd[MVT_MAX_LEVELS];
    long long idx_off[MVT_MAX_LEVELS];
    int n_levels, B, M, k, n_shards, Ms;
    unsigned budget;             // != 0: bounded-effort search (phase A)
    const u64 *caps;             // [n_shards, n_levels, B, Ms] or null
    unsigned char *exact_out;    // same layout, or null: 1 = the emitted list is the exact local answer
    const unsigned char *skip;   // same layout, or null: 1 = leave this query's keys as they are
    const float *query;
    u64 *out;
    PeersDev idx_peers;          // n != 0: store int32 index lists into idx_all of every rank instead of keys into `out`
    int out_shard, out_shards;   // ... at track shard out_shard of out_shards
    unsigned n_items;            // n_levels * B * n_shards * Ms
};

// one query (level lvl, frame b, padded track m of n_shards*Ms) by one warp
template <bool RASTER_ONLY>
__device__ __forceinline__ void keys_item(const KeysParams &p, const unsigned char *win_tab, const int lvl, const int b, const int m) {
    const int lane = threadIdx.x & 31;
    const int k = p.k;
    const int shard = m / p.Ms, mm = m - shard * p.Ms;
    const size_t qslot = (((size_t)shard * p.n_levels + lvl) * p.B + b) * p.Ms + mm;
    u64 *o = p.out ? p.out + qslot * k : nullptr;
    u64 key = kEmptyKey64;
    if (p.skip && p.skip[qslot]) return;          // warp-uniform
    bool exact = true;
    if (m < p.M) {
        const long long q = (long long)b * p.M + m;
        const float qx = __ldg(p.query + q * 3), qy = __ldg(p.query + q * 3 + 1), qz = __ldg(p.query + q * 3 + 2);
        const Group<32> g(lane);
        TopK tk;
        const SearchDev &sd = p.sd[lvl];
        // cap: distance-only bound from phase A -- keep every point with d2 <= the bound's d2 (any index: ties stay in)
        u64 cap = kEmptyKey64;
        if (p.caps) {
            const u64 c = __ldg(p.caps + qslot);
            if ((c >> 32) < 0x7f800000ull) cap = c | 0xffffffffull;      // finite distance; else no bound is known
        }
        if (RASTER_ONLY || sd.raster) exact = raster_search<32, true>(g, sd.rv, win_tab, b, true, qx, qy, qz, k, tk, sd.g.stats, (int)q, cap, p.budget);
        else knn_search(sd.g, b, qx, qy, qz, k, lane, tk);
        key = tk.t;
        if (key_idx(key) != kNoIdx) key += (u64)p.idx_off[lvl];      // local -> global index (no carry: idx < 2^31)
        else key = kEmptyKey64;
    }
    if (o && lane < k) o[lane] = key;
    if (p.exact_out && lane == 0) p.exact_out[qslot] = exact ? 1 : 0;
    if (p.idx_peers.n) {
        // four indices per 16-byte store; lane j < k/4 stores neighbours 4j..4j+3 to every rank
        const int i = (key == kEmptyKey64) ? -1 : key_idx(key);
        const int src = (4 * lane) & 31;
        const int i0 = __shfl_sync(FULL, i, src), i1 = __shfl_sync(FULL, i, src + 1);
        const int i2 = __shfl_sync(FULL, i, src + 2), i3 = __shfl_sync(FULL, i, src + 3);
        if (lane < (k >> 2)) {
            const size_t slot = (((size_t)p.out_shard * p.n_levels + lvl) * p.B + b) * p.Ms + mm;
            store_all(p.idx_peers, slot * (size_t)(k >> 2) + lane,
                      make_float4(__int_as_float(i0), __int_as_float(i1), __int_as_float(i2), __int_as_float(i3)));
        }
    }
}

// Work distribution as in knn_corr_kernel (mvt_knn.cu): one wave of blocks, block j owns items j, j + gridDim.x, ...,
// its warps draw from that list through a shared counter.
template <bool RASTER_ONLY>
#ifndef MVT_KEYS_MINB
#define MVT_KEYS_MINB 4
#endif
#ifndef MVT_KEYS_THREADS
#define MVT_KEYS_THREADS 256
#endif
__global__ void __launch_bounds__(MVT_KEYS_THREADS, MVT_KEYS_MINB) knn_keys_kernel(const KeysParams p) {
    __shared__ unsigned next_item;
    if (threadIdx.x == 0) next_item = 0u;
    MVT_WARP_SCRATCH(win_tab, MVT_KEYS_THREADS / 32);
    __syncthreads();                     // next_item
    const int lane = threadIdx.x & 31;
    const unsigned Mp = (unsigned)(p.n_shards * p.Ms);
    for (;;) {
        unsigned i = 0u;
        if (lane == 0) i = atomicAdd(&next_item, 1u);
        i = __shfl_sync(FULL, i, 0);
        const unsigned long long item64 = (unsigned long long)i * gridDim.x + blockIdx.x;
        if (item64 >= (unsigned long long)p.n_items) break;
        const unsigned item = (unsigned)item64;
        const unsigned fr = item / Mp, m = item - fr * Mp;
        const unsigned lvl = fr / (unsigned)p.B, b = fr - lvl * (unsigned)p.B;
        keys_item<RASTER_ONLY>(p, win_tab, (int)lvl, (int)b, (int)m);
    }
}

// one warp per query: merge n_lists sorted lists of k keys
__global__ void __launch_bounds__(256) topk_merge_kernel(const u64 *__restrict__ keys, int n_lists, long long Q, int k,
                                                         int *__restrict__ idx_out, u64 *__restrict__ keys_out) {
    const int lane = threadIdx.x & 31;
    const long long q = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (q >= Q) return;
    const Group<32> g(lane);
    TopK tk;
    tk.init();
    const int total = n_lists * k;
    for (int t0 = 0; t0 < total; t0 += 32) {
        const int t = t0 + lane;
        u64 key = kEmptyKey64;
        if (t < total) {
            const int list = t / k, j = t - list * k;
            key = keys[((size_t)list * Q + q) * k + j];
        }
        topk_offer(g, tk, key, k);
    }
    if (lane < k) {
        const int i = key_idx(tk.t);
        if (idx_out) idx_out[q * k + lane] = (i != kNoIdx) ? i : -1;
        if (keys_out) keys_out[q * k + lane] = tk.t;
    }
}

struct OwnedLevel {
    const float4 *xyz4;     // [B, Nloc] records of the owned points
    const void *fvec;       // [B, Nloc, 128]
    long long lo, hi;       // owned global index range [lo, hi), Nloc = hi - lo
    long long out_off;
    const float4 *all_xyz4; // [B, n_all] records of ALL views (this rank's replica), or null -- see corr_from_idx_kernel
    long long n_all;
};
struct CorrIdxParams {
    OwnedLevel lv[MVT_MAX_LEVELS];
    int n_levels, B, M, k, n_shards, Ms;
    const int *idx;         // [n_shards, n_levels, B, Ms, k] global indices (-1 = missing)
    const void *targets;    // [B, M, 128]
    const float *coords;    // [B, M, 3]
    float *out;
    long long row_stride;
    float scale;
    PeersDev out_peers;     // n != 0: records go to `out` of every rank (row_stride in floats, same on every rank)
    int shard;              // this rank's track shard (it stores the zero records of missing neighbours)
};

// One warp per (level, frame, track).  Only the neighbours whose rows live on this rank are gathered, four rows in flight
// per round (with N view shards a query owns 16 / N of its neighbours on average, often none).  Arithmetic per neighbour
// is that of knn_corr_kernel -- the same per-lane fma chain and the same butterfly order over the lanes of a row -- so
// the records are bit-identical to the fused single-GPU kernel's whatever the sharding.
#ifndef MVT_CIDX_MINB
#define MVT_CIDX_MINB 8
#endif
#ifndef MVT_CIDX_ROUND
#define MVT_CIDX_ROUND 4       // rows in flight per round (fp32; bf16 rows are half as long: ROUND / 2 pairs)
#endif
template <bool BF16>
__global__ void __launch_bounds__(256, MVT_CIDX_MINB) corr_from_idx_kernel(const CorrIdxParams p) {
    const int lane = threadIdx.x & 31;
    const int m = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (m >= p.M) return;
    const int b = blockIdx.y, lvl = blockIdx.z, k = p.k;
    const OwnedLevel &L = p.lv[lvl];
    const long long q = (long long)b * p.M + m;
    const int shard = m / p.Ms, mm = m - shard * p.Ms;
    const int *ip = p.idx + ((((size_t)shard * p.n_levels + lvl) * p.B + b) * p.Ms + mm) * k;
    const long long gi = (lane < k) ? (long long)__ldg(ip + lane) : -1ll;
    const bool own = gi >= L.lo && gi < L.hi;
    const int row = own ? (int)(gi - L.lo) : 0;        // local row of neighbour `lane`
    const size_t nloc = (size_t)(L.hi - L.lo);
    unsigned om = __ballot_sync(FULL, own);
    // Split records (shared output + a replica of every view's records on each rank): the offsets of ALL k neighbours are
    // computed here from the replica and written to this rank's own block with plain stores; only the correlation values
    // (4 of a record's 16 bytes, known to the rank that holds the feature row) cross NVLink.
    const bool split = L.all_xyz4 != nullptr && p.out_peers.n != 0;
    if (split) {
        if (lane < k) {
            float *o = p.out + q * p.row_stride + L.out_off + (size_t)lane * 4;
            if (gi >= 0) {
                const float qx = __ldg(p.coords + q * 3), qy = __ldg(p.coords + q * 3 + 1), qz = __ldg(p.coords + q * 3 + 2);
                const float4 n = __ldg(L.all_xyz4 + (size_t)b * L.n_all + gi);
                o[1] = n.x - qx;
                *reinterpret_cast<float2 *>(o + 2) = make_float2(n.y - qy, n.z - qz);
            } else {
                *reinterpret_cast<float4 *>(o) = make_float4(0.f, 0.f, 0.f, 0.f);      // missing neighbour: nobody owns it
            }
        }
        if (om == 0u) return;
    } else if (om == 0u) {      // none of this query's neighbours lives here: no target row, no gather
        if (lane < k) {
            const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
            if (p.out_peers.n == 0) *reinterpret_cast<float4 *>(p.out + q * p.row_stride + L.out_off + (size_t)lane * 4) = z;
            else if (gi < 0 && shard == p.shard) store_all(p.out_peers, (size_t)((q * p.row_stride + L.out_off) >> 2) + lane, z);
        }
        return;
    }
    // everything that depends only on the index list is requested before the feature rows: one round trip, not three
    const float qx = __ldg(p.coords + q * 3), qy = __ldg(p.coords + q * 3 + 1), qz = __ldg(p.coords + q * 3 + 2);
    float4 n = make_float4(0.f, 0.f, 0.f, 0.f);
    if (own) n = __ldg(L.xyz4 + (size_t)b * nloc + row);
    float corr_mine = 0.f;
    if (!BF16) {
        const float4 tg = __ldg(reinterpret_cast<const float4 *>(p.targets) + q * 32 + lane);
        const float4 *fv = reinterpret_cast<const float4 *>(L.fvec) + (size_t)b * nloc * 32 + lane;
        while (om) {
            int j[MVT_CIDX_ROUND];
            float4 f[MVT_CIDX_ROUND];
#pragma unroll
            for (int u = 0; u < MVT_CIDX_ROUND; ++u) {
                j[u] = om ? __ffs(om) - 1 : -1;
                om &= om - 1;
            }
#pragma unroll
            for (int u = 0; u < MVT_CIDX_ROUND; ++u) {
                const int rj = __shfl_sync(FULL, row, j[u] & 31);
                f[u] = make_float4(0.f, 0.f, 0.f, 0.f);
                if (j[u] >= 0) f[u] = __ldg(fv + (size_t)rj * 32);
            }
#pragma unroll
            for (int u = 0; u < MVT_CIDX_ROUND; ++u) {
                if (j[u] < 0) break;
                float d = fmaf(tg.w, f[u].w, fmaf(tg.z, f[u].z, fmaf(tg.y, f[u].y, tg.x * f[u].x)));
#pragma unroll
                for (int o = 16; o; o >>= 1) d += __shfl_xor_sync(FULL, d, o);
                if (lane == j[u]) corr_mine = d;
            }
        }
    } else {
        // 16 lanes x 8 channels per row: each half-warp takes one neighbour of a pair
        const uint4 tgh = __ldg(reinterpret_cast<const uint4 *>(p.targets) + q * 16 + (lane & 15));
        const uint4 *fv = reinterpret_cast<const uint4 *>(L.fvec) + (size_t)b * nloc * 16 + (lane & 15);
        const int half = lane >> 4;
        while (om) {
            int ja[MVT_CIDX_ROUND / 2], jb[MVT_CIDX_ROUND / 2];
            uint4 f[MVT_CIDX_ROUND / 2];
#pragma unroll
            for (int u = 0; u < MVT_CIDX_ROUND / 2; ++u) {
                ja[u] = om ? __ffs(om) - 1 : -1;
                om &= om - 1;
                jb[u] = om ? __ffs(om) - 1 : -1;
                om &= om - 1;
            }
#pragma unroll
            for (int u = 0; u < MVT_CIDX_ROUND / 2; ++u) {
                const int jm = half ? jb[u] : ja[u];
                const int rj = __shfl_sync(FULL, row, jm & 31);
                f[u] = make_uint4(0u, 0u, 0u, 0u);
                if (jm >= 0) f[u] = __ldg(fv + (size_t)rj * 16);
            }
#pragma unroll
            for (int u = 0; u < MVT_CIDX_ROUND / 2; ++u) {
                if (ja[u] < 0) break;
                const __nv_bfloat162 *x = reinterpret_cast<const __nv_bfloat162 *>(&tgh);
                const __nv_bfloat162 *y = reinterpret_cast<const __nv_bfloat162 *>(&f[u]);
                float acc = 0.f;
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    float2 uu = __bfloat1622float2(x[e]), vv = __bfloat1622float2(y[e]);
                    acc = fmaf(uu.x, vv.x, acc);
                    acc = fmaf(uu.y, vv.y, acc);
                }
#pragma unroll
                for (int o = 8; o; o >>= 1) acc += __shfl_xor_sync(FULL, acc, o);
                const float va = __shfl_sync(FULL, acc, 0), vb = __shfl_sync(FULL, acc, 16);
                if (lane == ja[u]) corr_mine = va;
                if (lane == jb[u]) corr_mine = vb;
            }
        }
    }
    if (split) {
        if (own) store_all_f32(p.out_peers, (size_t)(q * p.row_stride + L.out_off) + (size_t)lane * 4, __fdiv_rn(corr_mine, p.scale));
        return;
    }
    if (lane < k) {
        float4 rec = make_float4(0.f, 0.f, 0.f, 0.f);
        if (own) rec = make_float4(__fdiv_rn(corr_mine, p.scale), n.x - qx, n.y - qy, n.z - qz);
        if (p.out_peers.n == 0) {
            *reinterpret_cast<float4 *>(p.out + q * p.row_stride + L.out_off + (size_t)lane * 4) = rec;
        } else if (own || (gi < 0 && shard == p.shard)) {     // exactly one rank stores each record
            store_all(p.out_peers, (size_t)((q * p.row_stride + L.out_off) >> 2) + lane, rec);
        }
    }
}


// ---------------------------------------------------------------------------------------------------------------
// One kernel per refinement iteration: search -> records to the owners' inboxes -> correlation -> result records to all
// ---------------------------------------------------------------------------------------------------------------
struct ShardLocal {             // this rank's private counters (zero between launches)
    unsigned prod_done;         // producer warps (one query each) that have pushed their records or deferred their query
    unsigned hard_count;        // queries deferred to the tile search (large pixel boxes), hard_cursor taken, hard_done finished
    unsigned hard_cursor;
    unsigned err;               // set when a wait timed out (a peer never showed up): results are then invalid
    unsigned hard_done;
    unsigned published;         // 1 once this rank's totals have been sent to the owners
    unsigned cons_done;         // consumer warps that have left
    unsigned pad;
    unsigned tail[MVT_MAX_PEERS];     // records I have pushed to owner o
    unsigned cursor[MVT_MAX_PEERS];   // next record of source s to take from my inbox
};
struct ShardHard { unsigned qid, pad; u64 cap; };   // a deferred query: ((lvl * B + b) * Ms + mm), k-th key found so far
struct ShardLevelDev {
    SearchDev sd;
    const float4 *xyz4_local;
    const void *fvec_local;
    long long n_local;
    int hw;
};
struct ShardParams {
    ShardLevelDev lv[MVT_MAX_LEVELS];
    int n_levels, B, M, Ms, k;
    const void *targets;
    const float *coords;
    u64 *inbox[MVT_MAX_PEERS];        // every rank's inbox as mapped here ([n_ranks][cap] each)
    u64 *ctrl[MVT_MAX_PEERS];         // every rank's control words: [0..7] total of source s (+1), [8..15] epoch of source s
    PeersDev result;
    ShardLocal *local;
    long long cap, row_stride;
    int rank, n;
    int view_first[MVT_MAX_PEERS + 1];
    u64 epoch;
    float scale;
    unsigned n_items;           // real queries of my shard: n_levels * B * (tracks of my shard)
    unsigned n_cons_blocks;     // consumer blocks: n_early in the grid's first row (extra columns), the rest in its last z slice
    unsigned n_early, nx;       // nx: producer columns of the grid
    int no_local;               // debug (MVT_SHARD_NO_LOCAL): locally owned neighbours go through the inbox as well
    ShardHard *hard;            // [n_levels * B * Ms] deferred-query list (local)
};
constexpr u64 kRecValid = 1ull << 63;
constexpr unsigned long long kShardTimeoutNs = 4000000000ull;

__device__ __forceinline__ void st_sys(u64 *p, u64 v) { asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory"); }
__device__ __forceinline__ void st_release_sys(u64 *p, u64 v) { asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory"); }
__device__ __forceinline__ u64 ld_sys(const u64 *p) {
    u64 v;
    asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ u64 ld_acquire_sys(const u64 *p) {
    u64 v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned ld_volatile_u32(const unsigned *p) { return *reinterpret_cast<const volatile unsigned *>(p); }
__device__ __forceinline__ unsigned long long now_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// One 8-byte record per neighbour of a finished search into the inbox of the rank that owns the neighbour's feature row:
//   [63] valid | [62:58] slot j | [57:56] level | [55:50] frame b | [49:32] track m | [31:0] index among the owner's points
// tkt: lane j holds the j-th neighbour's key.
// The neighbours whose rows THIS rank owns are not pushed: the searching warp correlates them itself (shard_local_corr).
template <bool BF16>
__device__ __forceinline__ void shard_local_corr(const ShardParams &p, int lvl, int b, int m, bool local, unsigned loc);

template <bool BF16>
__device__ __forceinline__ void shard_push_records(const ShardParams &p, int lvl, int b, int m, u64 tkt) {
    const int lane = threadIdx.x & 31;
    const ShardLevelDev &L = p.lv[lvl];
    const long long q = (long long)b * p.M + m;
    const int idx = key_idx(tkt);
    const bool mine = lane < p.k;
    const bool none = mine && idx == kNoIdx;
    const unsigned qid = ((unsigned)lvl << 24) | ((unsigned)b << 18) | (unsigned)m;
    if (none)       // fewer than k searchable points: nobody owns this record -- zeros, from here
        store_all(p.result, (size_t)((q * p.row_stride + (long long)lvl * p.k * 4) >> 2) + lane, make_float4(0.f, 0.f, 0.f, 0.f));
    int owner = 0x100 + lane;                              // lanes without a record: a group of their own
    unsigned loc = 0u;
    if (mine && !none) {
        const int v = idx / L.hw;
        owner = 0;
#pragma unroll
        for (int r = 1; r < MVT_MAX_PEERS; ++r)
            if (r < p.n && v >= p.view_first[r]) owner = r;
        loc = (unsigned)(idx - p.view_first[owner] * L.hw);
    }
    const bool local = mine && !none && owner == p.rank && !p.no_local;
    if (__any_sync(FULL, local)) shard_local_corr<BF16>(p, lvl, b, m, local, loc);      // (uniform)
    if (local) owner = 0x100 + lane;                       // ... and is not sent anywhere
    const unsigned grp = __match_any_sync(FULL, owner);
    const bool has = mine && !none && !local;
    const int leader = __ffs(grp) - 1;
    unsigned base = 0u;
    if (has && lane == leader) base = atomicAdd(&p.local->tail[owner], (unsigned)__popc(grp));
    base = __shfl_sync(FULL, base, leader);                // every lane reads its own group's leader
    if (has) {
        const unsigned slot = base + (unsigned)__popc(grp & ((1u << lane) - 1u));
        const u64 rec = kRecValid | ((u64)lane << 58) | ((u64)qid << 32) | (u64)loc;
        st_sys(p.inbox[owner] + (size_t)p.rank * p.cap + slot, rec);
    }
}

// The searching warp's own share of the correlation: lane j < k holds neighbour j; `local` = this rank owns its feature row
// (index `loc` among the rank's points).  Same gather, per-row arithmetic and lane order as the fused single-GPU kernel
// (mvt_knn.cu: corr_epilogue), with the rows of the other ranks' neighbours left out.
template <bool BF16>
__device__ __forceinline__ void shard_local_corr(const ShardParams &p, int lvl, int b, int m, bool local, unsigned loc) {
    const int lane = threadIdx.x & 31;
    const ShardLevelDev &L = p.lv[lvl];
    const size_t bm = (size_t)b * p.M + m;
    float corr_mine;
    if (!BF16) {
        const float4 tg = __ldg(reinterpret_cast<const float4 *>(p.targets) + bm * 32 + lane);
        const float4 *fv = reinterpret_cast<const float4 *>(L.fvec_local) + (size_t)b * L.n_local * 32 + lane;
        float pr[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            const unsigned ij = __shfl_sync(FULL, loc, j);
            const bool on = __shfl_sync(FULL, (int)local, j) != 0;
            float4 f = make_float4(0.f, 0.f, 0.f, 0.f);
            if (on) f = __ldg(fv + (size_t)ij * 32);
            pr[j] = dot4(tg, f);
        }
        const float tot = transpose_reduce16(pr, lane);
        corr_mine = __shfl_sync(FULL, tot, (2 * lane) & 31);
    } else {
        const uint4 tgh = __ldg(reinterpret_cast<const uint4 *>(p.targets) + bm * 16 + (lane & 15));
        const uint4 *fv = reinterpret_cast<const uint4 *>(L.fvec_local) + (size_t)b * L.n_local * 16 + (lane & 15);
        const int half = lane >> 4;
        float pr[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int j = 2 * i + half;
            const unsigned ij = __shfl_sync(FULL, loc, j);
            const bool on = __shfl_sync(FULL, (int)local, j) != 0;
            uint4 f = make_uint4(0u, 0u, 0u, 0u);
            if (on) f = __ldg(fv + (size_t)ij * 16);
            pr[i] = dot8_bf16(tgh, f);
        }
        const float tot = transpose_reduce8_half(pr, lane);
        corr_mine = __shfl_sync(FULL, tot, ((lane & 1) << 4) | ((lane >> 1) << 1));
    }
    if (local) {
        const float4 n = __ldg(L.xyz4_local + (size_t)b * L.n_local + loc);
        const float qx = __ldg(p.coords + bm * 3), qy = __ldg(p.coords + bm * 3 + 1), qz = __ldg(p.coords + bm * 3 + 2);
        store_all(p.result, (size_t)(((long long)bm * p.row_stride + (long long)lvl * p.k * 4) >> 2) + lane,
                  make_float4(__fdiv_rn(corr_mine, p.scale), n.x - qx, n.y - qy, n.z - qz));
    }
}

// Phase B, one batch of 32 consecutive records of source `src` (lane i holds record i, 0 = none) by one warp: the feature
// rows named by the records are gathered sixteen at a time, correlated with their queries' target rows (per-row arithmetic
// and lane order of the fused single-GPU kernel), and each lane stores its record's [corr, dx, dy, dz] to every rank.
template <bool BF16>
__device__ __forceinline__ void shard_corr_batch(const ShardParams &p, const u64 rec) {
    const int lane = threadIdx.x & 31;
    const unsigned valid = __ballot_sync(FULL, (rec & kRecValid) != 0ull);
    if (valid == 0u) return;
    float corr_mine = 0.f;
#pragma unroll 1
    for (int r = 0; r < 2; ++r) {
        if (((valid >> (16 * r)) & 0xffffu) == 0u) continue;            // (uniform)
        float tot;
        if (!BF16) {
            float pr[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                const u64 rj = __shfl_sync(FULL, rec, 16 * r + j);
                pr[j] = 0.f;
                if (rj & kRecValid) {                                     // (uniform)
                    const unsigned qid = (unsigned)(rj >> 32), li = (unsigned)rj;
                    const unsigned lvl = (qid >> 24) & 3u, b = (qid >> 18) & 63u, bm = b * (unsigned)p.M + (qid & 0x3ffffu);
                    const ShardLevelDev &L = p.lv[lvl];
                    const float4 f = __ldg(reinterpret_cast<const float4 *>(L.fvec_local) + ((size_t)b * L.n_local + li) * 32 + lane);
                    const float4 tg = __ldg(reinterpret_cast<const float4 *>(p.targets) + (size_t)bm * 32 + lane);
                    pr[j] = dot4(tg, f);
                }
            }
            tot = transpose_reduce16(pr, lane);                           // lane L: total of record L >> 1 of this round
            tot = __shfl_sync(FULL, tot, (2 * lane) & 31);                // lane l: total of record l & 15
        } else {
            const int half = lane >> 4;
            float pr[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const u64 rj = __shfl_sync(FULL, rec, 16 * r + 2 * i + half);
                pr[i] = 0.f;
                if (rj & kRecValid) {
                    const unsigned qid = (unsigned)(rj >> 32), li = (unsigned)rj;
                    const unsigned lvl = (qid >> 24) & 3u, b = (qid >> 18) & 63u, bm = b * (unsigned)p.M + (qid & 0x3ffffu);
                    const ShardLevelDev &L = p.lv[lvl];
                    const uint4 f = __ldg(reinterpret_cast<const uint4 *>(L.fvec_local) + ((size_t)b * L.n_local + li) * 16 + (lane & 15));
                    const uint4 tg = __ldg(reinterpret_cast<const uint4 *>(p.targets) + (size_t)bm * 16 + (lane & 15));
                    pr[i] = dot8_bf16(tg, f);
                }
            }
            tot = transpose_reduce8_half(pr, lane);
            // record j of the round lives in half (j & 1), lanes 2 * (j >> 1) (+1) of that half
            tot = __shfl_sync(FULL, tot, ((lane & 1) << 4) | (((lane & 15) >> 1) << 1));
        }
        if ((lane >> 4) == r) corr_mine = tot;
    }
    if (rec & kRecValid) {
        const unsigned qid = (unsigned)(rec >> 32), li = (unsigned)rec, j = (unsigned)(rec >> 58) & 31u;
        const unsigned lvl = (qid >> 24) & 3u, b = (qid >> 18) & 63u, bm = b * (unsigned)p.M + (qid & 0x3ffffu);
        const ShardLevelDev &L = p.lv[lvl];
        const float4 n = __ldg(L.xyz4_local + (size_t)b * L.n_local + li);
        const float qx = __ldg(p.coords + (size_t)bm * 3), qy = __ldg(p.coords + (size_t)bm * 3 + 1), qz = __ldg(p.coords + (size_t)bm * 3 + 2);
        store_all(p.result, (size_t)(((long long)bm * p.row_stride + (long long)lvl * p.k * 4) >> 2) + j,
                  make_float4(__fdiv_rn(corr_mine, p.scale), n.x - qx, n.y - qy, n.z - qz));
    }
}

// Housekeeping of a consumer warp (warp-uniform): (1) finish one deferred query of this rank, if there is one -- tile search,
// then its records go to the owners; tickets, not compare-and-swap: thousands of warps contend for a few hundred entries;
// (2) once every producer warp and every deferred query of this rank is through, tell each owner how many records to expect.
// Returns true when it did a query.  Called from every wait of the consumer, so that a warp that waits for records never
// keeps the deferred queries (on which the end of the record streams depends) from being done.
template <bool BF16>
__device__ __forceinline__ bool shard_service(const ShardParams &p, int &ticket) {
    const int lane = threadIdx.x & 31;
    ShardLocal *loc = p.local;
    bool did = false;
    if (ticket < 0) {
        const unsigned cur = ld_volatile_u32(&loc->hard_cursor), cnt = ld_volatile_u32(&loc->hard_count);
        if (cur < cnt) {
            unsigned t = 0u;
            if (lane == 0) t = atomicAdd(&loc->hard_cursor, 1u);
            ticket = (int)__shfl_sync(FULL, t, 0);
        }
    }
    if (ticket >= 0) {
        if ((unsigned)ticket < ld_volatile_u32(&loc->hard_count)) {
            volatile ShardHard *e = p.hard + ticket;
            const unsigned long long t0 = now_ns();
            while (e->pad != (unsigned)p.epoch) {                         // counted before written: wait for the writer
                __nanosleep(32);
                if (now_ns() - t0 > kShardTimeoutNs) { loc->err = 1u; break; }
            }
            const unsigned qid = e->qid;
            const u64 cap = e->cap;
            const unsigned per_level = (unsigned)p.B * (unsigned)p.Ms;
            const unsigned lvl = qid / per_level, rem = qid - lvl * per_level, b = rem / (unsigned)p.Ms, mm = rem - b * (unsigned)p.Ms;
            const int m = p.rank * p.Ms + (int)mm;
            const long long q = (long long)b * p.M + m;
            const float qx = __ldg(p.coords + q * 3), qy = __ldg(p.coords + q * 3 + 1), qz = __ldg(p.coords + q * 3 + 2);
            TopK tk;
            tk.init();
            tk.th = cap;
            tk = tile_search(p.lv[lvl].sd.rv, (int)b, qx, qy, qz, p.k, tk, p.lv[lvl].sd.g.stats);
            shard_push_records<BF16>(p, (int)lvl, (int)b, m, tk.t);
            __threadfence();
            if (lane == 0) atomicAdd(&loc->hard_done, 1u);
            ticket = -1;
            did = true;
        } else if (ld_volatile_u32(&loc->prod_done) == p.n_items) {
            __threadfence();                                              // every producer is through: the count is final
            if ((unsigned)ticket >= ld_volatile_u32(&loc->hard_count)) ticket = -1;          // a ticket past the end: void
        }
    }
    if (!did && ld_volatile_u32(&loc->published) == 0u && ld_volatile_u32(&loc->prod_done) == p.n_items) {
        __threadfence();                                                  // prod_done complete => hard_count is final
        const unsigned cnt = ld_volatile_u32(&loc->hard_count);
        if (ld_volatile_u32(&loc->hard_done) == cnt && lane == 0 && atomicCAS(&loc->published, 0u, 1u) == 0u) {
            __threadfence();
            for (int o = 0; o < p.n; ++o) st_sys(p.ctrl[o] + p.rank, (u64)atomicAdd(&loc->tail[o], 0u) + 1ull);
        }
        __syncwarp();
    }
    return did;
}

// A consumer warp: housekeeping (above), and batches of 32 records from this rank's inboxes, until every source has delivered
// its total and all of it has been taken.
template <bool BF16>
__device__ __forceinline__ bool shard_consume(const ShardParams &p, const int start) {
    const int lane = threadIdx.x & 31;
    ShardLocal *loc = p.local;
    u64 *inbox = p.inbox[p.rank];
    const u64 *total = p.ctrl[p.rank];
    unsigned exhausted = 0u;
    const unsigned all = (1u << p.n) - 1u;
    int src = start % p.n, idle = 0, ticket = -1;
    unsigned long long t_last = now_ns();                                 // last time this warp found work
    while (exhausted != all) {
        bool took = shard_service<BF16>(p, ticket);
        if (!took && !((exhausted >> src) & 1u)) {
            const unsigned c = ld_volatile_u32(&loc->cursor[src]);
            const u64 tot = ld_sys(total + src);                          // records of this source + 1, 0 = not known yet
            if (tot != 0ull && (u64)c + 1ull >= tot) {
                exhausted |= 1u << src;
            } else if ((long long)c < p.cap && (ld_sys(inbox + (size_t)src * p.cap + c) & kRecValid)) {
                unsigned i0 = 0u;
                if (lane == 0) i0 = atomicAdd(&loc->cursor[src], 32u);
                i0 = __shfl_sync(FULL, i0, 0);
                const long long pos = (long long)i0 + lane;
                u64 *slot = inbox + (size_t)src * p.cap + (pos < p.cap ? pos : 0);
                u64 rec = 0ull;
                bool have = pos >= p.cap;
                const unsigned long long t0 = now_ns();
                for (;;) {                                                // (warp-uniform loop: the housekeeping needs all lanes)
                    if (!have) {
                        rec = ld_sys(slot);
                        have = (rec & kRecValid) != 0ull;
                        if (!have) {
                            const u64 tt = ld_sys(total + src);
                            if (tt != 0ull && (u64)pos + 1ull >= tt) { rec = 0ull; have = true; }    // beyond the source's last record
                        }
                    }
                    if (__all_sync(FULL, have)) break;
                    if (!shard_service<BF16>(p, ticket)) __nanosleep(64);
                    if (now_ns() - t0 > kShardTimeoutNs) { loc->err = 1u; if (!have) rec = 0ull; break; }
                }
                if (rec & kRecValid) *slot = 0ull;                        // the inbox is empty again for the next launch
                shard_corr_batch<BF16>(p, rec);
                took = true;
            }
        }
        if (took) { idle = 0; t_last = now_ns(); continue; }
        src = (src + 1 == p.n) ? 0 : src + 1;
        if (++idle > p.n) {                                               // a whole round without work: back off
            __nanosleep(128);
            if (now_ns() - t_last > kShardTimeoutNs || ld_volatile_u32(&loc->err)) { loc->err = 1u; return false; }
        }
    }
    return true;
}

#ifndef MVT_SHARD_THREADS
#define MVT_SHARD_THREADS 64
#endif
#ifndef MVT_SHARD_MINB
#define MVT_SHARD_MINB 18
#endif
// Grid (ceil(Ms / warps per block), B, n_levels + 1).  z < n_levels: PRODUCER blocks, the shape of the fused single-GPU kernel
// -- one warp per query of this rank's track shard: window search, then one record per neighbour into the owner's inbox; a
// query with a large pixel box is deferred to the consumers.  z == n_levels: CONSUMER blocks (the first n_cons_blocks of the
// slice; one wave).  Blocks are dispatched in order, so a consumer starts only when every producer block of this rank has at
// least been placed: consumers wait for producers (here and on the peers), producers wait for nobody.
template <bool BF16>
__global__ void __launch_bounds__(MVT_SHARD_THREADS, MVT_SHARD_MINB) shard_iter_kernel(const ShardParams p) {
    MVT_WARP_SCRATCH(win_tab, MVT_SHARD_THREADS / 32);
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    ShardLocal *loc = p.local;
    unsigned cb = 0xffffffffu;                  // consumer block id, if this block is one
    if (blockIdx.x >= p.nx) {
        // extra columns: the first row's are the EARLY consumers -- dispatched right behind the first producers, they overlap
        // the correlation with the searches (a bounded share of the resident blocks: producers always find room)
        if (blockIdx.y != 0u || blockIdx.z != 0u) return;
        cb = blockIdx.x - p.nx;
    } else if ((int)blockIdx.z < p.n_levels) {
        const int mm = (int)blockIdx.x * nwarps + wib;
        const int m = p.rank * p.Ms + mm;
        if (mm >= p.Ms || m >= p.M) return;
        const int b = blockIdx.y, lvl = blockIdx.z;
        const ShardLevelDev &L = p.lv[lvl];
        const long long q = (long long)b * p.M + m;
        const float qx = __ldg(p.coords + q * 3), qy = __ldg(p.coords + q * 3 + 1), qz = __ldg(p.coords + q * 3 + 2);
        const Group<32> g(lane);
        TopK tk;
        const bool exact = raster_windows<32, true>(g, L.sd.rv, win_tab, b, true, qx, qy, qz, p.k, tk, L.sd.g.stats, (int)q, kEmptyKey64, 0u);
        if (exact) {
            shard_push_records<BF16>(p, lvl, b, m, tk.t);
        } else if (lane == 0) {
            const unsigned i = atomicAdd(&loc->hard_count, 1u);
            ShardHard *e = p.hard + i;
            e->qid = (unsigned)((lvl * p.B + b) * p.Ms + mm);
            e->cap = tk.th | 0xffffffffull;
            __threadfence();
            e->pad = (unsigned)p.epoch;                                   // the tag last: the entry is complete when it shows
        }
        __threadfence();
        if (lane == 0) atomicAdd(&loc->prod_done, 1u);
        return;
    }
    if (cb == 0xffffffffu) cb = p.n_early + blockIdx.y * p.nx + blockIdx.x;      // last z slice: the late consumers
    if (cb >= p.n_cons_blocks) return;
    shard_consume<BF16>(p, p.rank + 1 + (int)(cb * nwarps + wib));
    // ---- end: the last consumer warp resets this rank's state, signals the epoch and waits for every peer's ----
    __threadfence_system();            // this warp's result stores are performed before it counts as done
    if (lane == 0 && atomicAdd(&loc->cons_done, 1u) == p.n_cons_blocks * (unsigned)nwarps - 1u) {
        __threadfence_system();
        loc->prod_done = 0u; loc->hard_count = 0u; loc->hard_cursor = 0u; loc->hard_done = 0u; loc->published = 0u; loc->cons_done = 0u;
        for (int r = 0; r < MVT_MAX_PEERS; ++r) { loc->tail[r] = 0u; loc->cursor[r] = 0u; }
        u64 *mine = p.ctrl[p.rank];
        for (int r = 0; r < p.n; ++r) mine[r] = 0ull;                     // next launch's totals start unknown
        __threadfence_system();
        for (int r = 0; r < p.n; ++r) st_release_sys(p.ctrl[r] + MVT_MAX_PEERS + p.rank, p.epoch);
        const unsigned long long t0 = now_ns();
        for (int r = 0; r < p.n; ++r)
            while (ld_acquire_sys(mine + MVT_MAX_PEERS + r) < p.epoch) {
                __nanosleep(64);
                if (now_ns() - t0 > kShardTimeoutNs) { loc->err = 1u; break; }
            }
    }
}

}  // namespace mvt

// ---------------------------------------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------------------------------------
MVT_EXPORT int mvt_knn_keys(const mvt_search *searches, const long long *index_offsets, int n_levels, int B,
                            const float *query, int M, int k, int n_shards, const unsigned long long *cap_keys,
                            int budget, unsigned char *exact_out, const unsigned char *skip,
                            unsigned long long *keys_out, void *stream) {
    using namespace mvt;
    MVT_CHECK_ARG(searches && index_offsets && query && keys_out, "mvt_knn_keys: null pointer");
    MVT_CHECK_ARG(n_levels >= 1 && n_levels <= MVT_MAX_LEVELS, "mvt_knn_keys: n_levels must be 1..%d (got %d)", MVT_MAX_LEVELS, n_levels);
    MVT_CHECK_ARG(B > 0 && B <= 65535 && M > 0, "mvt_knn_keys: bad shape B=%d M=%d", B, M);
    MVT_CHECK_ARG(k >= 1 && k <= MVT_MAX_K, "mvt_knn_keys: k must be 1..%d (got %d)", MVT_MAX_K, k);
    MVT_CHECK_ARG(n_shards >= 1, "mvt_knn_keys: n_shards=%d", n_shards);
    KeysParams p;
    p.n_levels = n_levels; p.B = B; p.M = M; p.k = k; p.n_shards = n_shards; p.Ms = (M + n_shards - 1) / n_shards;
    p.query = query; p.out = (u64 *)keys_out;
    MVT_CHECK_ARG(budget >= 0, "mvt_knn_keys: budget=%d", budget);
    p.caps = (const u64 *)cap_keys; p.budget = (unsigned)budget; p.exact_out = exact_out; p.skip = skip;
    p.idx_peers = peers_dev(nullptr); p.out_shard = 0; p.out_shards = 1;
    for (int l = 0; l < n_levels; ++l) {
        int rc = make_search(&searches[l], B, 0, p.sd[l], "mvt_knn_keys");
        MVT_CHECK_ARG(!(cap_keys || budget) || p.sd[l].raster, "mvt_knn_keys: caps / budget need the raster search");
        if (rc) return rc;
        MVT_CHECK_ARG(index_offsets[l] >= 0 && index_offsets[l] + searches[l].N < 0x7fffffffLL, "mvt_knn_keys: index offset %lld out of range", index_offsets[l]);
        p.idx_off[l] = index_offsets[l];
    }
    for (int l = n_levels; l < MVT_MAX_LEVELS; ++l) { p.sd[l] = p.sd[0]; p.idx_off[l] = 0; }
    const int padded = p.n_shards * p.Ms;
    const long long n_items = (long long)n_levels * B * padded;
    MVT_CHECK_ARG(n_items < (1ll << 31), "mvt_knn_keys: too many queries (%lld)", n_items);
    p.n_items = (unsigned)n_items;
    const unsigned grid = one_wave_grid(n_items, MVT_KEYS_THREADS / 32, MVT_KEYS_MINB);
    bool all_raster = true;
    for (int l = 0; l < n_levels; ++l) all_raster = all_raster && p.sd[l].raster;
    if (all_raster) knn_keys_kernel<true><<<grid, MVT_KEYS_THREADS, 0, (cudaStream_t)stream>>>(p);
    else knn_keys_kernel<false><<<grid, MVT_KEYS_THREADS, 0, (cudaStream_t)stream>>>(p);
    MVT_CHECK_LAUNCH("knn_keys");
    return MVT_OK;
}

MVT_EXPORT int mvt_topk_merge(const unsigned long long *keys_in, int n_lists, long long Q, int k, int *idx_out,
                              unsigned long long *keys_out, void *stream) {
    using namespace mvt;
    MVT_CHECK_ARG(keys_in && (idx_out || keys_out), "mvt_topk_merge: null pointer");
    MVT_CHECK_ARG(n_lists >= 1 && Q >= 0 && k >= 1 && k <= MVT_MAX_K, "mvt_topk_merge: bad n_lists=%d Q=%lld k=%d", n_lists, Q, k);
    if (Q == 0) return MVT_OK;
    MVT_CHECK_ARG((Q + 7) / 8 < 0x7fffffffLL, "mvt_topk_merge: too many queries");
    topk_merge_kernel<<<(unsigned)((Q + 7) / 8), 256, 0, (cudaStream_t)stream>>>((const u64 *)keys_in, n_lists, Q, k, idx_out, (u64 *)keys_out);
    MVT_CHECK_LAUNCH("topk_merge");
    return MVT_OK;
}

static int corr_from_idx_impl(const mvt_owned_level *levels, int n_levels, int B, int M, int C, int feat_dtype, int k,
                              int n_shards, const int *idx, const void *targets, const float *coords, float *out,
                              long long out_row_stride, const mvt_peers *out_peers, int shard, void *stream,
                              const float *const *all_xyz4 = nullptr, const long long *n_all = nullptr) {
    using namespace mvt;
    MVT_CHECK_ARG(levels && idx && targets && coords && (out || out_peers), "mvt_corr_from_idx: null pointer");
    MVT_CHECK_ARG(!out_peers || (out_peers->n >= 1 && out_peers->n <= MVT_MAX_PEERS && shard >= 0 && shard < n_shards),
                  "mvt_corr_from_idx_shared: bad peers / shard");
    MVT_CHECK_ARG(n_levels >= 1 && n_levels <= MVT_MAX_LEVELS, "mvt_corr_from_idx: n_levels must be 1..%d (got %d)", MVT_MAX_LEVELS, n_levels);
    MVT_CHECK_ARG(B > 0 && B <= 65535 && M > 0, "mvt_corr_from_idx: bad shape B=%d M=%d", B, M);
    MVT_CHECK_ARG(C == 128, "mvt_corr_from_idx: the sharded path supports C == 128, groups == 1 (got C=%d)", C);
    MVT_CHECK_ARG(k >= 1 && k <= 16, "mvt_corr_from_idx: k must be 1..16 (got %d)", k);
    MVT_CHECK_ARG(feat_dtype == MVT_F32 || feat_dtype == MVT_BF16, "mvt_corr_from_idx: feat_dtype %d", feat_dtype);
    MVT_CHECK_ARG(n_shards >= 1, "mvt_corr_from_idx: n_shards=%d", n_shards);
    MVT_CHECK_ARG((((uintptr_t)out | (uintptr_t)targets) & 15) == 0 && out_row_stride % 4 == 0, "mvt_corr_from_idx: out/targets must be 16-byte aligned");
    CorrIdxParams p;
    p.n_levels = n_levels; p.B = B; p.M = M; p.k = k; p.n_shards = n_shards; p.Ms = (M + n_shards - 1) / n_shards;
    p.idx = idx; p.targets = targets; p.coords = coords; p.out = out; p.row_stride = out_row_stride;
    p.scale = sqrtf((float)C);
    p.out_peers = peers_dev(out_peers); p.shard = shard;
    for (int l = 0; l < n_levels; ++l) {
        const mvt_owned_level &u = levels[l];
        MVT_CHECK_ARG(u.xyz4 && u.fvec && u.hi > u.lo && u.lo >= 0, "mvt_corr_from_idx: level %d: bad shard", l);
        MVT_CHECK_ARG((((uintptr_t)u.xyz4 | (uintptr_t)u.fvec) & 15) == 0 && u.out_offset % 4 == 0, "mvt_corr_from_idx: level %d: 16-byte alignment", l);
        p.lv[l].xyz4 = (const float4 *)u.xyz4; p.lv[l].fvec = u.fvec; p.lv[l].lo = u.lo; p.lv[l].hi = u.hi; p.lv[l].out_off = u.out_offset;
        p.lv[l].all_xyz4 = nullptr; p.lv[l].n_all = 0;
        if (all_xyz4 && all_xyz4[l]) {
            MVT_CHECK_ARG(n_all && n_all[l] >= u.hi && ((uintptr_t)all_xyz4[l] & 15) == 0, "mvt_corr_from_idx_shared: level %d: bad replica", l);
            p.lv[l].all_xyz4 = (const float4 *)all_xyz4[l]; p.lv[l].n_all = n_all[l];
        }
    }
    for (int l = n_levels; l < MVT_MAX_LEVELS; ++l) p.lv[l] = p.lv[0];
    dim3 grid((unsigned)((M + 7) / 8), (unsigned)B, (unsigned)n_levels);
    if (feat_dtype == MVT_BF16) corr_from_idx_kernel<true><<<grid, 256, 0, (cudaStream_t)stream>>>(p);
    else corr_from_idx_kernel<false><<<grid, 256, 0, (cudaStream_t)stream>>>(p);
    MVT_CHECK_LAUNCH("corr_from_idx");
    return MVT_OK;
}

MVT_EXPORT int mvt_corr_from_idx(const mvt_owned_level *levels, int n_levels, int B, int M, int C, int feat_dtype, int k,
                                 int n_shards, const int *idx, const void *targets, const float *coords, float *out,
                                 long long out_row_stride, void *stream) {
    return corr_from_idx_impl(levels, n_levels, B, M, C, feat_dtype, k, n_shards, idx, targets, coords, out, out_row_stride,
                              nullptr, 0, stream);
}

MVT_EXPORT int mvt_corr_from_idx_shared(const mvt_owned_level *levels, int n_levels, int B, int M, int C, int feat_dtype, int k,
                                        int n_shards, int shard, const int *idx, const void *targets, const float *coords,
                                        const mvt_peers *out, long long out_row_stride, const float *const *all_xyz4,
                                        const long long *n_all, void *stream) {
    MVT_CHECK_ARG(out, "mvt_corr_from_idx_shared: null out");
    return corr_from_idx_impl(levels, n_levels, B, M, C, feat_dtype, k, n_shards, idx, targets, coords,
                              (float *)out->ptr[shard < MVT_MAX_PEERS && shard >= 0 ? shard : 0], out_row_stride, out, shard, stream,
                              all_xyz4, n_all);
}

MVT_EXPORT int mvt_knn_idx_shared(const mvt_search *searches, int n_levels, int B, const float *query, int Ms, int k,
                                  int shard, int n_shards, const mvt_peers *idx_all, void *stream) {
    using namespace mvt;
    MVT_CHECK_ARG(searches && query && idx_all, "mvt_knn_idx_shared: null pointer");
    MVT_CHECK_ARG(n_levels >= 1 && n_levels <= MVT_MAX_LEVELS, "mvt_knn_idx_shared: n_levels must be 1..%d (got %d)", MVT_MAX_LEVELS, n_levels);
    MVT_CHECK_ARG(B > 0 && B <= 65535 && Ms > 0, "mvt_knn_idx_shared: bad shape B=%d Ms=%d", B, Ms);
    MVT_CHECK_ARG(k >= 4 && k <= MVT_MAX_K && k % 4 == 0, "mvt_knn_idx_shared: k must be a multiple of 4 in 4..%d (got %d)", MVT_MAX_K, k);
    MVT_CHECK_ARG(idx_all->n >= 1 && idx_all->n <= MVT_MAX_PEERS && shard >= 0 && shard < n_shards, "mvt_knn_idx_shared: bad peers / shard");
    KeysParams p;
    p.n_levels = n_levels; p.B = B; p.M = Ms; p.k = k; p.n_shards = 1; p.Ms = Ms;
    p.query = query; p.out = nullptr;
    p.caps = nullptr; p.budget = 0u; p.exact_out = nullptr; p.skip = nullptr;
    p.idx_peers = peers_dev(idx_all); p.out_shard = shard; p.out_shards = n_shards;
    for (int l = 0; l < n_levels; ++l) {
        int rc = make_search(&searches[l], B, 0, p.sd[l], "mvt_knn_idx_shared");
        if (rc) return rc;
        MVT_CHECK_ARG(p.sd[l].raster, "mvt_knn_idx_shared: needs the raster search");
        p.idx_off[l] = 0;
    }
    for (int l = n_levels; l < MVT_MAX_LEVELS; ++l) { p.sd[l] = p.sd[0]; p.idx_off[l] = 0; }
    const long long n_items = (long long)n_levels * B * Ms;
    MVT_CHECK_ARG(n_items < (1ll << 31), "mvt_knn_idx_shared: too many queries (%lld)", n_items);
    p.n_items = (unsigned)n_items;
    const unsigned grid = one_wave_grid(n_items, MVT_KEYS_THREADS / 32, MVT_KEYS_MINB);
    knn_keys_kernel<true><<<grid, MVT_KEYS_THREADS, 0, (cudaStream_t)stream>>>(p);
    MVT_CHECK_LAUNCH("knn_idx_shared");
    return MVT_OK;
}

MVT_EXPORT size_t mvt_shard_local_bytes(int n_levels, int B, int M, int n_ranks) {
    if (n_levels <= 0 || B <= 0 || M <= 0 || n_ranks <= 0) return 0;
    const long long Ms = ((long long)M + n_ranks - 1) / n_ranks;
    return 256 + sizeof(mvt::ShardHard) * (size_t)n_levels * (size_t)B * (size_t)Ms;
}

MVT_EXPORT int mvt_shard_iter(const mvt_shard_level *levels, int n_levels, int B, int M, int C, int feat_dtype, int k,
                              const void *targets, const float *coords, const mvt_shard_ctx *ctx, long long result_row_stride,
                              unsigned long long epoch, void *stream) {
    using namespace mvt;
    MVT_CHECK_ARG(levels && targets && coords && ctx && ctx->local, "mvt_shard_iter: null pointer");
    MVT_CHECK_ARG(n_levels >= 1 && n_levels <= MVT_MAX_LEVELS, "mvt_shard_iter: n_levels must be 1..%d (got %d)", MVT_MAX_LEVELS, n_levels);
    MVT_CHECK_ARG(B > 0 && B <= 65535 && M > 0, "mvt_shard_iter: bad shape B=%d M=%d", B, M);
    MVT_CHECK_ARG(C == 128, "mvt_shard_iter: the sharded path supports C == 128, groups == 1 (got C=%d)", C);
    MVT_CHECK_ARG(k >= 4 && k <= 16 && k % 4 == 0, "mvt_shard_iter: k must be 4, 8, 12 or 16 (got %d)", k);
    MVT_CHECK_ARG(feat_dtype == MVT_F32 || feat_dtype == MVT_BF16, "mvt_shard_iter: feat_dtype %d", feat_dtype);
    MVT_CHECK_ARG(ctx->n_ranks >= 1 && ctx->n_ranks <= MVT_MAX_PEERS && ctx->rank >= 0 && ctx->rank < ctx->n_ranks,
                  "mvt_shard_iter: rank %d of %d", ctx->rank, ctx->n_ranks);
    MVT_CHECK_ARG(ctx->inbox.n == ctx->n_ranks && ctx->ctrl.n == ctx->n_ranks && ctx->result.n == ctx->n_ranks,
                  "mvt_shard_iter: the symmetric buffers must name all %d ranks", ctx->n_ranks);
    MVT_CHECK_ARG(B <= 64 && M <= (1 << 18), "mvt_shard_iter: the record format holds B <= 64 frames, M <= 262144 tracks (got %d, %d)", B, M);
    MVT_CHECK_ARG(epoch > 0 && result_row_stride % 4 == 0 && (((uintptr_t)targets) & 15) == 0, "mvt_shard_iter: epoch / alignment");
    ShardParams p;
    p.n_levels = n_levels; p.B = B; p.M = M; p.k = k;
    p.n = ctx->n_ranks; p.rank = ctx->rank;
    p.Ms = (M + p.n - 1) / p.n;
    p.targets = targets; p.coords = coords;
    p.local = (ShardLocal *)ctx->local;
    p.cap = ctx->inbox_cap; p.row_stride = result_row_stride;
    p.epoch = epoch; p.scale = sqrtf((float)C);
    MVT_CHECK_ARG(p.cap >= (long long)n_levels * B * p.Ms * k && p.cap < (1ll << 31), "mvt_shard_iter: inbox_cap %lld too small", p.cap);
    for (int r = 0; r < MVT_MAX_PEERS; ++r) {
        p.inbox[r] = r < p.n ? (u64 *)ctx->inbox.ptr[r] : nullptr;
        p.ctrl[r] = r < p.n ? (u64 *)ctx->ctrl.ptr[r] : nullptr;
        MVT_CHECK_ARG(r >= p.n || (p.inbox[r] && p.ctrl[r] && ctx->result.ptr[r]), "mvt_shard_iter: rank %d: null peer mapping", r);
    }
    p.result = peers_dev(&ctx->result);
    for (int r = 0; r <= MVT_MAX_PEERS; ++r) p.view_first[r] = ctx->view_first[r <= p.n ? r : p.n];
    MVT_CHECK_ARG(p.view_first[0] == 0, "mvt_shard_iter: view_first[0] must be 0");
    for (int r = 0; r < p.n; ++r) MVT_CHECK_ARG(p.view_first[r + 1] >= p.view_first[r], "mvt_shard_iter: view_first must ascend");
    for (int l = 0; l < n_levels; ++l) {
        const mvt_shard_level &u = levels[l];
        int rc = make_search(&u.search, B, 0, p.lv[l].sd, "mvt_shard_iter");
        if (rc) return rc;
        MVT_CHECK_ARG(p.lv[l].sd.raster, "mvt_shard_iter: needs the raster search");
        MVT_CHECK_ARG(u.hw == u.search.H * u.search.W && p.view_first[p.n] == u.search.V, "mvt_shard_iter: level %d: views / raster mismatch", l);
        const long long nloc = (long long)(p.view_first[p.rank + 1] - p.view_first[p.rank]) * u.hw;
        MVT_CHECK_ARG(u.n_local == nloc && (nloc == 0 || (u.xyz4_local && u.fvec_local)), "mvt_shard_iter: level %d: n_local %lld != %lld", l, u.n_local, nloc);
        MVT_CHECK_ARG((((uintptr_t)u.xyz4_local | (uintptr_t)u.fvec_local) & 15) == 0, "mvt_shard_iter: level %d: 16-byte alignment", l);
        p.lv[l].xyz4_local = (const float4 *)u.xyz4_local; p.lv[l].fvec_local = u.fvec_local; p.lv[l].n_local = u.n_local; p.lv[l].hw = u.hw;
    }
    for (int l = n_levels; l < MVT_MAX_LEVELS; ++l) p.lv[l] = p.lv[0];
    const long long mine = (long long)p.M - (long long)p.rank * p.Ms;
    const long long tracks = mine < 0 ? 0 : (mine > p.Ms ? p.Ms : mine);
    p.n_items = (unsigned)((long long)n_levels * B * tracks);
    static_assert(sizeof(ShardLocal) <= 256, "ShardLocal must fit the first 256 bytes of the local buffer");
    p.hard = reinterpret_cast<ShardHard *>(reinterpret_cast<unsigned char *>(ctx->local) + 256);
    int dev = 0, sms = 0;
    cudaGetDevice(&dev);
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms < 1) sms = kNumSMs;
    const int wpb = MVT_SHARD_THREADS / 32;
    p.nx = (unsigned)((p.Ms + wpb - 1) / wpb);
    const long long slice = (long long)p.nx * B, wave = (long long)sms * MVT_SHARD_MINB;
    static int early_pct = -1;
    if (early_pct < 0) {
        const char *e = getenv("MVT_SHARD_EARLY_PCT");             // share of one wave that consumes from the start (tuning knob)
        early_pct = e ? atoi(e) : 0;
        if (early_pct < 0 || early_pct > 75) early_pct = 0;
    }
    p.n_early = (unsigned)(wave * early_pct / 100);
    {
        const char *e = getenv("MVT_SHARD_NO_LOCAL");
        p.no_local = (e && e[0] == '1') ? 1 : 0;
    }
    const long long late = slice < wave - p.n_early ? slice : wave - p.n_early;
    p.n_cons_blocks = p.n_early + (unsigned)(late > 0 ? late : 0);
    if (p.n_cons_blocks == 0u) { p.n_early = 1u; p.n_cons_blocks = 1u; }
    const dim3 grid(p.nx + p.n_early, (unsigned)B, (unsigned)n_levels + 1u);
    if (feat_dtype == MVT_BF16) shard_iter_kernel<true><<<grid, MVT_SHARD_THREADS, 0, (cudaStream_t)stream>>>(p);
    else shard_iter_kernel<false><<<grid, MVT_SHARD_THREADS, 0, (cudaStream_t)stream>>>(p);
    MVT_CHECK_LAUNCH("shard_iter");
    return MVT_OK;
}
