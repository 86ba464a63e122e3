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
