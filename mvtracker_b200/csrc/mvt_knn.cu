// kNN query kernels: plain kNN, fused kNN + gather + correlation + offset encoding, and k=1 feature gather.
// Replaces knn / PointcloudCorrBlock.corr_sample / the feature-init block of the reference
// (mvtracker/models/core/mvtracker/mvtracker.py:26-90, 800-846, 627-643).
//
// Two exact search back ends behind one interface: the raster search (mvt_raster.cuh; clouds built by cloud_build,
// no index to build) and the hashed-grid search (mvt_knn.cuh; any cloud).
//
// Roofline of knn_corr: HBM bound on the neighbour-feature gather.  Algorithmic bytes per query (k=16, C=128, fp32):
//   read 12 (xyz) + 512 (target) + 16*(512 + 12) (neighbour rows + xyz), write 16*16  = 9 164 B  (SURVEY.md 8d).
// The neighbour rows go HBM -> registers -> dot product; they are never written back.
#include <mutex>
#include "mvt_search.cuh"
#include "mvt_corr.cuh"

namespace mvt {

// Lanes per query of the fused raster kernel: 32 (default) or 16 (two queries per warp in lock step; measured slightly
// slower on B200 so far -- kept selectable through mvt_debug_set_group_width for experiments and tests).
static int g_group_width = 32;

// RASTER_ONLY: the caller guarantees sd.raster (the grid code is then not even compiled into the kernel).
template <bool RASTER_ONLY = false>
__device__ __forceinline__ void search(const SearchDev &sd, const unsigned char *win_tab, int b, float qx, float qy, float qz,
                                       int k, int lane, float &bd, int &bi, int qtag = 0) {
    TopK tk;
    const Group<32> g(lane);
    if (RASTER_ONLY || sd.raster) raster_search<32, true>(g, sd.rv, win_tab, b, true, qx, qy, qz, k, tk, sd.g.stats, qtag);
    else knn_search(sd.g, b, qx, qy, qz, k, lane, tk);
    bd = key_d2(tk.t);
    bi = key_idx(tk.t);
}

// ---------------------------------------------------------------------------------------------------------------
// plain kNN
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) knn_kernel(const SearchDev sd, int B, int M, const float *__restrict__ query, int k,
                                                  int *__restrict__ idx32, long long *__restrict__ idx64,
                                                  float *__restrict__ dist) {
    MVT_WARP_SCRATCH(win_tab, 8);
    const int lane = threadIdx.x & 31;
    const long long w = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (w >= (long long)B * M) return;
    const int b = (int)(w / M);
    const float qx = __ldg(query + w * 3), qy = __ldg(query + w * 3 + 1), qz = __ldg(query + w * 3 + 2);
    float bd; int bi;
    search(sd, win_tab, b, qx, qy, qz, k, lane, bd, bi);
    if (lane < k) {
        bool ok = bi != kNoIdx;
        if (idx32) idx32[w * k + lane] = ok ? bi : -1;
        if (idx64) idx64[w * k + lane] = ok ? (long long)bi : -1ll;
        if (dist) dist[w * k + lane] = ok ? __fsqrt_rn(bd) : INFINITY;
    }
}

// ---------------------------------------------------------------------------------------------------------------
// fused kNN + gather + correlation
// ---------------------------------------------------------------------------------------------------------------
struct CorrLevelDev {
    SearchDev sd;
    const void *fvec;
    int k;
    long long out_off;
};

struct CorrParams {
    CorrLevelDev lv[MVT_MAX_LEVELS];
    int n_levels, B, M, C, G;
    int add_offset, add_xyz;
    const void *targets;
    const float *coords;
    float *out;
    long long row_stride, frame_stride;
    int *idx_out;
    const int *idx_hint;   // [n_levels,B,M,k] or null: any k points of the same clouds (e.g. last iteration's neighbours)
    float scale;      // sqrt(C/G)
    int vec_out;      // 1: every neighbour record is one aligned float4 (D == 4); 2: ... stored through the NVSwitch multicast
                      // address `out` (multimem.st: one store lands in every rank's block); 3: ... stored to every rank's mapping
    long long q_frame_stride;   // queries between two frames of targets / coords (M when they are dense [B, M, .])
    float *peer[MVT_MAX_PEERS]; // vec_out == 3: the output block of every rank (peer[0] == out)
    int n_peers;
    unsigned *hard;   // deferred queries (see knn_corr_hard_kernel): [0] count, [1] cursor, entries (uint4) from byte 256; or null
    int hard_mode;    // 1: defer the large-box queries; 2: only count them in hard[0] and finish them in place
};

// A query whose ball projects onto a large pixel box in some view (it floats in free space, far from every surface --
// e.g. a track of a moving object in the frames where the object has moved on: the reference starts every window with the
// same 3-D point in all frames, mvtracker.py:510 -- or sits next to a camera) would keep its warp scanning thousands of
// pixels.  The fused kernel gives such a query up as soon as a pass would read kTileMinPixels new pixels of one view and
// appends it here with the k-th distance it has found so far; knn_corr_hard_kernel finishes it with the depth-culled tile
// search (mvt_raster.cuh: tile_search) and the same gather / correlation code.  Costs the fused kernel no registers.
struct HardEntry { unsigned qid, pad; u64 cap; };

// MODE 0: generic fp32 (any C, G).   MODE 1: fp32, C == 128, G == 1, k <= 16.   MODE 2: bf16, C == 128, G == 1, k <= 16.
// GW = 32: one warp per query.   GW = 16 (raster search, MODE 1/2, V <= 16): one half warp per query.
// Grid: x = query blocks of one frame, y = frame b, z = pyramid level.
// Two warps (two queries) per block, 18 blocks per SM: 56 registers per thread, 36 warps per SM.  Small blocks hand a warp
// slot to the next query as soon as its own query is done (128 x 8: 0.734 ms on the panoptic window, 64 x 16: 0.718,
// 32 x 32: 0.699; 256 x 4: 0.802); with the per-view projection constants parked in shared memory between passes
// (MVT_CAM_SMEM, mvt_raster.cuh) 56 registers spill 12 bytes and 36 resident warps beat 32: 32 x 32 -> 0.689, 64 x 18 -> 0.681,
// 64 x 20 (48 registers, 40 warps) -> 0.726.
#ifndef MVT_CORR_MINB
#define MVT_CORR_MINB 18
#endif
#ifndef MVT_CORR_TILES_INLINE
#define MVT_CORR_TILES_INLINE 0     // 1 (A/B): the fused kernel finishes its own large-box queries with tile_search
#endif
#ifndef MVT_HARD_MINB
#define MVT_HARD_MINB 8             // resident 64-thread blocks per SM of knn_corr_hard_kernel
#endif
#ifndef MVT_CORR_THREADS
#define MVT_CORR_THREADS 64
#endif
// Everything after the search: neighbour lists out, gather of the k feature rows, correlation, offsets, store.
// tkt: the group's sorted result list (group lane j holds the j-th neighbour's key).
template <int MODE, bool RASTER_ONLY, int GW>
__device__ __forceinline__ void corr_epilogue(const CorrParams &p, const CorrLevelDev &L, const Group<GW> &g, const int lane, const bool live,
                                              const int b, const int m, const int lvl, const long long q, const float qx, const float qy,
                                              const float qz, const u64 tkt) {
    const int gl = g.gl;
    const int k = L.k, C = p.C;
    const int bi = key_idx(tkt);        // group lane j < k: the j-th neighbour

    if (p.idx_out && live && gl < k) p.idx_out[(((long long)lvl * p.B + b) * p.M + m) * k + gl] = (bi != kNoIdx) ? bi : -1;

    const int D = p.G + 3 * p.add_offset + 3 * p.add_xyz;
    float *orow = p.out + (long long)b * p.frame_stride + (long long)m * p.row_stride + L.out_off;
    const size_t seg = (size_t)b * L.sd.N;
    float corr_mine = 0.f;   // MODE 1/2: correlation of neighbour `gl` (GW == 16) / `lane` (GW == 32)

    if (MODE == 1 && GW == 32) {
        const float4 tg = __ldg(reinterpret_cast<const float4 *>(p.targets) + q * 32 + lane);
        const float4 *fv = reinterpret_cast<const float4 *>(L.fvec) + seg * 32 + lane;
        float pr[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            int ij = __shfl_sync(FULL, bi, j);
            float4 f = make_float4(0.f, 0.f, 0.f, 0.f);
            if (j < k && ij != kNoIdx) f = __ldg(fv + (size_t)ij * 32);
            pr[j] = dot4(tg, f);
        }
        float tot = transpose_reduce16(pr, lane);
        corr_mine = __shfl_sync(FULL, tot, (2 * lane) & 31);
    } else if (MODE == 1) {
        const float4 tg0 = __ldg(reinterpret_cast<const float4 *>(p.targets) + q * 32 + gl);
        const float4 tg1 = __ldg(reinterpret_cast<const float4 *>(p.targets) + q * 32 + 16 + gl);
        const float4 *fv = reinterpret_cast<const float4 *>(L.fvec) + seg * 32 + gl;
        float pr[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            int ij = g.shfl(bi, j);
            float4 f0 = make_float4(0.f, 0.f, 0.f, 0.f), f1 = f0;
            if (j < k && ij != kNoIdx) {
                f0 = __ldg(fv + (size_t)ij * 32);
                f1 = __ldg(fv + (size_t)ij * 32 + 16);
            }
            pr[j] = dot4(tg0, f0) + dot4(tg1, f1);
        }
        corr_mine = transpose_reduce16_half(g, pr);
    } else if (MODE == 2 && GW == 32) {
        const uint4 tgh = __ldg(reinterpret_cast<const uint4 *>(p.targets) + q * 16 + (lane & 15));
        const uint4 *fv = reinterpret_cast<const uint4 *>(L.fvec) + seg * 16 + (lane & 15);
        const int half = lane >> 4;
        float pr[8];   // pr[i]: partial of neighbour 2i + half
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            int j = 2 * i + half;
            int ij = __shfl_sync(FULL, bi, j);
            uint4 f = make_uint4(0u, 0u, 0u, 0u);
            if (j < k && ij != kNoIdx) f = __ldg(fv + (size_t)ij * 16);
            pr[i] = dot8_bf16(tgh, f);
        }
        float tot = transpose_reduce8_half(pr, lane);
        // neighbour j lives in half (j & 1), lanes 2*(j>>1), 2*(j>>1)+1 of that half
        corr_mine = __shfl_sync(FULL, tot, ((lane & 1) << 4) | ((lane >> 1) << 1));
    } else if (MODE == 2) {
        const uint4 tgh = __ldg(reinterpret_cast<const uint4 *>(p.targets) + q * 16 + gl);
        const uint4 *fv = reinterpret_cast<const uint4 *>(L.fvec) + seg * 16 + gl;
        float pr[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            int ij = g.shfl(bi, j);
            uint4 f = make_uint4(0u, 0u, 0u, 0u);
            if (j < k && ij != kNoIdx) f = __ldg(fv + (size_t)ij * 16);
            pr[j] = dot8_bf16(tgh, f);
        }
        corr_mine = transpose_reduce16_half(g, pr);
    }

    if (MODE != 0) {
        if (live && gl < k) {
            bool ok = bi != kNoIdx;
            float3 n = ok ? nbr_xyz<RASTER_ONLY>(L.sd, b, bi) : make_float3(0.f, 0.f, 0.f);
            float cv = ok ? __fdiv_rn(corr_mine, p.scale) : 0.f;
            float *o = orow + (size_t)gl * D;
            if (p.vec_out) {
                const float4 rec = ok ? make_float4(cv, n.x - qx, n.y - qy, n.z - qz) : make_float4(0.f, 0.f, 0.f, 0.f);
                if (p.vec_out == 1) {
                    *reinterpret_cast<float4 *>(o) = rec;
                } else if (p.vec_out == 2) {
                    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(o), "f"(rec.x), "f"(rec.y),
                                 "f"(rec.z), "f"(rec.w) : "memory");
                } else {
                    const long long off = o - p.out;
                    for (int r = 0; r < p.n_peers; ++r) *reinterpret_cast<float4 *>(p.peer[r] + off) = rec;
                }
            } else {
                int wpos = 0;
                o[wpos++] = cv;
                if (p.add_offset) { o[wpos++] = ok ? n.x - qx : 0.f; o[wpos++] = ok ? n.y - qy : 0.f; o[wpos++] = ok ? n.z - qz : 0.f; }
                if (p.add_xyz) { o[wpos++] = n.x; o[wpos++] = n.y; o[wpos++] = n.z; }
            }
        }
        return;
    }

    // ---- MODE 0: generic path (GW == 32), one neighbour and one group at a time ----
    const float *tgt = reinterpret_cast<const float *>(p.targets) + q * C;
    const float *fvf = reinterpret_cast<const float *>(L.fvec);
    const int Cg = C / p.G;
    for (int j = 0; j < k; ++j) {
        int ij = __shfl_sync(FULL, bi, j);
        bool ok = ij != kNoIdx;
        float *o = orow + (size_t)j * D;
        const float *row = fvf + (seg + (size_t)(ok ? ij : 0)) * C;
        for (int gi = 0; gi < p.G; ++gi) {
            float acc = 0.f;
            if (ok)
                for (int c = lane; c < Cg; c += 32) acc = fmaf(__ldg(tgt + gi * Cg + c), __ldg(row + gi * Cg + c), acc);
#pragma unroll
            for (int o2 = 16; o2; o2 >>= 1) acc += __shfl_xor_sync(FULL, acc, o2);
            if (lane == 0) o[gi] = ok ? __fdiv_rn(acc, p.scale) : 0.f;
        }
        if (p.add_offset || p.add_xyz) {
            float3 n = ok ? nbr_xyz<RASTER_ONLY>(L.sd, b, ij) : make_float3(0.f, 0.f, 0.f);
            if (lane < 3) {
                float pv = lane == 0 ? n.x : (lane == 1 ? n.y : n.z);
                float qv = lane == 0 ? qx : (lane == 1 ? qy : qz);
                int wpos = p.G;
                if (p.add_offset) { o[wpos + lane] = ok ? pv - qv : 0.f; wpos += 3; }
                if (p.add_xyz) o[wpos + lane] = pv;
            }
        }
    }
}

template <int MODE, bool RASTER_ONLY, int GW, bool DEFER = false>
__global__ void __launch_bounds__(MVT_CORR_THREADS, MVT_CORR_MINB) knn_corr_kernel(const CorrParams p) {
    MVT_WARP_SCRATCH(win_tab, MVT_CORR_THREADS / 32);
    const int lane = threadIdx.x & 31;
    const Group<GW> g(lane);
    const int gl = g.gl;
    const int m_raw = (int)((blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * (32 / GW)) + (GW == 32 ? 0 : (lane >> 4));
    const bool live = m_raw < p.M;       // half-warp groups: the last group of a frame may be padding
    if (GW == 32 && !live) return;
    const int m = live ? m_raw : p.M - 1;
    const int b = blockIdx.y, lvl = blockIdx.z;
    const long long q = (long long)b * p.q_frame_stride + m;
    const CorrLevelDev &L = p.lv[lvl];
    const int k = L.k;
    const float qx = __ldg(p.coords + q * 3), qy = __ldg(p.coords + q * 3 + 1), qz = __ldg(p.coords + q * 3 + 2);

    TopK tk;
    // warm start: the largest distance to k known points of this cloud bounds the k-th best -> cap of the search
    u64 cap = kEmptyKey64;
    if (p.idx_hint && (RASTER_ONLY || L.sd.raster)) {
        unsigned bits = 0xffffffffu;
        // plain (coherent) load: idx_hint may alias idx_out, which this launch writes -- each query reads its own k
        // entries before it writes them, but ld.global.nc promises that NOBODY writes the location during the kernel
        int hi = -1 - gl;                       // distinct dummies for the lanes beyond k
        if (gl < k) hi = p.idx_hint[(((long long)lvl * p.B + b) * p.M + m) * k + gl];
        // the cap is the largest distance to k DISTINCT points; a repeated index would make it too small: no cap then
        const unsigned same = __match_any_sync(FULL, GW == 32 ? hi : (hi ^ ((lane & 16) << 26)));
        const bool dup = gl < k && hi >= 0 && (same & (same - 1u)) != 0u;
        if (gl < k) {
            if (hi >= 0 && hi < L.sd.N && !dup) {
                const float4 hp = __ldg(L.sd.rv.xyz4 + (size_t)b * L.sd.N + hi);
                const float hd = dist2(hp.x, hp.y, hp.z, qx, qy, qz);
                const bool usable = L.sd.rv.skip_invalid ? (hp.w > 0.f) : !(hp.w < 0.f);
                if (usable && hd == hd) bits = __float_as_uint(hd);
            }
        } else {
            bits = 0u;
        }
        // group maximum (all lanes take part)
#pragma unroll
        for (int o = GW / 2; o; o >>= 1) bits = max(bits, g.shfl_xor(bits, o));
        if (bits < 0x7f800000u) cap = ((u64)bits << 32) | 0xffffffffull;
    }
    if (RASTER_ONLY || L.sd.raster) {
        if constexpr (DEFER) {
            // windows only; the search gives up on a large pixel box (the level has tile tables, rv.ztab != null)
            const bool exact = raster_windows<32, true>(g, L.sd.rv, win_tab, b, live, qx, qy, qz, k, tk, L.sd.g.stats, (int)q, cap, 0u,
                                                        p.hard_mode == 2 ? p.hard : nullptr);
#if MVT_CORR_TILES_INLINE
            if (!exact) tk = tile_search(L.sd.rv, b, qx, qy, qz, k, tk, L.sd.g.stats);      // (A/B variant: finish it here)
#else
            if (!exact) {       // (uniform) -> knn_corr_hard_kernel
                if (lane == 0) {
                    const unsigned i = atomicAdd(p.hard, 1u);
                    HardEntry e;
                    e.qid = (unsigned)(((long long)lvl * p.B + b) * p.M + m); e.pad = 0u;
                    e.cap = tk.th | 0xffffffffull;
                    reinterpret_cast<HardEntry *>(reinterpret_cast<unsigned char *>(p.hard) + 256)[i] = e;
                }
                return;
            }
#endif
        } else {
            raster_search<GW>(g, L.sd.rv, win_tab, b, live, qx, qy, qz, k, tk, L.sd.g.stats, (int)q, cap);
        }
    } else if (GW == 32) {
        knn_search(L.sd.g, b, qx, qy, qz, k, lane, tk);
    }
    corr_epilogue<MODE, RASTER_ONLY, GW>(p, L, g, lane, live, b, m, lvl, q, qx, qy, qz, tk.t);
}

// The deferred queries of knn_corr_kernel: one wave of blocks, every warp draws entries until the list is empty.
template <int MODE>
__global__ void __launch_bounds__(64, MVT_HARD_MINB) knn_corr_hard_kernel(const CorrParams p) {
    const int lane = threadIdx.x & 31;
    const Group<32> g(lane);
    const unsigned count = p.hard[0];
    const HardEntry *list = reinterpret_cast<const HardEntry *>(reinterpret_cast<const unsigned char *>(p.hard) + 256);
    for (;;) {
        unsigned i = 0u;
        if (lane == 0) i = atomicAdd(p.hard + 1, 1u);
        i = __shfl_sync(FULL, i, 0);
        if (i >= count) return;
        const HardEntry e = list[i];
        const int m = (int)(e.qid % (unsigned)p.M);
        const unsigned fr = e.qid / (unsigned)p.M;
        const int b = (int)(fr % (unsigned)p.B), lvl = (int)(fr / (unsigned)p.B);
        const long long q = (long long)b * p.q_frame_stride + m;
        const CorrLevelDev &L = p.lv[lvl];
        const float qx = __ldg(p.coords + q * 3), qy = __ldg(p.coords + q * 3 + 1), qz = __ldg(p.coords + q * 3 + 2);
        TopK tk;
        tk.init();
        tk.th = e.cap;
        tk = tile_search(L.sd.rv, b, qx, qy, qz, L.k, tk, L.sd.g.stats);
        corr_epilogue<MODE, true, 32>(p, L, g, lane, true, b, m, lvl, q, qx, qy, qz, tk.t);
    }
}

// ---------------------------------------------------------------------------------------------------------------
// k = 1 search + feature row copy (feature init)
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) knn_gather_kernel(const SearchDev sd, const void *__restrict__ fvec, int elem_bytes,
                                                         int B, int M, int C, const float *__restrict__ query,
                                                         void *__restrict__ feat_out, int *__restrict__ idx_out) {
    MVT_WARP_SCRATCH(win_tab, 8);
    const int lane = threadIdx.x & 31;
    const long long w = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (w >= (long long)B * M) return;
    const int b = (int)(w / M);
    const float qx = __ldg(query + w * 3), qy = __ldg(query + w * 3 + 1), qz = __ldg(query + w * 3 + 2);
    float bd; int bi;
    search(sd, win_tab, b, qx, qy, qz, 1, lane, bd, bi);
    int i0 = __shfl_sync(FULL, bi, 0);
    if (idx_out && lane == 0) idx_out[w] = (i0 != kNoIdx) ? i0 : -1;
    const size_t row_bytes = (size_t)C * elem_bytes;
    const unsigned char *src = reinterpret_cast<const unsigned char *>(fvec) + ((size_t)b * sd.N + (size_t)(i0 != kNoIdx ? i0 : 0)) * row_bytes;
    unsigned char *dst = reinterpret_cast<unsigned char *>(feat_out) + (size_t)w * row_bytes;
    if ((row_bytes & 15) == 0 && (((uintptr_t)src | (uintptr_t)dst) & 15) == 0) {
        for (size_t o = (size_t)lane * 16; o < row_bytes; o += 512) {
            uint4 v = (i0 != kNoIdx) ? __ldg(reinterpret_cast<const uint4 *>(src + o)) : make_uint4(0u, 0u, 0u, 0u);
            *reinterpret_cast<uint4 *>(dst + o) = v;
        }
    } else {
        for (size_t o = lane; o < row_bytes; o += 32) dst[o] = (i0 != kNoIdx) ? src[o] : 0;
    }
}

// ---------------------------------------------------------------------------------------------------------------
// tile depth tables (raster search: depth culling of large pixel boxes, tile_scan in mvt_raster.cuh)
// ---------------------------------------------------------------------------------------------------------------
struct TileJob {
    const float4 *xyz4;     // [B, V*H*W]
    const float *cam;
    float2 *out;            // [B, V, TH, TW]
    int V, S, H, W, TW, TH;
    long long first;        // first warp of this job
};
struct TileParams {
    TileJob job[8];
    int n, B;
    long long warps;
};

// one warp per 8x8 tile: {min, max} over its points of Z = third row of the view matrix . p, evaluated exactly as the
// search evaluates the query's Z (same fma chain), so that the cull test compares like with like
__global__ void __launch_bounds__(256) tile_depth_kernel(const TileParams p) {
    const int lane = threadIdx.x & 31;
    const long long w = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (w >= p.warps) return;
    int ji = 0;
#pragma unroll
    for (int i = 1; i < 8; ++i)
        if (i < p.n && w >= p.job[i].first) ji = i;
    const TileJob &J = p.job[ji];
    long long t = w - J.first;
    const int tx = (int)(t % J.TW); t /= J.TW;
    const int ty = (int)(t % J.TH); t /= J.TH;
    const int v = (int)(t % J.V);
    const int b = (int)(t / J.V);
    const float *cm = J.cam + ((size_t)v * J.S + b) * kCamStride;
    const float r0 = cm[30], r1 = cm[31], r2 = cm[32], t3 = cm[35];
    const float4 *vf = J.xyz4 + ((size_t)b * J.V + v) * J.H * J.W;
    float zmin = INFINITY, zmax = -INFINITY;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const int px = tx * kTile + (lane & 7), py = ty * kTile + (lane >> 3) + 4 * h;
        if (px < J.W && py < J.H) {
            const float4 q = __ldg(vf + (size_t)py * J.W + px);
            if (!(q.w < 0.f)) {       // every point any search may return (skip_invalid only narrows this)
                const float z = fmaf(r0, q.x, fmaf(r1, q.y, fmaf(r2, q.z, t3)));
                if (z == z) { zmin = fminf(zmin, z); zmax = fmaxf(zmax, z); }
            }
        }
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) {
        zmin = fminf(zmin, __shfl_xor_sync(FULL, zmin, o));
        zmax = fmaxf(zmax, __shfl_xor_sync(FULL, zmax, o));
    }
    if (lane == 0) J.out[(((size_t)b * J.V + v) * J.TH + ty) * J.TW + tx] = make_float2(zmin, zmax);
}

}  // namespace mvt

// ---------------------------------------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------------------------------------
MVT_EXPORT long long mvt_tile_depth_floats(int B, int V, int H, int W) {
    if (B <= 0 || V <= 0 || H <= 0 || W <= 0) return 0;
    return 2ll * B * V * ((H + MVT_TILE - 1) / MVT_TILE) * ((W + MVT_TILE - 1) / MVT_TILE);
}

MVT_EXPORT int mvt_tile_depth_build(const mvt_search *searches, int n, int B, float *const *tile_z_out, void *stream) {
    using namespace mvt;
    MVT_CHECK_ARG(searches && tile_z_out, "mvt_tile_depth_build: null pointer");
    MVT_CHECK_ARG(n >= 1 && n <= 8 && B > 0, "mvt_tile_depth_build: n must be 1..8, B > 0 (got n=%d B=%d)", n, B);
    TileParams p;
    p.n = n; p.B = B;
    long long warps = 0;
    for (int i = 0; i < n; ++i) {
        const mvt_search &u = searches[i];
        MVT_CHECK_ARG(!u.grid && u.xyz4 && u.cam && tile_z_out[i], "mvt_tile_depth_build: descriptor %d is not a raster search / null output", i);
        MVT_CHECK_ARG(u.V > 0 && u.V <= MVT_MAX_VIEWS && u.H > 0 && u.W > 0 && (long long)u.V * u.H * u.W == (long long)u.N,
                      "mvt_tile_depth_build: descriptor %d: bad raster V=%d H=%d W=%d N=%d", i, u.V, u.H, u.W, u.N);
        MVT_CHECK_ARG(u.cam_view_stride >= B, "mvt_tile_depth_build: descriptor %d: cam_view_stride %d < B", i, u.cam_view_stride);
        MVT_CHECK_ARG((((uintptr_t)u.xyz4 & 15) | ((uintptr_t)tile_z_out[i] & 7)) == 0, "mvt_tile_depth_build: descriptor %d: alignment", i);
        TileJob &J = p.job[i];
        J.xyz4 = (const float4 *)u.xyz4; J.cam = u.cam; J.out = (float2 *)tile_z_out[i];
        J.V = u.V; J.S = u.cam_view_stride; J.H = u.H; J.W = u.W;
        J.TW = (u.W + kTile - 1) / kTile; J.TH = (u.H + kTile - 1) / kTile;
        J.first = warps;
        warps += (long long)B * u.V * J.TH * J.TW;
    }
    for (int i = n; i < 8; ++i) p.job[i] = p.job[0];
    p.warps = warps;
    MVT_CHECK_ARG((warps + 7) / 8 < 0x7fffffffLL, "mvt_tile_depth_build: too many tiles");
    tile_depth_kernel<<<(unsigned)((warps + 7) / 8), 256, 0, (cudaStream_t)stream>>>(p);
    MVT_CHECK_LAUNCH("tile_depth_build");
    return MVT_OK;
}

static inline unsigned warp_blocks(long long warps) { return (unsigned)((warps + 7) / 8); }

namespace mvt {
int make_search(const mvt_search *u, int B, int need_xyz, SearchDev &sd, const char *who) {
    MVT_CHECK_ARG(u, "%s: null search descriptor", who);
    MVT_CHECK_ARG(u->N > 0, "%s: bad N=%d", who, u->N);
    sd.N = u->N;
    sd.xyz = u->xyz;
    if (u->grid) {
        sd.raster = 0;
        sd.g = grid_view(u->grid, B, u->N);
        sd.rv = RasterView{};
        MVT_CHECK_ARG(!need_xyz || u->xyz, "%s: grid mode needs xyz", who);
    } else {
        MVT_CHECK_ARG(u->xyz4 && u->cam, "%s: neither a grid nor raster records (xyz4 + cam) were given", who);
        MVT_CHECK_ARG(u->V > 0 && u->V <= MVT_MAX_VIEWS && u->H > 0 && u->W > 0, "%s: bad raster V=%d H=%d W=%d (V <= %d)", who,
                      u->V, u->H, u->W, MVT_MAX_VIEWS);
        MVT_CHECK_ARG((long long)u->V * u->H * u->W == (long long)u->N, "%s: N=%d != V*H*W=%lld", who, u->N,
                      (long long)u->V * u->H * u->W);
        MVT_CHECK_ARG(u->stride > 0.f && u->cam_view_stride >= B, "%s: bad stride %f / cam_view_stride %d", who, u->stride,
                      u->cam_view_stride);
        MVT_CHECK_ARG(((uintptr_t)u->xyz4 & 15) == 0, "%s: xyz4 must be 16-byte aligned", who);
        sd.raster = 1;
        sd.g = GridView{};
        sd.g.stats = debug_stats_ptr();
        sd.rv.xyz4 = reinterpret_cast<const float4 *>(u->xyz4);
        sd.rv.cam = u->cam;
        sd.rv.V = u->V; sd.rv.S = u->cam_view_stride; sd.rv.H = u->H; sd.rv.W = u->W;
        sd.rv.stride = u->stride;
        sd.rv.skip_invalid = u->skip_invalid ? 1 : 0;
        MVT_CHECK_ARG(((uintptr_t)u->tile_z & 7) == 0, "%s: tile_z must be 8-byte aligned", who);
        sd.rv.ztab = reinterpret_cast<const float2 *>(u->tile_z);
        sd.rv.TW = (u->W + kTile - 1) / kTile; sd.rv.TH = (u->H + kTile - 1) / kTile;
        sd.rv.neg = u->neg_depth;
    }
    return MVT_OK;
}
}  // namespace mvt
using mvt::make_search;

MVT_EXPORT int mvt_knn(const mvt_search *search, int B, const float *query, int M, int k,
                       void *idx_out, int idx_is_i64, float *dist_out, void *stream) {
    using namespace mvt;
    MVT_CHECK_ARG(search && query, "mvt_knn: null pointer");
    MVT_CHECK_ARG(B > 0 && M >= 0, "mvt_knn: bad shape B=%d M=%d", B, M);
    MVT_CHECK_ARG(k >= 1 && k <= MVT_MAX_K, "mvt_knn: k must be 1..%d (got %d)", MVT_MAX_K, k);
    if (M == 0) return MVT_OK;
    SearchDev sd;
    int rc = make_search(search, B, 0, sd, "mvt_knn");
    if (rc) return rc;
    long long warps = (long long)B * M;
    MVT_CHECK_ARG(warps / 8 < 0x7fffffffLL, "mvt_knn: too many queries");
    knn_kernel<<<warp_blocks(warps), 256, 0, (cudaStream_t)stream>>>(
        sd, B, M, query, k, idx_is_i64 ? nullptr : (int *)idx_out, idx_is_i64 ? (long long *)idx_out : nullptr, dist_out);
    MVT_CHECK_LAUNCH("knn");
    return MVT_OK;
}

MVT_EXPORT size_t mvt_knn_corr_workspace_bytes(int n_levels, int B, int M) {
    if (n_levels <= 0 || B <= 0 || M <= 0) return 0;
    return 256 + sizeof(mvt::HardEntry) * (size_t)n_levels * (size_t)B * (size_t)M;
}

// out_peers != null: `out` is ignored; the records go to every rank's block (mvt_knn_corr_shared).
static int knn_corr_impl(const mvt_corr_level *levels, int n_levels, int B, int M, int C, int G, int feat_dtype,
                         const void *targets, const float *coords, long long query_frame_stride, int add_offset, int add_xyz,
                         float *out, const mvt_peers *out_peers, long long out_row_stride, long long out_frame_stride,
                         const int *idx_hint, int *idx_out, void *workspace, int hard_query_mode, void *stream) {
    using namespace mvt;
    if (out_peers) {
        MVT_CHECK_ARG(out_peers->n >= 1 && out_peers->n <= MVT_MAX_PEERS, "mvt_knn_corr_shared: out has %d peers", out_peers->n);
        out = (float *)(out_peers->mc ? out_peers->mc : out_peers->ptr[0]);
    }
    MVT_CHECK_ARG(levels && targets && coords && out, "mvt_knn_corr: null pointer");
    MVT_CHECK_ARG(query_frame_stride >= M, "mvt_knn_corr: query frame stride %lld < M", query_frame_stride);
    MVT_CHECK_ARG(n_levels >= 1 && n_levels <= MVT_MAX_LEVELS, "mvt_knn_corr: n_levels must be 1..%d (got %d)", MVT_MAX_LEVELS, n_levels);
    MVT_CHECK_ARG(B > 0 && M >= 0 && C > 0, "mvt_knn_corr: bad shape B=%d M=%d C=%d", B, M, C);
    MVT_CHECK_ARG(G >= 1 && G <= C && C % G == 0, "mvt_knn_corr: groups=%d must divide C=%d", G, C);   // mvtracker.py:786-787
    MVT_CHECK_ARG(feat_dtype == MVT_F32 || feat_dtype == MVT_BF16, "mvt_knn_corr: feat_dtype %d", feat_dtype);
    if (M == 0) return MVT_OK;
    CorrParams p;
    p.n_levels = n_levels; p.B = B; p.M = M; p.C = C; p.G = G;
    p.add_offset = add_offset ? 1 : 0; p.add_xyz = add_xyz ? 1 : 0;
    p.targets = targets; p.coords = coords; p.out = out; p.row_stride = out_row_stride;
    p.frame_stride = out_frame_stride ? out_frame_stride : (long long)M * out_row_stride; p.idx_out = idx_out; p.idx_hint = idx_hint;
    p.scale = sqrtf((float)C / (float)G);
    p.q_frame_stride = query_frame_stride;
    p.n_peers = 0;
    for (int r = 0; r < MVT_MAX_PEERS; ++r) p.peer[r] = nullptr;
    const int D = G + 3 * p.add_offset + 3 * p.add_xyz;
    bool fast = (C == 128 && G == 1);
    bool vec = (D == 4) && (((uintptr_t)out & 15) == 0) && (out_row_stride % 4 == 0) && (out_frame_stride % 4 == 0);
    bool aligned = (((uintptr_t)targets & 15) == 0);
    bool all_raster = true;
    for (int l = 0; l < n_levels; ++l) {
        const mvt_corr_level &u = levels[l];
        MVT_CHECK_ARG(u.fvec, "mvt_knn_corr: level %d has a null fvec", l);
        MVT_CHECK_ARG(u.k >= 1 && u.k <= MVT_MAX_K, "mvt_knn_corr: level %d bad k=%d", l, u.k);
        MVT_CHECK_ARG(u.k <= u.search.N, "mvt_knn_corr: k should be less than or equal to N (level %d: k=%d N=%d)", l, u.k,
                      u.search.N);   // mvtracker.py:785
        MVT_CHECK_ARG((idx_out == nullptr && idx_hint == nullptr) || u.k == levels[0].k, "mvt_knn_corr: idx_out / idx_hint need the same k on every level");
        int rc = make_search(&u.search, B, p.add_offset || p.add_xyz, p.lv[l].sd, "mvt_knn_corr");
        if (rc) return rc;
        p.lv[l].fvec = u.fvec; p.lv[l].k = u.k; p.lv[l].out_off = u.out_offset;
        all_raster = all_raster && p.lv[l].sd.raster;
        fast = fast && u.k <= 16;
        vec = vec && (u.out_offset % 4 == 0);
        aligned = aligned && (((uintptr_t)u.fvec & 15) == 0);
    }
    for (int l = n_levels; l < MVT_MAX_LEVELS; ++l) p.lv[l] = p.lv[0];
    p.vec_out = vec ? 1 : 0;
    if (out_peers) {
        MVT_CHECK_ARG(vec, "mvt_knn_corr_shared: needs 16-byte records (groups = 1, add_offset = 1, add_xyz = 0) and 16-byte aligned rows");
        if (out_peers->mc) {
            p.vec_out = 2;
        } else {
            p.vec_out = 3;
            p.n_peers = out_peers->n;
            for (int r = 0; r < out_peers->n; ++r) {
                MVT_CHECK_ARG(out_peers->ptr[r] && ((uintptr_t)out_peers->ptr[r] & 15) == 0, "mvt_knn_corr_shared: peer %d pointer", r);
                p.peer[r] = (float *)out_peers->ptr[r];
            }
        }
    }
    MVT_CHECK_ARG(feat_dtype == MVT_F32 || (fast && aligned), "mvt_knn_corr: bf16 features need C=128, groups=1, k<=16, 16-byte aligned rows");
    MVT_CHECK_ARG(B <= 65535, "mvt_knn_corr: B=%d too large", B);
    cudaStream_t st = (cudaStream_t)stream;
    bool half = all_raster && fast && aligned;      // half-warp groups: k <= 16 (part of `fast`) and V <= 16
    for (int l = 0; l < n_levels; ++l) half = half && levels[l].search.V <= 16;
    if (mvt::g_group_width != 16) half = false;
    const int qpb = (half ? 2 : 1) * (MVT_CORR_THREADS / 32);      // queries per block
    const dim3 grid((unsigned)((M + qpb - 1) / qpb), (unsigned)B, (unsigned)n_levels);
    // deferral of the queries with large pixel boxes: needs the caller's scratch and every level's tile tables
    MVT_CHECK_ARG(hard_query_mode >= 0 && hard_query_mode <= 2, "mvt_knn_corr: hard_query_mode %d", hard_query_mode);
    bool defer = workspace != nullptr && hard_query_mode != 0 && all_raster && fast && aligned && !half && (long long)n_levels * B * M < (1ll << 31);
    for (int l = 0; l < n_levels; ++l) defer = defer && levels[l].search.tile_z != nullptr;
    p.hard = nullptr;
    p.hard_mode = hard_query_mode;
    if (defer) {
        MVT_CHECK_ARG(((uintptr_t)workspace & 15) == 0, "mvt_knn_corr: workspace must be 16-byte aligned");
        p.hard = (unsigned *)workspace;
        if (cudaMemsetAsync(workspace, 0, 16, st) != cudaSuccess) {
            set_error("mvt_knn_corr: cudaMemsetAsync(workspace) failed");
            return MVT_E_CUDA;
        }
    }
    if (feat_dtype == MVT_BF16) {
        if (half) knn_corr_kernel<2, true, 16><<<grid, MVT_CORR_THREADS, 0, st>>>(p);
        else if (defer) knn_corr_kernel<2, true, 32, true><<<grid, MVT_CORR_THREADS, 0, st>>>(p);
        else if (all_raster) knn_corr_kernel<2, true, 32><<<grid, MVT_CORR_THREADS, 0, st>>>(p);
        else knn_corr_kernel<2, false, 32><<<grid, MVT_CORR_THREADS, 0, st>>>(p);
    } else if (fast && aligned) {
        if (half) knn_corr_kernel<1, true, 16><<<grid, MVT_CORR_THREADS, 0, st>>>(p);
        else if (defer) knn_corr_kernel<1, true, 32, true><<<grid, MVT_CORR_THREADS, 0, st>>>(p);
        else if (all_raster) knn_corr_kernel<1, true, 32><<<grid, MVT_CORR_THREADS, 0, st>>>(p);
        else knn_corr_kernel<1, false, 32><<<grid, MVT_CORR_THREADS, 0, st>>>(p);
    } else {
        knn_corr_kernel<0, false, 32><<<grid, MVT_CORR_THREADS, 0, st>>>(p);
    }
    MVT_CHECK_LAUNCH("knn_corr");
    if (defer && hard_query_mode == 1) {
        const unsigned hgrid = one_wave_grid((long long)n_levels * B * M, 2, MVT_HARD_MINB);
        if (feat_dtype == MVT_BF16) knn_corr_hard_kernel<2><<<hgrid, 64, 0, st>>>(p);
        else knn_corr_hard_kernel<1><<<hgrid, 64, 0, st>>>(p);
        MVT_CHECK_LAUNCH("knn_corr_hard");
    }
    return MVT_OK;
}

// ---- mvt_knn_corr without any allocation: the deferred-query list lives in a static device array (include/mvt.h) ------------
// The array's name is a host-side stand-in in host code, not a device address: the address a kernel may use comes from
// cudaGetSymbolAddress, asked once per device with that device current.
static __device__ __align__(256) unsigned char g_noalloc_scratch[MVT_KNN_CORR_NOALLOC_BYTES];
static_assert(sizeof(mvt::HardEntry) == 16, "MVT_KNN_CORR_NOALLOC_QUERIES assumes 16-byte entries");

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_noalloc_mutex;
void *g_noalloc_ptr[kMaxDevices] = {};

int noalloc_scratch(void **ptr) {
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
        mvt::set_error("mvt_knn_corr_noalloc: cudaGetDevice failed or device %d out of range", dev);
        return MVT_E_CUDA;
    }
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (g_noalloc_ptr[dev] == nullptr) {
        void *p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);      // this device's copy of the array
        if (e != cudaSuccess || p == nullptr) {
            mvt::set_error("mvt_knn_corr_noalloc: cudaGetSymbolAddress on device %d: %s", dev, cudaGetErrorString(e));
            return MVT_E_CUDA;
        }
        g_noalloc_ptr[dev] = p;
    }
    *ptr = g_noalloc_ptr[dev];
    return MVT_OK;
}
}   // namespace

MVT_EXPORT int mvt_knn_corr_noalloc(const mvt_corr_level *levels, int n_levels, int B, int M, int C, int G, int feat_dtype,
                                    const void *targets, const float *coords, int add_offset, int add_xyz,
                                    float *out, long long out_row_stride, long long out_frame_stride, const int *idx_hint,
                                    int *idx_out, int hard_query_mode, void *stream) {
    MVT_CHECK_ARG(hard_query_mode == 1 || hard_query_mode == 2,
                  "mvt_knn_corr_noalloc: hard_query_mode must be 1 or 2 (got %d; mode 0 uses no scratch: call mvt_knn_corr)", hard_query_mode);
    MVT_CHECK_ARG(n_levels >= 1 && n_levels <= MVT_MAX_LEVELS && B > 0 && M >= 0, "mvt_knn_corr_noalloc: bad shape n_levels=%d B=%d M=%d",
                  n_levels, B, M);
    const unsigned long long queries = (unsigned long long)n_levels * (unsigned long long)B * (unsigned long long)M;
    if (queries > MVT_KNN_CORR_NOALLOC_QUERIES || mvt_knn_corr_workspace_bytes(n_levels, B, M) > MVT_KNN_CORR_NOALLOC_BYTES) {
        mvt::set_error("mvt_knn_corr_noalloc: n_levels*B*M = %llu exceeds the static scratch (%u queries, %u bytes); use mvt_knn_corr "
                       "with a caller-owned workspace", queries, (unsigned)MVT_KNN_CORR_NOALLOC_QUERIES, (unsigned)MVT_KNN_CORR_NOALLOC_BYTES);
        return MVT_E_WORKSPACE;
    }
    void *scratch = nullptr;
    int rc = noalloc_scratch(&scratch);
    if (rc) return rc;
    // mvt_knn_corr clears the counters at the head of the scratch on `stream` before its kernels run, and launches only
    return mvt_knn_corr(levels, n_levels, B, M, C, G, feat_dtype, targets, coords, add_offset, add_xyz, out, out_row_stride,
                        out_frame_stride, idx_hint, idx_out, scratch, hard_query_mode, stream);
}

MVT_EXPORT int mvt_knn_corr(const mvt_corr_level *levels, int n_levels, int B, int M, int C, int G, int feat_dtype,
                            const void *targets, const float *coords, int add_offset, int add_xyz,
                            float *out, long long out_row_stride, long long out_frame_stride, const int *idx_hint, int *idx_out,
                            void *workspace, int hard_query_mode, void *stream) {
    MVT_CHECK_ARG(out, "mvt_knn_corr: null pointer");
    return knn_corr_impl(levels, n_levels, B, M, C, G, feat_dtype, targets, coords, M, add_offset, add_xyz, out, nullptr, out_row_stride,
                         out_frame_stride, idx_hint, idx_out, workspace, hard_query_mode, stream);
}

MVT_EXPORT int mvt_knn_corr_shared(const mvt_corr_level *levels, int n_levels, int B, int M, int C, int feat_dtype,
                                   const void *targets, const float *coords, long long query_frame_stride,
                                   const mvt_peers *out, long long out_row_stride, long long out_frame_stride,
                                   void *workspace, int hard_query_mode, void *stream) {
    MVT_CHECK_ARG(out, "mvt_knn_corr_shared: null pointer");
    return knn_corr_impl(levels, n_levels, B, M, C, 1, feat_dtype, targets, coords, query_frame_stride, 1, 0, nullptr, out, out_row_stride,
                         out_frame_stride, nullptr, nullptr, workspace, hard_query_mode, stream);
}

MVT_EXPORT int mvt_debug_set_group_width(int lanes) {
    MVT_CHECK_ARG(lanes == 16 || lanes == 32, "mvt_debug_set_group_width: 16 or 32 (got %d)", lanes);
    mvt::g_group_width = lanes;
    return MVT_OK;
}

MVT_EXPORT int mvt_knn_gather(const mvt_search *search, const void *fvec, int feat_dtype, int B, int C,
                              const float *query, int M, void *feat_out, int *idx_out, void *stream) {
    using namespace mvt;
    MVT_CHECK_ARG(search && fvec && query && feat_out, "mvt_knn_gather: null pointer");
    MVT_CHECK_ARG(B > 0 && M >= 0 && C > 0, "mvt_knn_gather: bad shape B=%d M=%d C=%d", B, M, C);
    MVT_CHECK_ARG(feat_dtype == MVT_F32 || feat_dtype == MVT_BF16, "mvt_knn_gather: feat_dtype %d", feat_dtype);
    if (M == 0) return MVT_OK;
    SearchDev sd;
    int rc = make_search(search, B, 0, sd, "mvt_knn_gather");
    if (rc) return rc;
    long long warps = (long long)B * M;
    knn_gather_kernel<<<warp_blocks(warps), 256, 0, (cudaStream_t)stream>>>(sd, fvec, feat_dtype == MVT_F32 ? 4 : 2, B, M, C, query,
                                                                          feat_out, idx_out);
    MVT_CHECK_LAUNCH("knn_gather");
    return MVT_OK;
}
