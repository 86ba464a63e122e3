// kNN query kernels: plain kNN, fused kNN + gather + correlation + offset encoding, and k=1 feature gather.
// Replaces knn / PointcloudCorrBlock.corr_sample / the feature-init block of the reference
// (mvtracker/models/core/mvtracker/mvtracker.py:26-90, 800-846, 627-643).
//
// Roofline of knn_corr: HBM bound on the neighbour-feature gather.  Algorithmic bytes per query (k=16, C=128, fp32):
//   read 12 (xyz) + 512 (target) + 16*(512 + 12) (neighbour rows + xyz), write 16*16  = 9 164 B  (SURVEY.md 8d).
// The neighbour rows go HBM -> registers -> dot product; they are never written back.
#include "mvt_knn.cuh"

namespace mvt {

// ---------------------------------------------------------------------------------------------------------------
// plain kNN
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) knn_kernel(const GridView g, int B, int M, const float *__restrict__ query, int k,
                                                  int *__restrict__ idx32, long long *__restrict__ idx64,
                                                  float *__restrict__ dist) {
    const int lane = threadIdx.x & 31;
    const long long w = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (w >= (long long)B * M) return;
    const int b = (int)(w / M);
    const float qx = __ldg(query + w * 3), qy = __ldg(query + w * 3 + 1), qz = __ldg(query + w * 3 + 2);
    float bd; int bi;
    knn_search(g, b, qx, qy, qz, k, lane, bd, bi);
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
    GridView g;
    const float *xyz;
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
    long long row_stride;
    int *idx_out;
    float scale;      // sqrt(C/G)
    int vec_out;      // 1: every neighbour record is one aligned float4 (D == 4)
};

// MODE 0: generic fp32 (any C, G).   MODE 1: fp32, C == 128, G == 1, k <= 16 (one float4 of the row per lane).
// MODE 2: bf16, C == 128, G == 1, k <= 16 (16 lanes x 16 B per row, two rows per warp load).
template <int MODE>
__global__ void __launch_bounds__(256) knn_corr_kernel(const CorrParams p) {
    const int lane = threadIdx.x & 31;
    const long long per_level = (long long)p.B * p.M;
    const long long w = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (w >= per_level * p.n_levels) return;
    const int lvl = (int)(w / per_level);
    const long long q = w - (long long)lvl * per_level;   // b * M + m
    const int b = (int)(q / p.M);
    const CorrLevelDev &L = p.lv[lvl];
    const int k = L.k, C = p.C;
    const float qx = __ldg(p.coords + q * 3), qy = __ldg(p.coords + q * 3 + 1), qz = __ldg(p.coords + q * 3 + 2);

    // the target row is fetched before the search so its latency hides behind it
    float4 tg = make_float4(0.f, 0.f, 0.f, 0.f);
    uint4 tgh = make_uint4(0u, 0u, 0u, 0u);
    if (MODE == 1) tg = __ldg(reinterpret_cast<const float4 *>(p.targets) + q * 32 + lane);
    if (MODE == 2) tgh = __ldg(reinterpret_cast<const uint4 *>(p.targets) + q * 16 + (lane & 15));

    float bd; int bi;
    knn_search(L.g, b, qx, qy, qz, k, lane, bd, bi);

    if (p.idx_out && lane < k) p.idx_out[(w * k) + lane] = (bi != kNoIdx) ? bi : -1;

    const int D = p.G + 3 * p.add_offset + 3 * p.add_xyz;
    float *orow = p.out + q * p.row_stride + L.out_off;
    const size_t seg = (size_t)b * L.g.N;
    float corr_mine = 0.f;   // MODE 1/2: correlation of neighbour `lane`

    if (MODE == 1) {
        const float4 *fv = reinterpret_cast<const float4 *>(L.fvec);
        float pr[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            int ij = __shfl_sync(FULL, bi, j);
            float4 f = make_float4(0.f, 0.f, 0.f, 0.f);
            if (j < k && ij != kNoIdx) f = __ldg(fv + (seg + (size_t)ij) * 32 + lane);
            pr[j] = fmaf(tg.w, f.w, fmaf(tg.z, f.z, fmaf(tg.y, f.y, tg.x * f.x)));
        }
        // transpose-reduce: 16 partials x 32 lanes -> lane L holds the total of neighbour L>>1  (16 shuffles)
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            bool hi = lane & 16;
            float send = hi ? pr[i] : pr[i + 8], keep = hi ? pr[i + 8] : pr[i];
            pr[i] = keep + __shfl_xor_sync(FULL, send, 16);
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            bool hi = lane & 8;
            float send = hi ? pr[i] : pr[i + 4], keep = hi ? pr[i + 4] : pr[i];
            pr[i] = keep + __shfl_xor_sync(FULL, send, 8);
        }
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            bool hi = lane & 4;
            float send = hi ? pr[i] : pr[i + 2], keep = hi ? pr[i + 2] : pr[i];
            pr[i] = keep + __shfl_xor_sync(FULL, send, 4);
        }
        {
            bool hi = lane & 2;
            float send = hi ? pr[0] : pr[1], keep = hi ? pr[1] : pr[0];
            pr[0] = keep + __shfl_xor_sync(FULL, send, 2);
        }
        pr[0] += __shfl_xor_sync(FULL, pr[0], 1);
        corr_mine = __shfl_sync(FULL, pr[0], (2 * lane) & 31);
    } else if (MODE == 2) {
        const uint4 *fv = reinterpret_cast<const uint4 *>(L.fvec);
        const int half = lane >> 4, l16 = lane & 15;
        float pr[8];   // pr[i]: partial of neighbour 2i + half
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            int j = 2 * i + half;
            int ij = __shfl_sync(FULL, bi, j);
            uint4 f = make_uint4(0u, 0u, 0u, 0u);
            if (j < k && ij != kNoIdx) f = __ldg(fv + (seg + (size_t)ij) * 16 + l16);
            const __nv_bfloat162 *a = reinterpret_cast<const __nv_bfloat162 *>(&tgh);
            const __nv_bfloat162 *c = reinterpret_cast<const __nv_bfloat162 *>(&f);
            float acc = 0.f;
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                float2 x = __bfloat1622float2(a[e]), y = __bfloat1622float2(c[e]);
                acc = fmaf(x.x, y.x, acc);
                acc = fmaf(x.y, y.y, acc);
            }
            pr[i] = acc;
        }
        // reduce over the 16 lanes of each half: 8 partials -> lane (within half) l holds neighbour 2*(l>>1) + half
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            bool hi = lane & 8;
            float send = hi ? pr[i] : pr[i + 4], keep = hi ? pr[i + 4] : pr[i];
            pr[i] = keep + __shfl_xor_sync(FULL, send, 8);
        }
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            bool hi = lane & 4;
            float send = hi ? pr[i] : pr[i + 2], keep = hi ? pr[i + 2] : pr[i];
            pr[i] = keep + __shfl_xor_sync(FULL, send, 4);
        }
        {
            bool hi = lane & 2;
            float send = hi ? pr[0] : pr[1], keep = hi ? pr[1] : pr[0];
            pr[0] = keep + __shfl_xor_sync(FULL, send, 2);
        }
        pr[0] += __shfl_xor_sync(FULL, pr[0], 1);
        // neighbour j lives in half (j & 1), lanes 2*(j>>1), 2*(j>>1)+1 of that half
        corr_mine = __shfl_sync(FULL, pr[0], ((lane & 1) << 4) | ((lane >> 1) << 1));
    }

    if (MODE != 0) {
        if (lane < k) {
            bool ok = bi != kNoIdx;
            float nx = 0.f, ny = 0.f, nz = 0.f;
            if (ok) {
                const float *px = L.xyz + (seg + (size_t)bi) * 3;
                nx = __ldg(px); ny = __ldg(px + 1); nz = __ldg(px + 2);
            }
            float cv = ok ? __fdiv_rn(corr_mine, p.scale) : 0.f;
            float *o = orow + (size_t)lane * D;
            if (p.vec_out) {
                *reinterpret_cast<float4 *>(o) = ok ? make_float4(cv, nx - qx, ny - qy, nz - qz) : make_float4(0.f, 0.f, 0.f, 0.f);
            } else {
                int wpos = 0;
                o[wpos++] = cv;
                if (p.add_offset) { o[wpos++] = ok ? nx - qx : 0.f; o[wpos++] = ok ? ny - qy : 0.f; o[wpos++] = ok ? nz - qz : 0.f; }
                if (p.add_xyz) { o[wpos++] = nx; o[wpos++] = ny; o[wpos++] = nz; }
            }
        }
        return;
    }

    // ---- MODE 0: generic path, one neighbour and one group at a time ----
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
        if (lane < 3 && (p.add_offset || p.add_xyz)) {
            float pv = ok ? __ldg(L.xyz + (seg + (size_t)ij) * 3 + lane) : 0.f;
            float qv = lane == 0 ? qx : (lane == 1 ? qy : qz);
            int wpos = p.G;
            if (p.add_offset) { o[wpos + lane] = ok ? pv - qv : 0.f; wpos += 3; }
            if (p.add_xyz) o[wpos + lane] = pv;
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------
// k = 1 search + feature row copy (feature init)
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) knn_gather_kernel(const GridView g, const void *__restrict__ fvec, int elem_bytes,
                                                         int B, int M, int C, const float *__restrict__ query,
                                                         void *__restrict__ feat_out, int *__restrict__ idx_out) {
    const int lane = threadIdx.x & 31;
    const long long w = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (w >= (long long)B * M) return;
    const int b = (int)(w / M);
    const float qx = __ldg(query + w * 3), qy = __ldg(query + w * 3 + 1), qz = __ldg(query + w * 3 + 2);
    float bd; int bi;
    knn_search(g, b, qx, qy, qz, 1, lane, bd, bi);
    int i0 = __shfl_sync(FULL, bi, 0);
    if (idx_out && lane == 0) idx_out[w] = (i0 != kNoIdx) ? i0 : -1;
    const size_t row_bytes = (size_t)C * elem_bytes;
    const unsigned char *src = reinterpret_cast<const unsigned char *>(fvec) + ((size_t)b * g.N + (size_t)(i0 != kNoIdx ? i0 : 0)) * row_bytes;
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

}  // namespace mvt

// ---------------------------------------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------------------------------------
static inline unsigned warp_blocks(long long warps) { return (unsigned)((warps + 7) / 8); }

MVT_EXPORT int mvt_knn(const void *grid, int B, int N, const float *query, int M, int k,
                       void *idx_out, int idx_is_i64, float *dist_out, void *stream) {
    using namespace mvt;
    MVT_CHECK_ARG(grid && query, "mvt_knn: null pointer");
    MVT_CHECK_ARG(B > 0 && N > 0 && M >= 0, "mvt_knn: bad shape B=%d N=%d M=%d", B, N, M);
    MVT_CHECK_ARG(k >= 1 && k <= MVT_MAX_K, "mvt_knn: k must be 1..%d (got %d)", MVT_MAX_K, k);
    if (M == 0) return MVT_OK;
    long long warps = (long long)B * M;
    MVT_CHECK_ARG(warps / 8 < 0x7fffffffLL, "mvt_knn: too many queries");
    GridView g = grid_view(grid, B, N);
    knn_kernel<<<warp_blocks(warps), 256, 0, (cudaStream_t)stream>>>(
        g, B, M, query, k, idx_is_i64 ? nullptr : (int *)idx_out, idx_is_i64 ? (long long *)idx_out : nullptr, dist_out);
    MVT_CHECK_LAUNCH("knn");
    return MVT_OK;
}

MVT_EXPORT int mvt_knn_corr(const mvt_corr_level *levels, int n_levels, int B, int M, int C, int G, int feat_dtype,
                            const void *targets, const float *coords, int add_offset, int add_xyz,
                            float *out, long long out_row_stride, int *idx_out, void *stream) {
    using namespace mvt;
    MVT_CHECK_ARG(levels && targets && coords && out, "mvt_knn_corr: null pointer");
    MVT_CHECK_ARG(n_levels >= 1 && n_levels <= MVT_MAX_LEVELS, "mvt_knn_corr: n_levels must be 1..%d (got %d)", MVT_MAX_LEVELS, n_levels);
    MVT_CHECK_ARG(B > 0 && M >= 0 && C > 0, "mvt_knn_corr: bad shape B=%d M=%d C=%d", B, M, C);
    MVT_CHECK_ARG(G >= 1 && G <= C && C % G == 0, "mvt_knn_corr: groups=%d must divide C=%d", G, C);   // mvtracker.py:786-787
    MVT_CHECK_ARG(feat_dtype == MVT_F32 || feat_dtype == MVT_BF16, "mvt_knn_corr: feat_dtype %d", feat_dtype);
    if (M == 0) return MVT_OK;
    CorrParams p;
    p.n_levels = n_levels; p.B = B; p.M = M; p.C = C; p.G = G;
    p.add_offset = add_offset ? 1 : 0; p.add_xyz = add_xyz ? 1 : 0;
    p.targets = targets; p.coords = coords; p.out = out; p.row_stride = out_row_stride; p.idx_out = idx_out;
    p.scale = sqrtf((float)C / (float)G);
    const int D = G + 3 * p.add_offset + 3 * p.add_xyz;
    bool fast = (C == 128 && G == 1);
    bool vec = (D == 4) && (((uintptr_t)out & 15) == 0) && (out_row_stride % 4 == 0);
    bool aligned = (((uintptr_t)targets & 15) == 0);
    for (int l = 0; l < n_levels; ++l) {
        const mvt_corr_level &u = levels[l];
        MVT_CHECK_ARG(u.grid && u.xyz && u.fvec, "mvt_knn_corr: level %d has a null pointer", l);
        MVT_CHECK_ARG(u.N > 0 && u.k >= 1 && u.k <= MVT_MAX_K, "mvt_knn_corr: level %d bad N=%d k=%d", l, u.N, u.k);
        MVT_CHECK_ARG(u.k <= u.N, "mvt_knn_corr: k should be less than or equal to N (level %d: k=%d N=%d)", l, u.k, u.N);   // mvtracker.py:785
        MVT_CHECK_ARG(idx_out == nullptr || u.k == levels[0].k, "mvt_knn_corr: idx_out needs the same k on every level");
        p.lv[l].g = grid_view(u.grid, B, u.N);
        p.lv[l].xyz = u.xyz; p.lv[l].fvec = u.fvec; p.lv[l].k = u.k; p.lv[l].out_off = u.out_offset;
        fast = fast && u.k <= 16;
        vec = vec && (u.out_offset % 4 == 0);
        aligned = aligned && (((uintptr_t)u.fvec & 15) == 0);
    }
    for (int l = n_levels; l < MVT_MAX_LEVELS; ++l) p.lv[l] = p.lv[0];
    p.vec_out = vec ? 1 : 0;
    MVT_CHECK_ARG(feat_dtype == MVT_F32 || (fast && aligned), "mvt_knn_corr: bf16 features need C=128, groups=1, k<=16, 16-byte aligned rows");
    long long warps = (long long)B * M * n_levels;
    MVT_CHECK_ARG(warps / 8 < 0x7fffffffLL, "mvt_knn_corr: too many queries");
    cudaStream_t st = (cudaStream_t)stream;
    if (feat_dtype == MVT_BF16) knn_corr_kernel<2><<<warp_blocks(warps), 256, 0, st>>>(p);
    else if (fast && aligned) knn_corr_kernel<1><<<warp_blocks(warps), 256, 0, st>>>(p);
    else knn_corr_kernel<0><<<warp_blocks(warps), 256, 0, st>>>(p);
    MVT_CHECK_LAUNCH("knn_corr");
    return MVT_OK;
}

MVT_EXPORT int mvt_knn_gather(const void *grid, const void *fvec, int feat_dtype, int B, int N, int C,
                              const float *query, int M, void *feat_out, int *idx_out, void *stream) {
    using namespace mvt;
    MVT_CHECK_ARG(grid && fvec && query && feat_out, "mvt_knn_gather: null pointer");
    MVT_CHECK_ARG(B > 0 && N > 0 && M >= 0 && C > 0, "mvt_knn_gather: bad shape B=%d N=%d M=%d C=%d", B, N, M, C);
    MVT_CHECK_ARG(feat_dtype == MVT_F32 || feat_dtype == MVT_BF16, "mvt_knn_gather: feat_dtype %d", feat_dtype);
    if (M == 0) return MVT_OK;
    long long warps = (long long)B * M;
    GridView g = grid_view(grid, B, N);
    knn_gather_kernel<<<warp_blocks(warps), 256, 0, (cudaStream_t)stream>>>(g, fvec, feat_dtype == MVT_F32 ? 4 : 2, B, M, C, query,
                                                                          feat_out, idx_out);
    MVT_CHECK_LAUNCH("knn_gather");
    return MVT_OK;
}
