// Backward of the hot path: what torch autograd computes through the reference's
// PointcloudCorrBlock.corr_sample (mvtracker/models/core/mvtracker/mvtracker.py:827-835: gather + einsum / sqrt(C/G)) and
// through init_pointcloud_from_rgbd's feature branch (mvtracker/models/core/model_utils.py:439-441 chained avg_pool2d,
// :475-476 permute + reshape), as used by training (train.py:119-163).  Neighbour indices, positions and offsets carry no
// gradient (coords are detached, mvtracker.py:351; kNN is not differentiable).
//
//   corr_backward_kernel     grad_targets[q, c] (+)= sum_j gs_j * fvec[idx_j, c],   grad_fvec[idx_j, c] += gs_j * targets[q, c],
//                            gs_j = grad_out[q, j, group(c)] / sqrt(C/G).  One warp per query; neighbour rows are gathered once
//                            (as in the forward kernel) and the scatter is one 16-byte vector reduction per lane and
//                            neighbour (red.global.add.v4.f32).  Bound: HBM; algorithmic bytes per query (k = 16, C = 128, fp32)
//                            = 512 (targets) + 16*512 (rows) + 16*512 (reductions) + 512 (grad_targets) + 64 + 64.
//   feature_unpool_kernel    grad_fmaps[v, s, c, y, x] = grad_fvec_l[s, v*H_l*W_l + (y >> l)*W_l + (x >> l), c] / 4^l
//                            (0 outside the pooled area): point-major rows -> channel-major planes through a padded
//                            shared-memory tile, both sides coalesced.  Bound: HBM, (4 + 4 / 4^l) bytes per output element.
#include "mvt_common.cuh"

namespace mvt {

struct CorrBwdParams {
    const void *fvec;          // [B, N, C]
    const void *targets;       // [B, M, C]
    const int *idx;            // [B, M, k] (-1 = no neighbour)
    const float *grad_out;     // entry (b, m, j, g) at b*frame_stride + m*row_stride + j*D + g
    float *grad_targets;       // [B, M, C] fp32
    float *grad_fvec;          // [B, N, C] fp32, accumulated into
    long long row_stride, frame_stride;
    int B, N, M, C, G, k, D;
    float scale;               // sqrt(C / G)
    int accumulate;            // 1: grad_targets += (a further level of the same targets)
};

__device__ __forceinline__ float4 load4(const float *p) { return __ldg(reinterpret_cast<const float4 *>(p)); }
__device__ __forceinline__ float4 load4(const __nv_bfloat16 *p) {
    const uint2 r = __ldg(reinterpret_cast<const uint2 *>(p));
    const float2 a = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162 *>(&r.x));
    const float2 b = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162 *>(&r.y));
    return make_float4(a.x, a.y, b.x, b.y);
}

template <typename T>
__global__ void __launch_bounds__(256) corr_backward_kernel(const CorrBwdParams p) {
    const int lane = threadIdx.x & 31;
    const long long q = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (q >= (long long)p.B * p.M) return;
    const int b = (int)(q / p.M), m = (int)(q - (long long)b * p.M);
    const int C = p.C, Cg = C / p.G;
    const T *tg = reinterpret_cast<const T *>(p.targets) + q * C;
    const T *fv = reinterpret_cast<const T *>(p.fvec) + (size_t)b * p.N * C;
    float *gf = p.grad_fvec + (size_t)b * p.N * C;
    float *gt = p.grad_targets + q * C;
    const float *go = p.grad_out + (long long)b * p.frame_stride + (long long)m * p.row_stride;
    const int my_idx = lane < p.k ? __ldg(p.idx + q * p.k + lane) : -1;
    for (int c0 = lane * 4; c0 < C; c0 += 128) {
        const int gi = c0 / Cg;
        const float4 t = load4(tg + c0);
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int j = 0; j < p.k; ++j) {
            const int ij = __shfl_sync(0xffffffffu, my_idx, j);
            if (ij < 0) continue;                                     // (uniform)
            const float gs = __fdiv_rn(__ldg(go + (size_t)j * p.D + gi), p.scale);
            const float4 f = load4(fv + (size_t)ij * C + c0);
            acc.x = fmaf(gs, f.x, acc.x); acc.y = fmaf(gs, f.y, acc.y); acc.z = fmaf(gs, f.z, acc.z); acc.w = fmaf(gs, f.w, acc.w);
            atomicAdd(reinterpret_cast<float4 *>(gf + (size_t)ij * C + c0), make_float4(gs * t.x, gs * t.y, gs * t.z, gs * t.w));
        }
        float4 *o = reinterpret_cast<float4 *>(gt + c0);
        if (p.accumulate) {
            const float4 old = *o;
            acc.x += old.x; acc.y += old.y; acc.z += old.z; acc.w += old.w;
        }
        *o = acc;
    }
}

struct UnpoolParams {
    const float *grad_fvec;    // [S, V*Hl*Wl, C] fp32
    float *grad_fmaps;         // [V, S, C, H, W] fp32
    int V, S, C, H, W, level, Hl, Wl;
    float weight;              // 1 / 4^level
    int accumulate;
};

// block: 32 level-0 pixels of one row y of one (v, s) plane x all channels, 128 channels per barrier pair (16 loads in
// flight per thread); grid (ceil(W/32), H, V*S)
constexpr int kUnpoolCC = 128;
__global__ void __launch_bounds__(256) feature_unpool_kernel(const UnpoolParams p) {
    __shared__ float tile[kUnpoolCC][33];                        // [channel][pixel]
    const int x0 = blockIdx.x * 32, y = blockIdx.y, vs = blockIdx.z;
    const int v = vs / p.S, s = vs - v * p.S;
    const int l = p.level;
    const int yl = y >> l;
    const bool row_in = yl < p.Hl;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;      // 8 warps
    const size_t frame = ((size_t)s * p.V + v) * (size_t)p.Hl * p.Wl;
    for (int c0 = 0; c0 < p.C; c0 += kUnpoolCC) {
        // load: warp ty takes pixels ty, ty+8, ...; a pixel's 128 channels = 4 coalesced 128-byte pieces
#pragma unroll
        for (int ii = 0; ii < 4; ++ii) {
            const int i = ty + 8 * ii, x = x0 + i, xl = x >> l;
            const bool in = row_in && x < p.W && xl < p.Wl;
            const float *src = p.grad_fvec + (frame + (size_t)yl * p.Wl + (in ? xl : 0)) * p.C + c0;
#pragma unroll
            for (int cc = 0; cc < kUnpoolCC; cc += 32) {
                float g = 0.f;
                if (in && c0 + cc + tx < p.C) g = __ldg(src + cc + tx) * p.weight;
                tile[cc + tx][i] = g;
            }
        }
        __syncthreads();
        // store: warp ty takes channels ty, ty+8, ...: 32 consecutive pixels of row y per warp instruction
        const int x = x0 + tx;
#pragma unroll 4
        for (int i = ty; i < kUnpoolCC; i += 8) {
            const int c = c0 + i;
            if (c < p.C && x < p.W) {
                float *o = p.grad_fmaps + ((((size_t)v * p.S + s) * p.C + c) * p.H + y) * p.W + x;
                *o = p.accumulate ? *o + tile[i][tx] : tile[i][tx];
            }
        }
        __syncthreads();
    }
}

}  // namespace mvt

MVT_EXPORT int mvt_knn_corr_backward(const void *fvec, const void *targets, int feat_dtype, const int *idx, const float *grad_out,
                                     long long grad_row_stride, long long grad_frame_stride, int B, int N, int M, int C, int G,
                                     int k, int D, float *grad_targets, int accumulate_targets, float *grad_fvec, void *stream) {
    using namespace mvt;
    MVT_CHECK_ARG(fvec && targets && idx && grad_out && grad_targets && grad_fvec, "mvt_knn_corr_backward: null pointer");
    MVT_CHECK_ARG(B > 0 && N > 0 && M > 0 && k > 0 && k <= MVT_MAX_K, "mvt_knn_corr_backward: bad sizes B=%d N=%d M=%d k=%d", B, N, M, k);
    MVT_CHECK_ARG(G > 0 && C % G == 0 && C % 4 == 0 && (C / G) % 4 == 0,
                  "mvt_knn_corr_backward: C=%d, G=%d: channels per group must be a multiple of 4", C, G);
    MVT_CHECK_ARG(D >= G, "mvt_knn_corr_backward: D=%d < G=%d", D, G);
    MVT_CHECK_ARG(feat_dtype == MVT_F32 || feat_dtype == MVT_BF16, "mvt_knn_corr_backward: feat_dtype %d", feat_dtype);
    CorrBwdParams p;
    p.fvec = fvec; p.targets = targets; p.idx = idx; p.grad_out = grad_out; p.grad_targets = grad_targets; p.grad_fvec = grad_fvec;
    p.row_stride = grad_row_stride; p.frame_stride = grad_frame_stride;
    p.B = B; p.N = N; p.M = M; p.C = C; p.G = G; p.k = k; p.D = D;
    p.scale = sqrtf((float)(C / G));
    p.accumulate = accumulate_targets;
    const long long Q = (long long)B * M;
    const unsigned grid = (unsigned)((Q + 7) / 8);
    if (feat_dtype == MVT_F32) corr_backward_kernel<float><<<grid, 256, 0, (cudaStream_t)stream>>>(p);
    else corr_backward_kernel<__nv_bfloat16><<<grid, 256, 0, (cudaStream_t)stream>>>(p);
    MVT_CHECK_LAUNCH("corr_backward_kernel");
    return MVT_OK;
}

MVT_EXPORT int mvt_cloud_build_backward(const float *grad_fvec, int V, int S, int C, int H, int W, int level, float *grad_fmaps,
                                        int accumulate, void *stream) {
    using namespace mvt;
    MVT_CHECK_ARG(grad_fvec && grad_fmaps, "mvt_cloud_build_backward: null pointer");
    MVT_CHECK_ARG(V > 0 && S > 0 && C > 0 && H > 0 && W > 0 && level >= 0 && level < MVT_MAX_LEVELS,
                  "mvt_cloud_build_backward: bad sizes V=%d S=%d C=%d H=%d W=%d level=%d", V, S, C, H, W, level);
    MVT_CHECK_ARG((H >> level) > 0 && (W >> level) > 0, "mvt_cloud_build_backward: level %d of a %dx%d map is empty", level, H, W);
    MVT_CHECK_ARG(H <= 65535 && (long long)V * S <= 65535, "mvt_cloud_build_backward: grid too large (H=%d, V*S=%d)", H, V * S);
    UnpoolParams p;
    p.grad_fvec = grad_fvec; p.grad_fmaps = grad_fmaps;
    p.V = V; p.S = S; p.C = C; p.H = H; p.W = W; p.level = level; p.Hl = H >> level; p.Wl = W >> level;
    p.weight = 1.0f / (float)(1 << (2 * level));
    p.accumulate = accumulate;
    dim3 grid((unsigned)((W + 31) / 32), (unsigned)H, (unsigned)(V * S));
    feature_unpool_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(p);
    MVT_CHECK_LAUNCH("feature_unpool_kernel");
    return MVT_OK;
}
