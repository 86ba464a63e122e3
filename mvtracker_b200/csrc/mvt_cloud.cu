// Cloud build: fused depth unprojection + camera transform + multi-view fusion + feature pyramid.
// Replaces init_pointcloud_from_rgbd (reference mvtracker/models/core/model_utils.py:420-482): one pass over
// fmaps[V,S,C,H,W] emits every requested pyramid level's point-major xyz / feature records.
//
// Roofline: pure streaming, HBM bound.  Algorithmic bytes per level-0 pixel (fp32, all 4 levels):
//   read C*4 + 4, write (C*4 + 12 [+1 valid]) * (1 + 1/4 + 1/16 + 1/64)   
#include "mvt_common.cuh"

namespace mvt {

// ---------------------------------------------------------------------------------------------------------------
// camera_prep: fp64 inverse of K (3x3) and [R|t;0001] (affine), rounded to fp32.  One thread per (view, frame).
// ---------------------------------------------------------------------------------------------------------------
__device__ inline void inv3(const double *m, double *o) {
    double c00 = m[4] * m[8] - m[5] * m[7], c01 = m[5] * m[6] - m[3] * m[8], c02 = m[3] * m[7] - m[4] * m[6];
    double det = m[0] * c00 + m[1] * c01 + m[2] * c02;
    double id = 1.0 / det;
    o[0] = c00 * id; o[1] = (m[2] * m[7] - m[1] * m[8]) * id; o[2] = (m[1] * m[5] - m[2] * m[4]) * id;
    o[3] = c01 * id; o[4] = (m[0] * m[8] - m[2] * m[6]) * id; o[5] = (m[2] * m[3] - m[0] * m[5]) * id;
    o[6] = c02 * id; o[7] = (m[1] * m[6] - m[0] * m[7]) * id; o[8] = (m[0] * m[4] - m[1] * m[3]) * id;
}

__global__ void camera_prep_kernel(const float *__restrict__ intrs, const float *__restrict__ extrs, int n,
                                   float *__restrict__ cam) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double K[9], Ki[9], A[9], Ai[9], t[3];
    for (int j = 0; j < 9; ++j) K[j] = (double)intrs[i * 9 + j];
    for (int r = 0; r < 3; ++r) {
        for (int c = 0; c < 3; ++c) A[r * 3 + c] = (double)extrs[i * 12 + r * 4 + c];
        t[r] = (double)extrs[i * 12 + r * 4 + 3];
    }
    inv3(K, Ki);
    inv3(A, Ai);
    float *o = cam + (size_t)i * kCamStride;
    for (int j = 0; j < 9; ++j) o[j] = (float)Ki[j];
    for (int r = 0; r < 3; ++r) {
        for (int c = 0; c < 3; ++c) o[9 + r * 4 + c] = (float)Ai[r * 3 + c];
        o[9 + r * 4 + 3] = (float)(-(Ai[r * 3 + 0] * t[0] + Ai[r * 3 + 1] * t[1] + Ai[r * 3 + 2] * t[2]));
    }
    o[21] = o[22] = o[23] = 0.f;
    // forward camera for the raster search: R (9), t (3), fx, skew, cx, fy, cy, "K is upper triangular with K22 == 1"
    for (int r = 0; r < 3; ++r) {
        for (int c = 0; c < 3; ++c) o[24 + r * 3 + c] = extrs[i * 12 + r * 4 + c];
        o[33 + r] = extrs[i * 12 + r * 4 + 3];
    }
    o[36] = intrs[i * 9 + 0]; o[37] = intrs[i * 9 + 1]; o[38] = intrs[i * 9 + 2];
    o[39] = intrs[i * 9 + 4]; o[40] = intrs[i * 9 + 5];
    bool ok = intrs[i * 9 + 3] == 0.f && intrs[i * 9 + 6] == 0.f && intrs[i * 9 + 7] == 0.f && intrs[i * 9 + 8] == 1.f &&
              intrs[i * 9 + 0] > 0.f && intrs[i * 9 + 4] > 0.f;
    o[41] = ok ? 1.f : 0.f;
    // [42]: an upper bound of the largest singular value of the 3x3 block of extrs (1 for a rigid transform; the
    // reference also feeds similarity-scaled blocks, datasets/utils.py:403-415): the raster search projects a WORLD ball
    // of radius r as a VIEW-space ball of radius [42] * r.  sqrt of the Gershgorin bound of A^T A, in fp64, rounded up.
    double smax2 = 0.0;
    for (int r = 0; r < 3; ++r) {
        double row = 0.0;
        for (int c = 0; c < 3; ++c) row += fabs(A[0 * 3 + r] * A[0 * 3 + c] + A[1 * 3 + r] * A[1 * 3 + c] + A[2 * 3 + r] * A[2 * 3 + c]);
        smax2 = fmax(smax2, row);
    }
    const double smax = sqrt(smax2) * (1.0 + 1e-6);
    o[42] = (smax == smax && smax < 3.0e38) ? __double2float_ru(smax) : 3.0e38f;
    for (int j = 43; j < kCamStride; ++j) o[j] = 0.f;
}

// ---------------------------------------------------------------------------------------------------------------
// cloud_build
// ---------------------------------------------------------------------------------------------------------------
#ifndef MVT_CLOUD_BATCH
#define MVT_CLOUD_BATCH 32
#endif
#ifndef MVT_CLOUD_MINB
#define MVT_CLOUD_MINB 4
#endif
#ifndef MVT_CLOUD_PACKED
#define MVT_CLOUD_PACKED 1      // bf16 features: shared-memory tiles of channel PAIRS, 64-channel chunks (0: the fp32-tile path)
#endif
constexpr int TH = 8;     // tile rows   (level-0 pixels) = one level-3 pixel high
constexpr int TW = 32;    // tile cols   (level-0 pixels) = four level-3 pixels wide
constexpr int CC = 32;    // channels per smem chunk
constexpr int S0 = TH * TW + 1, S1 = 64 + 1, S2 = 16 + 1, S3 = 4 + 1;   // padded channel strides (== 1 mod 32 or odd)

struct CloudParams {
    const void *fmaps;
    const float *depths;
    const float *cam;
    int V, S, C, H, W;
    float stride;
    int depth_avg;
    int level_mask;
    int tiles_x;
    float *xyz[4];
    float4 *xyz4[4];
    void *fvec[4];
    unsigned char *valid[4];
    float *spacing[4];
    int *neg[4];
    PeersDev rec[4];       // replicated record buffers [S, V_total*H_l*W_l] of every rank (n == 0: not shared)
    int view_off, V_total;
    long long fm_vstride, dp_vstride;   // elements between consecutive views of fmaps / depths (dense: S*C*H*W, S*H*W)
    int cam_vstride;                    // camera records between consecutive views (dense: S)
};

template <typename T> __device__ __forceinline__ float to_f(T v);
template <> __device__ __forceinline__ float to_f<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_f<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <typename T> __device__ __forceinline__ T from_f(float v);
template <> __device__ __forceinline__ float from_f<float>(float v) { return v; }
template <> __device__ __forceinline__ __nv_bfloat16 from_f<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }
template <typename T> __device__ __forceinline__ float round_t(float v) { return to_f<T>(from_f<T>(v)); }

// 2x2 mean in the order of avg_pool2d's window loop: ((a+b)+c)+d, then /4 (exact scaling).
__device__ __forceinline__ float mean4(float a, float b, float c, float d) { return (((a + b) + c) + d) * 0.25f; }

// Store one level's tile (npx pixels x CC channels, smem layout tile[c*cs + px]) to the point-major feature array.
// Vector path: 8 lanes x 16 B cover one pixel's 32-channel chunk (128 B contiguous), 4 pixels per warp instruction.
template <typename T, int LOG_TW>
__device__ __forceinline__ void store_tile(const float *tile, int cs, int npx, T *out, size_t frame_base, int Wl, int Hl,
                                           int gx0, int gy0, int C, int c0, bool vec) {
    const int tid = threadIdx.x;
    if (vec) {
        const int c4 = tid & 7;
        for (int px = tid >> 3; px < npx; px += 32) {
            int ly = px >> LOG_TW, lx = px & ((1 << LOG_TW) - 1);
            int gy = gy0 + ly, gx = gx0 + lx;
            if (gy < Hl && gx < Wl && c0 + 4 * c4 < C) {
                float v0 = tile[(4 * c4 + 0) * cs + px], v1 = tile[(4 * c4 + 1) * cs + px];
                float v2 = tile[(4 * c4 + 2) * cs + px], v3 = tile[(4 * c4 + 3) * cs + px];
                T *dst = out + (frame_base + (size_t)gy * Wl + gx) * C + c0 + 4 * c4;
                if (sizeof(T) == 4) {
                    *reinterpret_cast<float4 *>(dst) = make_float4(v0, v1, v2, v3);
                } else {
                    __nv_bfloat162 lo = __floats2bfloat162_rn(v0, v1), hi = __floats2bfloat162_rn(v2, v3);
                    uint2 pk;
                    pk.x = *reinterpret_cast<unsigned *>(&lo);
                    pk.y = *reinterpret_cast<unsigned *>(&hi);
                    *reinterpret_cast<uint2 *>(dst) = pk;
                }
            }
        }
    } else {
        const int c = tid & 31;
        for (int px = tid >> 5; px < npx; px += 8) {
            int ly = px >> LOG_TW, lx = px & ((1 << LOG_TW) - 1);
            int gy = gy0 + ly, gx = gx0 + lx;
            if (gy < Hl && gx < Wl && c0 + c < C)
                out[(frame_base + (size_t)gy * Wl + gx) * C + c0 + c] = from_f<T>(tile[c * cs + px]);
        }
    }
}

// Reduce the 2x2 blocks of `src` (sw pixels wide, channel stride ss) into `dst` (dw = sw/2 wide, dnpx pixels).
template <typename T>
__device__ __forceinline__ void pool_tile(const float *src, int ss, int sw, float *dst, int ds, int dw, int dnpx) {
    const int c = threadIdx.x & 31;
    for (int px = threadIdx.x >> 5; px < dnpx; px += 8) {
        int y = px / dw, x = px - y * dw;
        const float *s = src + c * ss + (2 * y) * sw + 2 * x;
        dst[c * ds + px] = round_t<T>(mean4(s[0], s[1], s[sw], s[sw + 1]));
    }
}

// ---- 2-byte features, packed: the shared-memory tiles hold bf16 PAIRS (channels 2i, 2i+1 of one pixel in one 32-bit word), so a
// chunk is 64 channels in the space of 32 fp32 ones: half the chunk iterations (barriers, pooling passes), half the shared-memory
// stores per loaded byte, and a pixel's chunk is 128 contiguous bytes on the store side (8 lanes x 16 B) as in the fp32 path.
// Values are bf16 at every level, exactly as chained avg_pool2d on a bf16 tensor rounds them (level 0 is a bit copy).
__device__ __forceinline__ unsigned pool_pair(unsigned s0, unsigned s1, unsigned s2, unsigned s3) {
    const float a = mean4(__uint_as_float(s0 << 16), __uint_as_float(s1 << 16), __uint_as_float(s2 << 16), __uint_as_float(s3 << 16));
    const float b = mean4(__uint_as_float(s0 & 0xffff0000u), __uint_as_float(s1 & 0xffff0000u), __uint_as_float(s2 & 0xffff0000u),
                          __uint_as_float(s3 & 0xffff0000u));
    const __nv_bfloat162 r = __floats2bfloat162_rn(a, b);       // .x (low half) = a
    return *reinterpret_cast<const unsigned *>(&r);
}
__device__ __forceinline__ void pool_tile_packed(const unsigned *src, int ss, int sw, unsigned *dst, int ds, int dw, int dnpx) {
    const int c = threadIdx.x & 31;      // channel pair
    for (int px = threadIdx.x >> 5; px < dnpx; px += 8) {
        int y = px / dw, x = px - y * dw;
        const unsigned *s = src + c * ss + (2 * y) * sw + 2 * x;
        dst[c * ds + px] = pool_pair(s[0], s[1], s[sw], s[sw + 1]);
    }
}
// tile[cp * cs + px], 32 channel pairs -> out rows: 8 lanes x 16 B (4 pairs each) = the pixel's 64-channel chunk, 128 B contiguous
template <int LOG_TW>
__device__ __forceinline__ void store_tile_packed(const unsigned *tile, int cs, int npx, __nv_bfloat16 *out, size_t frame_base, int Wl,
                                                  int Hl, int gx0, int gy0, int C, int c0) {
    const int tid = threadIdx.x, c4 = tid & 7;
    for (int px = tid >> 3; px < npx; px += 32) {
        int ly = px >> LOG_TW, lx = px & ((1 << LOG_TW) - 1);
        int gy = gy0 + ly, gx = gx0 + lx;
        if (gy < Hl && gx < Wl) {
            uint4 w;
            w.x = tile[(4 * c4 + 0) * cs + px]; w.y = tile[(4 * c4 + 1) * cs + px];
            w.z = tile[(4 * c4 + 2) * cs + px]; w.w = tile[(4 * c4 + 3) * cs + px];
            *reinterpret_cast<uint4 *>(out + (frame_base + (size_t)gy * Wl + gx) * C + c0 + 8 * c4) = w;
        }
    }
}

__device__ __forceinline__ void emit_point(const CloudParams &p, const float *cm, int lvl, int s, int v, int gy, int gx,
                                           int Hl, int Wl, float depth, float &sp_sum, float &sp_cnt) {
    float st = p.stride * (float)(1 << lvl);
    float u = ((float)gx + 0.5f) * st - 0.5f;     // model_utils.py:462-466
    float w = ((float)gy + 0.5f) * st - 0.5f;
    float rx = fmaf(cm[0], u, fmaf(cm[1], w, cm[2]));
    float ry = fmaf(cm[3], u, fmaf(cm[4], w, cm[5]));
    float rz = fmaf(cm[6], u, fmaf(cm[7], w, cm[8]));
    float cx = rx * depth, cy = ry * depth, cz = rz * depth;           // :470
    float X = fmaf(cm[9], cx, fmaf(cm[10], cy, fmaf(cm[11], cz, cm[12])));   // :472 (w == 1 exactly: affine inverse)
    float Y = fmaf(cm[13], cx, fmaf(cm[14], cy, fmaf(cm[15], cz, cm[16])));
    float Z = fmaf(cm[17], cx, fmaf(cm[18], cy, fmaf(cm[19], cz, cm[20])));
    size_t P = (size_t)p.V * Hl * Wl;
    size_t pt = (size_t)s * P + (size_t)v * Hl * Wl + (size_t)gy * Wl + gx;
    if (p.xyz[lvl]) {
        float *o = p.xyz[lvl] + pt * 3;
        o[0] = X; o[1] = Y; o[2] = Z;
    }
    if (p.xyz4[lvl]) p.xyz4[lvl][pt] = make_float4(X, Y, Z, depth);
    if (p.rec[lvl].n)      // the search records of this view, straight into every rank's copy of the fused cloud
        store_all(p.rec[lvl], ((size_t)s * p.V_total + (size_t)(p.view_off + v)) * Hl * Wl + (size_t)gy * Wl + gx,
                  make_float4(X, Y, Z, depth));
    bool ok = depth > 0.f;
    if (depth < 0.f && p.neg[lvl]) p.neg[lvl][s] = 1;      // (benign race: every writer stores the same value)
    if (p.valid[lvl]) p.valid[lvl][pt] = ok ? 1 : 0;
    if (ok && depth < 3.0e38f) {
        sp_sum += depth * st * fabsf(cm[0]);
        sp_cnt += 1.f;
    }
}

template <typename T>
__global__ void __launch_bounds__(256, MVT_CLOUD_MINB) cloud_build_kernel(const CloudParams p) {
    __shared__ float t0[CC * S0];
    __shared__ float t1[CC * S1];
    __shared__ float t2[CC * S2];
    __shared__ float t3[CC * S3];
    __shared__ float dep[TH * TW + 64 + 16 + 4];
    __shared__ float cm[24];
    __shared__ float red[8][8];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int vs = blockIdx.y;                 // v * S + s
    const int v = vs / p.S, s = vs - v * p.S;
    const int ty = blockIdx.x / p.tiles_x, tx = blockIdx.x - ty * p.tiles_x;
    const int y0 = ty * TH, x0 = tx * TW;
    const int H = p.H, W = p.W, C = p.C;
    const size_t HW = (size_t)H * W;

    if (tid < 24) cm[tid] = p.cam[((size_t)v * p.cam_vstride + s) * kCamStride + tid];
    // depth pyramid of the tile
    {
        int yy = tid >> 5, xx = tid & 31;
        int gy = y0 + yy, gx = x0 + xx;
        dep[tid] = (gy < H && gx < W) ? p.depths[(size_t)v * p.dp_vstride + (size_t)s * HW + (size_t)gy * W + gx] : 0.f;
    }
    __syncthreads();
    float *d1 = dep + 256, *d2 = d1 + 64, *d3 = d2 + 16;
    if (tid < 64) {
        int y = tid >> 4, x = tid & 15;
        const float *q = dep + (2 * y) * 32 + 2 * x;
        d1[tid] = p.depth_avg ? mean4(q[0], q[1], q[32], q[33]) : q[0];
    }
    __syncthreads();
    if (tid < 16) {
        int y = tid >> 3, x = tid & 7;
        const float *q = d1 + (2 * y) * 16 + 2 * x;
        d2[tid] = p.depth_avg ? mean4(q[0], q[1], q[16], q[17]) : q[0];
    }
    __syncthreads();
    if (tid < 4) {
        const float *q = d2 + 2 * tid;
        d3[tid] = p.depth_avg ? mean4(q[0], q[1], q[8], q[9]) : q[0];
    }
    __syncthreads();

    // geometry: thread t emits level-l point t of the tile when t < 256 >> 2l
    float sp[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int l = 0; l < 4; ++l) {
        if (!((p.level_mask >> l) & 1)) continue;
        int Hl = H >> l, Wl = W >> l;
        int tw = TW >> l, npx = (TH >> l) * tw;
        if (tid < npx) {
            int ly = tid / tw, lx = tid - ly * tw;
            int gy = (y0 >> l) + ly, gx = (x0 >> l) + lx;
            if (gy < Hl && gx < Wl) {
                const float *dl = (l == 0) ? dep : (l == 1) ? d1 : (l == 2) ? d2 : d3;
                emit_point(p, cm, l, s, v, gy, gx, Hl, Wl, dl[tid], sp[2 * l], sp[2 * l + 1]);
            }
        }
    }
    if (p.spacing[0] || p.spacing[1] || p.spacing[2] || p.spacing[3]) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            float x = sp[i];
#pragma unroll
            for (int o = 16; o; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
            if (lane == 0) red[warp][i] = x;
        }
        __syncthreads();
        if (tid < 8) {
            float x = 0.f;
            for (int w2 = 0; w2 < 8; ++w2) x += red[w2][tid];
            int l = tid >> 1;
            if (p.spacing[l] && ((p.level_mask >> l) & 1) && x != 0.f) atomicAdd(p.spacing[l] + s * 2 + (tid & 1), x);
        }
    }

    // features: channel chunks through shared memory, transposed to point-major rows
    const bool any_f = p.fvec[0] || p.fvec[1] || p.fvec[2] || p.fvec[3];
    if (!any_f) return;
    const bool vec = (C % 4) == 0;
    const T *fm = reinterpret_cast<const T *>(p.fmaps) + (size_t)v * p.fm_vstride + (size_t)s * C * HW;
    const bool need1 = (p.level_mask & 0xE) != 0, need2 = (p.level_mask & 0xC) != 0, need3 = (p.level_mask & 0x8) != 0;
    const int yy = tid >> 5, xx = tid & 31;
    const bool inside = (y0 + yy < H) && (x0 + xx < W);
    const size_t pix = (size_t)(y0 + yy) * W + (x0 + xx);
    const bool pairs = ((W & 1) == 0) && ((reinterpret_cast<uintptr_t>(p.fmaps) & 3) == 0) && ((p.fm_vstride & 1) == 0);   // 4-byte pixel pairs are aligned

#if MVT_CLOUD_PACKED
    if (sizeof(T) == 2 && pairs && (C & 63) == 0 &&
        (((reinterpret_cast<uintptr_t>(p.fvec[0]) | reinterpret_cast<uintptr_t>(p.fvec[1]) | reinterpret_cast<uintptr_t>(p.fvec[2]) |
           reinterpret_cast<uintptr_t>(p.fvec[3])) & 15) == 0)) {
        unsigned *w0 = reinterpret_cast<unsigned *>(t0), *w1 = reinterpret_cast<unsigned *>(t1);
        unsigned *w2 = reinterpret_cast<unsigned *>(t2), *w3 = reinterpret_cast<unsigned *>(t3);
        __nv_bfloat16 *f0 = reinterpret_cast<__nv_bfloat16 *>(p.fvec[0]), *f1 = reinterpret_cast<__nv_bfloat16 *>(p.fvec[1]);
        __nv_bfloat16 *f2 = reinterpret_cast<__nv_bfloat16 *>(p.fvec[2]), *f3 = reinterpret_cast<__nv_bfloat16 *>(p.fvec[3]);
        const int q = tid >> 7, r2 = (tid >> 4) & 7, pp = tid & 15;
        const bool in2 = (y0 + r2 < H) && (x0 + 2 * pp < W);
        const size_t pix2 = (size_t)(y0 + r2) * W + (size_t)(x0 + 2 * pp);
        const bool swap = (tid >> 4) & 1;          // rows r2, r2+1 of a warp map to the same banks: alternate the order
        const __nv_bfloat16 *fmh = reinterpret_cast<const __nv_bfloat16 *>(fm);
        for (int c0 = 0; c0 < C; c0 += 2 * CC) {
            __syncthreads();   // previous chunk's readers of the tiles are done
            // thread (q, r2, pp): pixels (2pp, 2pp+1) of row r2, channel pairs q, q+2, ..., q+30: all 32 loads (4 B each: two pixels
            // of one channel) are in flight before the first shared-memory store
            unsigned val[CC];
#pragma unroll
            for (int i = 0; i < CC / 2; ++i) {
                const int ch = c0 + 2 * (2 * i + q);
                val[2 * i] = val[2 * i + 1] = 0u;
                if (in2) {
                    val[2 * i] = __ldg(reinterpret_cast<const unsigned *>(fmh + (size_t)ch * HW + pix2));
                    val[2 * i + 1] = __ldg(reinterpret_cast<const unsigned *>(fmh + (size_t)(ch + 1) * HW + pix2));
                }
            }
#pragma unroll
            for (int i = 0; i < CC / 2; ++i) {
                unsigned *d = w0 + (2 * i + q) * S0 + r2 * 32 + 2 * pp;
                const unsigned lo = __byte_perm(val[2 * i], val[2 * i + 1], 0x5410);     // pixel 2pp:   {ch, ch+1}
                const unsigned hi = __byte_perm(val[2 * i], val[2 * i + 1], 0x7632);     // pixel 2pp+1: {ch, ch+1}
                if (swap) { d[1] = hi; d[0] = lo; } else { d[0] = lo; d[1] = hi; }
            }
            __syncthreads();
            if (need1) pool_tile_packed(w0, S0, 32, w1, S1, 16, 64);
            if (f0 && (p.level_mask & 1)) store_tile_packed<5>(w0, S0, 256, f0, ((size_t)s * p.V + v) * HW, W, H, x0, y0, C, c0);
            if (!need1) continue;
            __syncthreads();
            if (need2) pool_tile_packed(w1, S1, 16, w2, S2, 8, 16);
            if (f1 && (p.level_mask & 2)) {
                int Hl = H >> 1, Wl = W >> 1;
                store_tile_packed<4>(w1, S1, 64, f1, ((size_t)s * p.V + v) * Hl * Wl, Wl, Hl, x0 >> 1, y0 >> 1, C, c0);
            }
            if (!need2) continue;
            __syncthreads();
            if (need3) pool_tile_packed(w2, S2, 8, w3, S3, 4, 4);
            if (f2 && (p.level_mask & 4)) {
                int Hl = H >> 2, Wl = W >> 2;
                store_tile_packed<3>(w2, S2, 16, f2, ((size_t)s * p.V + v) * Hl * Wl, Wl, Hl, x0 >> 2, y0 >> 2, C, c0);
            }
            if (!need3) continue;
            __syncthreads();
            if (f3 && (p.level_mask & 8)) {
                int Hl = H >> 3, Wl = W >> 3;
                store_tile_packed<2>(w3, S3, 4, f3, ((size_t)s * p.V + v) * Hl * Wl, Wl, Hl, x0 >> 3, y0 >> 3, C, c0);
            }
        }
        return;
    }
#endif
    for (int c0 = 0; c0 < C; c0 += CC) {
        __syncthreads();   // previous chunk's readers of t0..t3 are done
        if (sizeof(T) == 2 && pairs) {
            // 2-byte features: a thread loads TWO adjacent pixels of one channel (4 B), so that a warp instruction moves
            // 128 B like the fp32 path (half-warps on consecutive rows); threads 128..255 take the odd channels.  All 16
            // loads of the chunk are in flight before the first shared-memory store.
            const int hsel = tid >> 7, r2 = (tid >> 4) & 7, pp = tid & 15;
            const bool in2 = (y0 + r2 < H) && (x0 + 2 * pp < W);
            const size_t pix2 = (size_t)(y0 + r2) * W + (size_t)(x0 + 2 * pp);
            const bool swap = (tid >> 4) & 1;          // rows r2, r2+1 of a warp map to the same banks: alternate the order
            unsigned val[CC / 2];
#pragma unroll
            for (int c = 0; c < CC / 2; ++c) {
                const int ch = c0 + 2 * c + hsel;
                val[c] = 0u;
                if (in2 && ch < C) val[c] = __ldg(reinterpret_cast<const unsigned *>(fm + (size_t)ch * HW + pix2));
            }
#pragma unroll
            for (int c = 0; c < CC / 2; ++c) {
                float *d = t0 + (2 * c + hsel) * S0 + r2 * 32 + 2 * pp;
                const float lo = __uint_as_float(val[c] << 16), hi = __uint_as_float(val[c] & 0xffff0000u);
                if (swap) { d[1] = hi; d[0] = lo; } else { d[0] = lo; d[1] = hi; }
            }
        } else {
            // all loads of a batch are issued before the first shared-memory store: MVT_CLOUD_BATCH x 4 B per thread in flight
#pragma unroll
            for (int cb = 0; cb < CC; cb += MVT_CLOUD_BATCH) {
                float val[MVT_CLOUD_BATCH];
#pragma unroll
                for (int c = 0; c < MVT_CLOUD_BATCH; ++c) {
                    val[c] = 0.f;
                    if (inside && c0 + cb + c < C) val[c] = to_f<T>(__ldg(fm + (size_t)(c0 + cb + c) * HW + pix));
                }
#pragma unroll
                for (int c = 0; c < MVT_CLOUD_BATCH; ++c) t0[(cb + c) * S0 + tid] = val[c];
            }
        }
        __syncthreads();
        if (need1) pool_tile<T>(t0, S0, 32, t1, S1, 16, 64);
        if (p.fvec[0] && (p.level_mask & 1))
            store_tile<T, 5>(t0, S0, 256, reinterpret_cast<T *>(p.fvec[0]), ((size_t)s * p.V + v) * HW, W, H, x0, y0, C, c0, vec);
        if (!need1) continue;
        __syncthreads();
        if (need2) pool_tile<T>(t1, S1, 16, t2, S2, 8, 16);
        if (p.fvec[1] && (p.level_mask & 2)) {
            int Hl = H >> 1, Wl = W >> 1;
            store_tile<T, 4>(t1, S1, 64, reinterpret_cast<T *>(p.fvec[1]), ((size_t)s * p.V + v) * Hl * Wl, Wl, Hl, x0 >> 1, y0 >> 1, C, c0, vec);
        }
        if (!need2) continue;
        __syncthreads();
        if (need3) pool_tile<T>(t2, S2, 8, t3, S3, 4, 4);
        if (p.fvec[2] && (p.level_mask & 4)) {
            int Hl = H >> 2, Wl = W >> 2;
            store_tile<T, 3>(t2, S2, 16, reinterpret_cast<T *>(p.fvec[2]), ((size_t)s * p.V + v) * Hl * Wl, Wl, Hl, x0 >> 2, y0 >> 2, C, c0, vec);
        }
        if (!need3) continue;
        __syncthreads();
        if (p.fvec[3] && (p.level_mask & 8)) {
            int Hl = H >> 3, Wl = W >> 3;
            store_tile<T, 2>(t3, S3, 4, reinterpret_cast<T *>(p.fvec[3]), ((size_t)s * p.V + v) * Hl * Wl, Wl, Hl, x0 >> 3, y0 >> 3, C, c0, vec);
        }
    }
}

__global__ void zero_spacing_kernel(float *a, float *b, float *c, float *d, int n) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        if (a) a[i] = 0.f;
        if (b) b[i] = 0.f;
        if (c) c[i] = 0.f;
        if (d) d[i] = 0.f;
    }
}

}  // namespace mvt

// ---------------------------------------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------------------------------------
MVT_EXPORT int mvt_camera_prep(const float *intrs, const float *extrs, int n, float *cam, void *stream) {
    MVT_CHECK_ARG(intrs && extrs && cam, "mvt_camera_prep: null pointer");
    MVT_CHECK_ARG(n > 0, "mvt_camera_prep: n must be positive (got %d)", n);
    mvt::camera_prep_kernel<<<(n + 63) / 64, 64, 0, (cudaStream_t)stream>>>(intrs, extrs, n, cam);
    MVT_CHECK_LAUNCH("camera_prep");
    return MVT_OK;
}

static int cloud_build_impl(const void *fmaps, int feat_dtype, const float *depths, const float *cam,
                            int V, int S, int C, int H, int W, float stride, int depth_avg,
                            const mvt_level_out *levels, int level_mask, const mvt_peers *records, int view_offset,
                            int V_total, void *stream, long long fm_vstride = 0, long long dp_vstride = 0, int cam_vstride = 0) {
    using namespace mvt;
    MVT_CHECK_ARG(depths && cam && levels, "mvt_cloud_build: null pointer");
    MVT_CHECK_ARG(V > 0 && S > 0 && C > 0 && H > 0 && W > 0, "mvt_cloud_build: bad shape V=%d S=%d C=%d H=%d W=%d", V, S, C, H, W);
    MVT_CHECK_ARG(feat_dtype == MVT_F32 || feat_dtype == MVT_BF16, "mvt_cloud_build: feat_dtype %d", feat_dtype);
    MVT_CHECK_ARG(level_mask > 0 && level_mask < (1 << MVT_MAX_LEVELS), "mvt_cloud_build: level_mask 0x%x (levels 0..3 supported)", level_mask);
    MVT_CHECK_ARG((long long)V * S <= 65535, "mvt_cloud_build: V*S too large");
    CloudParams p;
    p.fmaps = fmaps; p.depths = depths; p.cam = cam;
    p.V = V; p.S = S; p.C = C; p.H = H; p.W = W;
    p.fm_vstride = fm_vstride ? fm_vstride : (long long)S * C * H * W;
    p.dp_vstride = dp_vstride ? dp_vstride : (long long)S * H * W;
    p.cam_vstride = cam_vstride ? cam_vstride : S;
    MVT_CHECK_ARG(p.fm_vstride >= (long long)S * C * H * W && p.dp_vstride >= (long long)S * H * W && p.cam_vstride >= S,
                  "mvt_cloud_build: view strides smaller than one view's frames");
    p.stride = stride; p.depth_avg = depth_avg; p.level_mask = level_mask;
    bool any_f = false, any_sp = false;
    for (int l = 0; l < 4; ++l) {
        bool on = (level_mask >> l) & 1;
        p.xyz[l] = on ? levels[l].xyz : nullptr;
        p.xyz4[l] = on ? (float4 *)levels[l].xyz4 : nullptr;
        p.fvec[l] = on ? levels[l].fvec : nullptr;
        p.valid[l] = on ? levels[l].valid : nullptr;
        p.spacing[l] = on ? levels[l].spacing : nullptr;
        p.neg[l] = on ? levels[l].neg_depth : nullptr;
        if (on) {
            MVT_CHECK_ARG((H >> l) > 0 && (W >> l) > 0, "mvt_cloud_build: level %d is empty for %dx%d", l, H, W);
            any_f |= p.fvec[l] != nullptr;
            any_sp |= p.spacing[l] != nullptr;
            MVT_CHECK_ARG(((uintptr_t)p.xyz4[l] & 15) == 0, "mvt_cloud_build: xyz4[%d] must be 16-byte aligned", l);
            if (p.fvec[l] && C % 4 == 0)
                MVT_CHECK_ARG(((uintptr_t)p.fvec[l] & 15) == 0, "mvt_cloud_build: fvec[%d] must be 16-byte aligned", l);
        }
    }
    MVT_CHECK_ARG(!any_f || fmaps, "mvt_cloud_build: fmaps is null but a feature output was requested");
    p.view_off = view_offset; p.V_total = V_total;
    for (int l = 0; l < 4; ++l) {
        p.rec[l] = peers_dev(nullptr);
        if (records && ((level_mask >> l) & 1)) {
            MVT_CHECK_ARG(records[l].n >= 1 && records[l].n <= MVT_MAX_PEERS, "mvt_cloud_build_shared: level %d: %d ranks", l, records[l].n);
            MVT_CHECK_ARG(view_offset >= 0 && view_offset + V <= V_total, "mvt_cloud_build_shared: views %d..%d outside 0..%d", view_offset, view_offset + V, V_total);
            p.rec[l] = peers_dev(&records[l]);
        }
    }
    cudaStream_t st = (cudaStream_t)stream;
    for (int l = 0; l < 4; ++l)
        if (p.neg[l] && cudaMemsetAsync(p.neg[l], 0, sizeof(int) * (size_t)S, st) != cudaSuccess) {
            mvt::set_error("mvt_cloud_build: cudaMemsetAsync(neg_depth[%d]) failed", l);
            return MVT_E_CUDA;
        }
    if (any_sp) {
        int n = S * 2;
        zero_spacing_kernel<<<(n + 127) / 128, 128, 0, st>>>(p.spacing[0], p.spacing[1], p.spacing[2], p.spacing[3], n);
        MVT_CHECK_LAUNCH("zero_spacing");
    }
    p.tiles_x = (W + TW - 1) / TW;
    int tiles_y = (H + TH - 1) / TH;
    dim3 grid((unsigned)(p.tiles_x * tiles_y), (unsigned)(V * S));
    if (feat_dtype == MVT_F32) cloud_build_kernel<float><<<grid, 256, 0, st>>>(p);
    else cloud_build_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>(p);
    MVT_CHECK_LAUNCH("cloud_build");
    return MVT_OK;
}

MVT_EXPORT int mvt_cloud_build(const void *fmaps, int feat_dtype, const float *depths, const float *cam,
                               int V, int S, int C, int H, int W, float stride, int depth_avg,
                               const mvt_level_out *levels, int level_mask, void *stream) {
    return cloud_build_impl(fmaps, feat_dtype, depths, cam, V, S, C, H, W, stride, depth_avg, levels, level_mask, nullptr, 0, V,
                            stream);
}

MVT_EXPORT int mvt_cloud_build_frames(const void *fmaps, long long fmaps_view_stride, int feat_dtype, const float *depths,
                                      long long depths_view_stride, const float *cam, int cam_view_stride,
                                      int V, int S, int C, int H, int W, float stride, int depth_avg,
                                      const mvt_level_out *levels, int level_mask, const mvt_peers *records, int view_offset,
                                      int V_total, void *stream) {
    return cloud_build_impl(fmaps, feat_dtype, depths, cam, V, S, C, H, W, stride, depth_avg, levels, level_mask, records,
                            records ? view_offset : 0, records ? V_total : V, stream, fmaps_view_stride, depths_view_stride,
                            cam_view_stride);
}

MVT_EXPORT int mvt_cloud_build_shared(const void *fmaps, int feat_dtype, const float *depths, const float *cam,
                                      int V, int S, int C, int H, int W, float stride, int depth_avg,
                                      const mvt_level_out *levels, int level_mask, const mvt_peers *records,
                                      int view_offset, int V_total, void *stream) {
    MVT_CHECK_ARG(records, "mvt_cloud_build_shared: null records");
    return cloud_build_impl(fmaps, feat_dtype, depths, cam, V, S, C, H, W, stride, depth_avg, levels, level_mask, records,
                            view_offset, V_total, stream);
}
