// Update-transformer input rows: everything around the correlation block that
// MVTracker.forward_iteration assembles with permute + cat + two broadcast adds
// (mvtracker/models/core/mvtracker/mvtracker.py:374-386; flow embedding: mvtracker/models/core/embeddings.py:134-161).
//
//   x[n, s, :] = cat( sincos(flow), flow,          3E + 3 columns,  flow = coords[s, n] - coords[0, n]
//                     corr block,                  Lc columns, written beforehand by mvt_knn_corr through its strides
//                     track features ffeats[s, n], C columns
//                     track_mask[s, n], vis[s, n] )
//                + pos_embed[n, :] + time_embed[s, :]
//
// One block per track, one pass: every column of the [N, S, D] buffer is written exactly once (the correlation columns
// are read-modify-written in place: + pos + time), instead of the reference's cat (read all parts, write D) followed by
// two out-of-place adds (read D, write D, twice).  Bound: HBM; algorithmic bytes per row = D*4 written + (C*e + 12 + 8)
// read + Lc*4 re-read; the embedding tables (N*D and S*D floats) are read once per track / stay in L2.
// Arithmetic follows the reference's order in fp32: div_j = float(2j) * float(1000/E); sin / cos (full-precision
// sinf / cosf) of flow * div_j at columns axis*E + 2j / + 2j + 1; x = (row + pos) + time.
#include "mvt_common.cuh"

namespace mvt {

struct InputParams {
    const float *coords;       // [S, N, 3]
    const void *ffeats;        // [S, N, C]
    const float *track_mask;   // [S, N]
    const float *vis;          // [S, N]
    const float *pos;          // [N, D] or null
    const float *tim;          // [S, D] or null
    float *x;                  // row (n, s) at n * track_stride + s * frame_stride
    long long track_stride, frame_stride;
    int S, N, C, E, Lc, D;
    float div_scale;           // float(1000 / E)
    int corr_mode;             // 0: leave the correlation columns alone, 1: add the embeddings to them in place
};

constexpr int kInputMaxS = 64;      // frames per window the kernel keeps flows for (the model uses S = 12)

// One block per track.  The S*D elements of the track's rows are independent: thread t takes elements t, t + 256, ...
// (several per loop trip, so that their loads are in flight together); the S flow vectors are computed once into shared memory.
template <typename T>
__global__ void __launch_bounds__(256) track_input_kernel(const InputParams p) {
    __shared__ float flow[3 * kInputMaxS];
    const int n = blockIdx.x;
    const int E3 = 3 * p.E, c_corr = E3 + 3, c_feat = c_corr + p.Lc, c_mask = c_feat + p.C, D = p.D;
    if (threadIdx.x < 3 * p.S) {
        const int s = threadIdx.x / 3, a = threadIdx.x - 3 * s;
        flow[threadIdx.x] = __fsub_rn(__ldg(p.coords + ((size_t)s * p.N + n) * 3 + a), __ldg(p.coords + (size_t)n * 3 + a));
    }
    __syncthreads();
    const float *pos = p.pos ? p.pos + (size_t)n * D : nullptr;
    const T *ff = reinterpret_cast<const T *>(p.ffeats);
    float *base = p.x + (long long)n * p.track_stride;
    auto magic_of = [](int d) { return (unsigned)((0x100000000ull + (unsigned)d - 1) / (unsigned)d); };   // e / d == (e * magic) >> 32 while e * d < 2^32 (checked on the host)
    auto div_by = [](int e, unsigned magic) { return (int)(((unsigned long long)(unsigned)e * magic) >> 32); };
    // (a) the sin / cos pairs: columns axis*E + 2i (sin) and + 2i + 1 (cos) share the angle -> one sincosf per pair
    {
        const int half = p.E >> 1, ppr = 3 * half, total = p.S * ppr;        // pairs per row
        const unsigned m_row = magic_of(ppr), m_axis = magic_of(half);
#pragma unroll 2
        for (int e = threadIdx.x; e < total; e += 256) {
            const int s = div_by(e, m_row), pr = e - s * ppr;
            const int axis = div_by(pr, m_axis), i = pr - axis * half;
            const int c = axis * p.E + 2 * i;
            const float a = __fmul_rn(flow[3 * s + axis], __fmul_rn((float)(2 * i), p.div_scale));
            float sv, cv;
            sincosf(a, &sv, &cv);
            if (pos) { sv = __fadd_rn(sv, __ldg(pos + c)); cv = __fadd_rn(cv, __ldg(pos + c + 1)); }
            if (p.tim) { sv = __fadd_rn(sv, __ldg(p.tim + (size_t)s * D + c)); cv = __fadd_rn(cv, __ldg(p.tim + (size_t)s * D + c + 1)); }
            float *row = base + (long long)s * p.frame_stride;
            row[c] = sv;
            row[c + 1] = cv;
        }
    }
    // (b) the copied columns [3E, D): flow, correlation block (in place), track features, masks
    {
        const int w = D - E3, total = p.S * w;
        const unsigned m_row = magic_of(w);
#pragma unroll 4
        for (int e = threadIdx.x; e < total; e += 256) {
            const int s = div_by(e, m_row), c = E3 + (e - s * w);
            const size_t q = (size_t)s * p.N + n;
            float *row = base + (long long)s * p.frame_stride;
            float v;
            if (c < c_corr) {
                v = flow[3 * s + (c - E3)];
            } else if (c < c_feat) {
                if (p.corr_mode == 0) continue;
                v = row[c];
            } else if (c < c_mask) {
                v = (float)ff[q * p.C + (c - c_feat)];
            } else {
                v = c == c_mask ? __ldg(p.track_mask + q) : __ldg(p.vis + q);
            }
            if (pos) v = __fadd_rn(v, __ldg(pos + c));
            if (p.tim) v = __fadd_rn(v, __ldg(p.tim + (size_t)s * D + c));
            row[c] = v;
        }
    }
}

// Lifts pixels of one depth map to world space: bilinear_sample2d of the depth at (x, y)
// (mvtracker/models/core/model_utils.py:81-165: floor / +1 corners clamped into the image, weights from the unclamped
// corners) followed by the same unprojection as the cloud build, camera = Kinv * [x, y, 1] * z, world = Einv * [camera, 1]
// (mvtracker/models/evaluation_predictor_3dpt.py:103-116).  One thread per pixel.
__global__ void __launch_bounds__(256) lift_pixels_kernel(const float *__restrict__ depth, int H, int W, const float *__restrict__ cam,
                                                          const float *__restrict__ pix, int n, float t_value,
                                                          float *__restrict__ out, int out_stride, int out_off) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float x = __ldg(pix + 2 * i), y = __ldg(pix + 2 * i + 1);
    const float xf = floorf(x), yf = floorf(y);
    const float x0f = fminf(fmaxf(xf, -2.0e9f), 2.0e9f), y0f = fminf(fmaxf(yf, -2.0e9f), 2.0e9f);
    const int x0 = (int)x0f, y0 = (int)y0f;
    const int x0c = min(max(x0, 0), W - 1), x1c = min(max(x0 + 1, 0), W - 1);
    const int y0c = min(max(y0, 0), H - 1), y1c = min(max(y0 + 1, 0), H - 1);
    const float x1f = xf + 1.f, y1f = yf + 1.f;
    const float w00 = __fmul_rn(x1f - x, y1f - y), w01 = __fmul_rn(x - xf, y1f - y);
    const float w10 = __fmul_rn(x1f - x, y - yf), w11 = __fmul_rn(x - xf, y - yf);
    const float d00 = __ldg(depth + (size_t)y0c * W + x0c), d01 = __ldg(depth + (size_t)y0c * W + x1c);
    const float d10 = __ldg(depth + (size_t)y1c * W + x0c), d11 = __ldg(depth + (size_t)y1c * W + x1c);
    // (w00*i00 + w01*i01 + w10*i10 + w11*i11), left to right, products rounded (model_utils.py:149)
    const float z = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(w00, d00), __fmul_rn(w01, d01)), __fmul_rn(w10, d10)), __fmul_rn(w11, d11));
    const float rx = fmaf(cam[0], x, fmaf(cam[1], y, cam[2])), ry = fmaf(cam[3], x, fmaf(cam[4], y, cam[5])),
                rz = fmaf(cam[6], x, fmaf(cam[7], y, cam[8]));
    const float cx = rx * z, cy = ry * z, cz = rz * z;
    float *o = out + (size_t)i * out_stride + out_off;
    if (out_off > 0) o[-1] = t_value;
    o[0] = fmaf(cam[9], cx, fmaf(cam[10], cy, fmaf(cam[11], cz, cam[12])));
    o[1] = fmaf(cam[13], cx, fmaf(cam[14], cy, fmaf(cam[15], cz, cam[16])));
    o[2] = fmaf(cam[17], cx, fmaf(cam[18], cy, fmaf(cam[19], cz, cam[20])));
}

}  // namespace mvt

MVT_EXPORT int mvt_lift_pixels(const float *depth, int H, int W, const float *cam, const float *pixels_xy, int n, int with_t,
                               float t_value, float *out, void *stream) {
    using namespace mvt;
    MVT_CHECK_ARG(depth && cam && pixels_xy && out, "mvt_lift_pixels: null pointer");
    MVT_CHECK_ARG(H > 0 && W > 0 && n >= 0, "mvt_lift_pixels: bad sizes H=%d W=%d n=%d", H, W, n);
    if (n == 0) return MVT_OK;
    lift_pixels_kernel<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(depth, H, W, cam, pixels_xy, n, t_value, out,
                                                                           with_t ? 4 : 3, with_t ? 1 : 0);
    MVT_CHECK_LAUNCH("lift_pixels_kernel");
    return MVT_OK;
}

MVT_EXPORT int mvt_track_input(const float *coords, const void *ffeats, int feat_dtype, const float *track_mask, const float *vis,
                               const float *pos_embed, const float *time_embed, int S, int N, int C, int E, int Lc, float *x,
                               long long track_stride, long long frame_stride, int corr_mode, void *stream) {
    using namespace mvt;
    MVT_CHECK_ARG(coords && ffeats && track_mask && vis && x, "mvt_track_input: null pointer");
    MVT_CHECK_ARG(S > 0 && N > 0 && C > 0 && Lc >= 0, "mvt_track_input: bad sizes S=%d N=%d C=%d Lc=%d", S, N, C, Lc);
    MVT_CHECK_ARG(S <= kInputMaxS, "mvt_track_input: at most %d frames per window (got %d)", kInputMaxS, S);
    MVT_CHECK_ARG(E > 0 && (E & 1) == 0, "mvt_track_input: flow embedding size must be even (got %d)", E);
    MVT_CHECK_ARG(feat_dtype == MVT_F32 || feat_dtype == MVT_BF16, "mvt_track_input: feat_dtype %d", feat_dtype);
    MVT_CHECK_ARG(corr_mode == 0 || corr_mode == 1, "mvt_track_input: corr_mode %d", corr_mode);
    InputParams p;
    p.coords = coords; p.ffeats = ffeats; p.track_mask = track_mask; p.vis = vis; p.pos = pos_embed; p.tim = time_embed;
    p.x = x; p.track_stride = track_stride; p.frame_stride = frame_stride;
    p.S = S; p.N = N; p.C = C; p.E = E; p.Lc = Lc; p.D = 3 * E + 3 + Lc + C + 2;
    p.div_scale = (float)(1000.0 / (double)E);
    p.corr_mode = corr_mode;
    MVT_CHECK_ARG((long long)S * p.D * p.D < (1ll << 32), "mvt_track_input: S*D*D = %lld too large", (long long)S * p.D * p.D);
    MVT_CHECK_ARG(track_stride >= p.D || N == 1, "mvt_track_input: track stride %lld < row width %d", track_stride, p.D);
    if (feat_dtype == MVT_F32) track_input_kernel<float><<<N, 256, 0, (cudaStream_t)stream>>>(p);
    else track_input_kernel<__nv_bfloat16><<<N, 256, 0, (cudaStream_t)stream>>>(p);
    MVT_CHECK_LAUNCH("track_input_kernel");
    return MVT_OK;
}
