/*
 * mvt.h -- C ABI of libmvt.so: the B200 (sm_100a) implementation of MVTracker's
 * data-parallel hot path.  Plain pointers and sizes only; every pointer named
 * "device" must be CUDA device memory of the current device; `stream` is a
 * cudaStream_t passed as void* (NULL = legacy default stream).  All entry points
 * are asynchronous (enqueue only, no host sync, CUDA-graph capturable), never
 * allocate user-visible memory, and return 0 on success or a negative MVT_E_*
 * code; mvt_last_error() returns a thread-local message for the last failure.
 *
 * Reference interfaces replaced (paths under the reference repo):
 *   mvt_camera_prep + mvt_cloud_build  <- init_pointcloud_from_rgbd
 *        mvtracker/models/core/model_utils.py:420-482
 *   mvt_grid_build + mvt_knn            <- knn(k, xyz_ref, xyz_query) i.e. _knn_pointops / _knn_torch
 *        mvtracker/models/core/mvtracker/mvtracker.py:26-90  (native precedent: pointops.knn_query, :55-65)
 *   mvt_knn_corr                        <- PointcloudCorrBlock.corr_sample
 *        mvtracker/models/core/mvtracker/mvtracker.py:800-846 (and the level loop :355-369)
 *   mvt_knn_gather                      <- feature init, k = 1
 *        mvtracker/models/core/mvtracker/mvtracker.py:627-643
 *   mvt_knn_corr_backward, mvt_cloud_build_backward  <- torch autograd through the two functions above (training)
 *   mvt_lift_pixels                     <- support-grid unprojection of EvaluationPredictor.forward
 *        mvtracker/models/evaluation_predictor_3dpt.py:99-116 (bilinear_sample2d: model_utils.py:81-165)
 *   mvt_track_input                     <- assembly of the update transformer's input around the correlation block
 *        mvtracker/models/core/mvtracker/mvtracker.py:374-386, mvtracker/models/core/embeddings.py:134-161
 */
#ifndef MVT_H_
#define MVT_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MVT_VERSION 100

#define MVT_OK 0
#define MVT_E_ARG (-1)      /* bad argument (shape, alignment, unsupported combination) */
#define MVT_E_CUDA (-2)     /* a CUDA runtime call / launch failed */
#define MVT_E_WORKSPACE (-3) /* workspace too small */

#define MVT_F32 0
#define MVT_BF16 1

#define MVT_MAX_LEVELS 4    /* pyramid levels 0..3 (configs/model/mvtracker.yaml: corr_n_levels=4) */
#define MVT_MAX_K 32
#define MVT_CAM_STRIDE 48  /* floats per camera record written by mvt_camera_prep */
#define MVT_MAX_VIEWS 32
#define MVT_MAX_PEERS 8     /* GPUs of one node sharing a symmetric buffer */

int mvt_version(void);
const char *mvt_last_error(void);
/* Number of kernels this library has launched in the process so far (monotone; for benchmarks' gpu_launches). */
unsigned long long mvt_launch_count(void);

/* Inverse camera matrices, fp64 on device, rounded to fp32 (model_utils.py:454-457).
 * intrs [n,3,3], extrs [n,3,4] (world->camera), device fp32, n = V*S.
 * cam   [n,MVT_CAM_STRIDE] device fp32:  [0..8] Kinv | [9..20] Einv rows 0..2 (3x4) | [24..32] R | [33..35] t |
 *       [36..40] fx, skew, cx, fy, cy | [41] 1 if K is upper triangular with K22 == 1 (raster search usable) |
 *       [42] upper bound of the largest singular value of extrs[:, :3] (1 for a rigid transform; similarity-scaled
 *       extrinsics as datasets/utils.py:403-415 produces them are searched exactly). */
int mvt_camera_prep(const float *intrs, const float *extrs, int n, float *cam, void *stream);

/* One pyramid level's outputs of the cloud build.  Any pointer may be NULL (skip that output). */
typedef struct mvt_level_out {
    float *xyz;             /* [S, P_l, 3] fp32, P_l = V*(H>>l)*(W>>l), point = v*H_l*W_l + y*W_l + x (model_utils.py:475) */
    float *xyz4;            /* [S, P_l, 4] fp32 {x, y, z, depth}: the same points as 16-byte raster records, the
                               search structure of the raster kNN (no build step) */
    void *fvec;             /* [S, P_l, C] feature dtype (model_utils.py:476) */
    unsigned char *valid;   /* [S, P_l]   depth > 0 (model_utils.py:478-480) */
    float *spacing;         /* [S, 2]     sum and count of the pixel footprint depth*stride_l/fx over valid points
                                          (kNN cell-size hint; accumulated, zeroed by the call) */
    int *neg_depth;         /* [S]        set to 1 when frame s of this level holds a point with depth < 0 (zeroed by the
                                          call).  Such points lie BEHIND their camera on the pixel's line; the reference's
                                          kNN (mvtracker.py:816, filter_invalid=False) still returns them, so a raster
                                          search that is handed this array scans those frames' rasters exhaustively */
} mvt_level_out;

/* Fused unprojection + view fusion + feature pyramid, all requested levels in ONE pass over fmaps.
 * fmaps  [V,S,C,H,W] device, feat_dtype MVT_F32|MVT_BF16;  depths [V,S,H,W] device fp32 (already strided by the caller,
 * mvtracker.py:558-563);  cam from mvt_camera_prep;  stride = pixel stride of level 0 (model_utils.py:424).
 * depth_avg: 0 = 'nearest' (depth_l = depth_0[y<<l, x<<l]), 1 = 'avg' (chained 2x2 means) (model_utils.py:441-446).
 * levels[l] describes level l; level_mask bit l set = emit level l. */
int mvt_cloud_build(const void *fmaps, int feat_dtype, const float *depths, const float *cam,
                    int V, int S, int C, int H, int W, float stride, int depth_avg,
                    const mvt_level_out *levels, int level_mask, void *stream);

/* mvt_cloud_build for a RUN OF FRAMES inside larger arrays -- what a sliding window needs when the clouds of the frames it
 * shares with the previous window are kept (the reference rebuilds all S frames of every window although S/2 of them
 * overlap, mvtracker.py:297-306 vs :549-555): fmaps / depths are frame runs of views that are fmaps_view_stride /
 * depths_view_stride ELEMENTS apart (e.g. frames 6..11 of a [V,12,C,H,W] tensor: pointer at frame 6, view stride 12*C*H*W),
 * the camera record of (view v, frame s) is cam + (v*cam_view_stride + s)*MVT_CAM_STRIDE, and the output pointers of
 * `levels` point at the first frame to write inside the caller's per-frame arrays.  0 strides = dense.
 * records (optional, [MVT_MAX_LEVELS], see mvt_cloud_build_shared below): also store these views' 16-byte records into the
 * replicated per-frame arrays of every rank, at views view_offset.. of V_total (pointers at the first frame to write). */
struct mvt_peers;
int mvt_cloud_build_frames(const void *fmaps, long long fmaps_view_stride, int feat_dtype, const float *depths,
                           long long depths_view_stride, const float *cam, int cam_view_stride,
                           int V, int S, int C, int H, int W, float stride, int depth_avg,
                           const mvt_level_out *levels, int level_mask, const struct mvt_peers *records, int view_offset,
                           int V_total, void *stream);

/* Hashed uniform grid over B independent clouds ("segments", one per frame) of N points each. */
size_t mvt_grid_workspace_bytes(int B, int N);

typedef struct mvt_grid_job {
    const float *xyz;           /* [B,N,3] device fp32 */
    const unsigned char *mask;  /* [B,N] or NULL: points with mask==0 are left out of the search */
    const float *spacing;       /* [B,2] (sum,count) cell-size hint from mvt_cloud_build, or NULL (volumetric estimate) */
    void *workspace;            /* device, >= mvt_grid_workspace_bytes(B,N), 256-byte aligned */
    int B, N;
    int k_hint;                 /* the k the grid will mostly be queried with (cell size ~ k-NN radius) */
    int index_mode;             /* 0: neighbours are reported by original index; 1: by rank among mask-valid points
                                   (bit-compatible with the reference's filter_invalid path, mvtracker.py:820-829) */
    float cell_scale;           /* multiplies the estimated cell size; <= 0 selects the default */
} mvt_grid_job;

/* Builds up to 8 grids (e.g. the 4 pyramid levels) with one set of launches. */
int mvt_grid_build(const mvt_grid_job *jobs, int n_jobs, void *stream);

/* What a query kernel searches in.  Two structures give identical (exact) results:
 *   grid mode   (grid != NULL): the hashed grid built by mvt_grid_build over xyz [B,N,3]; any cloud.
 *   raster mode (grid == NULL): the {x,y,z,depth} records xyz4 [B,N,4] written by mvt_cloud_build plus the camera
 *                records of mvt_camera_prep; N == V*H*W.  No build step; needs a cloud that came from depth rasters. */
typedef struct mvt_search {
    const void *grid;
    const float *xyz;        /* grid mode: [B,N,3] (neighbour coordinates for the offsets) */
    const float *xyz4;       /* raster mode: [B,N,4] */
    const float *cam;        /* raster mode: record of (view v, batch element b) at cam + (v*cam_view_stride + b)*MVT_CAM_STRIDE */
    int cam_view_stride;     /* records between consecutive views (= S of the mvt_camera_prep call) */
    int N;
    int V, H, W;             /* raster mode: rasters of this pyramid level */
    float stride;            /* raster mode: input-image pixels per raster pixel (stride * 2^level) */
    int skip_invalid;        /* raster mode: 1 = points with depth <= 0 are not searchable */
    const float *tile_z;     /* raster mode, optional: [B, V, ceil(H/8), ceil(W/8), 2] {min, max} view-space depth of the points of
                                every MVT_TILE x MVT_TILE pixel tile (mvt_tile_depth_build); lets a search whose ball projects onto
                                a large pixel box skip the tiles whose depth range the ball cannot reach.  NULL: no culling.
                                Never changes a result. */
    const int *neg_depth;    /* raster mode, optional: [B] flags of mvt_level_out.neg_depth.  NULL: the caller promises that
                                no searchable point has depth < 0 (such points would not be found) */
} mvt_search;

#define MVT_TILE 8
/* Floats in the tile_z table of one raster level. */
long long mvt_tile_depth_floats(int B, int V, int H, int W);
/* Fills tile_z_out[i] for n (<= 8) raster-mode descriptors (e.g. the 4 pyramid levels) in one launch.  The tile_z fields
 * of the descriptors themselves are ignored.  No reference counterpart (the reference's kNN, mvtracker.py:26-90, is brute
 * force / pointops); this is part of the search structure, like mvt_grid_build. */
int mvt_tile_depth_build(const mvt_search *searches, int n, int B, float *const *tile_z_out, void *stream);

/* Exact kNN.  query [B,M,3] device fp32.  idx_out [B,M,k] int32 or int64 (idx_is_i64), dist_out [B,M,k] fp32 Euclidean,
 * either may be NULL.  Order: ascending (d2, index) with d2 = fma(dz,dz,fma(dy,dy,dx*dx)) in fp32.
 * Missing neighbours (fewer than k searchable points): index -1, distance +inf. */
int mvt_knn(const mvt_search *search, int B, const float *query, int M, int k,
            void *idx_out, int idx_is_i64, float *dist_out, void *stream);

typedef struct mvt_corr_level {
    mvt_search search;
    const void *fvec;       /* [B,N,C] */
    int k;
    long long out_offset;   /* float offset of this level's k*D block inside an output row */
} mvt_corr_level;

/* Fused kNN + neighbour gather + grouped dot-product correlation + offset encoding for up to 4 levels in one launch.
 * targets [B,M,C] (feature dtype), coords [B,M,3] fp32.
 * out: fp32; row (b,m) starts at out + b*out_frame_stride + m*out_row_stride (out_frame_stride 0 = M*out_row_stride, the
 * dense [B,M,.] block); with out_frame_stride = D_in, out_row_stride = S*D_in and out pointing at its column the block
 * lands directly in the [N,S,D_in] transformer input that mvtracker.py:374-384 otherwise assembles with permute + cat.
 * Level l writes k*D floats at its out_offset, neighbour j
 * at j*D:  [corr_0..corr_{G-1}, dx,dy,dz (add_offset), x,y,z (add_xyz)],  D = G + 3*add_offset + 3*add_xyz,
 * corr_g = <target_g, nbr_g> / sqrt(C/G)  (mvtracker.py:832-846).
 * idx_out: optional [n_levels,B,M,k] int32.
 * idx_hint: optional [n_levels,B,M,k] int32 -- ANY k points of the same clouds per query, typically the idx_out of the
 * previous refinement iteration (tracks move by centimetres between iterations, mvtracker.py:392): the largest distance
 * to them bounds the k-th best, so the raster search starts with a tight radius.  A warm start only: results are exact
 * and identical with or without it (entries < 0, repeated entries or entries of unsearchable points disable it for that
 * query).  May alias idx_out.
 * workspace: optional device scratch of >= mvt_knn_corr_workspace_bytes(n_levels, B, M) bytes, 16-byte aligned.  With it
 * (and tile_z tables in every level's search descriptor, C == 128, G == 1, k <= 16) the fused kernel hands the queries whose
 * ball projects onto a large pixel box -- tracks floating in free space, far from every surface: the reference starts each
 * window with the same 3-D point in all frames (mvtracker.py:510), so tracks of moving objects do -- to a second kernel
 * that finishes them with the depth-culled tile search; the results are identical, the slow queries no longer hold the
 * fused kernel's warps (hard_query_mode 1).  hard_query_mode 2: one kernel does everything, but the queries that mode 1 would
 * have handed over are counted; either way the count is left in the first 4 bytes of the workspace (a caller can read it back
 * asynchronously and pick the mode of its next call: deferral pays when several per cent of the queries are such, it costs
 * a little when there are only a few).  NULL / mode 0: one kernel, nothing counted. */
size_t mvt_knn_corr_workspace_bytes(int n_levels, int B, int M);
int mvt_knn_corr(const mvt_corr_level *levels, int n_levels, int B, int M, int C, int G, int feat_dtype,
                 const void *targets, const float *coords, int add_offset, int add_xyz,
                 float *out, long long out_row_stride, long long out_frame_stride, const int *idx_hint, int *idx_out,
                 void *workspace, int hard_query_mode, void *stream);

/* mvt_knn_corr for callers that cannot allocate device memory (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc).  No counterpart in the reference.
 * Same arguments, arithmetic and results as mvt_knn_corr with a workspace, but the scratch (the list of deferred large-box
 * queries) is a file-scope `static __device__` array of MVT_KNN_CORR_NOALLOC_BYTES inside libmvt.so -- every GPU that loads
 * the library holds that much until the process exits.  The call allocates nothing (no cudaMalloc / cudaFree, no allocating
 * library routine); the caller passes inputs and outputs only.
 * Limit: mvt_knn_corr_workspace_bytes(n_levels, B, M) <= MVT_KNN_CORR_NOALLOC_BYTES, i.e. n_levels * B * M <=
 * MVT_KNN_CORR_NOALLOC_QUERIES (BASELINE.json config 2: 4 * 12 * 2048 = 98 304); larger shapes are refused with
 * MVT_E_WORKSPACE and a message in mvt_last_error().  hard_query_mode 1 or 2 as above (0 would not use the scratch: refused).
 * ALL calls on one GPU share the array (per process): it keeps the previous call's list, the counters at its head are
 * cleared on the call's stream before the kernels run; two calls that overlap on the GPU (different streams without an
 * ordering between them) race on it -- serialise them or use mvt_knn_corr with one workspace per stream. */
#define MVT_KNN_CORR_NOALLOC_BYTES (4u << 20)
#define MVT_KNN_CORR_NOALLOC_QUERIES ((MVT_KNN_CORR_NOALLOC_BYTES - 256u) / 16u)
int mvt_knn_corr_noalloc(const mvt_corr_level *levels, int n_levels, int B, int M, int C, int G, int feat_dtype,
                         const void *targets, const float *coords, int add_offset, int add_xyz,
                         float *out, long long out_row_stride, long long out_frame_stride, const int *idx_hint, int *idx_out,
                         int hard_query_mode, void *stream);

/* k = 1 search returning the nearest point's feature row (feature init, mvtracker.py:636-643).
 * fvec [B,N,C]; feat_out [B,M,C] feature dtype; idx_out optional [B,M] int32. */
int mvt_knn_gather(const mvt_search *search, const void *fvec, int feat_dtype, int B, int C,
                   const float *query, int M, void *feat_out, int *idx_out, void *stream);

/* Update-transformer input rows (mvtracker.py:374-386 builds them with permute + cat + two adds):
 *   x[n, s, :] = cat( get_3d_embedding(coords[s,n] - coords[0,n], E) (3E sin/cos columns, then the 3 flow components;
 *                     embeddings.py:134-161), the Lc correlation columns, ffeats[s,n,:] (C), track_mask[s,n], vis[s,n] )
 *                + pos_embed[n,:] + time_embed[s,:]                      D = 3E + 3 + Lc + C + 2 columns
 * coords [S,N,3] fp32; ffeats [S,N,C] feature dtype; track_mask, vis [S,N] fp32; pos_embed [N,D], time_embed [S,D] fp32
 * (either may be NULL = not added).  x: row (n, s) at x + n*track_stride + s*frame_stride (floats), unit column stride.
 * corr_mode 1: the correlation columns already hold what mvt_knn_corr wrote there (out = x + 3E+3, row stride =
 * track_stride, frame stride = frame_stride) and get the two embeddings added in place; 0: they are left untouched. */
int mvt_track_input(const float *coords, const void *ffeats, int feat_dtype, const float *track_mask, const float *vis,
                    const float *pos_embed, const float *time_embed, int S, int N, int C, int E, int Lc, float *x,
                    long long track_stride, long long frame_stride, int corr_mode, void *stream);

/* Pixels of ONE depth map lifted to world space (the support grid of EvaluationPredictor.forward,
 * mvtracker/models/evaluation_predictor_3dpt.py:99-116): z = bilinear_sample2d(depth, x, y) (model_utils.py:81-165), then
 * world = Einv * [Kinv * [x, y, 1] * z, 1] with the camera record of mvt_camera_prep.
 * depth [H,W] device fp32; cam [MVT_CAM_STRIDE] (one record); pixels_xy [n,2] (x, y) in pixels of this map.
 * out [n,3] (with_t 0) or [n,4] = (t_value, x, y, z) (with_t 1: the query / support point layout). */
int mvt_lift_pixels(const float *depth, int H, int W, const float *cam, const float *pixels_xy, int n, int with_t,
                    float t_value, float *out, void *stream);

/* ---- backward of the path (training, train.py:119-163) ---------------------------------------------
 * What autograd computes through corr_sample's gather + einsum / sqrt(C/G) (mvtracker.py:827-835) for ONE level, given the
 * neighbour lists of the forward pass (mvt_knn_corr's idx_out, [B,M,k] int32, -1 = none):
 *   grad_targets[b,m,c] (+)= sum_j gs * fvec[b, idx_j, c]      gs = grad_out[b,m,j,group(c)] / sqrt(C/G)
 *   grad_fvec[b,idx_j,c]  += gs * targets[b,m,c]               (vector reductions; accumulation order is not fixed)
 * grad_out entry (b,m,j,g) at grad_out[b*grad_frame_stride + m*grad_row_stride + j*D + g] (the level's block of the
 * forward output; offset / xyz columns carry no gradient).  grad_targets [B,M,C] fp32 (accumulate_targets 1: added to,
 * for the further levels of the same targets), grad_fvec [B,N,C] fp32, zeroed by the caller.  C/G % 4 == 0. */
int mvt_knn_corr_backward(const void *fvec, const void *targets, int feat_dtype, const int *idx, const float *grad_out,
                          long long grad_row_stride, long long grad_frame_stride, int B, int N, int M, int C, int G,
                          int k, int D, float *grad_targets, int accumulate_targets, float *grad_fvec, void *stream);

/* Feature branch of init_pointcloud_from_rgbd backwards (model_utils.py:439-441 chained avg_pool2d, :475-476 permute):
 *   grad_fmaps[v,s,c,y,x] (+)= grad_fvec[s, v*Hl*Wl + (y>>level)*Wl + (x>>level), c] / 4^level, 0 outside the pooled area.
 * grad_fvec [S, V*Hl*Wl, C] fp32 (Hl = H>>level); grad_fmaps [V,S,C,H,W] fp32. */
int mvt_cloud_build_backward(const float *grad_fvec, int V, int S, int C, int H, int W, int level, float *grad_fmaps,
                             int accumulate, void *stream);

/* ---- view-sharded kNN correlation across GPUs (mvtracker_b200/csrc/mvt_shard.cu; driven by mvtracker_b200/dist.py) --------
 * Each rank keeps only the clouds of the views it unprojected; 8-byte (d2, index) keys and 16-byte result records are
 * all that crosses NVLink.  The reference has no multi-GPU inference; the three calls below are
 * the per-rank halves of PointcloudCorrBlock.corr_sample (mvtracker.py:800-846) split at the two exchange points. */

/* Exact top-k of all queries among THIS rank's points, as sorted packed keys: (bits(d2) << 32) | (index + index_offset),
 * missing = 0x7f8000007fffffff.  searches[l] / index_offsets[l] describe level l (same k on every level).
 * keys_out [n_shards, n_levels, B, Ms, k] with Ms = ceil(M / n_shards): track m goes to shard m / Ms, padded tracks are
 * filled with missing keys -- the layout an all-to-all over track shards sends as is.
 * cap_keys (optional, [n_shards, n_levels, B, Ms] in the same track layout): only points whose squared distance is <= the
 * cap key's distance part are searched for (an upper bound of the GLOBAL k-th distance obtained from other ranks, so that a
 * rank whose views do not see the query's surface stops early); fewer than k keys may then come back.
 * budget > 0: bounded effort -- a query's search gives up once it knows k points and has looked at `budget` candidates; its
 * list is then k real points (a valid bound), not the exact answer.  exact_out (optional, same layout, bytes): 1 where the list
 * is exact.  skip (optional, same layout): queries flagged 1 are not searched and their keys in keys_out are left untouched
 * (second phase over the buffer of the first). */
int mvt_knn_keys(const mvt_search *searches, const long long *index_offsets, int n_levels, int B,
                 const float *query, int M, int k, int n_shards, const unsigned long long *cap_keys, int budget,
                 unsigned char *exact_out, const unsigned char *skip, unsigned long long *keys_out, void *stream);

/* Merges n_lists sorted candidate lists per query: keys_in [n_lists, Q, k] -> idx_out [Q, k] int32 (-1 = missing) and /
 * or keys_out [Q, k], ascending (d2, index). */
int mvt_topk_merge(const unsigned long long *keys_in, int n_lists, long long Q, int k, int *idx_out,
                   unsigned long long *keys_out, void *stream);

typedef struct mvt_owned_level {
    const float *xyz4;      /* [B, hi-lo, 4] raster records of the points this rank owns at this level */
    const void *fvec;       /* [B, hi-lo, C] */
    long long lo, hi;       /* owned global index range [lo, hi) */
    long long out_offset;   /* float offset of the level's k*4 block inside an output row */
} mvt_owned_level;

/* Correlation + offsets for GIVEN neighbour lists, for the neighbours this rank owns; records of neighbours owned by other
 * ranks are written as zeros, so the sum over ranks is the full mvt_knn_corr output (add_offset = 1, add_xyz = 0, C = 128,
 * groups = 1).  idx [n_shards, n_levels, B, Ms, k] global indices in the layout of mvt_knn_keys. */
int mvt_corr_from_idx(const mvt_owned_level *levels, int n_levels, int B, int M, int C, int feat_dtype, int k,
                      int n_shards, const int *idx, const void *targets, const float *coords, float *out,
                      long long out_row_stride, void *stream);

/* ---- fused compute + exchange over NVLink (symmetric memory; csrc/mvt_shard.cu, csrc/mvt_cloud.cu) ----------------------
 * The three exchanges of the track-sharded path are folded into the kernels that produce the data: every 16-byte unit
 * (a point record, four neighbour indices, a [corr,dx,dy,dz] record) has exactly ONE producer, which stores it into the
 * same buffer on every rank -- through the NVSwitch multicast address (one multimem.st) when the buffer has one, else
 * through the per-rank peer mappings.  No NCCL call, no reduction, no zero fill; the caller puts a cross-rank barrier
 * (symmetric-memory signal pads, ~6 us) between producer and consumer kernels. */
typedef struct mvt_peers {
    void *mc;                    /* multicast address of the buffer, or NULL */
    void *ptr[MVT_MAX_PEERS];    /* the buffer as mapped for rank 0..n-1 (ptr[own rank] = the local address) */
    int n;                       /* ranks; 16-byte aligned addresses */
} mvt_peers;

/* mvt_cloud_build that ALSO stores the {x,y,z,depth} records of this rank's views into the replicated
 * [S, V_total*H_l*W_l, 4] buffers records[l] of every rank (views view_offset .. view_offset+V-1). */
int mvt_cloud_build_shared(const void *fmaps, int feat_dtype, const float *depths, const float *cam,
                           int V, int S, int C, int H, int W, float stride, int depth_avg,
                           const mvt_level_out *levels, int level_mask, const mvt_peers *records /* [MVT_MAX_LEVELS] */,
                           int view_offset, int V_total, void *stream);

/* Exact kNN of THIS rank's track shard (query [B, Ms, 3]) over the replicated records; the int32 index lists are stored
 * into idx_all [n_shards, n_levels, B, Ms, k] of every rank at shard `shard`.  k must be a multiple of 4. */
int mvt_knn_idx_shared(const mvt_search *searches, int n_levels, int B, const float *query, int Ms, int k,
                       int shard, int n_shards, const mvt_peers *idx_all, void *stream);

/* mvt_corr_from_idx whose records go to out [B, M, n_levels*k*4] of every rank: a record is stored by the rank that owns
 * the neighbour's features; records of missing neighbours (index < 0) are stored as zeros by rank `shard` for its own
 * track shard.  After the barrier every rank holds the complete block.
 * all_xyz4 / n_all (optional, [n_levels]): this rank's replica of every view's records per level ([B, n_all[l], 4], what
 * mvt_cloud_build_shared filled).  Then each rank writes the three offsets of every neighbour into its own block itself
 * and only the correlation value of a record (4 of its 16 bytes) is stored to the other ranks: a quarter of the NVLink
 * traffic, same bits. */
int mvt_corr_from_idx_shared(const mvt_owned_level *levels, int n_levels, int B, int M, int C, int feat_dtype, int k,
                             int n_shards, int shard, const int *idx, const void *targets, const float *coords,
                             const mvt_peers *out, long long out_row_stride, const float *const *all_xyz4,
                             const long long *n_all, void *stream);

/* ---- frames sharded over the GPUs: the fused kernel storing into every rank's result block ------------------------------------
 * The frames of a window are the batch axis of the reference's kNN correlation (corr_sample runs on [B*S, ...] clouds,
 * mvtracker.py:800-846; frame s of the window only ever meets frame s of the cloud): they are independent units.  A rank that
 * owns a subset of the window's frames holds those frames' clouds of ALL views, runs mvt_knn_corr's fused kernel for its
 * frames x all tracks, and stores each 16-byte [corr, dx, dy, dz] record into the [S, M, n_levels*k*4] block of EVERY rank
 * (one multimem.st through `out->mc`, or one store per peer mapping without multicast) -- the all-gather of the results is the
 * kernel's own epilogue; a cross-rank barrier after the launch completes the block.  No index lists, no inboxes, no second
 * kernel.  groups = 1, add_offset = 1, add_xyz = 0, C and k as mvt_knn_corr's fast path or not (any C works).
 * B: this rank's frames (levels[].search describe the clouds of exactly these frames, in order); targets / coords point at
 * the rows of the FIRST local frame inside the replicated [S, M, .] tensors and consecutive local frames are
 * query_frame_stride queries apart (round-robin ownership over N ranks: N*M); `out` (ptr[] / mc already offset to the first
 * local frame's rows) with out_frame_stride floats between local frames, out_row_stride between tracks.
 * workspace / hard_query_mode: as mvt_knn_corr. */
int mvt_knn_corr_shared(const mvt_corr_level *levels, int n_levels, int B, int M, int C, int feat_dtype,
                        const void *targets, const float *coords, long long query_frame_stride,
                        const mvt_peers *out, long long out_row_stride, long long out_frame_stride,
                        void *workspace, int hard_query_mode, void *stream);

/* ---- one kernel per refinement iteration: search + exchange + correlation + exchange (csrc/mvt_shard.cu) ----------------------
 * The compute-then-collective chain of the calls above (mvt_knn_idx_shared -> barrier -> mvt_corr_from_idx_shared ->
 * barrier: four stages, every rank walking every query's list) as ONE kernel launch per rank:
 *   producers  (the block shape of the fused single-GPU kernel) one warp per query of THIS rank's track shard: window search
 *              over the replicated records; the search knows the owner of every neighbour (index -> view -> rank) and pushes
 *              an 8-byte record {query, slot, owner-local index} straight into the OWNER's inbox (peer store over NVLink);
 *              records are self-validating (top bit), no fence or counter per record; queries with large pixel boxes are
 *              deferred to the consumers (tile search);
 *   consumers  one wave of blocks at the end of the same grid: as records arrive in this rank's inboxes they gather the
 *              feature rows it owns, correlate them with the queries' target rows and store the [corr, dx, dy, dz] records
 *              into the result block of every rank (multicast store) -- overlapping the search tails of all ranks;
 *   end        the last consumer warp of a rank resets its counters, signals `epoch` to every peer and waits for theirs: when
 *              the kernel has completed on a rank, that rank's result block is complete.
 * No host-visible barrier, no NCCL, one launch.  Records are bit-identical to mvt_knn_corr's (same per-row arithmetic).
 * C == 128, groups == 1, k <= 16 (a multiple of 4), add_offset = 1, add_xyz = 0. */
typedef struct mvt_shard_level {
    mvt_search search;        /* raster mode over the replicated records of ALL views at this level (+ tile tables) */
    const float *xyz4_local;  /* [B, n_local, 4] records of the views this rank owns */
    const void *fvec_local;   /* [B, n_local, C] feature rows of the views this rank owns */
    long long n_local;        /* owned points per frame at this level */
    int hw;                   /* H_l * W_l: points per view */
} mvt_shard_level;

typedef struct mvt_shard_ctx {
    mvt_peers inbox;          /* symmetric: per rank [n_ranks][inbox_cap] 8-byte records, zeroed once by the caller */
    mvt_peers ctrl;           /* symmetric: per rank 2*MVT_MAX_PEERS 8-byte words (totals, epochs), zeroed once */
    mvt_peers result;         /* symmetric: per rank [B, M, n_levels*k*4] fp32; mc = its multicast address or NULL */
    void *local;              /* this rank's private scratch: mvt_shard_local_bytes(...) bytes, 16-byte aligned, zeroed once */
    long long inbox_cap;      /* records per (destination, source) pair: >= n_levels * B * ceil(M / n_ranks) * k */
    int rank, n_ranks;
    int view_first[MVT_MAX_PEERS + 1];   /* rank r owns views view_first[r] .. view_first[r+1]-1 (contiguous, ascending) */
} mvt_shard_ctx;

size_t mvt_shard_local_bytes(int n_levels, int B, int M, int n_ranks);
/* targets [B, M, C] and coords [B, M, 3]: ALL tracks (replicated inputs, as the replicated transformer holds them); rank r
 * searches tracks r*Ms .. (r+1)*Ms-1, Ms = ceil(M / n_ranks).  epoch: a number that grows by one per call on every rank. */
int mvt_shard_iter(const mvt_shard_level *levels, int n_levels, int B, int M, int C, int feat_dtype, int k,
                   const void *targets, const float *coords, const mvt_shard_ctx *ctx, long long result_row_stride,
                   unsigned long long epoch, void *stream);

/* Debug / profiling: per-segment grid parameters copied out for inspection.
 * info [B,16] device fp32: ox,oy,oz,h,nx,ny,nz,n_valid,n_cells_used,table_slots,... */
int mvt_grid_info(const void *grid, int B, int N, float *info, void *stream);

/* Debug: device array of 16 uint64 counters the search kernels add to (NULL = off, the default):
 * grid mode:   [0] queries [1..3] finished after ring 1..3 [4] exhaustive fallbacks [5] candidates scanned
 *              [6] top-k insertions [7] cell lookups [8] small-segment exhaustive scans;
 * raster mode: [0] queries [1..4] finished after 1, 2, 3, >=4 passes [5] candidates scanned [6] insertions. */
int mvt_debug_set_stats(void *dev_counters);

/* Debug / experiments: lanes that own one query in mvt_knn_corr's raster kernels, 32 (default) or 16 (two queries per
 * warp in lock step; needs k <= 16, V <= 16, C == 128, groups == 1, else 32 is used).  Results are identical. */
int mvt_debug_set_group_width(int lanes);

#ifdef __cplusplus
}
#endif
#endif /* MVT_H_ */
