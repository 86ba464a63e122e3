// Hashed uniform grid over B independent clouds ("segments"): the search structure behind mvt_knn / mvt_knn_corr.
// Cells are keyed by packed integer coordinates (10 bits/axis) in an open-addressing table per segment; the points of
// a cell are contiguous in `sorted` as {x, y, z, bit-cast index} records.  Layout inside a cell and the order of cells
// are NOT deterministic (atomics) -- results are, because the search orders candidates by the total order (d2, index).
//
// Build = 5 launches for up to 8 grids (all pyramid levels at once): init, aabb, plan, count, alloc+scatter.
#include "mvt_common.cuh"

namespace mvt {

struct GridJobDev {
    const float *xyz;
    const unsigned char *mask;
    const float *spacing;
    GridSeg *segs;
    float *aabb;          // [B,8] min xyz, max xyz, 2 pad  (float bits, atomically reduced)
    unsigned *keys, *cnt;
    uint2 *vals;
    float4 *sorted;
    unsigned *pt_slot, *pt_rank, *cidx;
    int B, N;
    unsigned T;
    int tshift;
    int k_hint, index_mode;
    float cell_scale;
};

struct GridJobs {
    GridJobDev j[kMaxJobs];
    int n;
};

// float atomic min/max through the sign-aware integer trick (values are finite here)
__device__ __forceinline__ void atomic_min_f(float *a, float v) {
    v += 0.0f;   // -0 -> +0 (the integer trick orders -0 below every negative)
    if (v >= 0.f) atomicMin(reinterpret_cast<int *>(a), __float_as_int(v));
    else atomicMax(reinterpret_cast<unsigned *>(a), __float_as_uint(v));
}
__device__ __forceinline__ void atomic_max_f(float *a, float v) {
    v += 0.0f;
    if (v >= 0.f) atomicMax(reinterpret_cast<int *>(a), __float_as_int(v));
    else atomicMin(reinterpret_cast<unsigned *>(a), __float_as_uint(v));
}

__device__ __forceinline__ bool finite3(float x, float y, float z) {
    return fabsf(x) < 3.0e38f && fabsf(y) < 3.0e38f && fabsf(z) < 3.0e38f;   // false for NaN / Inf
}

// ---- 1. init: clear keys / counts, reset aabb + cursors -------------------------------------------------------
__global__ void grid_init_kernel(const GridJobs jobs) {
    const GridJobDev &J = jobs.j[blockIdx.y];
    size_t total = (size_t)J.B * J.T;
    size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        J.keys[i] = kEmptyKey;
        J.cnt[i] = 0u;
    }
    size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (gid < (size_t)J.B) {
        float *a = J.aabb + gid * 8;
        a[0] = a[1] = a[2] = INFINITY;
        a[3] = a[4] = a[5] = -INFINITY;
        a[6] = a[7] = 0.f;
        GridSeg *s = J.segs + gid;
        s->n_valid = 0; s->cursor = 0; s->n_cells = 0; s->pad = 0;
    }
}

// ---- 2. aabb: per-segment bounds over searchable (finite, mask-valid) points -----------------------------------
__global__ void grid_aabb_kernel(const GridJobs jobs) {
    const GridJobDev &J = jobs.j[blockIdx.z];
    const int b = blockIdx.y;
    if (b >= J.B) return;
    const float *xyz = J.xyz + (size_t)b * J.N * 3;
    const unsigned char *mk = J.mask ? J.mask + (size_t)b * J.N : nullptr;
    float mn[3] = {INFINITY, INFINITY, INFINITY}, mx[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < J.N; i += gridDim.x * blockDim.x) {
        float x = xyz[i * 3], y = xyz[i * 3 + 1], z = xyz[i * 3 + 2];
        if ((mk && !mk[i]) || !finite3(x, y, z)) continue;
        mn[0] = fminf(mn[0], x); mn[1] = fminf(mn[1], y); mn[2] = fminf(mn[2], z);
        mx[0] = fmaxf(mx[0], x); mx[1] = fmaxf(mx[1], y); mx[2] = fmaxf(mx[2], z);
    }
#pragma unroll
    for (int a = 0; a < 3; ++a) {
#pragma unroll
        for (int o = 16; o; o >>= 1) {
            mn[a] = fminf(mn[a], __shfl_xor_sync(0xffffffffu, mn[a], o));
            mx[a] = fmaxf(mx[a], __shfl_xor_sync(0xffffffffu, mx[a], o));
        }
    }
    if ((threadIdx.x & 31) == 0 && mn[0] <= mx[0]) {
        float *A = J.aabb + (size_t)b * 8;
#pragma unroll
        for (int a = 0; a < 3; ++a) { atomic_min_f(A + a, mn[a]); atomic_max_f(A + 3 + a, mx[a]); }
    }
}

// ---- 3. plan: cell size and grid dims per segment ---------------------------------------------------------------
// With the pixel-footprint hint (mean of depth*stride/fx over valid pixels, from cloud_build):
//     h = cell_scale * footprint * sqrt(k/16)        (k-NN radius on a surface sampled at that footprint)
// otherwise a measure-based estimate from the bounding box: h = cell_scale * (measure * k / n)^(1/d) over the
// non-degenerate axes.  h is raised so that no axis needs more than 1023 cells.
__global__ void grid_plan_kernel(const GridJobs jobs) {
    const GridJobDev &J = jobs.j[blockIdx.y];
    int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= J.B) return;
    const float *A = J.aabb + (size_t)b * 8;
    GridSeg *s = J.segs + b;
    float kk = (float)(J.k_hint > 0 ? J.k_hint : 16);
    if (!(A[0] <= A[3])) {   // no searchable point
        s->ox = s->oy = s->oz = 0.f; s->h = 1.f; s->inv_h = 1.f; s->nx = s->ny = s->nz = 1;
        return;
    }
    float ex = A[3] - A[0], ey = A[4] - A[1], ez = A[5] - A[2];
    float emax = fmaxf(ex, fmaxf(ey, ez));
    float h = 0.f;
    float scale = J.cell_scale > 0.f ? J.cell_scale : 1.f;
    if (J.spacing && J.spacing[b * 2 + 1] > 0.f) {
        float fp = J.spacing[b * 2] / J.spacing[b * 2 + 1];
        h = 2.0f * scale * fp * sqrtf(fmaxf(kk, 4.f) / 16.f);
    } else {
        float measure = 1.f; int d = 0;
        float eps = 1e-6f * emax;
        if (ex > eps) { measure *= ex; ++d; }
        if (ey > eps) { measure *= ey; ++d; }
        if (ez > eps) { measure *= ez; ++d; }
        float n = fmaxf((float)J.N, 1.f);
        if (d > 0) h = 0.75f * scale * powf(measure * kk / n, 1.f / (float)d);
    }
    float hmin = emax / (float)(kMaxGridDim - 1);
    if (!(h > hmin)) h = hmin;
    if (!(h > 0.f)) h = 1.f;   // all points identical
    s->ox = A[0]; s->oy = A[1]; s->oz = A[2];
    s->h = h; s->inv_h = 1.f / h;
    int nx = (int)floorf(ex / h) + 1, ny = (int)floorf(ey / h) + 1, nz = (int)floorf(ez / h) + 1;
    s->nx = min(max(nx, 1), kMaxGridDim); s->ny = min(max(ny, 1), kMaxGridDim); s->nz = min(max(nz, 1), kMaxGridDim);
}

// ---- 4. count: insert each point's cell into the hash table, remember (slot, rank in cell) ----------------------
__global__ void grid_count_kernel(const GridJobs jobs) {
    const GridJobDev &J = jobs.j[blockIdx.z];
    const int b = blockIdx.y;
    if (b >= J.B) return;
    const GridSeg sg = J.segs[b];
    const float *xyz = J.xyz + (size_t)b * J.N * 3;
    const unsigned char *mk = J.mask ? J.mask + (size_t)b * J.N : nullptr;
    unsigned *keys = J.keys + (size_t)b * J.T, *cnt = J.cnt + (size_t)b * J.T;
    const unsigned tmask = J.T - 1;
    unsigned local_valid = 0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < J.N; i += gridDim.x * blockDim.x) {
        float x = xyz[i * 3], y = xyz[i * 3 + 1], z = xyz[i * 3 + 2];
        unsigned slot = kEmptyKey, rank = 0;
        if (!(mk && !mk[i]) && finite3(x, y, z)) {
            unsigned key = cell_key(cell_coord(x, sg.ox, sg.inv_h, sg.nx), cell_coord(y, sg.oy, sg.inv_h, sg.ny),
                                    cell_coord(z, sg.oz, sg.inv_h, sg.nz));
            slot = cell_hash(key, J.tshift);
            while (true) {
                unsigned cur = reinterpret_cast<volatile unsigned *>(keys)[slot];
                if (cur == key) break;
                if (cur == kEmptyKey) {
                    unsigned old = atomicCAS(keys + slot, kEmptyKey, key);
                    if (old == kEmptyKey || old == key) break;
                }
                slot = (slot + 1) & tmask;
            }
            rank = atomicAdd(cnt + slot, 1u);
            ++local_valid;
        }
        J.pt_slot[(size_t)b * J.N + i] = slot;
        J.pt_rank[(size_t)b * J.N + i] = rank;
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) local_valid += __shfl_xor_sync(0xffffffffu, local_valid, o);
    if ((threadIdx.x & 31) == 0 && local_valid) atomicAdd(&J.segs[b].n_valid, local_valid);
}

// ---- 4b. (index_mode 1 only) rank of every point among the mask-valid points of its segment, original order ------
__global__ void __launch_bounds__(1024) grid_cidx_kernel(const GridJobs jobs) {
    const GridJobDev &J = jobs.j[blockIdx.y];
    const int b = blockIdx.x;
    if (b >= J.B || J.index_mode != 1) return;
    __shared__ unsigned wsum[32];
    __shared__ unsigned carry_s;
    const unsigned char *mk = J.mask ? J.mask + (size_t)b * J.N : nullptr;
    unsigned *out = J.cidx + (size_t)b * J.N;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) carry_s = 0;
    __syncthreads();
    for (int base = 0; base < J.N; base += 1024) {
        int i = base + threadIdx.x;
        const unsigned carry = carry_s;   // written before the previous iteration's last barrier
        unsigned v = (i < J.N) ? (mk ? (mk[i] != 0) : 1u) : 0u;
        unsigned inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            unsigned t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        if (lane == 31) wsum[warp] = inc;
        __syncthreads();
        if (warp == 0) {
            unsigned w = wsum[lane], winc = w;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                unsigned t = __shfl_up_sync(0xffffffffu, winc, o);
                if (lane >= o) winc += t;
            }
            wsum[lane] = winc - w;   // exclusive warp offsets
            if (lane == 31) carry_s = carry + winc;
        }
        __syncthreads();
        if (i < J.N) out[i] = carry + wsum[warp] + inc - v;
        __syncthreads();
    }
}

// ---- 5. alloc: give every occupied cell a contiguous range of `sorted` ------------------------------------------
__global__ void __launch_bounds__(256) grid_alloc_kernel(const GridJobs jobs) {
    const GridJobDev &J = jobs.j[blockIdx.z];
    const int b = blockIdx.y;
    if (b >= J.B) return;
    __shared__ unsigned wsum[8], wocc[8];
    __shared__ unsigned base_s;
    const unsigned *cnt = J.cnt + (size_t)b * J.T;
    uint2 *vals = J.vals + (size_t)b * J.T;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (unsigned base = blockIdx.x * 256u; base < J.T; base += gridDim.x * 256u) {
        unsigned slot = base + threadIdx.x;
        unsigned c = slot < J.T ? cnt[slot] : 0u;
        unsigned inc = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            unsigned t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        unsigned occ = __popc(__ballot_sync(0xffffffffu, c > 0));
        if (lane == 31) { wsum[warp] = inc; wocc[warp] = occ; }
        __syncthreads();
        if (threadIdx.x == 0) {
            unsigned tot = 0, to = 0;
            for (int w = 0; w < 8; ++w) { unsigned t = wsum[w]; wsum[w] = tot; tot += t; to += wocc[w]; }
            base_s = tot ? atomicAdd(&J.segs[b].cursor, tot) : 0u;
            if (to) atomicAdd(&J.segs[b].n_cells, to);
        }
        __syncthreads();
        if (c) vals[slot] = make_uint2(base_s + wsum[warp] + inc - c, c);
        __syncthreads();
    }
}

// ---- 6. scatter: write {x,y,z,index} records cell by cell -------------------------------------------------------
__global__ void grid_scatter_kernel(const GridJobs jobs) {
    const GridJobDev &J = jobs.j[blockIdx.z];
    const int b = blockIdx.y;
    if (b >= J.B) return;
    const float *xyz = J.xyz + (size_t)b * J.N * 3;
    const uint2 *vals = J.vals + (size_t)b * J.T;
    float4 *sorted = J.sorted + (size_t)b * J.N;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < J.N; i += gridDim.x * blockDim.x) {
        unsigned slot = J.pt_slot[(size_t)b * J.N + i];
        if (slot == kEmptyKey) continue;
        unsigned pos = vals[slot].x + J.pt_rank[(size_t)b * J.N + i];
        int idx = (J.index_mode == 1) ? (int)J.cidx[(size_t)b * J.N + i] : i;
        sorted[pos] = make_float4(xyz[i * 3], xyz[i * 3 + 1], xyz[i * 3 + 2], __int_as_float(idx));
    }
}

__global__ void grid_info_kernel(const GridSeg *segs, int B, unsigned T, float *info) {
    int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= B) return;
    const GridSeg s = segs[b];
    float *o = info + (size_t)b * 16;
    o[0] = s.ox; o[1] = s.oy; o[2] = s.oz; o[3] = s.h;
    o[4] = (float)s.nx; o[5] = (float)s.ny; o[6] = (float)s.nz;
    o[7] = (float)s.n_valid; o[8] = (float)s.n_cells; o[9] = (float)T; o[10] = (float)s.cursor;
    for (int i = 11; i < 16; ++i) o[i] = 0.f;
}

}  // namespace mvt

// ---------------------------------------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------------------------------------
MVT_EXPORT size_t mvt_grid_workspace_bytes(int B, int N) {
    if (B <= 0 || N <= 0) return 0;
    return mvt::grid_layout(B, N).bytes;
}

MVT_EXPORT int mvt_grid_build(const mvt_grid_job *jobs, int n_jobs, void *stream) {
    using namespace mvt;
    MVT_CHECK_ARG(jobs && n_jobs > 0 && n_jobs <= kMaxJobs, "mvt_grid_build: n_jobs must be 1..%d (got %d)", kMaxJobs, n_jobs);
    GridJobs G;
    G.n = n_jobs;
    int maxB = 0;
    size_t maxN = 0, maxBT = 0;
    bool any_cidx = false;
    for (int i = 0; i < n_jobs; ++i) {
        const mvt_grid_job &u = jobs[i];
        MVT_CHECK_ARG(u.xyz && u.workspace, "mvt_grid_build: job %d has a null pointer", i);
        MVT_CHECK_ARG(u.B > 0 && u.N > 0 && u.B <= 65535, "mvt_grid_build: job %d bad shape B=%d N=%d", i, u.B, u.N);
        MVT_CHECK_ARG((size_t)u.N < (1u << 30), "mvt_grid_build: N too large");
        MVT_CHECK_ARG(((uintptr_t)u.workspace & 255) == 0, "mvt_grid_build: workspace must be 256-byte aligned");
        MVT_CHECK_ARG(u.index_mode == 0 || (u.index_mode == 1 && u.mask), "mvt_grid_build: index_mode %d needs a mask", u.index_mode);
        GridLayout L = grid_layout(u.B, u.N);
        char *p = (char *)u.workspace;
        GridJobDev &d = G.j[i];
        d.xyz = u.xyz; d.mask = u.mask; d.spacing = u.spacing;
        d.segs = (GridSeg *)(p + L.off_seg);
        d.aabb = (float *)(p + L.off_aabb);
        d.keys = (unsigned *)(p + L.off_keys);
        d.cnt = (unsigned *)(p + L.off_cnt);
        d.vals = (uint2 *)(p + L.off_vals);
        d.sorted = (float4 *)(p + L.off_sorted);
        d.pt_slot = (unsigned *)(p + L.off_slot);
        d.pt_rank = (unsigned *)(p + L.off_rank);
        d.cidx = (unsigned *)(p + L.off_cidx);
        d.B = u.B; d.N = u.N; d.T = L.T; d.tshift = L.tshift;
        d.k_hint = u.k_hint; d.index_mode = u.index_mode; d.cell_scale = u.cell_scale;
        maxB = u.B > maxB ? u.B : maxB;
        maxN = (size_t)u.N > maxN ? (size_t)u.N : maxN;
        size_t bt = (size_t)u.B * L.T;
        maxBT = bt > maxBT ? bt : maxBT;
        any_cidx |= u.index_mode == 1;
    }
    for (int i = n_jobs; i < kMaxJobs; ++i) G.j[i] = G.j[0];
    cudaStream_t st = (cudaStream_t)stream;
    auto blocks = [](size_t items, int per_block, int cap) {
        size_t b = (items + per_block - 1) / per_block;
        return (unsigned)(b < 1 ? 1 : (b > (size_t)cap ? (size_t)cap : b));
    };
    {
        dim3 g(blocks(maxBT, 256 * 4, kNumSMs * 8), n_jobs);
        if (g.x * 256u < (unsigned)maxB) g.x = (maxB + 255) / 256;
        grid_init_kernel<<<g, 256, 0, st>>>(G);
        MVT_CHECK_LAUNCH("grid_init");
    }
    // per-segment point loops: x blocks cover N, y = segment, z = job
    unsigned xb = blocks(maxN, 256 * 4, (kNumSMs * 8 + maxB - 1) / maxB);
    {
        dim3 g(xb, maxB, n_jobs);
        grid_aabb_kernel<<<g, 256, 0, st>>>(G);
        MVT_CHECK_LAUNCH("grid_aabb");
        grid_plan_kernel<<<dim3((maxB + 63) / 64, n_jobs), 64, 0, st>>>(G);
        MVT_CHECK_LAUNCH("grid_plan");
        dim3 gc(blocks(maxN, 256, (kNumSMs * 16 + maxB - 1) / maxB), maxB, n_jobs);
        grid_count_kernel<<<gc, 256, 0, st>>>(G);
        MVT_CHECK_LAUNCH("grid_count");
        if (any_cidx) {
            grid_cidx_kernel<<<dim3(maxB, n_jobs), 1024, 0, st>>>(G);
            MVT_CHECK_LAUNCH("grid_cidx");
        }
        size_t maxT = maxBT / (size_t)1;   // upper bound on T of any job (B>=1)
        dim3 ga(blocks(maxT, 256, (kNumSMs * 16 + maxB - 1) / maxB), maxB, n_jobs);
        grid_alloc_kernel<<<ga, 256, 0, st>>>(G);
        MVT_CHECK_LAUNCH("grid_alloc");
        grid_scatter_kernel<<<gc, 256, 0, st>>>(G);
        MVT_CHECK_LAUNCH("grid_scatter");
    }
    return MVT_OK;
}

MVT_EXPORT int mvt_grid_info(const void *grid, int B, int N, float *info, void *stream) {
    using namespace mvt;
    MVT_CHECK_ARG(grid && info && B > 0 && N > 0, "mvt_grid_info: bad argument");
    GridView g = grid_view(grid, B, N);
    grid_info_kernel<<<(B + 63) / 64, 64, 0, (cudaStream_t)stream>>>(g.segs, B, g.T, info);
    MVT_CHECK_LAUNCH("grid_info");
    return MVT_OK;
}
