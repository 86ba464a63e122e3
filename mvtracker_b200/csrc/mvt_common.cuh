// Shared internals of libmvt.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include "../../include/mvt.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "libmvt is written for sm_100a (B200) only"
#endif

namespace mvt {

void set_error(const char *fmt, ...);
void count_launch();

#define MVT_EXPORT extern "C" __attribute__((visibility("default")))

#define MVT_CHECK_ARG(cond, ...)                     \
    do {                                             \
        if (!(cond)) {                               \
            mvt::set_error(__VA_ARGS__);             \
            return MVT_E_ARG;                        \
        }                                            \
    } while (0)

#define MVT_CHECK_LAUNCH(what)                                                        \
    do {                                                                              \
        cudaError_t e__ = cudaGetLastError();                                         \
        if (e__ != cudaSuccess) {                                                     \
            mvt::set_error("%s: %s", what, cudaGetErrorString(e__));                  \
            return MVT_E_CUDA;                                                        \
        }                                                                             \
        mvt::count_launch();                                                          \
    } while (0)

constexpr int kNumSMs = 148;   // B200
constexpr unsigned kEmptyKey = 0xFFFFFFFFu;
constexpr int kMaxGridDim = 1023;   // 10 bits per axis in the packed cell key
constexpr int kMaxJobs = 8;
constexpr int kCamStride = MVT_CAM_STRIDE;   // floats per (view, frame) camera record

// The same 16-byte-record buffer on every rank of the node (symmetric memory): stores go to all ranks at once through
// the NVSwitch multicast address when there is one (one multimem.st per record), else to each rank's mapping in turn.
struct PeersDev {
    float4 *mc;
    float4 *ptr[MVT_MAX_PEERS];
    int n;      // 0 = not shared (nothing is stored)
};
__device__ __forceinline__ void store_all(const PeersDev &P, size_t i, float4 v) {
    if (P.mc) {
        asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(P.mc + i), "f"(v.x), "f"(v.y),
                     "f"(v.z), "f"(v.w)
                     : "memory");
    } else {
        for (int r = 0; r < P.n; ++r) P.ptr[r][i] = v;
    }
}
__device__ __forceinline__ void store_all_f32(const PeersDev &P, size_t i, float v) {      // i in floats
    if (P.mc) {
        asm volatile("multimem.st.relaxed.sys.global.f32 [%0], %1;" ::"l"(reinterpret_cast<float *>(P.mc) + i), "f"(v) : "memory");
    } else {
        for (int r = 0; r < P.n; ++r) reinterpret_cast<float *>(P.ptr[r])[i] = v;
    }
}
__host__ inline PeersDev peers_dev(const mvt_peers *u) {
    PeersDev d;
    d.mc = nullptr; d.n = 0;
    for (int r = 0; r < MVT_MAX_PEERS; ++r) d.ptr[r] = nullptr;
    if (u) {
        d.mc = (float4 *)u->mc; d.n = u->n;
        for (int r = 0; r < u->n && r < MVT_MAX_PEERS; ++r) d.ptr[r] = (float4 *)u->ptr[r];
    }
    return d;
}

// Per-segment grid parameters, written on device by grid_plan.
struct __align__(16) GridSeg {
    float ox, oy, oz, h;
    float inv_h;
    int nx, ny, nz;
    unsigned n_valid;    // points inserted into the grid
    unsigned cursor;     // range allocator (== n_valid when the build is done)
    unsigned n_cells;    // occupied cells
    unsigned pad;
};

// Workspace layout of one grid (B segments x N points); the same function is used by build and query.
struct GridLayout {
    unsigned T;          // hash slots per segment (power of two, >= 2N)
    int tshift;          // 32 - log2(T)
    size_t off_seg, off_aabb, off_keys, off_cnt, off_vals, off_sorted, off_slot, off_rank, off_cidx, bytes;
};

__host__ __device__ inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

__host__ inline GridLayout grid_layout(int B, int N) {
    GridLayout L;
    unsigned T = 64;
    int lg = 6;
    while ((size_t)T < 2 * (size_t)N) { T <<= 1; ++lg; }
    L.T = T;
    L.tshift = 32 - lg;
    size_t o = 0;
    L.off_seg = o;    o = align_up(o + sizeof(GridSeg) * (size_t)B, 256);
    L.off_aabb = o;   o = align_up(o + sizeof(float) * 8 * (size_t)B, 256);
    L.off_keys = o;   o = align_up(o + sizeof(unsigned) * (size_t)B * T, 256);
    L.off_cnt = o;    o = align_up(o + sizeof(unsigned) * (size_t)B * T, 256);
    L.off_vals = o;   o = align_up(o + sizeof(uint2) * (size_t)B * T, 256);
    L.off_sorted = o; o = align_up(o + sizeof(float4) * (size_t)B * N, 256);
    L.off_slot = o;   o = align_up(o + sizeof(unsigned) * (size_t)B * N, 256);
    L.off_rank = o;   o = align_up(o + sizeof(unsigned) * (size_t)B * N, 256);
    L.off_cidx = o;   o = align_up(o + sizeof(unsigned) * (size_t)B * N, 256);
    L.bytes = o;
    return L;
}

// Device view of one built grid (all pointers into the workspace).
struct GridView {
    const GridSeg *segs;
    const unsigned *keys;
    const uint2 *vals;
    const float4 *sorted;   // {x, y, z, bit-cast index}
    unsigned T;
    int tshift;
    int N;
    unsigned long long *stats;   // optional debug counters (mvt_debug_set_stats), else nullptr
};

unsigned long long *debug_stats_ptr();

__host__ inline GridView grid_view(const void *ws, int B, int N) {
    GridLayout L = grid_layout(B, N);
    const char *p = (const char *)ws;
    GridView g;
    g.segs = (const GridSeg *)(p + L.off_seg);
    g.keys = (const unsigned *)(p + L.off_keys);
    g.vals = (const uint2 *)(p + L.off_vals);
    g.sorted = (const float4 *)(p + L.off_sorted);
    g.T = L.T;
    g.tshift = L.tshift;
    g.N = N;
    g.stats = debug_stats_ptr();
    return g;
}

__device__ __forceinline__ unsigned cell_key(int x, int y, int z) { return (unsigned)x | ((unsigned)y << 10) | ((unsigned)z << 20); }
__device__ __forceinline__ unsigned cell_hash(unsigned key, int tshift) { return (key * 2654435761u) >> tshift; }

__device__ __forceinline__ int cell_coord(float x, float o, float inv_h, int n) {
    float f = floorf((x - o) * inv_h);
    // NaN -> comparisons false -> 0
    int c = (f > 0.f) ? (f < (float)(n - 1) ? (int)f : n - 1) : 0;
    return c;
}

// Grid of a work-drawing kernel (256-thread blocks, `per_block` items in flight per block): one wave of resident blocks
// (SMs x blocks_per_sm), fewer when there is less work than that.
__host__ inline unsigned one_wave_grid(long long n_items, int per_block, int blocks_per_sm) {
    static int sms[64] = {0};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64) dev = 0;
    if (sms[dev] == 0) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        sms[dev] = n;
    }
    long long wave = (long long)sms[dev] * blocks_per_sm, need = (n_items + per_block - 1) / per_block;
    return (unsigned)(need < wave ? (need > 0 ? need : 1) : wave);
}

}  // namespace mvt
