// Per-row arithmetic of the correlation, shared by the fused single-GPU kernel (mvt_knn.cu) and the sharded kernels
// (mvt_shard.cu): the same per-lane fma chain and the same butterfly order over the lanes of a row, so that a record is
// bit-identical whichever kernel -- and whichever rank -- computes it.
#pragma once
#include "mvt_knn.cuh"

namespace mvt {

// 16 partial sums per lane (one per neighbour) -> lane L ends with the warp total of neighbour L>>1 (16 shuffles).
__device__ __forceinline__ float transpose_reduce16(float (&pr)[16], int lane) {
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
    return pr[0] + __shfl_xor_sync(FULL, pr[0], 1);
}

// 8 partial sums per lane, reduced over the 16 lanes of each half-warp -> lane l of a half holds partial (l&15)>>1.
__device__ __forceinline__ float transpose_reduce8_half(float (&pr)[8], int lane) {
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
    return pr[0] + __shfl_xor_sync(FULL, pr[0], 1);
}

// 16 partial sums per lane, reduced over the 16 lanes of a half-warp group -> group lane l holds the total of partial l.
template <int GW>
__device__ __forceinline__ float transpose_reduce16_half(const Group<GW> &g, float (&pr)[16]) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        bool hi = g.gl & 8;
        float send = hi ? pr[i] : pr[i + 8], keep = hi ? pr[i + 8] : pr[i];
        pr[i] = keep + g.shfl_xor(send, 8);
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        bool hi = g.gl & 4;
        float send = hi ? pr[i] : pr[i + 4], keep = hi ? pr[i + 4] : pr[i];
        pr[i] = keep + g.shfl_xor(send, 4);
    }
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        bool hi = g.gl & 2;
        float send = hi ? pr[i] : pr[i + 2], keep = hi ? pr[i + 2] : pr[i];
        pr[i] = keep + g.shfl_xor(send, 2);
    }
    {
        bool hi = g.gl & 1;
        float send = hi ? pr[0] : pr[1], keep = hi ? pr[1] : pr[0];
        pr[0] = keep + g.shfl_xor(send, 1);
    }
    return pr[0];
}

__device__ __forceinline__ float dot4(const float4 &a, const float4 &b) {
    return fmaf(a.w, b.w, fmaf(a.z, b.z, fmaf(a.y, b.y, a.x * b.x)));
}
__device__ __forceinline__ float dot8_bf16(const uint4 &a, const uint4 &b) {
    const __nv_bfloat162 *x = reinterpret_cast<const __nv_bfloat162 *>(&a);
    const __nv_bfloat162 *y = reinterpret_cast<const __nv_bfloat162 *>(&b);
    float acc = 0.f;
#pragma unroll
    for (int e = 0; e < 4; ++e) {
        float2 u = __bfloat1622float2(x[e]), v = __bfloat1622float2(y[e]);
        acc = fmaf(u.x, v.x, acc);
        acc = fmaf(u.y, v.y, acc);
    }
    return acc;
}

}  // namespace mvt
