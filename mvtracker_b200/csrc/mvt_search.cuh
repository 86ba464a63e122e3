// What a query kernel searches in (device-side form of mvt_search), shared by the query kernels of mvt_knn.cu and
// the view-sharded kernels of mvt_shard.cu.
#pragma once
#include "mvt_raster.cuh"

namespace mvt {

struct SearchDev {
    int raster;          // 1: rv is valid, 0: g is valid
    GridView g;
    const float *xyz;    // grid mode neighbour coordinates [B,N,3]
    RasterView rv;
    int N;
};

// coordinates of neighbour `i` of batch element b
template <bool RASTER_ONLY = false>
__device__ __forceinline__ float3 nbr_xyz(const SearchDev &sd, int b, int i) {
    if (RASTER_ONLY || sd.raster) {
        float4 p = __ldg(sd.rv.xyz4 + (size_t)b * sd.N + i);
        return make_float3(p.x, p.y, p.z);
    }
    const float *px = sd.xyz + ((size_t)b * sd.N + i) * 3;
    return make_float3(__ldg(px), __ldg(px + 1), __ldg(px + 2));
}


// Validates a user descriptor and fills the device form (defined in mvt_knn.cu).
int make_search(const mvt_search *u, int B, int need_xyz, SearchDev &sd, const char *who);

}  // namespace mvt
