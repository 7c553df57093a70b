// Host entry of the GPU BVH8 builder (mrt_bvh_build.cu).
#pragma once
#include "mrt_types.h"

namespace mrt {

struct Bvh8 {
    Node8 *nodes = nullptr;       // device
    TriGeom *tri_geom = nullptr;  // device, leaf order
    uint32_t n_nodes = 0;
    uint32_t n_prims = 0;
    uint32_t levels = 0;
    float build_ms = 0.0f;        // device time of the whole chain (CUDA events on the build stream)
    size_t bytes() const { return static_cast<size_t>(n_nodes) * sizeof(Node8) + static_cast<size_t>(n_prims) * sizeof(TriGeom); }
};

// d_tris9: device array of n x 9 floats (A, AB, AC) in primitive-id order. n may be 0 (an empty root is produced).
// flags bit 0: rebuild the upper tree over LBVH clusters with binned SAH.
cudaError_t build_bvh8(const float *d_tris9, int n, uint32_t flags, cudaStream_t stream, Bvh8 *out);
void free_bvh8(Bvh8 *b);

}  // namespace mrt
