// rt_lbvh.cuh -- device-side BVH build (rt_lbvh.cu): Morton sort + Karras radix tree + bottom-up refit.
#pragma once

#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

#include "rt_layout.h"
#include "rt_scene.hpp"

namespace rt {

size_t lbvh_scratch_bytes(uint32_t n_leaves);

// n >= 2 leaf records (device) -> n - 1 nodes (node 0 = root), leaf_obj / leaf_cull in leaf (Morton) order.
// lo / hi: bounds of the leaf centres.  All work is enqueued on `stream`; nothing is synchronised.
cudaError_t build_lbvh_device(const LbvhLeaf* d_leaves, uint32_t n, const CullRec* d_cull, const float lo[3], const float hi[3],
                              float omax, void* scratch, size_t scratch_bytes, BvhNode* d_nodes, uint32_t* d_leaf_obj,
                              CullRec* d_leaf_cull, cudaStream_t stream);

}  // namespace rt
