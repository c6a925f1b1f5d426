// rt_bounds.cuh -- host interface of the device-side per-object scene preparation (rt_bounds.cu).
#pragma once

#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

#include "../../include/rt_b200.h"
#include "rt_layout.h"
#include "rt_scene.hpp"

namespace rt {

enum { RT_BOUNDS_NONFINITE = 1, RT_BOUNDS_PROJECTIVE = 2, RT_BOUNDS_SINGULAR = 3, RT_BOUNDS_DEGENERATE = 4, RT_BOUNDS_ANISOTROPY = 5 };

struct BoundsGlobals {
    double r0;                     // max over the live objects of |centre| + bounding radius
    double wdev_max;               // largest deviation of a matrix's w entry from 1
    unsigned long long first_bad;  // (object index << 8) | code of the first object outside the supported domain, ~0: none
    float h_min, f_max;            // smallest filter coefficient h, largest noise coefficient f
    float lo[3], hi[3];            // bounds of the leaf centres, stored as ordered integers (bounds_decode_ordered)
    float box_lo[3], box_hi[3];    // bounds of the noise-inflated boxes of the live objects (uniform grid, rt_grid3.cuh), same coding
    uint32_t live;                 // objects that can be hit
};

size_t bounds_scratch_bytes(uint32_t n_objs);
void bounds_globals_init(BoundsGlobals* host_value);
float bounds_decode_ordered(float stored);

// pass 1: AoS -> SoA repack (minv, mv: 4 float4 per object; mat: 3) and the geometry of every object; fills r0 / wdev_max
cudaError_t bounds_pass1_launch(const rt_object* d_objs, uint32_t n, float4* minv, float4* mv, float4* mat, void* scratch, BoundsGlobals* d_g,
                                cudaStream_t stream);
// pass 2 (needs omax, a function of pass 1's r0): filter records and the leaves of the device tree builder
// aabb: per object the half extents of its box including the noise inflation (HostScene::aabb), w < 0: never hit
cudaError_t bounds_pass2_launch(const rt_object* d_objs, uint32_t n, const void* scratch, double omax, CullRec* cull, LbvhLeaf* leaves, float4* aabb,
                                BoundsGlobals* d_g, cudaStream_t stream);

}  // namespace rt
