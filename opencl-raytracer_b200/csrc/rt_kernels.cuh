// rt_kernels.cuh -- launch interface between the C ABI (rt_api.cu) and the kernels (rt_kernels.cu).
#pragma once

#ifndef RT_HOST_EMULATION
#include <cuda_runtime.h>
#endif
#include <stdint.h>

#include "rt_device.cuh"

namespace rt {

struct Counters {  // device-side, filled only when counting is on (RT_OPT_COUNT_RAYS)
    unsigned long long ref_primary, ref_shadow, ref_reflect;  // rays the reference kernel traces for this frame
    unsigned long long traced_shadow, traced_reflect;         // rays this path actually traced (dead ones skipped)
    unsigned long long pair_tests;                            // ray x object filter tests (or, brute: exact tests)
    unsigned long long exact_tests;                           // pairs that ran the reference arithmetic
    unsigned long long node_visits;                           // BVH inner nodes visited
    unsigned long long hit_pixels;
};

struct Frame {
    // pixel source: either the pinhole camera or an explicit ray array
    const Ray* rays;  // nullptr -> camera
    uint32_t width, height;
    float zdir;
    // shard: row-blocks of block_rows rows dealt round-robin; this context owns b % n_shards == shard
    uint32_t block_rows, shard, n_shards;
    uint32_t row0;            // added to every row (a band of a banded whole-frame render; 0 otherwise)
    uint32_t lrow0;           // first local row of this launch (a band of a banded sharded render; 0 otherwise)
    uint32_t local_rows;      // rows of this shard (camera) handled by this launch -- local image is width x local_rows
    uint64_t n_local;         // pixels / rays handled by this launch
    uint32_t max_bounces;
    int mode;                 // rt_mode
    int skip_dead;            // 1: do not trace rays whose result cannot reach the output
    float4* out;              // full-frame framebuffer (index = global pixel index); may be a peer pointer
    int32_t* hit_id;          // optional AOVs (global pixel index)
    float* hit_t;
    Counters* counters;       // nullptr: no counting
};

// accel: 0 brute (exact arithmetic on every pair), 1 cull, 2 bvh
cudaError_t launch_megakernel(const Scene& sc, const Frame& fr, int accel, bool aov_only, cudaStream_t stream);
cudaError_t launch_trace(const Scene& sc, const Ray* rays, uint64_t n, int accel, int32_t* hit_id, float* hit_t,
                         Counters* counters, cudaStream_t stream);
cudaError_t launch_quantize(const float4* rgba, uint64_t n, uint8_t* rgb, unsigned long long* n_negative, cudaStream_t stream);
cudaError_t launch_fill(float4* out, uint64_t n, cudaStream_t stream);
// register-only FFMA loop; returns flops executed through *flops, kernel on `stream`
cudaError_t launch_fp32_peak(float* sink, int iters, int blocks, int threads, bool packed, cudaStream_t stream);

}  // namespace rt
