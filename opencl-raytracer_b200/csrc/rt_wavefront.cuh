// rt_wavefront.cuh -- host interface of the wavefront pipeline (rt_wavefront.cu).
#pragma once

#include "rt_kernels.cuh"
#include "rt_wave.cuh"

namespace rt {

struct WaveBuffers {
    PathArrays pa;      // capacity paths
    TaskQueue q[2];     // capacity tasks each
    float* res_t;       // capacity
    int32_t* res_id;    // capacity
    uint32_t* counts;   // 2 * n_counts words: queue sizes per generation, then fetch cursors per generation
    uint32_t n_counts;
    uint64_t capacity;
};

// Upper bound on the number of trace/advance generations of one frame.
uint32_t wave_max_generations(uint32_t max_bounces, uint32_t n_lights, int mode);

// Enqueue every kernel of one frame on `stream` (no host synchronisation inside).
// accel: 0 exact brute force, 1 filter (N-body tiles), 2 BVH.  *launches receives the number of kernels enqueued.
cudaError_t wave_render(const Scene& sc, const Frame& fr, const WaveBuffers& wb, int accel, uint32_t leaf_max, int sm_count,
                        uint32_t* launches, cudaStream_t stream);

}  // namespace rt
