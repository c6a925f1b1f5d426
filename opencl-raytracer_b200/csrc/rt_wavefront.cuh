// rt_wavefront.cuh -- host interface of the wavefront pipeline (rt_wavefront.cu).
#pragma once

#include "rt_kernels.cuh"
#include "rt_grid3.cuh"
#include "rt_pgrid.cuh"
#include "rt_wave.cuh"

namespace rt {

struct WaveBuffers {
    PathArrays pa;      // capacity paths
    TaskQueue q[2];     // capacity tasks each
    float* res_t;       // capacity
    int32_t* res_id;    // capacity
    uint32_t* counts;   // 4 * n_counts words: queue sizes (class B, class A) per generation, then the fetch cursors likewise
    uint32_t n_counts;
    uint64_t capacity;
    uint32_t qcap;      // entries of one task queue of this launch sequence (class A tasks are stored from the back)
};

// The perspective grids of the camera and of the lights (rt_pgrid.cuh); valid_mask == 0: none.
struct WaveGrids {
    GridSet cam, lit;
    Grid3 g3;  // uniform grid for the rays without a common apex (valid == 0: they walk the BVH)
    const float4* filter_tiles = nullptr;  // accel 1: the filter records packed into the tile layout (wave_pack_filter_tiles)
};

// accel 1 (wave_trace_filter): the scene's filter records in the layout its shared-memory tiles have, padded to whole tiles,
// so that a tile is one bulk copy.  `out` holds 2 * wave_filter_tile_pad(n_objs) float4.
uint32_t wave_filter_tile_pad(uint32_t n_objs);
cudaError_t wave_pack_filter_tiles(const Scene& sc, float4* out, cudaStream_t stream);

// Upper bound on the number of trace/advance generations of one frame.
uint32_t wave_max_generations(uint32_t max_bounces, uint32_t n_lights, int mode);

// Enqueue every kernel of one frame on `stream` (no host synchronisation inside).
// accel: 0 exact brute force, 1 filter (N-body tiles), 2 BVH, 3 perspective grids for camera rays and point-light shadow rays
// (grids != nullptr) + BVH for everything else.  *launches receives the number of kernels enqueued.  tail_gen > 0 (BVH /
// grid searches only): generations tail_gen, tail_gen + 1, ... are finished by one wave_tail launch.
cudaError_t wave_render(const Scene& sc, const Frame& fr, const WaveBuffers& wb, int accel, const WaveGrids* grids, int sm_count,
                        uint32_t* launches, cudaStream_t stream, int tail_gen = 0);

}  // namespace rt
