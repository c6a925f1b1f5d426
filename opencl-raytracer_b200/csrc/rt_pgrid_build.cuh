// rt_pgrid_build.cuh -- host interface of the device-side perspective-grid builder (rt_pgrid.cu).
#pragma once

#include <cuda_runtime.h>
#include <stddef.h>

#include "rt_grid3.cuh"
#include "rt_pgrid.cuh"

namespace rt {

size_t pgrid_scan_bytes(uint32_t n_cells);

// Pass 1: per-cell counts of the rectangles of every (object, grid) pair, then their exclusive scan into
// d_cell_start[0 .. n_cells] (d_cell_start[n_cells] = number of entries).  d_cursor: n_cells + 1 words of scratch.
cudaError_t pgrid_count(const PGrid* d_grids, uint32_t n_grids, uint32_t n_cells, const CullRec* d_cull, uint32_t n_objs, float omax,
                        uint32_t* d_cursor, uint32_t* d_cell_start, void* scan_tmp, size_t scan_bytes, cudaStream_t stream);
// Pass 2: the entries, every cell's list sorted by (dnear, obj).
cudaError_t pgrid_fill(const PGrid* d_grids, uint32_t n_grids, uint32_t n_cells, const CullRec* d_cull, uint32_t n_objs, float omax,
                       uint32_t* d_cursor, const uint32_t* d_cell_start, GridEntry* d_entries, cudaStream_t stream);

// The uniform 3-D grid (rt_grid3.cuh), same two passes; g.cell_start / g.entries are not read.  d_aabb: per object the
// noise-inflated half extents (HostScene::aabb).
cudaError_t grid3_count(const Grid3& g, const CullRec* d_cull, const float4* d_aabb, uint32_t n_objs, uint32_t* d_cursor, uint32_t* d_cell_start,
                        void* scan_tmp, size_t scan_bytes, cudaStream_t stream);
cudaError_t grid3_fill(const Grid3& g, const CullRec* d_cull, const float4* d_aabb, uint32_t n_objs, uint32_t* d_cursor, const uint32_t* d_cell_start,
                       uint32_t* d_entries, cudaStream_t stream);

}  // namespace rt
