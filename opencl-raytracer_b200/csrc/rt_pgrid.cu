// rt_pgrid.cu -- device-side construction of the perspective grids (rt_pgrid.cuh): count -> scan -> fill -> sort.
// Runs at scene upload (light grids) and when the camera changes (camera grid); about a millisecond for 10^5 objects.
#include "rt_pgrid_build.cuh"

#include <cub/cub.cuh>

namespace rt {
namespace {

constexpr int kB = 256;

// One thread per (object, grid): the cells of the object's rectangle are counted (FILL = false) or filled (FILL = true,
// cursor[] holds the next free slot of every cell).
template <bool FILL>
__global__ void __launch_bounds__(kB) pgrid_scatter(const PGrid* __restrict__ grids, const uint32_t n_grids, const CullRec* __restrict__ cull,
                                                    const uint32_t n_objs, const float omax, uint32_t* __restrict__ cursor,
                                                    const uint32_t* __restrict__ cell_start, GridEntry* __restrict__ entries) {
    const uint64_t idx = (uint64_t)blockIdx.x * kB + threadIdx.x;
    if (idx >= (uint64_t)n_objs * n_grids) return;
    const uint32_t obj = (uint32_t)(idx / n_grids), gi = (uint32_t)(idx % n_grids);
    const PGrid g = grids[gi];
    if (g.nu == 0u) return;
    const CullRec cr = cull[obj];
    float r, dnear;
    if (!pgrid_object_sphere(g, cr, omax, r, dnear)) return;
    int u0, u1, v0, v1;
    if (!pgrid_rect(g, cr.cx, cr.cy, cr.cz, r, u0, u1, v0, v1)) return;
    GridEntry e;
    e.obj = obj;
    e.dnear = dnear;
    for (int v = v0; v <= v1; ++v)
        for (int u = u0; u <= u1; ++u) {
            const uint32_t cell = g.cell_base + (uint32_t)v * g.nu + (uint32_t)u;
            const uint32_t k = atomicAdd(cursor + cell, 1u);
            if (FILL) entries[cell_start[cell] + k] = e;
        }
}

__device__ __forceinline__ bool entry_less(const GridEntry a, const GridEntry b) { return a.dnear < b.dnear || (a.dnear == b.dnear && a.obj < b.obj); }

// One thread per cell: order the list by (dnear, obj) -- nearest first, so that a ray can stop at the first entry beyond
// its current hit, and deterministic whatever order the atomics of the fill pass ran in.
__global__ void __launch_bounds__(kB) pgrid_sort(const uint32_t* __restrict__ cell_start, const uint32_t n_cells, GridEntry* __restrict__ entries) {
    const uint32_t c = blockIdx.x * kB + threadIdx.x;
    if (c >= n_cells) return;
    const uint32_t b = cell_start[c], n = cell_start[c + 1] - b;
    GridEntry* e = entries + b;
    if (n <= 32u) {  // insertion sort
        for (uint32_t i = 1; i < n; ++i) {
            const GridEntry x = e[i];
            uint32_t j = i;
            while (j > 0 && entry_less(x, e[j - 1])) {
                e[j] = e[j - 1];
                --j;
            }
            e[j] = x;
        }
        return;
    }
    // heap sort
    auto sift = [&](uint32_t start, uint32_t end) {
        uint32_t root = start;
        for (;;) {
            uint32_t child = 2 * root + 1;
            if (child >= end) break;
            if (child + 1 < end && entry_less(e[child], e[child + 1])) ++child;
            if (!entry_less(e[root], e[child])) break;
            const GridEntry t = e[root];
            e[root] = e[child];
            e[child] = t;
            root = child;
        }
    };
    for (uint32_t s = n / 2; s-- > 0;) sift(s, n);
    for (uint32_t end = n - 1; end > 0; --end) {
        const GridEntry t = e[0];
        e[0] = e[end];
        e[end] = t;
        sift(0, end);
    }
}

// ---- uniform 3-D grid (rt_grid3.cuh): one thread per object, the cells its conservative (noise-inflated) box touches
template <bool FILL>
__global__ void __launch_bounds__(kB) grid3_scatter(const Grid3 g, const CullRec* __restrict__ cull, const float4* __restrict__ aabb, const uint32_t n_objs,
                                                    uint32_t* __restrict__ cursor, const uint32_t* __restrict__ cell_start, uint32_t* __restrict__ entries) {
    const uint32_t obj = blockIdx.x * kB + threadIdx.x;
    if (obj >= n_objs) return;
    const CullRec cr = cull[obj];
    const float4 hb = aabb[obj];
    int lo[3], hi[3];
    if (!grid3_object_cells(g, cr, hb.x, hb.y, hb.z, hb.w, lo, hi)) return;
    for (int z = lo[2]; z <= hi[2]; ++z)
        for (int y = lo[1]; y <= hi[1]; ++y)
            for (int x = lo[0]; x <= hi[0]; ++x) {
                const uint32_t cell = ((uint32_t)z * g.ny + (uint32_t)y) * g.nx + (uint32_t)x;
                const uint32_t k = atomicAdd(cursor + cell, 1u);
                if (FILL) entries[cell_start[cell] + k] = obj;
            }
}

// One thread per cell: ascending object index (deterministic whatever order the atomics ran in).
__global__ void __launch_bounds__(kB) grid3_sort(const uint32_t* __restrict__ cell_start, const uint32_t n_cells, uint32_t* __restrict__ entries) {
    const uint32_t c = blockIdx.x * kB + threadIdx.x;
    if (c >= n_cells) return;
    const uint32_t b = cell_start[c], n = cell_start[c + 1] - b;
    uint32_t* e = entries + b;
    for (uint32_t i = 1; i < n; ++i) {  // lists are short by construction (the builder rejects grids with long ones)
        const uint32_t x = e[i];
        uint32_t j = i;
        while (j > 0 && x < e[j - 1]) {
            e[j] = e[j - 1];
            --j;
        }
        e[j] = x;
    }
}

}  // namespace

cudaError_t grid3_count(const Grid3& g, const CullRec* d_cull, const float4* d_aabb, uint32_t n_objs, uint32_t* d_cursor, uint32_t* d_cell_start,
                        void* scan_tmp, size_t scan_bytes, cudaStream_t stream) {
    const uint32_t n_cells = g.nx * g.ny * g.nz;
    cudaError_t e = cudaMemsetAsync(d_cursor, 0, sizeof(uint32_t) * ((size_t)n_cells + 1), stream);
    if (e != cudaSuccess) return e;
    if (n_objs) grid3_scatter<false><<<(n_objs + kB - 1) / kB, kB, 0, stream>>>(g, d_cull, d_aabb, n_objs, d_cursor, nullptr, nullptr);
    e = cub::DeviceScan::ExclusiveSum(scan_tmp, scan_bytes, d_cursor, d_cell_start, (int)(n_cells + 1), stream);
    if (e != cudaSuccess) return e;
    return cudaGetLastError();
}

cudaError_t grid3_fill(const Grid3& g, const CullRec* d_cull, const float4* d_aabb, uint32_t n_objs, uint32_t* d_cursor, const uint32_t* d_cell_start,
                       uint32_t* d_entries, cudaStream_t stream) {
    const uint32_t n_cells = g.nx * g.ny * g.nz;
    cudaError_t e = cudaMemsetAsync(d_cursor, 0, sizeof(uint32_t) * ((size_t)n_cells + 1), stream);
    if (e != cudaSuccess) return e;
    if (n_objs) grid3_scatter<true><<<(n_objs + kB - 1) / kB, kB, 0, stream>>>(g, d_cull, d_aabb, n_objs, d_cursor, d_cell_start, d_entries);
    if (n_cells) grid3_sort<<<(n_cells + kB - 1) / kB, kB, 0, stream>>>(d_cell_start, n_cells, d_entries);
    return cudaGetLastError();
}

size_t pgrid_scan_bytes(uint32_t n_cells) {
    size_t bytes = 0;
    cub::DeviceScan::ExclusiveSum(nullptr, bytes, (const uint32_t*)nullptr, (uint32_t*)nullptr, (int)(n_cells + 1));
    return bytes;
}

cudaError_t pgrid_count(const PGrid* d_grids, uint32_t n_grids, uint32_t n_cells, const CullRec* d_cull, uint32_t n_objs, float omax,
                        uint32_t* d_cursor, uint32_t* d_cell_start, void* scan_tmp, size_t scan_bytes, cudaStream_t stream) {
    cudaError_t e = cudaMemsetAsync(d_cursor, 0, sizeof(uint32_t) * ((size_t)n_cells + 1), stream);
    if (e != cudaSuccess) return e;
    const uint64_t jobs = (uint64_t)n_objs * n_grids;
    if (jobs) pgrid_scatter<false><<<(unsigned)((jobs + kB - 1) / kB), kB, 0, stream>>>(d_grids, n_grids, d_cull, n_objs, omax, d_cursor, nullptr, nullptr);
    e = cub::DeviceScan::ExclusiveSum(scan_tmp, scan_bytes, d_cursor, d_cell_start, (int)(n_cells + 1), stream);
    if (e != cudaSuccess) return e;
    return cudaGetLastError();
}

cudaError_t pgrid_fill(const PGrid* d_grids, uint32_t n_grids, uint32_t n_cells, const CullRec* d_cull, uint32_t n_objs, float omax,
                       uint32_t* d_cursor, const uint32_t* d_cell_start, GridEntry* d_entries, cudaStream_t stream) {
    cudaError_t e = cudaMemsetAsync(d_cursor, 0, sizeof(uint32_t) * ((size_t)n_cells + 1), stream);
    if (e != cudaSuccess) return e;
    const uint64_t jobs = (uint64_t)n_objs * n_grids;
    if (jobs) pgrid_scatter<true><<<(unsigned)((jobs + kB - 1) / kB), kB, 0, stream>>>(d_grids, n_grids, d_cull, n_objs, omax, d_cursor, d_cell_start, d_entries);
    if (n_cells) pgrid_sort<<<(n_cells + kB - 1) / kB, kB, 0, stream>>>(d_cell_start, n_cells, d_entries);
    return cudaGetLastError();
}

}  // namespace rt
