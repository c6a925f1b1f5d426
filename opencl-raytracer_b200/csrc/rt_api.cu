// rt_api.cu -- the extern "C" layer of include/rt_b200.h: device/context, scene upload, launch, readback.
// This is what OpenCLRaytracer does against Boost.Compute (OpenCLRaytracer.cpp:13-105), on the CUDA runtime.
// No CPU fallback: without an sm_100 device rt_create fails.
#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <chrono>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/rt_b200.h"
#include "rt_kernels.cuh"
#include "rt_bounds.cuh"
#include "rt_lbvh.cuh"
#include "rt_pgrid_build.cuh"
#include "rt_scene.hpp"
#include "rt_wavefront.cuh"

namespace {

thread_local std::string g_last_error;

int fail(int code, const std::string& msg) {
    g_last_error = msg;
    return code;
}

#define RT_CUDA(call)                                                                                          \
    do {                                                                                                       \
        cudaError_t e_ = (call);                                                                               \
        if (e_ != cudaSuccess)                                                                                 \
            return fail(RT_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_));                      \
    } while (0)

template <typename T>
struct DevBuf {
    T* p = nullptr;
    size_t n = 0;
    cudaError_t ensure(size_t count) {
        if (count <= n && p) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr;
        n = 0;
        cudaError_t e = cudaMalloc(&p, std::max<size_t>(count, 1) * sizeof(T));
        if (e == cudaSuccess) n = std::max<size_t>(count, 1);
        return e;
    }
    cudaError_t upload(const std::vector<T>& v, cudaStream_t s) {
        cudaError_t e = ensure(v.size());
        if (e != cudaSuccess || v.empty()) return e;
        return cudaMemcpyAsync(p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice, s);
    }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        n = 0;
    }
};

}  // namespace

struct rt_ctx {
    int device = 0;
    cudaStream_t stream = nullptr, copy_stream = nullptr, copy_stream2 = nullptr;  // copy_stream2: readback of the odd lane's bands (staggered plan)
    cudaEvent_t copy_done = nullptr, copy_done2 = nullptr, copy_t0 = nullptr, copy_t1 = nullptr;
    // RT_SEQ_TRACE=1: start / end of every launch sequence and the end of its band's copy, printed by rt_render (diagnostics)
    cudaEvent_t tr_start[16] = {}, tr_end[16] = {}, tr_copy[16] = {};
    int tr_n = 0;
    int last_n_seq = 0;  // launch sequences of the last rt_render (their band_done events time the end of the kernels)
    uint32_t tr_units[16] = {};
    // readback bands: kernel time : copy time per unit of the frame, measured on the previous frame (an idle PCIe link gives ~2.6
    // for the 8K cfg5 frame; eight ranks sharing the host's ingest bandwidth ~1) -- the bands shrink by this factor
    double band_ratio = 3.0;
    uint32_t band0_units = 0, band_units_total = 0;
    // staggered plan: every band's copy is timed (start = band finished and copy stream free), the kernel : copy ratio of the
    // next frame comes from all of them -- one band's copy time depends on what the other ranks happen to copy meanwhile
    cudaEvent_t cp_s[16] = {}, cp_e[16] = {};
    uint32_t cp_units[16] = {};
    int cp_n = 0;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev2 = nullptr;
    static const int kMaxBands = 16;
    cudaEvent_t band_done[kMaxBands] = {};
    cudaStream_t seq_stream[kMaxBands] = {};  // one per concurrent launch sequence, descending priority

    // host copy of the AoS scene (kept so that the bounds can be rebuilt when the ray-origin bound grows)
    std::vector<rt_object> objs;
    std::vector<rt_light> lights;
    bool have_scene = false, scene_dirty = true;
    rt::HostScene hs;
    DevBuf<rt::F4> d_minv, d_mv, d_mat, d_lights;
    DevBuf<rt::CullRec> d_cull, d_leaf_cull;
    DevBuf<rt::BvhNode> d_nodes;
    DevBuf<uint32_t> d_leaf_obj;
    DevBuf<rt::LbvhLeaf> d_lbvh_leaves;  // device tree builder (RT_OPT_BVH_BUILDER = 1): input records and scratch
    DevBuf<char> d_lbvh_scratch;
    DevBuf<rt_object> d_objs;  // raw AoS scene for the device-side preparation (rt_bounds.cu, RT_OPT_BVH_BUILDER = 1)
    DevBuf<char> d_bounds_scratch;
    DevBuf<rt::BoundsGlobals> d_bounds_globals;
    float last_upload_ms = 0.f;
    rt::Scene scene;

    // rays / camera
    bool use_camera = false, have_rays = false;
    uint32_t width = 0, height = 0;
    float zdir = 0.f;
    DevBuf<rt::Ray> d_rays;
    uint64_t n_rays = 0;
    float omax_user = 0.f;
    uint32_t block_rows = 1, shard = 0, n_shards = 1;

    DevBuf<float4> d_fb;
    DevBuf<int32_t> d_hit_id;
    DevBuf<float> d_hit_t;
    DevBuf<rt::Counters> d_counters;
    DevBuf<float> d_sink;

    // wavefront pipeline buffers (allocated on first use)
    DevBuf<float4> w_acc, w_refl, w_P, w_Nm, w_Rc, w_qs[2], w_qd[2];
    DevBuf<uint32_t> w_out, w_qpath[2], w_counts;
    DevBuf<float> w_res_t;
    DevBuf<int32_t> w_res_id;
    int sm_count = 148;
    int last_pipeline = 0;
    uint32_t last_launches = 1;

    // perspective grids (rt_pgrid.cuh): one set for the camera, one for the lights, built lazily on the device
    struct GridBufs {
        DevBuf<rt::PGrid> grids;
        DevBuf<uint32_t> cell_start;
        DevBuf<rt::GridEntry> entries;
        rt::GridSet set;
        bool dirty = true;
        uint64_t n_entries = 0;
        uint32_t n_cells = 0;
        float build_ms = 0.f;
    } g_cam, g_lit;
    DevBuf<uint32_t> g_cursor;
    DevBuf<char> g_scan;
    DevBuf<float4> d_filter_tiles;  // accel 1 on the wavefront pipeline: filter records in tile layout (rt_wavefront.cu)
    bool filter_tiles_dirty = true;
    // uniform 3-D grid (rt_grid3.cuh) for the rays without a common apex
    DevBuf<uint32_t> g3_cell_start, g3_entries;
    DevBuf<rt::F4> d_aabb;
    rt::Grid3 g3 = {};
    bool g3_dirty = true;
    // a scene prepared on the device (RT_OPT_BVH_BUILDER 1) leaves the inflated boxes in d_aabb and their bounds here
    bool dev_boxes = false;
    double dev_box_lo[3] = {0, 0, 0}, dev_box_hi[3] = {0, 0, 0};
    uint32_t dev_live = 0;
    uint64_t g3_n_entries = 0;
    int opt_grid3 = 400;  // cells per 100 objects (default 4 per object); 0: off, the BVH serves these rays
    uint32_t g_cam_w = 0, g_cam_h = 0;
    float g_cam_z = 0.f;
    int opt_grid_res = 512;

    // options
    int opt_detect_pinhole = 1;
    int opt_tail_gen = -1;  // -1 auto; 0 off; g > 0: generations >= g of a launch sequence are finished by one wave_tail launch
    int opt_zero_copy = 0;  // 0 (default) never, 1 whenever the target is mapped pinned memory, -1 on the wavefront pipeline only -- measured slower, see rt_render
    int opt_pipeline = -1, opt_count = 0, opt_leaf = 1, opt_skip_dead = 1, opt_bands = 0, opt_builder = 0, opt_lanes = 0;

    rt_stats stats;
    float fp32_peak_tf[2] = {0.f, 0.f};  // last rt_measure_fp32_peak: scalar FFMA loop, packed FFMA2 loop
};

namespace {

uint32_t shard_rows(const rt_ctx* c) {
    if (!c->use_camera) return 0;
    const uint32_t br = c->block_rows;
    const uint32_t n_blocks = (c->height + br - 1) / br;
    uint32_t rows = 0;
    for (uint32_t b = c->shard; b < n_blocks; b += c->n_shards) rows += std::min(br, c->height - b * br);
    return rows;
}

uint32_t shard_local_rows(const rt_ctx* c) {  // rows of the local image incl. the ragged last block
    if (!c->use_camera) return 0;
    const uint32_t br = c->block_rows;
    const uint32_t n_blocks = (c->height + br - 1) / br;
    uint32_t owned = 0;
    for (uint32_t b = c->shard; b < n_blocks; b += c->n_shards) ++owned;
    return owned * br;
}

void bind_scene(rt_ctx* c, uint32_t n_nodes) {
    rt::Scene& s = c->scene;
    s.minv = reinterpret_cast<const float4*>(c->d_minv.p);
    s.mv = reinterpret_cast<const float4*>(c->d_mv.p);
    s.mat = reinterpret_cast<const float4*>(c->d_mat.p);
    s.cull = c->d_cull.p;
    s.nodes = c->d_nodes.p;
    s.leaf_obj = c->d_leaf_obj.p;
    s.leaf_cull = c->d_leaf_cull.p;
    s.lights = reinterpret_cast<const rt::Light*>(c->d_lights.p);
    s.n_objs = c->hs.n_objs;
    s.n_lights = c->hs.n_lights;
    s.n_nodes = n_nodes;
    s.root = c->hs.root;
    s.root_f = c->hs.f_global;
    s.filter_h = c->hs.h_min;
    c->scene_dirty = false;
    c->g_cam.dirty = c->g_lit.dirty = c->g3_dirty = c->filter_tiles_dirty = true;
    c->g_cam.set = c->g_lit.set = rt::GridSet{};
    c->g3 = rt::Grid3{};
}

// RT_OPT_BVH_BUILDER = 1: everything per object happens on the device -- repack, bounds (rt_bounds.cu), tree (rt_lbvh.cu).
// The host uploads the raw AoS array and reads back two small records (the scene radius that fixes the ray-origin bound,
// then the leaf-centre box and the scene-wide coefficients).
int sync_scene_device(rt_ctx* c) {
    const uint32_t n = (uint32_t)c->objs.size(), L = (uint32_t)c->lights.size();
    RT_CUDA(c->d_objs.upload(c->objs, c->stream));
    RT_CUDA(c->d_minv.ensure(4 * (size_t)n));
    RT_CUDA(c->d_mv.ensure(4 * (size_t)n));
    RT_CUDA(c->d_mat.ensure(3 * (size_t)n));
    RT_CUDA(c->d_cull.ensure(n));
    RT_CUDA(c->d_lbvh_leaves.ensure(n));
    RT_CUDA(c->d_aabb.ensure(n));
    c->dev_boxes = false;
    RT_CUDA(c->d_bounds_scratch.ensure(rt::bounds_scratch_bytes(n)));
    RT_CUDA(c->d_bounds_globals.ensure(1));
    rt::BoundsGlobals g;
    rt::bounds_globals_init(&g);
    RT_CUDA(cudaMemcpyAsync(c->d_bounds_globals.p, &g, sizeof g, cudaMemcpyHostToDevice, c->stream));
    RT_CUDA(rt::bounds_pass1_launch(c->d_objs.p, n, reinterpret_cast<float4*>(c->d_minv.p), reinterpret_cast<float4*>(c->d_mv.p),
                                    reinterpret_cast<float4*>(c->d_mat.p), c->d_bounds_scratch.p, c->d_bounds_globals.p, c->stream));
    auto bad = [&](const rt::BoundsGlobals& gg) -> int {
        if (gg.first_bad == ~0ull) return RT_OK;
        static const char* what[] = {"", "non-finite matrix", "projective (non-affine) matrices are outside the supported domain", "singular mvInverse",
                                     "degenerate transform", "anisotropy too extreme for conservative bounds"};
        const unsigned code = (unsigned)(gg.first_bad & 0xffu);
        return fail(RT_ERR_UNSUPPORTED, "object " + std::to_string(gg.first_bad >> 8) + ": " + what[code < 6 ? code : 0]);
    };
    RT_CUDA(cudaMemcpyAsync(&g, c->d_bounds_globals.p, sizeof g, cudaMemcpyDeviceToHost, c->stream));
    RT_CUDA(cudaStreamSynchronize(c->stream));
    int rc = bad(g);
    if (rc != RT_OK) return rc;
    const double omax = std::max(1.05 * g.r0 + 1.0, (double)c->omax_user * 1.0001 + 1e-3);  // rt_scene.cpp
    RT_CUDA(rt::bounds_pass2_launch(c->d_objs.p, n, c->d_bounds_scratch.p, omax, c->d_cull.p, c->d_lbvh_leaves.p,
                                    reinterpret_cast<float4*>(c->d_aabb.p), c->d_bounds_globals.p, c->stream));
    RT_CUDA(cudaMemcpyAsync(&g, c->d_bounds_globals.p, sizeof g, cudaMemcpyDeviceToHost, c->stream));
    RT_CUDA(cudaStreamSynchronize(c->stream));
    rc = bad(g);
    if (rc != RT_OK) return rc;
    c->hs = rt::HostScene();
    c->hs.n_objs = n;
    c->hs.n_lights = L;
    c->hs.omax = (float)omax;
    c->hs.scene_radius = (float)g.r0;
    c->hs.h_min = g.h_min;
    c->hs.f_global = g.f_max;
    c->hs.root = 0;
    c->hs.lights.resize(4 * (size_t)L);
    for (uint32_t i = 0; i < L; ++i) {
        const rt_light& l = c->lights[i];
        c->hs.lights[4 * i + 0] = rt::F4{l.ambient[0], l.ambient[1], l.ambient[2], 0.f};
        c->hs.lights[4 * i + 1] = rt::F4{l.diffuse[0], l.diffuse[1], l.diffuse[2], 0.f};
        c->hs.lights[4 * i + 2] = rt::F4{l.specular[0], l.specular[1], l.specular[2], 0.f};
        c->hs.lights[4 * i + 3] = rt::F4{l.position[0], l.position[1], l.position[2], l.position[3]};
    }
    RT_CUDA(c->d_lights.upload(c->hs.lights, c->stream));
    for (int k = 0; k < 3; ++k) {
        c->hs.lbvh_lo[k] = rt::bounds_decode_ordered(g.lo[k]);
        c->hs.lbvh_hi[k] = rt::bounds_decode_ordered(g.hi[k]);
        c->dev_box_lo[k] = rt::bounds_decode_ordered(g.box_lo[k]);
        c->dev_box_hi[k] = rt::bounds_decode_ordered(g.box_hi[k]);
    }
    c->dev_live = g.live;
    c->dev_boxes = true;
    const size_t sb = rt::lbvh_scratch_bytes(n);
    RT_CUDA(c->d_lbvh_scratch.ensure(sb));
    RT_CUDA(c->d_nodes.ensure(n - 1));
    RT_CUDA(c->d_leaf_obj.ensure(n));
    RT_CUDA(c->d_leaf_cull.ensure(n));
    RT_CUDA(rt::build_lbvh_device(c->d_lbvh_leaves.p, n, c->d_cull.p, c->hs.lbvh_lo, c->hs.lbvh_hi, c->hs.omax, c->d_lbvh_scratch.p, sb, c->d_nodes.p,
                                  c->d_leaf_obj.p, c->d_leaf_cull.p, c->stream));
    RT_CUDA(cudaStreamSynchronize(c->stream));
    bind_scene(c, n - 1);
    return RT_OK;
}

int sync_scene(rt_ctx* c) {
    if (!c->have_scene) return fail(RT_ERR_INVALID, "rt_upload_scene has not been called");
    const float need_omax = c->use_camera ? 0.f : c->omax_user;
    if (!c->scene_dirty && need_omax <= c->hs.omax) return RT_OK;
    if (c->opt_builder == 1 && c->objs.size() >= 2 && c->objs.size() < (1u << 27)) return sync_scene_device(c);
    std::string err;
    c->dev_boxes = false;
    const int rc = rt::build_scene(c->objs.data(), (uint32_t)c->objs.size(), c->lights.data(), (uint32_t)c->lights.size(),
                                   c->omax_user, c->opt_builder == 1 ? 0 : c->opt_leaf, c->hs, err);
    if (rc != RT_OK) return fail(rc, err);
    const bool device_tree = !c->hs.lbvh_leaves.empty();
    RT_CUDA(c->d_minv.upload(c->hs.minv, c->stream));
    RT_CUDA(c->d_mv.upload(c->hs.mv, c->stream));
    RT_CUDA(c->d_mat.upload(c->hs.mat, c->stream));
    RT_CUDA(c->d_lights.upload(c->hs.lights, c->stream));
    RT_CUDA(c->d_cull.upload(c->hs.cull, c->stream));
    uint32_t n_nodes = (uint32_t)c->hs.nodes.size();
    if (device_tree) {  // bounds came from the host, the tree is built here: Morton sort, radix tree, refit (rt_lbvh.cu)
        const uint32_t n = (uint32_t)c->hs.lbvh_leaves.size();
        RT_CUDA(c->d_lbvh_leaves.upload(c->hs.lbvh_leaves, c->stream));
        const size_t sb = rt::lbvh_scratch_bytes(n);
        RT_CUDA(c->d_lbvh_scratch.ensure(sb));
        RT_CUDA(c->d_nodes.ensure(n - 1));
        RT_CUDA(c->d_leaf_obj.ensure(n));
        RT_CUDA(c->d_leaf_cull.ensure(n));
        RT_CUDA(rt::build_lbvh_device(c->d_lbvh_leaves.p, n, c->d_cull.p, c->hs.lbvh_lo, c->hs.lbvh_hi, c->hs.omax, c->d_lbvh_scratch.p, sb,
                                      c->d_nodes.p, c->d_leaf_obj.p, c->d_leaf_cull.p, c->stream));
        n_nodes = n - 1;
    } else {
        RT_CUDA(c->d_leaf_cull.upload(c->hs.leaf_cull, c->stream));
        RT_CUDA(c->d_nodes.upload(c->hs.nodes, c->stream));
        RT_CUDA(c->d_leaf_obj.upload(c->hs.leaf_obj, c->stream));
    }
    RT_CUDA(cudaStreamSynchronize(c->stream));
    bind_scene(c, n_nodes);
    return RT_OK;
}

// Build one set of perspective grids on the device (rt_pgrid.cu): count, scan, fill, sort.  `grids` arrive without
// cell_base.  A set whose lists would be out of proportion (an apex buried in large objects) is dropped -- its rays then
// use the BVH -- by returning with set.valid_mask == 0.
int build_grid_set(rt_ctx* c, rt_ctx::GridBufs& gb, std::vector<rt::PGrid>& grids, uint32_t valid_mask) {
    gb.set = rt::GridSet{};
    gb.n_entries = 0;
    gb.n_cells = 0;
    gb.dirty = false;
    uint64_t cells = 0;
    for (rt::PGrid& g : grids) {
        g.cell_base = (uint32_t)cells;
        cells += (uint64_t)g.nu * g.nv;
    }
    if (valid_mask == 0u || cells == 0 || cells >= (1ull << 28)) return RT_OK;
    const uint32_t n_cells = (uint32_t)cells, n_grids = (uint32_t)grids.size(), n_objs = c->hs.n_objs;
    const auto t0 = std::chrono::steady_clock::now();
    RT_CUDA(gb.grids.ensure(n_grids));
    RT_CUDA(cudaMemcpyAsync(gb.grids.p, grids.data(), n_grids * sizeof(rt::PGrid), cudaMemcpyHostToDevice, c->stream));
    RT_CUDA(gb.cell_start.ensure((size_t)n_cells + 1));
    RT_CUDA(c->g_cursor.ensure((size_t)n_cells + 1));
    const size_t sb = rt::pgrid_scan_bytes(n_cells);
    RT_CUDA(c->g_scan.ensure(sb));
    RT_CUDA(rt::pgrid_count(gb.grids.p, n_grids, n_cells, c->d_cull.p, n_objs, c->hs.omax, c->g_cursor.p, gb.cell_start.p, c->g_scan.p, sb, c->stream));
    uint32_t total = 0;
    RT_CUDA(cudaMemcpyAsync(&total, gb.cell_start.p + n_cells, sizeof total, cudaMemcpyDeviceToHost, c->stream));
    RT_CUDA(cudaStreamSynchronize(c->stream));
    if ((uint64_t)total > 96ull * n_objs + 4ull * n_cells + (1ull << 20)) return RT_OK;  // not worth it: BVH instead
    RT_CUDA(gb.entries.ensure(std::max<size_t>(total, 1)));
    RT_CUDA(rt::pgrid_fill(gb.grids.p, n_grids, n_cells, c->d_cull.p, n_objs, c->hs.omax, c->g_cursor.p, gb.cell_start.p, gb.entries.p, c->stream));
    RT_CUDA(cudaStreamSynchronize(c->stream));
    gb.set.grids = gb.grids.p;
    gb.set.cell_start = gb.cell_start.p;
    gb.set.entries = gb.entries.p;
    gb.set.n_grids = n_grids;
    gb.set.valid_mask = valid_mask;
    gb.n_entries = total;
    gb.n_cells = n_cells;
    gb.build_ms = (float)std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    return RT_OK;
}

// The uniform grid over the whole scene (rt_grid3.cuh): cubic cells, about opt_grid3 / 100 of them per object.  Not built --
// the BVH stays in charge of the rays without a common apex -- when the lists come out long: objects that are large against
// the cell (few big objects among many small ones, a clustered scene) would be entered into thousands of cells.
// Measured on the 8K cfg5 frame (profiles/r2_sweep_grid3.txt): the walk needs 18 cells and 23 filter tests per reflection ray
// where the BVH needs 52 node steps of two inflated slab tests each; frame 31.9 ms (BVH) -> 27.3 ms at 4 cells per object
// (1: 29.1, 2: 27.6, 8: 27.7, 16: 28.8) -> 24.9 ms with four resident CTAs per SM.  (A first version that entered objects by
// the cube around their bounding SPHERE and recomputed the cell boundaries in every step was no faster than the BVH: 33.2 ms.)
int build_grid3(rt_ctx* c) {
    c->g3 = rt::Grid3{};
    c->g3_dirty = false;
    c->g3_n_entries = 0;
    const uint32_t n_objs = c->hs.n_objs;
    if (c->opt_grid3 <= 0 || n_objs < 64) return RT_OK;
    const float ext0 = 2.f * c->hs.scene_radius + 1.f;
    const float eps = 2e-5f * ext0 + 1e-4f;
    double lo[3] = {1e30, 1e30, 1e30}, hi[3] = {-1e30, -1e30, -1e30};
    uint32_t live = 0;
    if (c->dev_boxes) {  // prepared on the device (rt_bounds.cu pass 2): boxes already in d_aabb, their bounds came back with the globals
        live = c->dev_live;
        for (int k = 0; k < 3; ++k) {
            lo[k] = c->dev_box_lo[k] - eps;
            hi[k] = c->dev_box_hi[k] + eps;
        }
    } else {
        if (c->hs.cull.size() < n_objs || c->hs.aabb.size() < n_objs) return RT_OK;
        for (uint32_t o = 0; o < n_objs; ++o) {
            const rt::F4& hb = c->hs.aabb[o];
            if (hb.w < 0.f || c->hs.cull[o].r2 < 0.f) continue;
            ++live;
            const float ctr[3] = {c->hs.cull[o].cx, c->hs.cull[o].cy, c->hs.cull[o].cz};
            const float half[3] = {hb.x, hb.y, hb.z};
            for (int k = 0; k < 3; ++k) {
                lo[k] = std::min(lo[k], (double)ctr[k] - half[k] - eps);
                hi[k] = std::max(hi[k], (double)ctr[k] + half[k] + eps);
            }
        }
    }
    if (live < 64) return RT_OK;
    double vol = 1;
    for (int k = 0; k < 3; ++k) vol *= std::max(hi[k] - lo[k], 1e-3);
    float h = (float)cbrt(vol / (c->opt_grid3 * 0.01 * live));
    for (int k = 0; k < 3; ++k) h = std::max(h, (float)((hi[k] - lo[k]) / 1024.0));
    rt::Grid3 g = {};
    g.h = h;
    g.inv_h = 1.f / h;
    g.lox = (float)lo[0] - 2 * eps;
    g.loy = (float)lo[1] - 2 * eps;
    g.loz = (float)lo[2] - 2 * eps;
    g.nx = (uint32_t)((hi[0] + 2 * eps - g.lox) / h) + 1;
    g.ny = (uint32_t)((hi[1] + 2 * eps - g.loy) / h) + 1;
    g.nz = (uint32_t)((hi[2] + 2 * eps - g.loz) / h) + 1;
    g.eps = eps;
    const uint64_t cells = (uint64_t)g.nx * g.ny * g.nz;
    if (cells >= (1ull << 28)) return RT_OK;
    const uint32_t n_cells = (uint32_t)cells;
    RT_CUDA(c->g3_cell_start.ensure((size_t)n_cells + 1));
    RT_CUDA(c->g_cursor.ensure((size_t)n_cells + 1));
    const size_t sb = rt::pgrid_scan_bytes(n_cells);
    RT_CUDA(c->g_scan.ensure(sb));
    if (!c->dev_boxes) RT_CUDA(c->d_aabb.upload(c->hs.aabb, c->stream));
    RT_CUDA(rt::grid3_count(g, c->d_cull.p, reinterpret_cast<const float4*>(c->d_aabb.p), n_objs, c->g_cursor.p, c->g3_cell_start.p, c->g_scan.p, sb, c->stream));
    uint32_t total = 0;
    RT_CUDA(cudaMemcpyAsync(&total, c->g3_cell_start.p + n_cells, sizeof total, cudaMemcpyDeviceToHost, c->stream));
    RT_CUDA(cudaStreamSynchronize(c->stream));
    if ((uint64_t)total > 64ull * live) return RT_OK;  // objects large against the cells: the BVH serves this scene better
    {   // ... and so it does when the objects cluster: the list length an entry sees on average (sum len^2 / sum len) must stay
        // small, or the walk degenerates into brute force inside the crowded cells
        std::vector<uint32_t> cs((size_t)n_cells + 1);
        RT_CUDA(cudaMemcpyAsync(cs.data(), c->g3_cell_start.p, cs.size() * sizeof(uint32_t), cudaMemcpyDeviceToHost, c->stream));
        RT_CUDA(cudaStreamSynchronize(c->stream));
        double s2 = 0.0;
        for (uint32_t k = 0; k < n_cells; ++k) {
            const double len = (double)(cs[k + 1] - cs[k]);
            s2 += len * len;
        }
        if (total > 0 && s2 / (double)total > 24.0) return RT_OK;
    }
    RT_CUDA(c->g3_entries.ensure(std::max<size_t>(total, 1)));
    RT_CUDA(rt::grid3_fill(g, c->d_cull.p, reinterpret_cast<const float4*>(c->d_aabb.p), n_objs, c->g_cursor.p, c->g3_cell_start.p, c->g3_entries.p, c->stream));
    RT_CUDA(cudaStreamSynchronize(c->stream));
    g.cell_start = c->g3_cell_start.p;
    g.entries = c->g3_entries.p;
    g.valid = 1;
    c->g3 = g;
    c->g3_n_entries = total;
    return RT_OK;
}

// The camera's grid: apex = the view-space origin every primary ray starts from (OpenCL-Raytracer.cpp:68-72), cells = tiles of
// TW x TH pixels (8 x 4, the patch one warp of wave_gen covers, unless the frame has more than 2 M of them).
int build_camera_grid(rt_ctx* c) {
    std::vector<rt::PGrid> grids(1);
    rt::PGrid& g = grids[0];
    memset(&g, 0, sizeof g);
    const float f = -c->zdir;
    uint32_t tw = 8, th = 4;
    while ((uint64_t)((c->width + tw - 1) / tw) * ((c->height + th) / th) > (2ull << 20)) (tw <= th ? tw : th) *= 2;
    const float half_w = (float)c->width / 2.0f, half_h = (float)c->height / 2.0f;
    uint32_t valid = 0;
    if (c->use_camera && f > 0.f && std::isfinite(f) && half_w / f < 30.f && half_h / f < 30.f && c->hs.n_objs > 0) {
        g.ax = g.ay = g.az = 0.f;
        g.axis = 2u | 4u;  // looks down -z
        g.u0 = -half_w / f;
        g.v0 = -half_h / f;
        g.inv_du = f / (float)tw;
        g.inv_dv = f / (float)th;
        g.nu = (c->width + tw - 1) / tw;
        g.nv = c->height / th + 1;  // dir.y = (H - jj) - H/2 runs from -H/2 + 1 to H/2 inclusive
        g.kind = 0;
        valid = 1;
    }
    c->g_cam_w = c->width;
    c->g_cam_h = c->height;
    c->g_cam_z = c->zdir;
    return build_grid_set(c, c->g_cam, grids, valid);
}

// Six grids (a cube map) around every point light; a light at infinity (position.w == 0, SAR:198) has parallel shadow rays
// and keeps the BVH.
int build_light_grids(rt_ctx* c) {
    const uint32_t L = std::min<uint32_t>((uint32_t)c->lights.size(), RT_GRID_MAX_LIGHTS);
    int res = c->opt_grid_res;
    for (int attempt = 0; attempt < 3; ++attempt, res /= 2) {
        std::vector<rt::PGrid> grids(6 * (size_t)L);
        uint32_t valid = 0;
        for (uint32_t li = 0; li < L; ++li) {
            const rt_light& lt = c->lights[li];
            const bool ok = lt.position[3] != 0.f && std::isfinite(lt.position[0]) && std::isfinite(lt.position[1]) && std::isfinite(lt.position[2]) &&
                            c->hs.n_objs > 0;
            for (uint32_t face = 0; face < 6; ++face) {
                rt::PGrid& g = grids[6 * li + face];
                memset(&g, 0, sizeof g);
                if (!ok) continue;
                g.ax = lt.position[0];
                g.ay = lt.position[1];
                g.az = lt.position[2];
                g.axis = (face % 3u) | (face >= 3u ? 4u : 0u);
                g.u0 = g.v0 = -1.f;
                g.inv_du = g.inv_dv = (float)res / 2.f;
                g.nu = g.nv = (uint32_t)res;
                g.kind = 1;
            }
            if (ok) valid |= 1u << li;
        }
        const int rc = build_grid_set(c, c->g_lit, grids, valid);
        if (rc != RT_OK) return rc;
        if (c->g_lit.set.valid_mask || valid == 0u || res <= 64) break;
    }
    return RT_OK;
}

// Is `rays` the W x H pinhole grid main() builds (OpenCL-Raytracer.cpp:18-26,68-72): start (0,0,0,1), direction
// (ii - W/2, (H - jj) - H/2, zdir, 0), row-major, zdir < 0 the same for all?  W is where direction.y first changes.
bool detect_pinhole(const rt_ray* rays, uint64_t n, uint32_t* w_out, uint32_t* h_out, float* z_out) {
    const float z = rays[0].direction[2], y0 = rays[0].direction[1];
    if (!(z < 0.f) || !std::isfinite(z)) return false;
    uint64_t w = n;
    for (uint64_t i = 1; i < n; ++i)
        if (rays[i].direction[1] != y0) {
            w = i;
            break;
        }
    if (w == 0 || n % w != 0 || w >= (1ull << 31) || n / w >= (1ull << 31)) return false;
    const uint64_t h = n / w;
    const float half_w = (float)w / 2.0f, half_h = (float)h / 2.0f;
    const int nt = (int)std::min<uint64_t>(std::max<uint64_t>(n >> 20, 1), std::min<unsigned>(std::max(1u, std::thread::hardware_concurrency()), 16u));
    std::vector<char> ok(nt, 1);
    auto check = [&](int t) {
        const uint64_t r0 = h * t / nt, r1 = h * (t + 1) / nt;
        for (uint64_t jj = r0; jj < r1; ++jj) {
            const float dy = (float)((int)h - (int)jj) - half_h;
            const rt_ray* row = rays + jj * w;
            for (uint64_t ii = 0; ii < w; ++ii) {
                const rt_ray& r = row[ii];
                if (r.start[0] != 0.f || r.start[1] != 0.f || r.start[2] != 0.f || r.start[3] != 1.f || r.direction[3] != 0.f ||
                    r.direction[0] != (float)ii - half_w || r.direction[1] != dy || r.direction[2] != z) {
                    ok[t] = 0;
                    return;
                }
            }
        }
    };
    if (nt == 1) {
        check(0);
    } else {
        std::vector<std::thread> pool;
        for (int t = 0; t < nt; ++t) pool.emplace_back(check, t);
        for (auto& th : pool) th.join();
    }
    for (char k : ok)
        if (!k) return false;
    *w_out = (uint32_t)w;
    *h_out = (uint32_t)h;
    *z_out = z;
    return true;
}

int pick_accel(const rt_ctx* c, int accel) {
    if (accel == RT_ACCEL_AUTO) {
        const size_t m = c->objs.size();
        // small scenes: the bounding-sphere filter (7 fused multiply-adds) in front of the exact arithmetic -- most pixels of the
        // reference's own scenes miss everything (98.7 % for simpleSphere) and should not pay ~130 unfused instructions per object
        if (m <= 48) return RT_ACCEL_CULL;
        return RT_ACCEL_GRID;
    }
    return accel;
}

int check_render_args(rt_ctx* c, int mode, int accel) {
    if (!c) return fail(RT_ERR_INVALID, "null context");
    if (mode < RT_MODE_HITTEST || mode > RT_MODE_SHADE_AND_REFLECT) return fail(RT_ERR_INVALID, "bad mode");
    if (accel < RT_ACCEL_AUTO || accel > RT_ACCEL_GRID) return fail(RT_ERR_INVALID, "bad accel");
    if (!c->use_camera && !c->have_rays) return fail(RT_ERR_INVALID, "neither rt_set_rays nor rt_set_camera was called");
    return RT_OK;
}

uint64_t n_pixels(const rt_ctx* c) { return c->use_camera ? (uint64_t)c->width * c->height : c->n_rays; }
uint64_t local_pixels(const rt_ctx* c) { return c->use_camera ? (uint64_t)shard_rows(c) * c->width : c->n_rays; }

// band >= 0 (camera renders on the wavefront pipeline): the context's part of the frame is cut into n_bands
// contiguous bands so that the readback of a finished band (copy stream) overlaps the next band's kernels.  The unit
// is one row (whole frame) or one owned row-block (sharded frame).  Every band costs the pipeline its ramp-up and its
// thin last generations again, and the copy is faster than the kernels of the same rows, so the bands shrink
// geometrically by band_ratio = kernel time : copy time, measured on the previous frame (7 : 1 on one GPU over an
// idle PCIe link, < 2 : 1 when 8 ranks share the host's ingest bandwidth): only the last, small band's copy is exposed.
uint32_t plan_units(uint32_t height, uint32_t block_rows, uint32_t shard, uint32_t n_shards) {
    if (n_shards <= 1) return height;
    const uint32_t n_blocks = (height + block_rows - 1) / block_rows;
    return n_blocks > shard ? (n_blocks - shard + n_shards - 1) / n_shards : 0u;
}
void plan_band(uint32_t units, int band, int n_bands, double ratio, uint32_t* u0, uint32_t* u1) {
    const double q = 1.0 / ratio;
    double total = 0.0, w = 1.0;
    for (int i = 0; i < n_bands; ++i, w *= q) total += w;
    double acc = 0.0;
    w = 1.0;
    uint32_t start = 0, end = 0;
    for (int i = 0; i <= band; ++i, w *= q) {
        start = end;
        acc += w;
        end = i == n_bands - 1 ? units : std::min(units, (uint32_t)(acc / total * units + 0.5));
        if (end < start) end = start;
    }
    *u0 = start;
    *u1 = end;
}
uint32_t band_units(const rt_ctx* c) { return plan_units(c->height, c->block_rows, c->shard, c->n_shards); }

// The staggered readback plan of rt_render: `units` (rows, or owned row-blocks) of `unit_px` pixels each are cut into launch
// sequences, each its own readback band.  k_req > 1 fixes their number, q_req > 0 the size ratio of consecutive sequences;
// otherwise both follow `ratio`, the measured kernel : copy time per unit of the previous frames.  Returns the number of
// sequences K (<= max_k) and bounds[0 .. K].
int plan_sequences(uint32_t units, uint64_t unit_px, int k_req, double q_req, double ratio, int max_k, uint32_t* bounds, double* q_out) {
    double q = q_req;
    if (!(q > 0.0)) {
        // each sequence is `q` of the one before.  Copy much faster than the kernels (one GPU on an idle link, kernel : copy
        // time 2.5): 0.7, so that only a small last copy is exposed.  Copy as slow as the kernels or slower (8 ranks share the
        // host's ~100 GB/s of ingest: 66 MB per rank take 5.4 ms, the kernels 4.3): equal sequences, and many -- the frame is
        // then home one copy time after the FIRST sequence has finished, so that one has to be small.
        q = std::min(1.0, std::max(0.7, 1.3 - 0.3 * ratio));
    }
    int k = k_req;
    if (k < 1) {
        // auto: six sequences (eight when they are about equal), fewer when the smallest one would fall below 0.4 Mpixel
        k = q > 0.85 ? 8 : 6;
        const double px = (double)units * (double)unit_px;
        auto smallest = [&](int n) {
            double sum = 0.0, w = 1.0;
            for (int i = 0; i < n; ++i, w *= q) sum += w;
            return px * (w / q) / sum;
        };
        while (k > 2 && smallest(k) < 400000.0) --k;
    }
    k = std::min(k, max_k);
    while (k > 1 && units < (uint32_t)(4 * k)) --k;
    k = std::max(k, 1);
    double total_w = 0.0, w = 1.0;
    for (int b = 0; b < k; ++b, w *= q) total_w += w;
    double acc = 0.0;
    w = 1.0;
    bounds[0] = 0;
    for (int b = 0; b < k; ++b, w *= q) {
        acc += w;
        bounds[b + 1] = b == k - 1 ? units : std::max(bounds[b], std::min(units, (uint32_t)(acc / total_w * units + 0.5)));
    }
    int n = 0;  // sequences that rounding left empty are dropped
    for (int b = 0; b < k; ++b)
        if (bounds[b + 1] > bounds[n]) bounds[++n] = bounds[b + 1];
    if (n == 0) bounds[++n] = units;
    if (q_out) *q_out = q;
    return n;
}

// Number of bands when the caller leaves it open: two.  One more band costs its ramp-up and thin last generations
// again (0.8 ms of an 8K frame on one GPU, 1.4 ms at 8 GPUs) and hides less than that in every configuration measured
// (1, 2, 4, 8 GPUs); a cost model on the measured copy time is not robust, because that time depends on how many ranks
// copy at the same moment, i.e. on the decision itself.


// Frame descriptor.  Camera frames are described by a range [u0, u1) of the context's units (band_units(): rows of the
// whole frame, or owned row-blocks of a shard); u1 == 0 means "everything" (and is the only form for explicit rays).
void fill_frame_desc(rt_ctx* c, rt::Frame& fr, uint32_t max_bounces, int mode, float4* out, uint32_t u0 = 0, uint32_t u1 = 0) {
    memset(&fr, 0, sizeof fr);
    fr.rays = c->use_camera ? nullptr : c->d_rays.p;
    fr.width = c->width;
    fr.height = c->height;
    fr.zdir = c->zdir;
    if (c->use_camera && u1 > 0 && c->n_shards <= 1) {  // rows u0 .. u1
        fr.block_rows = std::max(u1 - u0, 1u);
        fr.shard = 0;
        fr.n_shards = 1;
        fr.row0 = u0;
        fr.local_rows = u1 - u0;
        fr.n_local = (uint64_t)(u1 - u0) * c->width;
    } else if (c->use_camera && u1 > 0) {  // owned row-blocks u0 .. u1 of the shard
        fr.block_rows = c->block_rows;
        fr.shard = c->shard;
        fr.n_shards = c->n_shards;
        fr.lrow0 = u0 * c->block_rows;
        fr.local_rows = (u1 - u0) * c->block_rows;
        uint64_t rows = 0;
        for (uint32_t u = u0; u < u1; ++u) {
            const uint32_t g = (c->shard + u * c->n_shards) * c->block_rows;
            rows += g < c->height ? std::min(c->block_rows, c->height - g) : 0u;
        }
        fr.n_local = rows * c->width;
    } else {
        fr.block_rows = c->use_camera ? c->block_rows : 1;
        fr.shard = c->use_camera ? c->shard : 0;
        fr.n_shards = c->use_camera ? c->n_shards : 1;
        fr.local_rows = shard_local_rows(c);
        fr.n_local = c->use_camera ? (uint64_t)shard_rows(c) * c->width : c->n_rays;
    }
    fr.max_bounces = max_bounces;
    fr.mode = mode;
    fr.skip_dead = c->opt_skip_dead;
    fr.out = out;
    fr.counters = c->opt_count ? c->d_counters.p : nullptr;
}

int collect_stats(rt_ctx* c, int accel, uint64_t pixels, uint32_t launches) {
    rt_stats& st = c->stats;
    memset(&st, 0, sizeof st);
    st.pixels = pixels;
    st.kernel_launches = c->last_launches ? c->last_launches : launches;
    st.accel_used = (uint32_t)accel;
    st.pipeline_used = (uint32_t)c->last_pipeline;
    st.grid_walk_used = (accel == RT_ACCEL_GRID && c->last_pipeline == 1 && c->g3.valid) ? 1u : 0u;
    if (c->opt_count) {
        rt::Counters h;
        RT_CUDA(cudaMemcpyAsync(&h, c->d_counters.p, sizeof h, cudaMemcpyDeviceToHost, c->stream));
        RT_CUDA(cudaStreamSynchronize(c->stream));
        st.rays_primary = h.ref_primary;
        st.rays_shadow = h.ref_shadow;
        st.rays_reflect = h.ref_reflect;
        st.traced_shadow = h.traced_shadow;
        st.traced_reflect = h.traced_reflect;
        st.pair_tests = h.pair_tests;
        st.exact_tests = h.exact_tests;
        st.node_visits = h.node_visits;
        st.hit_pixels = h.hit_pixels;
    }
    return RT_OK;
}

// 0 megakernel, 1 wavefront.  The wavefront always skips dead rays, so asking for every reference ray
// (RT_OPT_SKIP_DEAD_RAYS = 0) selects the megakernel; tiny scenes are launch-bound and stay on the megakernel too.
constexpr uint32_t kMaxWaveGenerations = 4096;
int pick_pipeline(const rt_ctx* c, int accel, uint32_t max_bounces, int mode) {
    if (!c->opt_skip_dead) return 0;
    if (c->opt_pipeline == 0 || c->opt_pipeline == 1) return c->opt_pipeline;
    // many lights x many bounces: thousands of (mostly empty) generations -- the megakernel handles those in-thread
    if (rt::wave_max_generations(max_bounces, (uint32_t)c->lights.size(), mode) > kMaxWaveGenerations) return 0;
    return (accel != RT_ACCEL_BRUTE && c->objs.size() > 48) ? 1 : 0;
}

// Wavefront buffers for `capacity` paths in total, cut into one slice per concurrent launch sequence: sequence k owns the
// entries [offset[k], offset[k] + cap[k]).
int ensure_wave_buffers(rt_ctx* c, uint64_t capacity, uint32_t gens, int n_seq, const uint64_t* offset, const uint64_t* cap, rt::WaveBuffers* wb) {
    RT_CUDA(c->w_acc.ensure(capacity));
    RT_CUDA(c->w_refl.ensure(capacity));
    RT_CUDA(c->w_P.ensure(capacity));
    RT_CUDA(c->w_Nm.ensure(capacity));
    RT_CUDA(c->w_Rc.ensure(capacity));
    RT_CUDA(c->w_out.ensure(capacity));
    for (int i = 0; i < 2; ++i) {
        RT_CUDA(c->w_qs[i].ensure(capacity));
        RT_CUDA(c->w_qd[i].ensure(capacity));
        RT_CUDA(c->w_qpath[i].ensure(capacity));
    }
    RT_CUDA(c->w_res_t.ensure(capacity));
    RT_CUDA(c->w_res_id.ensure(capacity));
    const size_t per_seq = 4 * (size_t)(gens + 2);
    RT_CUDA(c->w_counts.ensure((size_t)n_seq * per_seq));
    for (int l = 0; l < n_seq; ++l) {
        if (cap[l] > RT_TASK_PATH_MASK) return fail(RT_ERR_INVALID, "too many rays for one launch sequence (2^27 - 1): more lanes / shards needed");
        const uint64_t o = offset[l];
        wb[l].pa.acc = c->w_acc.p + o;
        wb[l].pa.refl = c->w_refl.p + o;
        wb[l].pa.P = c->w_P.p + o;
        wb[l].pa.Nm = c->w_Nm.p + o;
        wb[l].pa.Rc = c->w_Rc.p + o;
        wb[l].pa.out = c->w_out.p + o;
        for (int i = 0; i < 2; ++i) {
            wb[l].q[i].s = c->w_qs[i].p + o;
            wb[l].q[i].d = c->w_qd[i].p + o;
            wb[l].q[i].path = c->w_qpath[i].p + o;
        }
        wb[l].res_t = c->w_res_t.p + o;
        wb[l].res_id = c->w_res_id.p + o;
        wb[l].counts = c->w_counts.p + l * per_seq;
        wb[l].n_counts = (uint32_t)(per_seq / 4);
        wb[l].capacity = cap[l];
        wb[l].qcap = (uint32_t)cap[l];
    }
    return RT_OK;
}

// How a camera render is cut into launch sequences (wavefront pipeline).
//  * Frame stays on the device (rt_render_device): 2 concurrent sequences (lanes) on two streams -- the blocks of one fill the
//    thin generation tails of the other.
//  * Frame goes to the host (rt_render): the frame (or the shard's row-blocks) is cut into launch sequences, each its own readback
//    band with its own copy, dealt round-robin to three streams (two copy streams); a finished band goes home under the kernels
//    of the others.  Sizes: each sequence 0.7 of the one before where the copy is faster than the kernels -- the thin last
//    generations of a sequence run under the heavy first ones of the next, only the small last copy is exposed; equal sizes where
//    the copy is the slower side (8 ranks share ~100 GB/s of host ingest: the frame is home one copy time after the FIRST band
//    finished).  Measured on the 8K cfg5 frame, kernels 23.5 ms, copy 9.4 ms (profiles/r2_sweep_stagger.txt,
//    r2_sweep_shard_stagger.txt, r2_sweep_bands.txt): **26.2 ms** end to end with 6 sequences x 0.7 on 3 streams; bands of two
//    equal lanes one after the other (the lanes reach their thin generations together, every band pays them again): 28.2 ms;
//    all bands concurrent on their own streams (RT_BAND_MODE=0): 29.1 ms -- stream priorities do not order persistent kernels,
//    the bands finish together; an SM copy kernel into the mapped frame instead of the copy engine: 35.7 ms; an event chain
//    that starts band k after the heavy generations of band k - 1: slower still.  8 GPUs: 8.9 -> 7.1 ms, 4: 12.9 -> 10.6 ms.
int auto_sequences(const rt_ctx* c, bool to_host, bool aov_only) {
    if (!c->use_camera || aov_only) return 1;
    const uint64_t px = local_pixels(c);
    // (concurrent layout) to the host: 4 bands for 12 Mpixel and more, 2 below
    int n = to_host && c->opt_bands > 0 ? c->opt_bands : (c->opt_lanes > 0 ? c->opt_lanes : (to_host && px >= (12ull << 20) ? 4 : 2));
    n = std::min(n, rt_ctx::kMaxBands);
    while (n > 1 && (px / n < (1ull << 19) || band_units(c) < (uint32_t)(4 * n))) --n;  // at least 0.5 Mpixel and 4 units per band
    return std::max(n, 1);
}

// Kernels of one render, bracketed by ev0/ev1 on c->stream; host_out != nullptr: plus the device->host copy of every band
// (copy stream), ev2 after the last one.  Nothing is synchronised here.
int render_on_device(rt_ctx* c, uint32_t max_bounces, int mode, int accel, float4* out, int* accel_used, bool aov_only = false,
                     float* host_out = nullptr) {
    int rc = sync_scene(c);
    if (rc != RT_OK) return rc;
    accel = pick_accel(c, accel);
    *accel_used = accel;
    const int pipeline = pick_pipeline(c, accel, max_bounces, mode);
    c->last_pipeline = pipeline;
    if (accel == RT_ACCEL_GRID && pipeline == 1) {  // perspective grids, (re)built on the device when the scene / the camera changed
        if (c->g_lit.dirty) {
            rc = build_light_grids(c);
            if (rc != RT_OK) return rc;
        }
        if (c->g3_dirty) {
            rc = build_grid3(c);
            if (rc != RT_OK) return rc;
        }
        if (c->use_camera && (c->g_cam.dirty || c->g_cam_w != c->width || c->g_cam_h != c->height || c->g_cam_z != c->zdir)) {
            rc = build_camera_grid(c);
            if (rc != RT_OK) return rc;
        }
    } else if (accel == RT_ACCEL_GRID) {
        *accel_used = RT_ACCEL_BVH;  // the megakernel has no grid search
    }
    if (accel == RT_ACCEL_CULL && pipeline == 1 && c->filter_tiles_dirty) {
        RT_CUDA(c->d_filter_tiles.ensure(2 * (size_t)std::max(rt::wave_filter_tile_pad(c->scene.n_objs), 1u)));
        RT_CUDA(rt::wave_pack_filter_tiles(c->scene, c->d_filter_tiles.p, c->stream));
        c->filter_tiles_dirty = false;
    }
    if (c->opt_count) {
        RT_CUDA(c->d_counters.ensure(1));
        RT_CUDA(cudaMemsetAsync(c->d_counters.p, 0, sizeof(rt::Counters), c->stream));
    }
    c->last_launches = 0;

    if (pipeline == 1) {
        const uint32_t gens = rt::wave_max_generations(max_bounces, c->scene.n_lights, mode);
        if (gens > kMaxWaveGenerations) return fail(RT_ERR_INVALID, "max_bounces x lights too large for the wavefront pipeline; use RT_OPT_PIPELINE 0");
        // ---- the launch plan: bands (units [b0, b1) of the frame, copied to the host as a whole) made of lanes (launch sequences)
        struct Seq {
            uint32_t u0, u1;
            int stream, band;
        };
        Seq seq[rt_ctx::kMaxBands];
        uint32_t band_u[rt_ctx::kMaxBands + 1];
        int n_seq = 0;
        const uint32_t units = c->use_camera ? band_units(c) : 0u;
        const uint64_t unit_px = (uint64_t)c->width * (c->n_shards <= 1 ? 1u : c->block_rows);
        static const int band_mode_env = [] {
            const char* v = getenv("RT_BAND_MODE");
            return v && *v ? atoi(v) : -1;
        }();
        static const double band_ratio_env = [] {
            const char* v = getenv("RT_BAND_RATIO");
            return v && *v ? atof(v) : 0.0;
        }();
        // experiment knobs, read on every render so that one process can sweep them
        auto env_num = [](const char* name, double dflt) {
            const char* v = getenv(name);
            return v && *v ? atof(v) : dflt;
        };
        int stagger_k = (int)env_num("RT_STAGGER_K", -1);  // -1: auto; 0 / 1: off (bands of two equal lanes); K > 1: K sequences
        double stagger_q = env_num("RT_STAGGER_Q", 0.0);
        const int stagger_s = std::max(1, std::min(4, (int)env_num("RT_STAGGER_S", 3)));
        if (c->opt_bands > 0 && stagger_k < 0) stagger_k = c->opt_bands;  // RT_OPT_READBACK_BANDS = number of sequences
        const bool sequential = host_out != nullptr && c->use_camera && !aov_only && (band_mode_env < 0 ? true : band_mode_env >= 1);
        const bool staggered = sequential && (stagger_k < 0 || stagger_k > 1);
        if (staggered) {
            // every launch sequence is a band of its own; sizes shrink geometrically, streams alternate: a sequence's thin last
            // generations run under the heavy first generations of the next sequence on another stream
            const int nb = plan_sequences(units, unit_px, stagger_k, stagger_q, c->band_ratio, (int)rt_ctx::kMaxBands, band_u, nullptr);
            for (int b = 0; b < nb; ++b) {
                seq[n_seq].u0 = band_u[b];
                seq[n_seq].u1 = band_u[b + 1];
                seq[n_seq].stream = b % stagger_s;
                seq[n_seq].band = b;
                ++n_seq;
            }
        } else if (sequential) {
            // bands one after the other, each as two lanes on two streams; the copy of band b runs under the kernels of band
            // b + 1, so the bands shrink geometrically by (copy time : kernel time) and only the small last copy is exposed
            // 4 bands for a large frame, and for a shard whose copy is as slow as its kernels (ranks sharing the host's ingest)
            int nb = c->opt_bands > 0 ? c->opt_bands : ((local_pixels(c) >= (12ull << 20) || c->band_ratio < 1.5) ? 4 : 2);
            nb = std::min(nb, rt_ctx::kMaxBands / 2);
            while (nb > 1 && units < (uint32_t)(4 * nb)) --nb;
            const double ratio = band_ratio_env > 0.0 ? band_ratio_env : std::min(8.0, std::max(1.0, std::max(1.0, c->band_ratio) * (nb >= 4 ? 1.4 : 1.0)));
            band_u[0] = 0;
            for (int b = 0; b < nb; ++b) {
                uint32_t u0, u1;
                plan_band(units, b, nb, ratio, &u0, &u1);
                band_u[b + 1] = u1;
                const int lanes = c->opt_lanes == 1 ? 1 : ((u1 - u0) >= 8 && (uint64_t)(u1 - u0) * unit_px >= (1ull << 20) ? 2 : 1);
                for (int l = 0; l < lanes; ++l) {
                    seq[n_seq].u0 = u0 + (uint32_t)((uint64_t)(u1 - u0) * l / lanes);
                    seq[n_seq].u1 = u0 + (uint32_t)((uint64_t)(u1 - u0) * (l + 1) / lanes);
                    seq[n_seq].stream = l;
                    seq[n_seq].band = b;
                    ++n_seq;
                }
            }
        } else {
            // concurrent equal bands, one stream each
            n_seq = auto_sequences(c, host_out != nullptr, aov_only);
            for (int k = 0; k < n_seq; ++k) {
                seq[k].u0 = (uint32_t)((uint64_t)units * k / n_seq);
                seq[k].u1 = (uint32_t)((uint64_t)units * (k + 1) / n_seq);
                seq[k].stream = k;
                seq[k].band = k;
                band_u[k] = seq[k].u0;
                band_u[k + 1] = seq[k].u1;
            }
        }
        static const int tail_env = [] {
            const char* v = getenv("RT_TAIL_GEN");
            return v && *v ? atoi(v) : -1;
        }();
        rt::Frame fr[rt_ctx::kMaxBands];
        uint64_t offset[rt_ctx::kMaxBands], cap[rt_ctx::kMaxBands];
        uint64_t total = 0;
        for (int k = 0; k < n_seq; ++k) {
            if (c->use_camera)
                fill_frame_desc(c, fr[k], max_bounces, mode, out, seq[k].u0, seq[k].u1);
            else
                fill_frame_desc(c, fr[k], max_bounces, mode, out);
            if (c->use_camera && seq[k].u1 == seq[k].u0) fr[k].n_local = 0;
            if (aov_only) {
                fr[k].hit_id = c->d_hit_id.p;
                fr[k].hit_t = c->d_hit_t.p;
            }
            offset[k] = total;
            cap[k] = (fr[k].n_local + 255u) & ~(uint64_t)255;
            total += cap[k];
        }
        rt::WaveBuffers wb[rt_ctx::kMaxBands];
        rc = ensure_wave_buffers(c, std::max<uint64_t>(total, 256), gens, n_seq, offset, cap, wb);
        if (rc != RT_OK) return rc;
        rt::WaveGrids wg;
        wg.cam = c->use_camera ? c->g_cam.set : rt::GridSet{};
        wg.lit = c->g_lit.set;
        wg.g3 = c->g3;
        wg.filter_tiles = c->d_filter_tiles.p;
        RT_CUDA(cudaEventRecord(c->ev0, c->stream));  // every sequence starts after what was queued before this render
        cudaStream_t cps = c->copy_stream;  // the copy stream of the band in hand
        auto copy_rows = [&](uint64_t first, uint64_t cnt) -> cudaError_t {
            return cudaMemcpyAsync(host_out + 4 * first, c->d_fb.p + first, cnt * sizeof(float4), cudaMemcpyDeviceToHost, cps);
        };
        bool waited[rt_ctx::kMaxBands] = {};
        for (int k = 0; k < n_seq; ++k) {
            cudaStream_t st = n_seq == 1 ? c->stream : c->seq_stream[seq[k].stream];
            if (n_seq > 1 && !waited[seq[k].stream]) {
                RT_CUDA(cudaStreamWaitEvent(st, c->ev0, 0));
                waited[seq[k].stream] = true;
            }
            // auto tail: the smaller a launch sequence, the earlier its generations are thin.  Swept on the 8K cfg5 frame and its
            // 1/2, 1/4, 1/8 shards (tools/sweep_tail.py, profiles/r2_sweep_tail.txt, two lights: 3 generations per bounce level):
            // whole frame 25.2 (chain only) -> 24.8 ms with the tail from generation 9, half 13.6 -> 13.2 (9), quarter 7.9 -> 7.35 (7),
            // eighth 4.87 -> 4.31 (4..7); too early a tail costs more than it saves (whole frame from generation 3: 30.7 ms) because
            // the per-thread searches of wave_tail do not have the warp scheduling of the trace kernels.
            int tail_gen = c->opt_tail_gen >= 0 ? c->opt_tail_gen : tail_env;
            if (tail_gen < 0) {
                const uint64_t px_seq = fr[k].n_local;
                const int per_level = 1 + (int)c->scene.n_lights;
                tail_gen = px_seq >= (6ull << 20) ? 3 * per_level : px_seq >= (1ull << 20) ? 2 * per_level + 1 : per_level + 1;
            }
            static const bool trace_env = getenv("RT_SEQ_TRACE") != nullptr;
            if (trace_env && host_out) {
                if (!c->tr_start[k]) {
                    RT_CUDA(cudaEventCreate(&c->tr_start[k]));
                    RT_CUDA(cudaEventCreate(&c->tr_end[k]));
                    RT_CUDA(cudaEventCreate(&c->tr_copy[k]));
                }
                RT_CUDA(cudaEventRecord(c->tr_start[k], st));
                c->tr_units[k] = seq[k].u1 - seq[k].u0;
                c->tr_n = n_seq;
            }
            uint32_t nl = 0;
            if (fr[k].n_local > 0) RT_CUDA(rt::wave_render(c->scene, fr[k], wb[k], accel, &wg, c->sm_count, &nl, st, tail_gen));
            c->last_launches += nl;
            if (n_seq > 1 || host_out) RT_CUDA(cudaEventRecord(c->band_done[k], st));
            if (trace_env && host_out) RT_CUDA(cudaEventRecord(c->tr_end[k], st));
            const bool band_complete = k + 1 == n_seq || seq[k + 1].band != seq[k].band;
            if (host_out && band_complete) {  // the band's rows go home under the kernels of the bands behind it
                cps = staggered && (seq[k].stream & 1) ? c->copy_stream2 : c->copy_stream;
                for (int j = k; j >= 0 && seq[j].band == seq[k].band; --j) RT_CUDA(cudaStreamWaitEvent(cps, c->band_done[j], 0));
                const bool time_it = sequential && seq[k].band == 0;
                if (time_it) RT_CUDA(cudaEventRecord(c->copy_t0, cps));
                if (staggered) {
                    if (!c->cp_s[k]) {
                        RT_CUDA(cudaEventCreate(&c->cp_s[k]));
                        RT_CUDA(cudaEventCreate(&c->cp_e[k]));
                    }
                    RT_CUDA(cudaEventRecord(c->cp_s[k], cps));
                }
                const uint32_t b0 = c->use_camera ? band_u[seq[k].band] : 0u, b1 = c->use_camera ? band_u[seq[k].band + 1] : 0u;
                if (!c->use_camera) {
                    if (c->n_rays) RT_CUDA(cudaMemcpyAsync(host_out, c->d_fb.p, c->n_rays * sizeof(float4), cudaMemcpyDeviceToHost, cps));
                } else if (b1 == b0) {
                    // empty band
                } else if (c->n_shards <= 1) {
                    const uint64_t first = (uint64_t)b0 * c->width, cnt = (uint64_t)(b1 - b0) * c->width;
                    RT_CUDA(copy_rows(first, cnt));
                } else {
                    for (uint32_t u = b0; u < b1; ++u) {
                        const uint32_t row = (c->shard + u * c->n_shards) * c->block_rows;
                        if (row >= c->height) break;
                        const uint64_t first = (uint64_t)row * c->width, cnt = (uint64_t)std::min(c->block_rows, c->height - row) * c->width;
                        RT_CUDA(copy_rows(first, cnt));
                    }
                }
                if (trace_env) RT_CUDA(cudaEventRecord(c->tr_copy[k], cps));
                if (staggered) {
                    RT_CUDA(cudaEventRecord(c->cp_e[k], cps));
                    c->cp_units[k] = band_u[seq[k].band + 1] - band_u[seq[k].band];
                }
                if (time_it) {
                    RT_CUDA(cudaEventRecord(c->copy_t1, cps));
                    c->band0_units = b1 - b0;
                    c->band_units_total = units;
                }
            }
        }
        if (!sequential) c->band0_units = 0;
        c->cp_n = staggered ? n_seq : 0;
        c->last_n_seq = host_out ? n_seq : 0;
        if (n_seq > 1)
            for (int k = 0; k < n_seq; ++k) RT_CUDA(cudaStreamWaitEvent(c->stream, c->band_done[k], 0));
        RT_CUDA(cudaEventRecord(c->ev1, c->stream));
        if (host_out) {
            RT_CUDA(cudaEventRecord(c->copy_done, c->copy_stream));
            RT_CUDA(cudaStreamWaitEvent(c->stream, c->copy_done, 0));
            if (staggered) {
                RT_CUDA(cudaEventRecord(c->copy_done2, c->copy_stream2));
                RT_CUDA(cudaStreamWaitEvent(c->stream, c->copy_done2, 0));
            }
            RT_CUDA(cudaEventRecord(c->ev2, c->stream));
        }
        return RT_OK;
    }
    rt::Frame fr;
    fill_frame_desc(c, fr, max_bounces, mode, out);
    if (aov_only) {
        fr.hit_id = c->d_hit_id.p;
        fr.hit_t = c->d_hit_t.p;
    }
    c->last_launches = 1;
    RT_CUDA(cudaEventRecord(c->ev0, c->stream));
    RT_CUDA(rt::launch_megakernel(c->scene, fr, accel == RT_ACCEL_GRID ? (int)RT_ACCEL_BVH : accel, aov_only, c->stream));
    RT_CUDA(cudaEventRecord(c->ev1, c->stream));
    if (host_out) {  // readback: the shard's row-blocks (or everything)
        if (c->use_camera && c->n_shards > 1) {
            const uint32_t br = c->block_rows, n_blocks = (c->height + br - 1) / br;
            for (uint32_t b = c->shard; b < n_blocks; b += c->n_shards) {
                const uint64_t first = (uint64_t)b * br * c->width;
                const uint64_t cnt = (uint64_t)std::min(br, c->height - b * br) * c->width;
                RT_CUDA(cudaMemcpyAsync(host_out + 4 * first, c->d_fb.p + first, cnt * sizeof(float4), cudaMemcpyDeviceToHost, c->stream));
            }
        } else if (n_pixels(c)) {
            RT_CUDA(cudaMemcpyAsync(host_out, c->d_fb.p, n_pixels(c) * sizeof(float4), cudaMemcpyDeviceToHost, c->stream));
        }
        RT_CUDA(cudaEventRecord(c->ev2, c->stream));
    }
    return RT_OK;
}

}  // namespace

extern "C" {

const char* rt_last_error(void) { return g_last_error.c_str(); }
int rt_abi_version(void) { return RT_B200_ABI_VERSION; }

int rt_device_count(int* count) {
    if (!count) return fail(RT_ERR_INVALID, "null argument");
    *count = 0;
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) return RT_OK;
    for (int d = 0; d < n; ++d) {
        cudaDeviceProp prop;
        if (cudaGetDeviceProperties(&prop, d) == cudaSuccess && prop.major == 10) ++*count;
    }
    return RT_OK;
}

int rt_create(rt_ctx** out, int device) {
    if (!out) return fail(RT_ERR_INVALID, "null out pointer");
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0)
        return fail(RT_ERR_NO_DEVICE, std::string("no CUDA device (there is no CPU fallback): ") + cudaGetErrorString(e));
    if (device < 0) {
        if (cudaGetDevice(&device) != cudaSuccess) device = 0;
    }
    if (device >= n) return fail(RT_ERR_NO_DEVICE, "device index out of range");
    cudaDeviceProp prop;
    RT_CUDA(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10)
        return fail(RT_ERR_NO_DEVICE, std::string("device '") + prop.name + "' is not compute capability 10.x; this library holds sm_100a code only");
    RT_CUDA(cudaSetDevice(device));
    rt_ctx* c = new rt_ctx();
    c->device = device;
    c->sm_count = prop.multiProcessorCount;
    memset(&c->stats, 0, sizeof c->stats);
    memset(&c->scene, 0, sizeof c->scene);
    bool bands_ok = cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking) == cudaSuccess &&
                    cudaStreamCreateWithFlags(&c->copy_stream2, cudaStreamNonBlocking) == cudaSuccess;
    int prio_lo = 0, prio_hi = 0;  // numerically lower = higher priority
    if (cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi) != cudaSuccess) prio_lo = prio_hi = 0;
    for (int i = 0; i < rt_ctx::kMaxBands && bands_ok; ++i) {
        bands_ok = cudaEventCreate(&c->band_done[i]) == cudaSuccess &&  // timed: the end of the kernels is the latest of them
                   cudaStreamCreateWithPriority(&c->seq_stream[i], cudaStreamNonBlocking, std::min(prio_hi + i, prio_lo)) == cudaSuccess;
    }
    bands_ok = bands_ok && cudaEventCreateWithFlags(&c->copy_done, cudaEventDisableTiming) == cudaSuccess &&
               cudaEventCreateWithFlags(&c->copy_done2, cudaEventDisableTiming) == cudaSuccess &&
               cudaEventCreate(&c->copy_t0) == cudaSuccess && cudaEventCreate(&c->copy_t1) == cudaSuccess;
    if (!bands_ok || cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess || cudaEventCreate(&c->ev0) != cudaSuccess ||
        cudaEventCreate(&c->ev1) != cudaSuccess || cudaEventCreate(&c->ev2) != cudaSuccess) {
        delete c;
        return fail(RT_ERR_CUDA, "stream/event creation failed");
    }
    *out = c;
    return RT_OK;
}

int rt_destroy(rt_ctx* c) {
    if (!c) return RT_OK;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    c->d_minv.release(); c->d_mv.release(); c->d_mat.release(); c->d_lights.release();
    c->d_cull.release(); c->d_leaf_cull.release(); c->d_nodes.release(); c->d_lbvh_leaves.release(); c->d_lbvh_scratch.release(); c->d_leaf_obj.release();
    c->d_objs.release(); c->d_bounds_scratch.release(); c->d_bounds_globals.release();
    c->d_rays.release(); c->d_fb.release(); c->d_hit_id.release(); c->d_hit_t.release(); c->d_counters.release();
    c->d_sink.release();
    c->g_cam.grids.release(); c->g_cam.cell_start.release(); c->g_cam.entries.release();
    c->g_lit.grids.release(); c->g_lit.cell_start.release(); c->g_lit.entries.release();
    c->g_cursor.release(); c->g_scan.release(); c->d_filter_tiles.release(); c->g3_cell_start.release(); c->g3_entries.release(); c->d_aabb.release();
    c->w_acc.release(); c->w_refl.release(); c->w_P.release(); c->w_Nm.release(); c->w_Rc.release(); c->w_out.release();
    for (int i = 0; i < 2; ++i) { c->w_qs[i].release(); c->w_qd[i].release(); c->w_qpath[i].release(); }
    c->w_res_t.release(); c->w_res_id.release(); c->w_counts.release();
    if (c->ev0) cudaEventDestroy(c->ev0);
    if (c->ev1) cudaEventDestroy(c->ev1);
    if (c->ev2) cudaEventDestroy(c->ev2);
    for (int i = 0; i < rt_ctx::kMaxBands; ++i) {
        if (c->band_done[i]) cudaEventDestroy(c->band_done[i]);
        if (c->seq_stream[i]) cudaStreamDestroy(c->seq_stream[i]);
    }
    if (c->copy_done) cudaEventDestroy(c->copy_done);
    if (c->copy_done2) cudaEventDestroy(c->copy_done2);
    if (c->copy_t0) cudaEventDestroy(c->copy_t0);
    if (c->copy_t1) cudaEventDestroy(c->copy_t1);
    for (int i = 0; i < 16; ++i) {
        if (c->cp_s[i]) cudaEventDestroy(c->cp_s[i]);
        if (c->cp_e[i]) cudaEventDestroy(c->cp_e[i]);
        if (c->tr_start[i]) cudaEventDestroy(c->tr_start[i]);
        if (c->tr_end[i]) cudaEventDestroy(c->tr_end[i]);
        if (c->tr_copy[i]) cudaEventDestroy(c->tr_copy[i]);
    }
    if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
    if (c->copy_stream2) cudaStreamDestroy(c->copy_stream2);
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
    return RT_OK;
}

int rt_set_option(rt_ctx* c, int option, int64_t value) {
    if (!c) return fail(RT_ERR_INVALID, "null context");
    switch (option) {
        case RT_OPT_PIPELINE: c->opt_pipeline = (int)value; return RT_OK;
        case RT_OPT_COUNT_RAYS: c->opt_count = value ? 1 : 0; return RT_OK;
        case RT_OPT_BVH_LEAF:
            if (value < 1 || value > 16) return fail(RT_ERR_INVALID, "leaf size must be 1..16");
            c->opt_leaf = (int)value;
            c->scene_dirty = true;
            return RT_OK;
        case RT_OPT_SKIP_DEAD_RAYS: c->opt_skip_dead = value ? 1 : 0; return RT_OK;
        case RT_OPT_BVH_BUILDER:
            if (value != 0 && value != 1) return fail(RT_ERR_INVALID, "builder must be 0 (host SAH) or 1 (device LBVH)");
            c->opt_builder = (int)value;
            c->scene_dirty = true;
            return RT_OK;
        case RT_OPT_READBACK_BANDS:
            if (value < 0 || value > rt_ctx::kMaxBands) return fail(RT_ERR_INVALID, "readback bands must be 0 (auto) or 1..16");
            c->opt_bands = (int)value;
            return RT_OK;
        case RT_OPT_TAIL_GEN:
            if (value < -1 || value > 4096) return fail(RT_ERR_INVALID, "tail generation must be -1 (auto), 0 (off) or 1..4096");
            c->opt_tail_gen = (int)value;
            return RT_OK;
        case RT_OPT_DETECT_PINHOLE:
            c->opt_detect_pinhole = value ? 1 : 0;
            return RT_OK;
        case RT_OPT_ZERO_COPY:
            if (value < -1 || value > 1) return fail(RT_ERR_INVALID, "zero copy must be -1 (auto), 0 or 1");
            c->opt_zero_copy = (int)value;
            return RT_OK;
        case RT_OPT_GRID3_DENSITY:
            if (value < 0 || value > 100000) return fail(RT_ERR_INVALID, "grid density must be 0 (off) .. 100000 cells per 100 objects");
            c->opt_grid3 = (int)value;
            c->g3_dirty = true;
            return RT_OK;
        case RT_OPT_GRID_RES:
            if (value < 32 || value > 4096) return fail(RT_ERR_INVALID, "grid resolution must be 32..4096");
            c->opt_grid_res = (int)value;
            c->g_lit.dirty = true;
            return RT_OK;
        case RT_OPT_LANES:
            if (value < 0 || value > rt_ctx::kMaxBands) return fail(RT_ERR_INVALID, "lanes must be 0 (auto) or 1..16");
            c->opt_lanes = (int)value;
            return RT_OK;
        default: return fail(RT_ERR_INVALID, "unknown option");
    }
}

int rt_upload_scene(rt_ctx* c, const rt_object* objs, uint32_t n_objs, const rt_light* lights, uint32_t n_lights) {
    if (!c) return fail(RT_ERR_INVALID, "null context");
    if ((n_objs && !objs) || (n_lights && !lights)) return fail(RT_ERR_INVALID, "null scene arrays");
    RT_CUDA(cudaSetDevice(c->device));
    c->objs.assign(objs, objs + n_objs);
    c->lights.assign(lights, lights + n_lights);
    c->have_scene = true;
    c->scene_dirty = true;
    const auto t0 = std::chrono::steady_clock::now();
    const int rc = sync_scene(c);
    c->last_upload_ms = (float)std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    return rc;
}

int rt_set_rays(rt_ctx* c, const rt_ray* host_rays, uint64_t n) {
    if (!c) return fail(RT_ERR_INVALID, "null context");
    if (n && !host_rays) return fail(RT_ERR_INVALID, "null rays");
    RT_CUDA(cudaSetDevice(c->device));
    if (c->opt_detect_pinhole && n >= 2) {
        // The reference's main() hands its tracer exactly one kind of ray vector: the pinhole grid of OpenCL-Raytracer.cpp:68-72.
        // If that is what arrived, nothing is uploaded (1.06 GB at 8K): the rays are generated on the device, bit for bit the
        // same, and the frame gets the camera's perspective grid, bands and lanes -- the unmodified drop-in is the fast path.
        uint32_t w = 0, h = 0;
        float z = 0.f;
        if (detect_pinhole(host_rays, n, &w, &h, &z)) {
            const int rc = rt_set_camera(c, w, h, z);
            if (rc == RT_OK) {
                c->n_rays = n;
                c->have_rays = false;
                c->d_rays.release();
            }
            return rc;
        }
    }
    double omax2 = 0.0;
    for (uint64_t i = 0; i < n; ++i) {
        const float* s = host_rays[i].start;
        const float* d = host_rays[i].direction;
        if (!(fabsf(s[3] - 1.f) <= 1e-4f) || d[3] != 0.f)
            return fail(RT_ERR_UNSUPPORTED, "ray " + std::to_string(i) + ": start.w must be 1 and direction.w 0 (Ray3D.hpp:9)");
        const double o2 = (double)s[0] * s[0] + (double)s[1] * s[1] + (double)s[2] * s[2];
        if (!(o2 < 1e30)) return fail(RT_ERR_UNSUPPORTED, "ray " + std::to_string(i) + ": non-finite origin");
        omax2 = std::max(omax2, o2);
    }
    RT_CUDA(c->d_rays.ensure(n));
    if (n) RT_CUDA(cudaMemcpyAsync(c->d_rays.p, host_rays, n * sizeof(rt_ray), cudaMemcpyHostToDevice, c->stream));
    RT_CUDA(cudaStreamSynchronize(c->stream));
    c->n_rays = n;
    c->omax_user = (float)sqrt(omax2);
    c->have_rays = true;
    c->use_camera = false;
    RT_CUDA(c->d_fb.ensure(n));
    return RT_OK;
}

int rt_set_camera(rt_ctx* c, uint32_t width, uint32_t height, float zdir) {
    if (!c) return fail(RT_ERR_INVALID, "null context");
    if (width == 0 || height == 0 || !(zdir == zdir)) return fail(RT_ERR_INVALID, "bad camera");
    if ((uint64_t)width * height >= (1ull << 40)) return fail(RT_ERR_INVALID, "frame too large");
    RT_CUDA(cudaSetDevice(c->device));
    c->width = width;
    c->height = height;
    c->zdir = zdir;
    c->use_camera = true;
    c->g_cam.dirty = true;
    RT_CUDA(c->d_fb.ensure((uint64_t)width * height));
    return RT_OK;
}

int rt_get_camera(rt_ctx* c, uint32_t* width, uint32_t* height, float* zdir) {
    if (!c) return fail(RT_ERR_INVALID, "null context");
    if (!c->use_camera) return fail(RT_ERR_INVALID, "the context renders caller-supplied rays, not the pinhole camera");
    if (width) *width = c->width;
    if (height) *height = c->height;
    if (zdir) *zdir = c->zdir;
    return RT_OK;
}

int rt_set_shard(rt_ctx* c, uint32_t block_rows, uint32_t shard, uint32_t n_shards) {
    if (!c) return fail(RT_ERR_INVALID, "null context");
    if (block_rows == 0 || n_shards == 0 || shard >= n_shards) return fail(RT_ERR_INVALID, "bad shard");
    if (n_shards != c->n_shards) {
        // first guess of the kernel : copy time ratio the readback plan starts from (rt_render measures it from the first frame
        // on): one GPU on an idle link 2.5-3; ranks of a 4- or 8-GPU job share the host's ingest bandwidth, measured 1.2-1.4
        c->band_ratio = n_shards >= 4 ? 1.5 : 3.0;
    }
    c->block_rows = block_rows;
    c->shard = shard;
    c->n_shards = n_shards;
    return RT_OK;
}

int rt_render_device(rt_ctx* c, uint32_t max_bounces, int mode, int accel, float* dev_out) {
    int rc = check_render_args(c, mode, accel);
    if (rc != RT_OK) return rc;
    RT_CUDA(cudaSetDevice(c->device));
    int used = 0;
    float4* out = dev_out ? reinterpret_cast<float4*>(dev_out) : c->d_fb.p;
    rc = render_on_device(c, max_bounces, mode, accel, out, &used);
    if (rc != RT_OK) return rc;
    RT_CUDA(cudaStreamSynchronize(c->stream));
    rc = collect_stats(c, used, c->use_camera ? (uint64_t)shard_rows(c) * c->width : c->n_rays, 1);
    if (rc != RT_OK) return rc;
    float ms = 0.f;
    RT_CUDA(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
    c->stats.kernel_ms = ms;
    c->stats.total_ms = ms;
    return RT_OK;
}

int rt_render(rt_ctx* c, uint32_t max_bounces, int mode, int accel, float* host_out) {
    int rc = check_render_args(c, mode, accel);
    if (rc != RT_OK) return rc;
    if (!host_out) return fail(RT_ERR_INVALID, "null output buffer");
    RT_CUDA(cudaSetDevice(c->device));
    int used = 0;
    // Zero-copy: when the caller's frame is page-locked memory the device can address (rt_host_alloc, rt_host_register,
    // cudaHostAlloc), the kernels store every finished pixel straight into it -- one 128-bit store per pixel over PCIe,
    // spread over the whole frame time instead of a copy at its end, no bands.  Render() semantics are kept: the call
    // returns when the stream is idle, i.e. when the frame is in host memory (OpenCLRaytracer.cpp:91-94).
    // MEASURED AND NOT THE DEFAULT (8K cfg5 frame, B200, PCIe 5): the 33 M scattered 16-byte stores stall wave_advance --
    // kernels 33.0 -> 45.0 ms, end to end 45.1 ms against 37.2 ms for framebuffer in HBM + banded copies.
    if (c->opt_zero_copy != 0 && c->have_scene) {
        const bool wave = pick_pipeline(c, pick_accel(c, accel), max_bounces, mode) == 1;
        void* dp = nullptr;
        if ((wave || c->opt_zero_copy == 1) && n_pixels(c) > 0) {
            if (cudaHostGetDevicePointer(&dp, host_out, 0) != cudaSuccess) {
                (void)cudaGetLastError();  // pageable memory: the copy path below
                dp = nullptr;
            }
        }
        if (dp) {
            rc = render_on_device(c, max_bounces, mode, accel, reinterpret_cast<float4*>(dp), &used);
            if (rc != RT_OK) return rc;
            RT_CUDA(cudaEventRecord(c->ev2, c->stream));
            RT_CUDA(cudaStreamSynchronize(c->stream));
            rc = collect_stats(c, used, local_pixels(c), 1);
            if (rc != RT_OK) return rc;
            float ms = 0.f, ms2 = 0.f;
            RT_CUDA(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
            RT_CUDA(cudaEventElapsedTime(&ms2, c->ev0, c->ev2));
            c->stats.kernel_ms = ms;
            c->stats.total_ms = ms2;
            return RT_OK;
        }
    }
    // kernels + copy: on the wavefront pipeline the frame is cut into concurrent bands whose copies overlap the kernels of
    // the bands behind them (render_on_device)
    const auto enq0 = std::chrono::steady_clock::now();
    rc = render_on_device(c, max_bounces, mode, accel, c->d_fb.p, &used, false, host_out);
    if (rc != RT_OK) return rc;
    const double enq_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - enq0).count();
    RT_CUDA(cudaStreamSynchronize(c->stream));
    rc = collect_stats(c, used, local_pixels(c), 1);
    if (rc != RT_OK) return rc;
    float ms = 0.f, ms2 = 0.f;
    RT_CUDA(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
    RT_CUDA(cudaEventElapsedTime(&ms2, c->ev0, c->ev2));
    // the end of the kernels = the latest sequence end (ev1 sits in a stream of its own and can be held up behind the copies
    // when the streams outnumber the hardware queues)
    if (c->last_pipeline == 1 && c->last_n_seq > 0) {
        float kmax = 0.f;
        for (int k = 0; k < c->last_n_seq; ++k) {
            float t = 0.f;
            if (cudaEventElapsedTime(&t, c->ev0, c->band_done[k]) == cudaSuccess) kmax = std::max(kmax, t);
            else (void)cudaGetLastError();
        }
        if (kmax > 0.f) ms = std::min(ms, kmax);
    }
    c->stats.kernel_ms = ms;  // all bands' kernels (copies of finished bands run underneath)
    c->stats.total_ms = ms2;
    if (c->tr_n > 0 && getenv("RT_SEQ_TRACE")) {
        fprintf(stderr, "[rt seq trace] dev %d shard %u/%u kernels %.2f total %.2f host enqueue %.2f ms, %u launches, ratio %.2f\n", c->device, c->shard,
                c->n_shards, ms, ms2, enq_ms, c->last_launches, c->band_ratio);
        for (int k = 0; k < c->tr_n; ++k) {
            float a = 0.f, b = 0.f, cp = -1.f;
            cudaEventElapsedTime(&a, c->ev0, c->tr_start[k]);
            cudaEventElapsedTime(&b, c->ev0, c->tr_end[k]);
            if (cudaEventElapsedTime(&cp, c->ev0, c->tr_copy[k]) != cudaSuccess) {
                (void)cudaGetLastError();
                cp = -1.f;
            }
            fprintf(stderr, "  seq %d units %u start %.2f end %.2f copy_end %.2f\n", k, c->tr_units[k], a, b, cp);
        }
    }
    if (c->cp_n > 0 && c->band_units_total && c->last_pipeline == 1) {  // next frame's plan: all bands' copy times against the kernels
        double dur = 0.0, un = 0.0;
        for (int k = 0; k < c->cp_n; ++k) {
            float t = 0.f;
            if (c->cp_units[k] && cudaEventElapsedTime(&t, c->cp_s[k], c->cp_e[k]) == cudaSuccess && t > 0.f) {
                dur += t;
                un += c->cp_units[k];
            } else {
                (void)cudaGetLastError();
            }
        }
        if (dur > 0.0 && un > 0.0 && ms > 0.f) {
            const double r = ((double)ms / c->band_units_total) / (dur / un);
            c->band_ratio = std::min(8.0, std::max(0.3, 0.5 * c->band_ratio + 0.5 * r));
        }
    } else if (c->band0_units && c->band_units_total && c->last_pipeline == 1) {  // next frame's band split
        float cp = 0.f;
        if (cudaEventElapsedTime(&cp, c->copy_t0, c->copy_t1) == cudaSuccess && cp > 0.f && ms > 0.f) {
            const double r = ((double)ms / c->band_units_total) / ((double)cp / c->band0_units);
            c->band_ratio = std::min(8.0, std::max(0.3, 0.5 * c->band_ratio + 0.5 * r));
        } else {
            (void)cudaGetLastError();
        }
    }
    return RT_OK;
}

int rt_framebuffer(rt_ctx* c, float** dev_ptr, uint64_t* n) {
    if (!c || !dev_ptr) return fail(RT_ERR_INVALID, "null argument");
    if (!c->use_camera && !c->have_rays) return fail(RT_ERR_INVALID, "no rays / camera set");
    *dev_ptr = reinterpret_cast<float*>(c->d_fb.p);
    if (n) *n = n_pixels(c);
    return RT_OK;
}

int rt_readback(rt_ctx* c, float* host_out) {
    if (!c || !host_out) return fail(RT_ERR_INVALID, "null argument");
    if (!c->use_camera && !c->have_rays) return fail(RT_ERR_INVALID, "no rays / camera set");
    RT_CUDA(cudaSetDevice(c->device));
    const uint64_t n = n_pixels(c);
    if (n) RT_CUDA(cudaMemcpyAsync(host_out, c->d_fb.p, n * sizeof(float4), cudaMemcpyDeviceToHost, c->stream));
    RT_CUDA(cudaStreamSynchronize(c->stream));
    return RT_OK;
}

int rt_render_aov(rt_ctx* c, int accel, int32_t* host_hit_id, float* host_hit_t) {
    int rc = check_render_args(c, RT_MODE_HITTEST, accel);
    if (rc != RT_OK) return rc;
    if (!host_hit_id || !host_hit_t) return fail(RT_ERR_INVALID, "null output buffer");
    RT_CUDA(cudaSetDevice(c->device));
    const uint64_t n = n_pixels(c);
    RT_CUDA(c->d_hit_id.ensure(n));
    RT_CUDA(c->d_hit_t.ensure(n));
    if (c->use_camera && c->n_shards > 1) {  // rows of other shards read as "miss"
        RT_CUDA(cudaMemsetAsync(c->d_hit_id.p, 0xff, n * sizeof(int32_t), c->stream));
        RT_CUDA(cudaMemsetAsync(c->d_hit_t.p, 0, n * sizeof(float), c->stream));
    }
    rc = render_on_device(c, 0, RT_MODE_HITTEST, accel, nullptr, &accel, true);  // AOVs only: the framebuffer is not touched
    if (rc != RT_OK) return rc;
    if (n) {
        RT_CUDA(cudaMemcpyAsync(host_hit_id, c->d_hit_id.p, n * sizeof(int32_t), cudaMemcpyDeviceToHost, c->stream));
        RT_CUDA(cudaMemcpyAsync(host_hit_t, c->d_hit_t.p, n * sizeof(float), cudaMemcpyDeviceToHost, c->stream));
    }
    RT_CUDA(cudaStreamSynchronize(c->stream));
    rc = collect_stats(c, accel, n, 1);
    if (rc != RT_OK) return rc;
    float ms = 0.f;
    RT_CUDA(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
    c->stats.kernel_ms = c->stats.total_ms = ms;
    return RT_OK;
}

int rt_trace(rt_ctx* c, const rt_ray* host_rays, uint64_t n, int accel, int32_t* host_hit_id, float* host_hit_t) {
    if (!c) return fail(RT_ERR_INVALID, "null context");
    if (accel < RT_ACCEL_AUTO || accel > RT_ACCEL_GRID) return fail(RT_ERR_INVALID, "bad accel");
    if (n && (!host_rays || !host_hit_id || !host_hit_t)) return fail(RT_ERR_INVALID, "null argument");
    RT_CUDA(cudaSetDevice(c->device));
    double omax2 = 0.0;
    for (uint64_t i = 0; i < n; ++i) {
        const float* s = host_rays[i].start;
        if (!(fabsf(s[3] - 1.f) <= 1e-4f) || host_rays[i].direction[3] != 0.f)
            return fail(RT_ERR_UNSUPPORTED, "ray " + std::to_string(i) + ": start.w must be 1 and direction.w 0 (Ray3D.hpp:9)");
        omax2 = std::max(omax2, (double)s[0] * s[0] + (double)s[1] * s[1] + (double)s[2] * s[2]);
    }
    if (!(omax2 < 1e30)) return fail(RT_ERR_UNSUPPORTED, "non-finite ray origin");
    const bool was_camera = c->use_camera;
    const float saved = c->omax_user;
    c->omax_user = std::max(c->omax_user, (float)sqrt(omax2));
    c->use_camera = false;  // make sync_scene honour omax_user
    int rc = sync_scene(c);
    c->use_camera = was_camera;
    if (rc != RT_OK) {
        c->omax_user = saved;
        return rc;
    }
    accel = pick_accel(c, accel);
    if (accel == RT_ACCEL_GRID) accel = RT_ACCEL_BVH;  // arbitrary rays have no common apex
    DevBuf<rt::Ray> rays;
    DevBuf<int32_t> ids;
    DevBuf<float> ts;
    cudaError_t e = rays.ensure(n);
    if (e == cudaSuccess) e = ids.ensure(n);
    if (e == cudaSuccess) e = ts.ensure(n);
    if (e == cudaSuccess && n) e = cudaMemcpyAsync(rays.p, host_rays, n * sizeof(rt_ray), cudaMemcpyHostToDevice, c->stream);
    rt::Counters* ctr = nullptr;
    if (e == cudaSuccess && c->opt_count) {
        e = c->d_counters.ensure(1);
        if (e == cudaSuccess) e = cudaMemsetAsync(c->d_counters.p, 0, sizeof(rt::Counters), c->stream);
        ctr = c->d_counters.p;
    }
    if (e == cudaSuccess) e = cudaEventRecord(c->ev0, c->stream);
    if (e == cudaSuccess) e = rt::launch_trace(c->scene, rays.p, n, accel, ids.p, ts.p, ctr, c->stream);
    if (e == cudaSuccess) e = cudaEventRecord(c->ev1, c->stream);
    if (e == cudaSuccess && n) e = cudaMemcpyAsync(host_hit_id, ids.p, n * sizeof(int32_t), cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess && n) e = cudaMemcpyAsync(host_hit_t, ts.p, n * sizeof(float), cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    rays.release();
    ids.release();
    ts.release();
    if (e != cudaSuccess) return fail(RT_ERR_CUDA, std::string("rt_trace: ") + cudaGetErrorString(e));
    rc = collect_stats(c, accel, n, 1);
    if (rc != RT_OK) return rc;
    float ms = 0.f;
    RT_CUDA(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
    c->stats.kernel_ms = c->stats.total_ms = ms;
    return RT_OK;
}

// ---- rt_trace without any allocation.  rt_trace stages the rays and the results in three buffers it cudaMallocs and frees on
// every call; callers that may not allocate get the same kernel with the staging in a fixed array that is part of the library's
// image: every GPU that loads librt_b200.so holds RT_TRACE_STATIC_BYTES (10 MiB) of it until the process exits.  The array is
// shared by every context and every call on a device and keeps the previous call's contents: the kernel overwrites every
// element the call reads back (nothing has to start at zero), but calls that overlap on one GPU race.
}  // extern "C"

namespace {

constexpr uint64_t kStaticRays = RT_TRACE_STATIC_MAX_RAYS;
constexpr uint64_t kStaticOffIds = kStaticRays * sizeof(rt::Ray);                 // rays first (32 B each) ...
constexpr uint64_t kStaticOffTs = kStaticOffIds + kStaticRays * sizeof(int32_t);  // ... then the hit ids, then t
constexpr uint64_t kStaticBytes = kStaticOffTs + kStaticRays * sizeof(float);
static_assert(kStaticBytes == RT_TRACE_STATIC_BYTES, "rt_b200.h states the size of the static scratch");
static_assert(kStaticOffIds % 256 == 0 && kStaticOffTs % 256 == 0, "sub-buffers keep the array's alignment");

static __device__ __align__(256) unsigned char g_trace_scratch[kStaticBytes];

// In host code `g_trace_scratch` is only a handle of the symbol: its device address differs per GPU and comes from
// cudaGetSymbolAddress with that GPU current.  Looked up once per device.
constexpr int kMaxStaticDevices = 64;
std::mutex g_static_mutex;
unsigned char* g_static_addr[kMaxStaticDevices] = {};

int static_scratch(int device, unsigned char** out) {
    if (device < 0 || device >= kMaxStaticDevices) return fail(RT_ERR_UNSUPPORTED, "rt_trace_static: device index beyond 63");
    std::lock_guard<std::mutex> lock(g_static_mutex);
    if (!g_static_addr[device]) {
        cudaError_t e = cudaSetDevice(device);
        void* p = nullptr;
        if (e == cudaSuccess) e = cudaGetSymbolAddress(&p, g_trace_scratch);
        if (e != cudaSuccess || !p)
            return fail(RT_ERR_CUDA, std::string("rt_trace_static: cudaGetSymbolAddress: ") + cudaGetErrorString(e));
        g_static_addr[device] = static_cast<unsigned char*>(p);
    }
    *out = g_static_addr[device];
    return RT_OK;
}

}  // namespace

extern "C" {

int rt_trace_static(rt_ctx* c, const rt_ray* host_rays, uint64_t n, int accel, int32_t* host_hit_id, float* host_hit_t) {
    if (!c) return fail(RT_ERR_INVALID, "null context");
    if (accel < RT_ACCEL_AUTO || accel > RT_ACCEL_GRID) return fail(RT_ERR_INVALID, "bad accel");
    if (n > kStaticRays)
        return fail(RT_ERR_UNSUPPORTED, "rt_trace_static: " + std::to_string(n) + " rays exceed the static scratch of " +
                                            std::to_string(kStaticRays) + " rays (RT_TRACE_STATIC_MAX_RAYS); split the batch or use rt_trace");
    if (n && (!host_rays || !host_hit_id || !host_hit_t)) return fail(RT_ERR_INVALID, "null argument");
    RT_CUDA(cudaSetDevice(c->device));
    unsigned char* scratch = nullptr;
    int rc = static_scratch(c->device, &scratch);  // n == 0: the per-device set-up alone
    if (rc != RT_OK || n == 0) return rc;
    double omax2 = 0.0;
    for (uint64_t i = 0; i < n; ++i) {
        const float* s = host_rays[i].start;
        if (!(fabsf(s[3] - 1.f) <= 1e-4f) || host_rays[i].direction[3] != 0.f)
            return fail(RT_ERR_UNSUPPORTED, "ray " + std::to_string(i) + ": start.w must be 1 and direction.w 0 (Ray3D.hpp:9)");
        omax2 = std::max(omax2, (double)s[0] * s[0] + (double)s[1] * s[1] + (double)s[2] * s[2]);
    }
    if (!(omax2 < 1e30)) return fail(RT_ERR_UNSUPPORTED, "non-finite ray origin");
    // The scene must already be resident and prepared for origins this far out: preparing it again would allocate.
    if (!c->have_scene) return fail(RT_ERR_INVALID, "rt_trace_static: no scene uploaded");
    if (c->scene_dirty || !((float)sqrt(omax2) <= c->hs.omax))
        return fail(RT_ERR_UNSUPPORTED, "rt_trace_static: the resident scene is not prepared for these rays (origins up to " +
                                            std::to_string(sqrt(omax2)) + " from the world origin, scene prepared for " +
                                            std::to_string(c->hs.omax) + "): call rt_trace or rt_set_rays once with such rays first");
    accel = pick_accel(c, accel);
    if (accel == RT_ACCEL_GRID) accel = RT_ACCEL_BVH;  // arbitrary rays have no common apex
    rt::Ray* d_rays = reinterpret_cast<rt::Ray*>(scratch);
    int32_t* d_ids = reinterpret_cast<int32_t*>(scratch + kStaticOffIds);
    float* d_ts = reinterpret_cast<float*>(scratch + kStaticOffTs);
    cudaError_t e = cudaMemcpyAsync(d_rays, host_rays, n * sizeof(rt_ray), cudaMemcpyHostToDevice, c->stream);
    if (e == cudaSuccess) e = rt::launch_trace(c->scene, d_rays, n, accel, d_ids, d_ts, nullptr, c->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(host_hit_id, d_ids, n * sizeof(int32_t), cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(host_hit_t, d_ts, n * sizeof(float), cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    if (e != cudaSuccess) return fail(RT_ERR_CUDA, std::string("rt_trace_static: ") + cudaGetErrorString(e));
    return RT_OK;
}

int rt_get_stats(rt_ctx* c, rt_stats* out) {
    if (!c || !out) return fail(RT_ERR_INVALID, "null argument");
    *out = c->stats;
    return RT_OK;
}

int rt_quantize_rgb8(rt_ctx* c, const float* host_rgba, uint64_t n, uint8_t* host_rgb_out, uint64_t* n_negative) {
    if (!c) return fail(RT_ERR_INVALID, "null context");
    if (n_negative) *n_negative = 0;
    const float4* src = nullptr;
    DevBuf<float4> in;
    RT_CUDA(cudaSetDevice(c->device));
    if (!host_rgba) {  // the frame the last render left in this context's device framebuffer: nothing is uploaded
        if (!c->use_camera && !c->have_rays) return fail(RT_ERR_INVALID, "no frame: neither rays nor a camera are set");
        n = n_pixels(c);
        src = c->d_fb.p;
    }
    if (n == 0) return RT_OK;
    if (!host_rgb_out) return fail(RT_ERR_INVALID, "null argument");
    DevBuf<uint8_t> outb;
    DevBuf<unsigned long long> neg;
    cudaError_t e = outb.ensure(3 * n);
    if (e == cudaSuccess) e = neg.ensure(1);
    if (e == cudaSuccess) e = cudaMemsetAsync(neg.p, 0, sizeof(unsigned long long), c->stream);
    if (e == cudaSuccess && host_rgba) {
        e = in.ensure(n);
        if (e == cudaSuccess) e = cudaMemcpyAsync(in.p, host_rgba, n * sizeof(float4), cudaMemcpyHostToDevice, c->stream);
        src = in.p;
    }
    unsigned long long h_neg = 0;
    if (e == cudaSuccess) e = rt::launch_quantize(src, n, outb.p, neg.p, c->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(host_rgb_out, outb.p, 3 * n, cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(&h_neg, neg.p, sizeof h_neg, cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    in.release();
    outb.release();
    neg.release();
    if (e != cudaSuccess) return fail(RT_ERR_CUDA, std::string("rt_quantize_rgb8: ") + cudaGetErrorString(e));
    if (n_negative) *n_negative = h_neg;
    return RT_OK;
}

int rt_plan_bands(uint32_t height, uint32_t block_rows, uint32_t shard, uint32_t n_shards, int n_bands, double ratio, uint32_t* bounds) {
    if (!bounds || n_bands < 1 || n_bands > rt_ctx::kMaxBands || !(ratio >= 1.0) || block_rows == 0 || n_shards == 0 || shard >= n_shards)
        return fail(RT_ERR_INVALID, "bad band plan arguments");
    const uint32_t units = plan_units(height, block_rows, shard, n_shards);
    bounds[0] = 0;
    for (int b = 0; b < n_bands; ++b) {
        uint32_t u0, u1;
        plan_band(units, b, n_bands, ratio, &u0, &u1);
        bounds[b + 1] = u1;
    }
    return RT_OK;
}

int rt_plan_sequences(uint32_t width, uint32_t height, uint32_t block_rows, uint32_t shard, uint32_t n_shards, int n_sequences, double q,
                      double kernel_copy_ratio, uint32_t* bounds, int* n_out) {
    if (!bounds || !n_out || n_sequences < 0 || n_sequences > rt_ctx::kMaxBands || !(q >= 0.0) || q > 1.0 || !(kernel_copy_ratio > 0.0) || width == 0 ||
        block_rows == 0 || n_shards == 0 || shard >= n_shards)
        return fail(RT_ERR_INVALID, "bad sequence plan arguments");
    const uint32_t units = plan_units(height, block_rows, shard, n_shards);
    const uint64_t unit_px = (uint64_t)width * (n_shards <= 1 ? 1u : block_rows);
    *n_out = plan_sequences(units, unit_px, n_sequences, q, kernel_copy_ratio, (int)rt_ctx::kMaxBands, bounds, nullptr);
    return RT_OK;
}

int rt_host_build_bvh(const rt_object* objs, uint32_t n_objs, const rt_light* lights, uint32_t n_lights, float omax_user, int leaf_size,
                      int threads, rt_bvh_info* info, void* nodes_out, uint64_t nodes_capacity_bytes, uint32_t* leaf_obj_out,
                      uint64_t leaf_capacity) {
    if (!info) return fail(RT_ERR_INVALID, "null info");
    if (leaf_size < 1 || leaf_size > 16) return fail(RT_ERR_INVALID, "leaf size must be 1..16");
    rt::HostScene hs;
    std::string err;
    const auto t0 = std::chrono::steady_clock::now();
    const int rc = rt::build_scene(objs, n_objs, lights, n_lights, omax_user, leaf_size, hs, err, threads);
    const auto t1 = std::chrono::steady_clock::now();
    if (rc != RT_OK) return fail(rc, err);
    memset(info, 0, sizeof *info);
    info->n_nodes = (uint32_t)hs.nodes.size();
    info->n_leaf_objs = (uint32_t)hs.leaf_obj.size();
    info->depth = hs.bvh_depth;
    info->root = hs.root;
    info->omax = hs.omax;
    info->scene_radius = hs.scene_radius;
    info->build_ms = std::chrono::duration<double, std::milli>(t1 - t0).count();
    if (nodes_out) {
        if (nodes_capacity_bytes < hs.nodes.size() * sizeof(rt::BvhNode)) return fail(RT_ERR_INVALID, "nodes_out too small");
        memcpy(nodes_out, hs.nodes.data(), hs.nodes.size() * sizeof(rt::BvhNode));
    }
    if (leaf_obj_out) {
        if (leaf_capacity < hs.leaf_obj.size()) return fail(RT_ERR_INVALID, "leaf_obj_out too small");
        memcpy(leaf_obj_out, hs.leaf_obj.data(), hs.leaf_obj.size() * sizeof(uint32_t));
    }
    return RT_OK;
}

int rt_measure_fp32_peak(rt_ctx* c, float* tflops, float* sm_mhz_hint) {
    if (!c || !tflops) return fail(RT_ERR_INVALID, "null argument");
    RT_CUDA(cudaSetDevice(c->device));
    cudaDeviceProp prop;
    RT_CUDA(cudaGetDeviceProperties(&prop, c->device));
    RT_CUDA(c->d_sink.ensure(1));
    const int blocks = prop.multiProcessorCount * 8, threads = 256, iters = 4096;
    float best = 0.f;
    // scalar FFMA and packed FFMA2 loops; the denominator is whichever the device runs faster
    for (int packed = 0; packed < 2; ++packed) {
        float b = 0.f;
        for (int rep = 0; rep < 5; ++rep) {
            RT_CUDA(cudaEventRecord(c->ev0, c->stream));
            RT_CUDA(rt::launch_fp32_peak(c->d_sink.p, iters, blocks, threads, packed != 0, c->stream));
            RT_CUDA(cudaEventRecord(c->ev1, c->stream));
            RT_CUDA(cudaStreamSynchronize(c->stream));
            float ms = 0.f;
            RT_CUDA(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
            const double flops = (packed ? 4.0 : 2.0) * 8.0 * 16.0 * (double)iters * (double)blocks * threads;
            if (rep > 0 && ms > 0.f) b = std::max(b, (float)(flops / (ms * 1e-3) / 1e12));
        }
        c->fp32_peak_tf[packed] = b;
        best = std::max(best, b);
    }
    *tflops = best;
    if (sm_mhz_hint) *sm_mhz_hint = best > 0.f ? (float)(best * 1e12 / (2.0 * 128.0 * prop.multiProcessorCount) / 1e6) : 0.f;
    return RT_OK;
}

int rt_fp32_peak_detail(rt_ctx* c, float* scalar_tflops, float* packed_tflops) {
    if (!c || !scalar_tflops || !packed_tflops) return fail(RT_ERR_INVALID, "null argument");
    *scalar_tflops = c->fp32_peak_tf[0];
    *packed_tflops = c->fp32_peak_tf[1];
    return RT_OK;
}

int rt_ipc_export_framebuffer(rt_ctx* c, void* handle) {
    if (!c || !handle) return fail(RT_ERR_INVALID, "null argument");
    if (!c->d_fb.p) return fail(RT_ERR_INVALID, "no framebuffer yet: call rt_set_camera / rt_set_rays first");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    RT_CUDA(cudaSetDevice(c->device));
    cudaIpcMemHandle_t h;
    RT_CUDA(cudaIpcGetMemHandle(&h, c->d_fb.p));
    memcpy(handle, &h, sizeof h);
    return RT_OK;
}

int rt_ipc_open(rt_ctx* c, const void* handle, void** dev_ptr) {
    if (!c || !handle || !dev_ptr) return fail(RT_ERR_INVALID, "null argument");
    RT_CUDA(cudaSetDevice(c->device));
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof h);
    RT_CUDA(cudaIpcOpenMemHandle(dev_ptr, h, cudaIpcMemLazyEnablePeerAccess));
    return RT_OK;
}

int rt_ipc_close(rt_ctx* c, void* dev_ptr) {
    if (!c) return fail(RT_ERR_INVALID, "null context");
    RT_CUDA(cudaSetDevice(c->device));
    if (dev_ptr) RT_CUDA(cudaIpcCloseMemHandle(dev_ptr));
    return RT_OK;
}

int rt_host_alloc(void** out, uint64_t bytes) {
    if (!out) return fail(RT_ERR_INVALID, "null argument");
    RT_CUDA(cudaHostAlloc(out, std::max<uint64_t>(bytes, 1), cudaHostAllocPortable | cudaHostAllocMapped));
    return RT_OK;
}

int rt_host_register(void* p, uint64_t bytes) {
    if (!p || !bytes) return fail(RT_ERR_INVALID, "null argument");
    RT_CUDA(cudaHostRegister(p, bytes, cudaHostRegisterPortable | cudaHostRegisterMapped));
    return RT_OK;
}

int rt_host_unregister(void* p) {
    if (p) RT_CUDA(cudaHostUnregister(p));
    return RT_OK;
}

int rt_host_free(void* p) {
    if (p) RT_CUDA(cudaFreeHost(p));
    return RT_OK;
}

}  // extern "C"
