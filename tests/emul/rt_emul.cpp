// TEST INFRASTRUCTURE -- not part of the product path.
//
// Host emulation of the CUDA path's device code: the same headers the kernels are made of, compiled for the
// CPU through cuda_shim.h, driven one pixel at a time.  Used by the CPU-only tests to check filters, BVH and
// the per-pixel state machine against the oracle before GPU time is spent.  Never loaded by the product.
#include "cuda_shim.h"

#include <string>
#include <thread>
#include <vector>
#include <atomic>

#include <algorithm>

#include "../../opencl-raytracer_b200/csrc/rt_pixel.cuh"
#include "../../opencl-raytracer_b200/csrc/rt_pgrid.cuh"
#include "../../opencl-raytracer_b200/csrc/rt_grid3.cuh"
#include "../../opencl-raytracer_b200/csrc/rt_wave.cuh"
#include "../../opencl-raytracer_b200/csrc/rt_scene.hpp"

// same formula as rt::wave_max_generations (rt_wavefront.cu, a CUDA translation unit)
static uint32_t rt_wave_max_generations_host(uint32_t max_bounces, uint32_t n_lights, int mode) {
    if (mode == 0) return 1;
    if (mode == 1) return 1 + n_lights;
    return 1u + n_lights + max_bounces * (1u + n_lights);
}

namespace {

void bind(const rt::HostScene& hs, rt::Scene& s) {
    memset(&s, 0, sizeof s);
    s.minv = reinterpret_cast<const float4*>(hs.minv.data());
    s.mv = reinterpret_cast<const float4*>(hs.mv.data());
    s.mat = reinterpret_cast<const float4*>(hs.mat.data());
    s.cull = hs.cull.data();
    s.nodes = hs.nodes.data();
    s.leaf_obj = hs.leaf_obj.data();
    s.leaf_cull = hs.leaf_cull.data();
    s.lights = reinterpret_cast<const rt::Light*>(hs.lights.data());
    s.n_objs = hs.n_objs;
    s.n_lights = hs.n_lights;
    s.n_nodes = (uint32_t)hs.nodes.size();
    s.root = hs.root;
    s.root_f = hs.f_global;
    s.filter_h = hs.h_min;
}

// Host stand-in for the device builder of the perspective grids (rt_pgrid.cu): the same per-object functions
// (pgrid_object_sphere, pgrid_rect), plain loops instead of count / scan / fill kernels.
struct HostGrids {
    std::vector<rt::PGrid> grids;
    std::vector<uint32_t> cell_start;
    std::vector<rt::GridEntry> entries;
    rt::GridSet set{};
    void build(const rt::HostScene& hs, uint32_t valid_mask) {
        uint32_t cells = 0;
        for (auto& g : grids) {
            g.cell_base = cells;
            cells += g.nu * g.nv;
        }
        std::vector<std::vector<rt::GridEntry>> lists(cells);
        for (const auto& g : grids) {
            if (!g.nu) continue;
            for (uint32_t o = 0; o < hs.n_objs; ++o) {
                float r, dnear;
                if (!rt::pgrid_object_sphere(g, hs.cull[o], hs.omax, r, dnear)) continue;
                int u0, u1, v0, v1;
                if (!rt::pgrid_rect(g, hs.cull[o].cx, hs.cull[o].cy, hs.cull[o].cz, r, u0, u1, v0, v1)) continue;
                for (int v = v0; v <= v1; ++v)
                    for (int u = u0; u <= u1; ++u) lists[g.cell_base + (uint32_t)v * g.nu + (uint32_t)u].push_back(rt::GridEntry{o, dnear});
            }
        }
        cell_start.assign(cells + 1, 0);
        for (uint32_t c = 0; c < cells; ++c) {
            std::sort(lists[c].begin(), lists[c].end(),
                      [](const rt::GridEntry& a, const rt::GridEntry& b) { return a.dnear < b.dnear || (a.dnear == b.dnear && a.obj < b.obj); });
            cell_start[c + 1] = cell_start[c] + (uint32_t)lists[c].size();
            entries.insert(entries.end(), lists[c].begin(), lists[c].end());
        }
        set.grids = grids.data();
        set.cell_start = cell_start.data();
        set.entries = entries.data();
        set.n_grids = (uint32_t)grids.size();
        set.valid_mask = valid_mask;
    }
};

// Host stand-in for the device builder of the uniform 3-D grid (rt_pgrid.cu), same per-object function.
struct HostGrid3 {
    std::vector<uint32_t> cell_start, entries;
    rt::Grid3 g{};
    void build(const rt::HostScene& hs, float cells_per_object) {
        double lo[3] = {1e30, 1e30, 1e30}, hi[3] = {-1e30, -1e30, -1e30};
        uint32_t live = 0;
        const float ext0 = 2.f * hs.scene_radius + 1.f;
        const float eps = 2e-5f * ext0 + 1e-4f;
        for (uint32_t o = 0; o < hs.n_objs; ++o) {
            const rt::F4& hb = hs.aabb[o];
            if (hb.w < 0.f || hs.cull[o].r2 < 0.f) continue;
            ++live;
            const float c[3] = {hs.cull[o].cx, hs.cull[o].cy, hs.cull[o].cz};
            const float half[3] = {hb.x, hb.y, hb.z};
            for (int k = 0; k < 3; ++k) {
                lo[k] = std::min(lo[k], (double)c[k] - half[k] - eps);
                hi[k] = std::max(hi[k], (double)c[k] + half[k] + eps);
            }
        }
        if (!live) return;
        double vol = 1;
        for (int k = 0; k < 3; ++k) vol *= std::max(hi[k] - lo[k], 1e-3);
        float h = (float)cbrt(vol / (cells_per_object * live));
        for (int k = 0; k < 3; ++k) h = std::max(h, (float)((hi[k] - lo[k]) / 1024.0));
        g.h = h;
        g.inv_h = 1.f / h;
        g.lox = (float)lo[0] - 2 * eps; g.loy = (float)lo[1] - 2 * eps; g.loz = (float)lo[2] - 2 * eps;
        g.nx = (uint32_t)((hi[0] + 2 * eps - g.lox) / h) + 1;
        g.ny = (uint32_t)((hi[1] + 2 * eps - g.loy) / h) + 1;
        g.nz = (uint32_t)((hi[2] + 2 * eps - g.loz) / h) + 1;
        g.eps = eps;
        const uint32_t cells = g.nx * g.ny * g.nz;
        std::vector<std::vector<uint32_t>> lists(cells);
        for (uint32_t o = 0; o < hs.n_objs; ++o) {
            int l[3], u[3];
            const rt::F4& hb = hs.aabb[o];
            if (!rt::grid3_object_cells(g, hs.cull[o], hb.x, hb.y, hb.z, hb.w, l, u)) continue;
            for (int z = l[2]; z <= u[2]; ++z)
                for (int y = l[1]; y <= u[1]; ++y)
                    for (int x = l[0]; x <= u[0]; ++x) lists[((uint32_t)z * g.ny + y) * g.nx + x].push_back(o);
        }
        cell_start.assign(cells + 1, 0);
        for (uint32_t c = 0; c < cells; ++c) {
            cell_start[c + 1] = cell_start[c] + (uint32_t)lists[c].size();
            entries.insert(entries.end(), lists[c].begin(), lists[c].end());
        }
        if (entries.empty()) entries.push_back(0);
        g.cell_start = cell_start.data();
        g.entries = entries.data();
        g.valid = 1;
    }
};

void camera_grid(HostGrids& hg, const rt::HostScene& hs, uint32_t W, uint32_t H, float zdir, uint32_t tw, uint32_t th) {
    rt::PGrid g;
    memset(&g, 0, sizeof g);
    const float f = -zdir;
    g.axis = 2u | 4u;
    g.u0 = -((float)W / 2.0f) / f;
    g.v0 = -((float)H / 2.0f) / f;
    g.inv_du = f / (float)tw;
    g.inv_dv = f / (float)th;
    g.nu = (W + tw - 1) / tw;
    g.nv = H / th + 1;
    g.kind = 0;
    hg.grids.assign(1, g);
    hg.build(hs, 1u);
}

void light_grids(HostGrids& hg, const rt::HostScene& hs, const rt_light* lights, uint32_t n_lights, uint32_t res) {
    const uint32_t L = n_lights < RT_GRID_MAX_LIGHTS ? n_lights : RT_GRID_MAX_LIGHTS;
    hg.grids.resize(6 * (size_t)L);
    uint32_t valid = 0;
    for (uint32_t li = 0; li < L; ++li) {
        const bool ok = lights[li].position[3] != 0.f;
        for (uint32_t face = 0; face < 6; ++face) {
            rt::PGrid& g = hg.grids[6 * li + face];
            memset(&g, 0, sizeof g);
            if (!ok) continue;
            g.ax = lights[li].position[0];
            g.ay = lights[li].position[1];
            g.az = lights[li].position[2];
            g.axis = (face % 3u) | (face >= 3u ? 4u : 0u);
            g.u0 = g.v0 = -1.f;
            g.inv_du = g.inv_dv = (float)res / 2.f;
            g.nu = g.nv = res;
            g.kind = 1;
        }
        if (ok) valid |= 1u << li;
    }
    hg.build(hs, valid);
}

template <int ACCEL>
void run(const rt::Scene& sc, const rt::Frame& fr, uint64_t n, int nthreads, rt::Counters* total) {
    if (nthreads <= 0) nthreads = (int)std::thread::hardware_concurrency();
    if (nthreads < 1) nthreads = 1;
    std::atomic<uint64_t> next{0};
    std::vector<rt::LocalCounters> lcs(nthreads);
    auto worker = [&](int t) {
        for (;;) {
            const uint64_t b = next.fetch_add(256);
            if (b >= n) return;
            const uint64_t e = b + 256 < n ? b + 256 : n;
            for (uint64_t i = b; i < e; ++i) {
                rt::Ray ray;
                ray.s = reinterpret_cast<const float4*>(fr.rays + i)[0];
                ray.d = reinterpret_cast<const float4*>(fr.rays + i)[1];
                rt::render_pixel<ACCEL, true>(sc, fr, ray, i, false, lcs[t]);
            }
        }
    };
    std::vector<std::thread> pool;
    for (int t = 0; t < nthreads; ++t) pool.emplace_back(worker, t);
    for (auto& t : pool) t.join();
    memset(total, 0, sizeof *total);
    for (auto& l : lcs) {
        total->ref_primary += l.ref_primary;
        total->ref_shadow += l.ref_shadow;
        total->ref_reflect += l.ref_reflect;
        total->traced_shadow += l.traced_shadow;
        total->traced_reflect += l.traced_reflect;
        total->pair_tests += l.pair_tests;
        total->exact_tests += l.exact_tests;
        total->node_visits += l.node_visits;
        total->hit_pixels += l.hit_pixels;
    }
}

}  // namespace

extern "C" {

// counters_out: 9 x uint64 in the order of rt::Counters.  Returns rt_status; message via emul_last_error().
static thread_local std::string g_err;
const char* emul_last_error(void) { return g_err.c_str(); }

int emul_render(const rt_object* objs, uint32_t n_objs, const rt_light* lights, uint32_t n_lights, const rt_ray* rays,
                uint64_t n_rays, uint32_t max_bounces, int mode, int accel, int skip_dead, int leaf, float omax_user,
                float* out_rgba, int32_t* hit_id, float* hit_t, uint64_t* counters_out, int nthreads) {
    rt::HostScene hs;
    const int rc = rt::build_scene(objs, n_objs, lights, n_lights, omax_user, leaf, hs, g_err);
    if (rc != RT_OK) return rc;
    rt::Scene sc;
    bind(hs, sc);
    rt::Frame fr;
    memset(&fr, 0, sizeof fr);
    fr.rays = reinterpret_cast<const rt::Ray*>(rays);
    fr.n_local = n_rays;
    fr.max_bounces = max_bounces;
    fr.mode = mode;
    fr.skip_dead = skip_dead;
    fr.out = reinterpret_cast<float4*>(out_rgba);
    fr.hit_id = hit_id;
    fr.hit_t = hit_t;
    rt::Counters total;
    if (accel == 0) run<0>(sc, fr, n_rays, nthreads, &total);
    else if (accel == 1) run<1>(sc, fr, n_rays, nthreads, &total);
    else run<2>(sc, fr, n_rays, nthreads, &total);
    if (counters_out) memcpy(counters_out, &total, sizeof total);
    return RT_OK;
}

// The wavefront pipeline's state machine (rt_wave.cuh) driven generation by generation on the host: queues,
// path state and compaction as in rt_wavefront.cu, the trace step done per ray with the functions above.
int emul_render_wave(const rt_object* objs, uint32_t n_objs, const rt_light* lights, uint32_t n_lights, const rt_ray* rays,
                     uint64_t n_rays, uint32_t max_bounces, int mode, int accel, int leaf, float omax_user, float* out_rgba,
                     uint64_t* counters_out, uint32_t* generations_out, uint32_t cam_w, uint32_t cam_h, float cam_zdir, uint32_t grid_res,
                     uint32_t tile_w, uint32_t tile_h, float grid3_density, uint64_t* cells_out) {
    rt::HostScene hs;
    const int rc = rt::build_scene(objs, n_objs, lights, n_lights, omax_user, leaf, hs, g_err);
    if (rc != RT_OK) return rc;
    rt::Scene sc;
    bind(hs, sc);
    const size_t n = (size_t)n_rays;
    std::vector<float4> acc(n), refl(n), P(n), Nm(n), Rc(n), qs[2], qd[2];
    std::vector<uint32_t> outi(n), qp[2];
    rt::PathArrays pa{acc.data(), refl.data(), P.data(), Nm.data(), Rc.data(), outi.data()};
    for (int i = 0; i < 2; ++i) {
        qs[i].resize(n);
        qd[i].resize(n);
        qp[i].resize(n);
    }
    for (size_t i = 0; i < n; ++i) {
        qs[0][i] = reinterpret_cast<const float4*>(rays + i)[0];
        qd[0][i] = reinterpret_cast<const float4*>(rays + i)[1];
        qp[0][i] = (uint32_t)i;
        outi[i] = (uint32_t)i;
        reinterpret_cast<float4*>(out_rgba)[i] = make_float4(-1.f, -1.f, -1.f, -1.f);
    }
    // accel 3: perspective grids for the camera rays (the rays must then be the cam_w x cam_h pinhole frame) and for the
    // shadow rays of point lights, the BVH for everything else -- as wave_render does (rt_wavefront.cu)
    HostGrids cam, lit;
    HostGrid3 g3;
    unsigned long long n_cells = 0;
    if (accel == 3 && grid3_density > 0.f && n_objs) g3.build(hs, grid3_density);
    if (accel == 3) {
        if (cam_w && n_objs) camera_grid(cam, hs, cam_w, cam_h, cam_zdir, tile_w, tile_h);
        if (n_objs) light_grids(lit, hs, lights, n_lights, grid_res);
    }
    rt::WaveParams wp{max_bounces, mode, lit.set.valid_mask, (uint32_t)n};
    rt::LocalCounters lc;
    size_t cnt = n;
    uint32_t g = 0;
    while (cnt > 0) {
        const int a = g & 1, b = (g + 1) & 1;
        size_t nout = 0;
        for (size_t i = 0; i < cnt; ++i) {
            rt::Ray ray{qs[a][i], qd[a][i]};
            const uint32_t pb = qp[a][i];
            rt::Hit h;
            unsigned long long ne = 0, nn = 0;
            if (accel == 0) h = rt::trace_brute(sc, ray, &ne);
            else if (accel == 1) h = rt::trace_cull(sc, ray, &ne);
            else if (accel == 3 && g == 0 && (cam.set.valid_mask & 1u)) {
                unsigned long long np = 0;
                h = rt::trace_grid<false>(sc, cam.set, 0, ray, &np, &ne);
                lc.pair_tests += np;
            } else if (accel == 3 && (pb & RT_TASK_ANYHIT) && ((pb >> RT_TASK_LIGHT_SHIFT) & 15u) != RT_TASK_NO_GRID) {
                unsigned long long np = 0;
                h = rt::trace_grid<true>(sc, lit.set, 6u * ((pb >> RT_TASK_LIGHT_SHIFT) & 15u), ray, &np, &ne);
                lc.pair_tests += np;
            } else if (accel == 3 && g3.g.valid) {
                unsigned long long np = 0;
                if (pb & RT_TASK_ANYHIT) h = rt::trace_grid3<true>(sc, g3.g, ray, 1.f, &np, &ne, &n_cells);
                else h = rt::trace_grid3<false>(sc, g3.g, ray, RT_MAX_FLOAT, &np, &ne, &n_cells);
                lc.pair_tests += np;
            } else if (pb & RT_TASK_ANYHIT) h = rt::trace_bvh<true>(sc, ray, 1.f, &ne, &nn);
            else h = rt::trace_bvh<false>(sc, ray, RT_MAX_FLOAT, &ne, &nn);
            lc.exact_tests += ne;
            lc.node_visits += nn;
            rt::StepOut o;
            const uint32_t p = pb & RT_TASK_PATH_MASK;
            rt::wave::step<true>(sc, wp, pa, p, g == 0, ray, h, o, lc);
            if (o.emit) {
                qs[b][nout] = o.ray.s;
                qd[b][nout] = o.ray.d;
                const bool to_a = o.anyhit && o.light < RT_TASK_NO_GRID && ((wp.grid_lights >> o.light) & 1u);
                qp[b][nout] = p | ((to_a ? o.light : RT_TASK_NO_GRID) << RT_TASK_LIGHT_SHIFT) | (o.anyhit ? RT_TASK_ANYHIT : 0u);
                ++nout;
            }
            if (o.write) reinterpret_cast<float4*>(out_rgba)[outi[p]] = o.px;
        }
        cnt = nout;
        ++g;
    }
    if (generations_out) *generations_out = g;
    if (cells_out) *cells_out = n_cells;
    if (g > rt_wave_max_generations_host(max_bounces, n_lights, mode)) {
        g_err = "more generations than wave_max_generations allows";
        return RT_ERR_INVALID;
    }
    if (counters_out) {
        rt::Counters total;
        memset(&total, 0, sizeof total);
        total.ref_primary = lc.ref_primary; total.ref_shadow = lc.ref_shadow; total.ref_reflect = lc.ref_reflect;
        total.traced_shadow = lc.traced_shadow; total.traced_reflect = lc.traced_reflect; total.pair_tests = lc.pair_tests;
        total.exact_tests = lc.exact_tests; total.node_visits = lc.node_visits; total.hit_pixels = lc.hit_pixels;
        memcpy(counters_out, &total, sizeof total);
    }
    return RT_OK;
}

// BVH facts for the host-logic tests
int emul_bvh_info(const rt_object* objs, uint32_t n_objs, int leaf, uint32_t* n_nodes, uint32_t* depth, float* omax) {
    rt::HostScene hs;
    const int rc = rt::build_scene(objs, n_objs, nullptr, 0, 0.f, leaf, hs, g_err);
    if (rc != RT_OK) return rc;
    *n_nodes = (uint32_t)hs.nodes.size();
    *depth = hs.bvh_depth;
    *omax = hs.omax;
    return RT_OK;
}
}
