// rt_grid3.cuh -- uniform 3-D grid with a DDA walk: the nearest-hit search for the rays that have NO common apex
// (reflection rays, shadow rays of directional lights, caller-supplied rays) when the objects are spread evenly enough.
//
// The reference tests every object for every ray (shade_and_reflect_kernel.cl:75).  The BVH (rt_scene.cpp) cuts that to
// ~50 node steps per ray, each two inflated slab tests and ~105 warp instructions of bookkeeping.  For a scene whose
// objects are small against its extent and evenly spread -- the headline workload places 104 096 of them uniformly in a
// slab -- a uniform grid with a few cells per object can do better: a DDA step is a dozen instructions, a ray leaves
// after a mean free path of ~a dozen cells, and each cell lists two or three objects that go through the same
// bounding-sphere filter + exact reference arithmetic as everywhere else.  An object is entered into every cell its
// conservative box touches -- the padded half extents of rt_scene.cpp plus the noise inflation f * (omax + |c|) a ray from
// anywhere in the scene can need, plus `eps` for the rounding of the walk -- so a hit at parameter t is always found in the
// cell that contains the hit point, and the walk may stop once the best t lies before the exit of the current cell.
// Built on the device (rt_pgrid.cu); the builder rejects a grid whose lists come out long (clustered scenes).
#pragma once

#include "rt_pgrid.cuh"

namespace rt {

struct Grid3 {
    float lox, loy, loz;  // lower corner
    float h, inv_h;       // cell edge and its reciprocal
    uint32_t nx, ny, nz;
    const uint32_t* cell_start;  // nx*ny*nz + 1
    const uint32_t* entries;     // object indices
    uint32_t valid;
    float eps;  // spatial slack every object was entered with (rounding of the walk)
};

RT_HD int grid3_clampi(int v, int n) { return v < 0 ? 0 : (v >= n ? n - 1 : v); }

// Cells touched by the box (c +- (half + eps)); false for an object that is never hit (half.w < 0).
RT_HD bool grid3_object_cells(const Grid3& g, const CullRec& cr, float hx, float hy, float hz, float hw, int* lo, int* hi) {
    if (hw < 0.f || cr.r2 < 0.f) return false;
    const float e = g.eps;
    lo[0] = grid3_clampi((int)floorf((cr.cx - hx - e - g.lox) * g.inv_h), (int)g.nx);
    hi[0] = grid3_clampi((int)floorf((cr.cx + hx + e - g.lox) * g.inv_h), (int)g.nx);
    lo[1] = grid3_clampi((int)floorf((cr.cy - hy - e - g.loy) * g.inv_h), (int)g.ny);
    hi[1] = grid3_clampi((int)floorf((cr.cy + hy + e - g.loy) * g.inv_h), (int)g.ny);
    lo[2] = grid3_clampi((int)floorf((cr.cz - hz - e - g.loz) * g.inv_h), (int)g.nz);
    hi[2] = grid3_clampi((int)floorf((cr.cz + hz + e - g.loz) * g.inv_h), (int)g.nz);
    return true;
}

// State of one ray's walk.  t is the ray's own parameter (direction not normalised, as in the reference).  Incremental:
// tm* = parameter of the next cell boundary per axis, dt* = parameter per cell, r* = cells left before the walk leaves the
// grid along the axis, l* = what a step adds to the linear cell index.  The accumulated rounding of tm += dt (< 100 steps,
// 6e-6 relative) is a few 1e-4 of a unit in space: inside eps.
struct Dda {
    float tmx, tmy, tmz;
    float dtx, dty, dtz;
    int rx, ry, rz;
    int lx, ly, lz;
    uint32_t cell;
    float t_end;  // parameter at which the ray leaves the grid (or its search limit)
};

// Returns false when the ray misses the grid (or its part before `tcap`) altogether.
RT_HD bool dda_start(const Grid3& g, const Ray& r, float tcap, Dda& w) {
    const float big = RT_MAX_FLOAT;
    float t0 = 0.f, t1 = tcap;
    const float s[3] = {r.s.x, r.s.y, r.s.z}, d[3] = {r.d.x, r.d.y, r.d.z};
    const float lo[3] = {g.lox, g.loy, g.loz};
    const int n[3] = {(int)g.nx, (int)g.ny, (int)g.nz};
    float inv[3];
    for (int k = 0; k < 3; ++k) {
        const float hi = lo[k] + (float)n[k] * g.h;
        if (fabsf(d[k]) < 1e-30f) {
            inv[k] = 0.f;
            if (s[k] < lo[k] || s[k] > hi) return false;  // parallel to the slab and outside it (objects keep eps away from the faces)
        } else {
            inv[k] = 1.f / d[k];
            const float ta = (lo[k] - s[k]) * inv[k], tb = (hi - s[k]) * inv[k];
            t0 = fmaxf(t0, fminf(ta, tb));
            t1 = fminf(t1, fmaxf(ta, tb));
        }
    }
    if (!(t0 <= t1)) return false;
    int c[3];
    float tm[3], dt[3];
    int rem[3], lin[3];
    const int stride[3] = {1, n[0], n[0] * n[1]};
    for (int k = 0; k < 3; ++k) {
        c[k] = grid3_clampi((int)floorf((s[k] + t0 * d[k] - lo[k]) * g.inv_h), n[k]);
        if (inv[k] == 0.f) {
            tm[k] = big;
            dt[k] = 0.f;
            rem[k] = 0x7fffffff;
            lin[k] = 0;
        } else if (inv[k] > 0.f) {
            tm[k] = (lo[k] + (float)(c[k] + 1) * g.h - s[k]) * inv[k];
            dt[k] = g.h * inv[k];
            rem[k] = n[k] - 1 - c[k];
            lin[k] = stride[k];
        } else {
            tm[k] = (lo[k] + (float)c[k] * g.h - s[k]) * inv[k];
            dt[k] = -g.h * inv[k];
            rem[k] = c[k];
            lin[k] = -stride[k];
        }
    }
    w.tmx = tm[0]; w.tmy = tm[1]; w.tmz = tm[2];
    w.dtx = dt[0]; w.dty = dt[1]; w.dtz = dt[2];
    w.rx = rem[0]; w.ry = rem[1]; w.rz = rem[2];
    w.lx = lin[0]; w.ly = lin[1]; w.lz = lin[2];
    w.cell = ((uint32_t)c[2] * g.ny + (uint32_t)c[1]) * g.nx + (uint32_t)c[0];
    w.t_end = t1;
    return true;
}

// One step: returns the parameter at which the ray leaves the cell it is in and moves into the next one; `out` is set when
// that next cell lies outside the grid.
RT_HD float dda_advance(Dda& w, bool& out) {
    float t_exit;
    if (w.tmx <= w.tmy && w.tmx <= w.tmz) {
        t_exit = w.tmx;
        w.tmx += w.dtx;
        w.cell += (uint32_t)w.lx;
        out = --w.rx < 0;
    } else if (w.tmy <= w.tmz) {
        t_exit = w.tmy;
        w.tmy += w.dty;
        w.cell += (uint32_t)w.ly;
        out = --w.ry < 0;
    } else {
        t_exit = w.tmz;
        w.tmz += w.dtz;
        w.cell += (uint32_t)w.lz;
        out = --w.rz < 0;
    }
    return t_exit;
}

// Per-ray search (host emulation; the wavefront kernel wave_trace_dda runs the same steps warp-scheduled).
// ANYHIT: only "is there an exact t < tcap" matters.
template <bool ANYHIT>
__device__ __forceinline__ Hit trace_grid3(const Scene& sc, const Grid3& g, const Ray& ray, float tcap, unsigned long long* n_pairs,
                                           unsigned long long* n_exact, unsigned long long* n_cells) {
    Hit h;
    h.t = RT_MAX_FLOAT;
    h.id = -1;
    uint32_t ktype = 0;
    {  // a ray that is not finite (a hit point at infinity) reaches the exact test on every object, as in the reference
        const float chk = ray.s.x + ray.s.y + ray.s.z + ray.d.x + ray.d.y + ray.d.z;
        if (!(fabsf(chk) < __int_as_float(0x7f800000))) return trace_brute(sc, ray, n_exact);
    }
    Dda w;
    if (!dda_start(g, ray, ANYHIT ? tcap : RT_MAX_FLOAT, w)) return h;
    const cv::RayK rk = cv::prepare(ray);
    const float4* cr = reinterpret_cast<const float4*>(sc.cull);
    uint32_t box[4] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu};  // the last objects that ran the exact test
    int nb = 0;
    unsigned long long np = 0, ne = 0, nc = 0;
    for (;;) {
        ++nc;
        for (uint32_t e = g.cell_start[w.cell], end = g.cell_start[w.cell + 1]; e < end; ++e) {
            const uint32_t obj = g.entries[e];
            const float4 a = __ldg(cr + 2 * (size_t)obj), b = __ldg(cr + 2 * (size_t)obj + 1);
            ++np;
            if (!cv::sphere_may_hit(rk, a.x, a.y, a.z, a.w, b.x)) continue;
            if (obj == box[0] || obj == box[1] || obj == box[2] || obj == box[3]) continue;
            box[nb] = obj;
            nb = (nb + 1) & 3;
            exact_pair(sc, ray, obj, __float_as_uint(b.w), h, ktype);
            ++ne;
            if (ANYHIT && h.t < tcap) goto done;
        }
        bool out;
        const float t_exit = dda_advance(w, out);
        // every object whose hit point lies before the exit of the cell just left has been seen (it is listed in the cell that
        // holds the point); 1e-5 relative covers the rounding of t_exit, the objects' eps the sliver that rounding can skip
        if (h.id >= 0 && h.t < t_exit * (1.f - 1e-5f)) break;
        if (t_exit > w.t_end || out) break;
    }
done:
    if (n_pairs) *n_pairs += np;
    if (n_exact) *n_exact += ne;
    if (n_cells) *n_cells += nc;
    return h;
}

}  // namespace rt
