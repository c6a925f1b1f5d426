// rt_grid3.cuh -- uniform 3-D grid with a DDA walk: the nearest-hit search for the rays that have NO common apex
// (reflection rays, shadow rays of directional lights, caller-supplied rays) when the objects are spread evenly enough.
//
// The reference tests every object for every ray (shade_and_reflect_kernel.cl:75).  The BVH (rt_scene.cpp) cuts that to
// ~50 node steps per ray, each two inflated slab tests and ~105 warp instructions of bookkeeping.  For a scene whose
// objects are small against its extent and evenly spread -- the headline workload places 104 096 of them uniformly in a
// slab -- a uniform grid with about one object per cell does better: a DDA step is a dozen instructions, a ray leaves
// after a mean free path of ~a dozen cells, and each cell lists two or three objects that go through the same
// bounding-sphere filter + exact reference arithmetic as everywhere else.  Objects are entered into every cell their
// conservative bounding sphere touches (same inflation for the reference's arithmetic noise as the light grids, rt_pgrid.cuh,
// plus `eps` for the rounding of the walk), so a hit at parameter t is always found in the cell that contains the hit point,
// and the walk may stop once the best t lies before the exit of the current cell.  Built on the device (rt_pgrid.cu); the
// builder rejects a grid whose lists come out long (clustered scenes) -- those scenes keep the BVH.
#pragma once

#include "rt_pgrid.cuh"

namespace rt {

// One list entry: the object's filter record inline (a walk step is then ONE dependent load, and the entries of a cell are
// consecutive, so the next one can be fetched while this one is tested) -- 32 bytes, two 128-bit loads.
struct alignas(16) Grid3Entry {
    float cx, cy, cz, r2;  // CullRec: centre, padded squared radius
    float h;               // CullRec: 1 - g
    uint32_t obj, type;
    uint32_t pad;
};

struct Grid3 {
    float lox, loy, loz;  // lower corner
    float h, inv_h;       // cell edge and its reciprocal
    uint32_t nx, ny, nz;
    const uint32_t* cell_start;  // nx*ny*nz + 1
    const Grid3Entry* entries;
    uint32_t valid;
    float eps;  // spatial slack every object was entered with (rounding of the walk)
};

// Conservative bounding sphere of an object for rays that start anywhere in the scene (|origin| <= omax): see
// pgrid_object_sphere, light kind.
RT_HD bool grid3_object_sphere(const CullRec& cr, float omax, float eps, float& r) {
    if (cr.r2 < 0.f) return false;
    const float cn = sqrtf(cr.cx * cr.cx + cr.cy * cr.cy + cr.cz * cr.cz);
    const float gq = fmaxf(1.f - cr.h, 0.f) * 1.0001f + 1e-9f;
    const float D = omax + cn;
    r = sqrtf(cr.r2 + gq * D * D) * 1.0001f + 1e-5f * (cn + omax + 1.f) + eps;
    return true;
}

RT_HD int grid3_clampi(int v, int n) { return v < 0 ? 0 : (v >= n ? n - 1 : v); }

RT_HD void grid3_object_cells(const Grid3& g, const CullRec& cr, float r, int* lo, int* hi) {
    lo[0] = grid3_clampi((int)floorf((cr.cx - r - g.lox) * g.inv_h), (int)g.nx);
    hi[0] = grid3_clampi((int)floorf((cr.cx + r - g.lox) * g.inv_h), (int)g.nx);
    lo[1] = grid3_clampi((int)floorf((cr.cy - r - g.loy) * g.inv_h), (int)g.ny);
    hi[1] = grid3_clampi((int)floorf((cr.cy + r - g.loy) * g.inv_h), (int)g.ny);
    lo[2] = grid3_clampi((int)floorf((cr.cz - r - g.loz) * g.inv_h), (int)g.nz);
    hi[2] = grid3_clampi((int)floorf((cr.cz + r - g.loz) * g.inv_h), (int)g.nz);
}

// State of one ray's walk.  t is the ray's own parameter (direction not normalised, as in the reference).
struct Dda {
    float ix, iy, iz;           // 1 / d (0 when the ray does not move along the axis)
    float bx, by, bz;           // (lo - s) * inv per axis: tmax_k = (cell_k + pos_k) * h * inv_k + b_k
    float hx, hy, hz;           // h * inv
    int cx, cy, cz;             // current cell
    int sx, sy, sz;             // step (+1 / -1 / 0)
    float t_end;                // parameter at which the ray leaves the grid (or its search limit)
};

// Returns false when the ray misses the grid (or its part before `tcap`) altogether.
RT_HD bool dda_start(const Grid3& g, const Ray& r, float tcap, Dda& w) {
    const float big = RT_MAX_FLOAT;
    float t0 = 0.f, t1 = tcap;
    const float s[3] = {r.s.x, r.s.y, r.s.z}, d[3] = {r.d.x, r.d.y, r.d.z};
    const float lo[3] = {g.lox, g.loy, g.loz};
    const float n[3] = {(float)g.nx, (float)g.ny, (float)g.nz};
    float inv[3];
    int st[3];
    for (int k = 0; k < 3; ++k) {
        const float hi = lo[k] + n[k] * g.h;
        if (fabsf(d[k]) < 1e-30f) {
            inv[k] = 0.f;
            st[k] = 0;
            if (s[k] < lo[k] || s[k] > hi) return false;  // parallel to the slab and outside it (objects keep eps away from the faces)
        } else {
            inv[k] = 1.f / d[k];
            st[k] = d[k] > 0.f ? 1 : -1;
            const float ta = (lo[k] - s[k]) * inv[k], tb = (hi - s[k]) * inv[k];
            t0 = fmaxf(t0, fminf(ta, tb));
            t1 = fminf(t1, fmaxf(ta, tb));
        }
    }
    if (!(t0 <= t1)) return false;  // also catches NaN
    w.ix = inv[0]; w.iy = inv[1]; w.iz = inv[2];
    w.sx = st[0]; w.sy = st[1]; w.sz = st[2];
    w.bx = (lo[0] - s[0]) * inv[0]; w.by = (lo[1] - s[1]) * inv[1]; w.bz = (lo[2] - s[2]) * inv[2];
    w.hx = g.h * inv[0]; w.hy = g.h * inv[1]; w.hz = g.h * inv[2];
    w.cx = grid3_clampi((int)floorf((s[0] + t0 * d[0] - lo[0]) * g.inv_h), (int)g.nx);
    w.cy = grid3_clampi((int)floorf((s[1] + t0 * d[1] - lo[1]) * g.inv_h), (int)g.ny);
    w.cz = grid3_clampi((int)floorf((s[2] + t0 * d[2] - lo[2]) * g.inv_h), (int)g.nz);
    w.t_end = t1;
    (void)big;
    return true;
}

// Parameter at which the ray leaves its current cell, and the axis it leaves through (0..2).
RT_HD float dda_exit(const Dda& w, int& axis) {
    const float big = RT_MAX_FLOAT;
    const float tx = w.sx ? fmaf((float)(w.cx + (w.sx > 0 ? 1 : 0)), w.hx, w.bx) : big;
    const float ty = w.sy ? fmaf((float)(w.cy + (w.sy > 0 ? 1 : 0)), w.hy, w.by) : big;
    const float tz = w.sz ? fmaf((float)(w.cz + (w.sz > 0 ? 1 : 0)), w.hz, w.bz) : big;
    axis = tx <= ty ? (tx <= tz ? 0 : 2) : (ty <= tz ? 1 : 2);
    return fminf(tx, fminf(ty, tz));
}

// Steps into the next cell; false when that leaves the grid.
RT_HD bool dda_step(const Grid3& g, Dda& w, int axis) {
    if (axis == 0) {
        w.cx += w.sx;
        return (unsigned)w.cx < g.nx;
    }
    if (axis == 1) {
        w.cy += w.sy;
        return (unsigned)w.cy < g.ny;
    }
    w.cz += w.sz;
    return (unsigned)w.cz < g.nz;
}

RT_HD uint32_t dda_cell(const Grid3& g, const Dda& w) { return ((uint32_t)w.cz * g.ny + (uint32_t)w.cy) * g.nx + (uint32_t)w.cx; }

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
    uint32_t box[4] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu};  // the last objects that ran the exact test
    int nb = 0;
    unsigned long long np = 0, ne = 0, nc = 0;
    for (;;) {
        const uint32_t cell = dda_cell(g, w);
        ++nc;
        for (uint32_t e = g.cell_start[cell], end = g.cell_start[cell + 1]; e < end; ++e) {
            const Grid3Entry en = g.entries[e];
            const uint32_t obj = en.obj;
            ++np;
            if (!cv::sphere_may_hit(rk, en.cx, en.cy, en.cz, en.r2, en.h)) continue;
            if (obj == box[0] || obj == box[1] || obj == box[2] || obj == box[3]) continue;
            box[nb] = obj;
            nb = (nb + 1) & 3;
            exact_pair(sc, ray, obj, en.type, h, ktype);
            ++ne;
            if (ANYHIT && h.t < tcap) goto done;
        }
        int axis;
        const float t_exit = dda_exit(w, axis);
        // every object whose hit point lies before the exit of this cell has been seen (it is listed in the cell that holds
        // the point); 1e-5 relative covers the rounding of t_exit, the objects' eps the sliver that rounding can skip
        if (h.id >= 0 && h.t < t_exit * (1.f - 1e-5f)) break;
        if (t_exit > w.t_end) break;
        if (!dda_step(g, w, axis)) break;
    }
done:
    if (n_pairs) *n_pairs += np;
    if (n_exact) *n_exact += ne;
    if (n_cells) *n_cells += nc;
    return h;
}

}  // namespace rt
