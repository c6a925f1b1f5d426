// rt_pgrid.cuh -- perspective grids: the nearest-hit / any-hit search for rays that pass through a common point.
//
// The reference searches every object for every ray (shade_and_reflect_kernel.cl:75).  Two of its three ray classes
// are not arbitrary rays, though:
//   * primary rays all start at the camera, the view-space origin (OpenCL-Raytracer.cpp:18-26,68-72);
//   * the shadow ray of a point light runs from the shaded point towards the light position, dir = pos - P
//     (SAR:196-209): its line passes through the light.
// For rays through a common apex A the search is two-dimensional: which objects a ray can touch depends only on its
// direction as seen from A.  A perspective grid cuts the directions of one half space (major axis +-x / +-y / +-z,
// coordinates u = w_b / |w_a|, v = w_c / |w_a| of the direction w) into nu x nv cells and stores per cell the objects
// whose conservative bounding sphere (inflated for the reference's arithmetic noise exactly like the bounding-sphere
// filter cv::sphere_may_hit, rt_scene.cpp) reaches into the cell's pyramid, sorted by their smallest possible
// distance from the apex.  A ray looks up ONE cell and walks its list: filter, then the exact reference arithmetic
// (ex::) on the survivors -- about half a dozen filter tests instead of ~50 BVH node steps.  The camera has one grid
// whose cells are pixel tiles; a light has six (a cube map).  Rays without a common apex (reflection rays, directional
// lights, caller-supplied rays) keep using the BVH.  Same (t, id) as every other search: the exact test decides.
//
// Everything in this header is plain float arithmetic usable on the host (tests/emul, host builder) and the device.
#pragma once

#include "rt_device.cuh"

#if defined(__CUDACC__)
#define RT_HD __host__ __device__ __forceinline__
#else
#define RT_HD inline
#endif

namespace rt {

struct PGrid {  // 48 B
    float ax, ay, az;      // apex
    uint32_t axis;         // major axis 0..2, | 4 when the grid looks down the negative axis
    float u0, v0;          // lower edge of the grid in u and v
    float inv_du, inv_dv;  // cells per unit of u / v
    uint32_t nu, nv;       // cells
    uint32_t cell_base;    // index of cell (0, 0) in GridSet::cell_start
    uint32_t kind;         // 0: camera (ray origins == apex), 1: light (ray origins anywhere in the scene)
};

struct GridEntry {  // 8 B
    uint32_t obj;   // object index
    float dnear;    // conservative lower bound of the distance apex -> any point of the (inflated) object
};

struct GridSet {
    const PGrid* grids;          // camera set: [0]; light set: [6 * light + face], face = axis | (negative ? 4 : 0) -> 0..2, 4..6 packed below
    const uint32_t* cell_start;  // n_cells + 1 offsets into entries
    const GridEntry* entries;
    uint32_t n_grids;
    uint32_t valid_mask;         // camera set: bit 0; light set: bit li = light li has its six grids (li < RT_GRID_MAX_LIGHTS)
};

#define RT_GRID_MAX_LIGHTS 15u  // light codes 0..14 travel in the task word, 15 = "no grid"

// face index 0..5 of a direction: +x +y +z -x -y -z  (major axis = largest |component|, lowest axis wins ties)
RT_HD uint32_t pgrid_face(float wx, float wy, float wz) {
    const float ax = fabsf(wx), ay = fabsf(wy), az = fabsf(wz);
    uint32_t a = 0;
    float m = ax, v = wx;
    if (ay > m) { a = 1; m = ay; v = wy; }
    if (az > m) { a = 2; m = az; v = wz; }
    return a + (v < 0.f ? 3u : 0u);
}

RT_HD void pgrid_swizzle(uint32_t axis, float x, float y, float z, float& a, float& b, float& c) {
    const uint32_t k = axis & 3u;
    a = k == 0 ? x : k == 1 ? y : z;
    b = k == 0 ? y : k == 1 ? z : x;
    c = k == 0 ? z : k == 1 ? x : y;
}

// Cell (index into cell_start) of the direction w as seen from the apex; w must point into the grid's half space and,
// for the camera grid, inside the frame.  The object side (pgrid_rect) pads every rectangle by 2e-3 cells, far more
// than the rounding of these few operations (< 1e-4 cells for grids of up to 4096 cells per side).
RT_HD uint32_t pgrid_cell(const PGrid& g, float wx, float wy, float wz) {
    float wa, wb, wc;
    pgrid_swizzle(g.axis, wx, wy, wz, wa, wb, wc);
    const float inv = 1.f / fabsf(wa);
    const float fu = (wb * inv - g.u0) * g.inv_du, fv = (wc * inv - g.v0) * g.inv_dv;
    int cu = (int)floorf(fu), cv = (int)floorf(fv);
    cu = cu < 0 ? 0 : (cu >= (int)g.nu ? (int)g.nu - 1 : cu);
    cv = cv < 0 ? 0 : (cv >= (int)g.nv ? (int)g.nv - 1 : cv);
    return g.cell_base + (uint32_t)cv * g.nu + (uint32_t)cu;
}

// One coordinate of pgrid_rect: the cells [lo, hi] of a grid axis (lower edge e0, `inv_d` cells per unit, n cells) that
// directions (x, y), x = component along the major axis (positive = in front), can fall into if their line from the apex
// comes within r of the point (x, y) -- the 3-D problem projected along the other minor axis, which can only widen it.
// Works in angles so that objects beside or partly behind the apex need no special cases: phi = atan2(y, x) +- asin(r / rho),
// intersected with the grid's angular range (always inside (-90, 90) degrees), mapped back with tan.
RT_HD bool pgrid_range(float x, float y, float r, float e0, float inv_d, uint32_t n, int& lo, int& hi) {
    const float rho2 = x * x + y * y;
    lo = 0;
    hi = (int)n - 1;
    if (!(rho2 > r * r * 1.001f + 1e-30f)) return true;  // the apex is inside (or at) the circle: every direction
    const float phi = atan2f(y, x);
    const float alpha = asinf(fminf(1.f, r / sqrtf(rho2) * 1.0001f)) + 2e-5f;
    const float g_lo = atanf(e0) - 2e-5f, g_hi = atanf(e0 + (float)n / inv_d) + 2e-5f;
    const float a_lo = fmaxf(phi - alpha, g_lo), a_hi = fminf(phi + alpha, g_hi);
    if (a_lo > a_hi) return false;
    const float f_lo = (tanf(a_lo) - e0) * inv_d - 2e-3f, f_hi = (tanf(a_hi) - e0) * inv_d + 2e-3f;
    if (f_lo > 0.f) lo = f_lo < (float)n ? (int)floorf(f_lo) : (int)n;
    if (f_hi < (float)n - 1.f) hi = f_hi >= 0.f ? (int)floorf(f_hi) : -1;
    return lo <= hi;
}

// Cells of grid g that a ray through (or, for shadow rays, within rounding of) its apex needs to know about an object with
// conservative bounding sphere (c, r): false = none.
RT_HD bool pgrid_rect(const PGrid& g, float cx, float cy, float cz, float r, int& u_lo, int& u_hi, int& v_lo, int& v_hi) {
    float x, y, z;
    pgrid_swizzle(g.axis, cx - g.ax, cy - g.ay, cz - g.az, x, y, z);
    if (g.axis & 4u) x = -x;
    if (!(x + r > 0.f)) return false;  // entirely behind the apex plane
    return pgrid_range(x, y, r, g.u0, g.inv_du, g.nu, u_lo, u_hi) && pgrid_range(x, z, r, g.v0, g.inv_dv, g.nv, v_lo, v_hi);
}

// The bounding sphere an object presents to the rays of a grid, from its filter record (rt_scene.cpp: a ray with origin o
// can only hit if its line passes within sqrt(r2 + g |oc|^2) of the centre, g = 1 - h), and the nearest distance from the
// apex at which such a ray can meet it.  Camera grid: o == apex, |oc| is known.  Light grids: o is a point of the scene,
// |oc| <= omax + |c| (omax bounds the norm of every secondary-ray origin).  The pad covers the rounding of c - apex and,
// for shadow rays, the distance by which the rounded ray start + direction misses the light position (a few ulp of the
// coordinates), 30x over.
RT_HD bool pgrid_object_sphere(const PGrid& g, const CullRec& cr, float omax, float& r, float& dnear) {
    if (cr.r2 < 0.f) return false;  // never hit (unknown type)
    const float vx = cr.cx - g.ax, vy = cr.cy - g.ay, vz = cr.cz - g.az;
    const float dist = sqrtf(vx * vx + vy * vy + vz * vz);
    const float cn = sqrtf(cr.cx * cr.cx + cr.cy * cr.cy + cr.cz * cr.cz);
    const float an = sqrtf(g.ax * g.ax + g.ay * g.ay + g.az * g.az);
    const float gq = fmaxf(1.f - cr.h, 0.f) * 1.0001f + 1e-9f;
    const float D = g.kind == 0u ? dist : omax + cn;
    r = sqrtf(cr.r2 + gq * D * D) * 1.0001f + 1e-5f * (cn + an + omax + 1.f);
    dnear = fmaxf(dist - r, 0.f) * 0.9999f;
    return true;
}

// ------------------------------------------------------------------ per-ray search (megakernel-style; the wavefront kernel
// in rt_wavefront.cu runs the same steps warp-scheduled)
// ANYHIT: shadow ray towards the light that owns the grids at gs.grids[first_grid .. +6): only "is there an exact t < 1"
// matters (SAR:229).  Otherwise: primary ray from the camera apex, nearest hit.
template <bool ANYHIT>
__device__ __forceinline__ Hit trace_grid(const Scene& sc, const GridSet& gs, uint32_t first_grid, const Ray& ray,
                                          unsigned long long* n_pairs, unsigned long long* n_exact) {
    Hit h;
    h.t = RT_MAX_FLOAT;
    h.id = -1;
    uint32_t ktype = 0;
    const float wx = ANYHIT ? -ray.d.x : ray.d.x, wy = ANYHIT ? -ray.d.y : ray.d.y, wz = ANYHIT ? -ray.d.z : ray.d.z;
    const PGrid g = gs.grids[first_grid + (ANYHIT ? pgrid_face(wx, wy, wz) : 0u)];
    const uint32_t cell = pgrid_cell(g, wx, wy, wz);
    uint32_t cur = gs.cell_start[cell];
    const uint32_t end = gs.cell_start[cell + 1];
    const cv::RayK rk = cv::prepare(ray);
    float cutoff;
    if (ANYHIT) {
        const float sx = ray.s.x - g.ax, sy = ray.s.y - g.ay, sz = ray.s.z - g.az;
        cutoff = sqrtf(sx * sx + sy * sy + sz * sz) * 1.0001f + 1e-4f;  // objects farther from the light than the ray start are behind it
    } else {
        cutoff = RT_MAX_FLOAT;
    }
    unsigned long long np = 0, ne = 0;
    const float4* cr = reinterpret_cast<const float4*>(sc.cull);
    for (; cur < end; ++cur) {
        const GridEntry e = gs.entries[cur];
        if (e.dnear > cutoff) break;
        const float4 a = __ldg(cr + 2 * (size_t)e.obj), b = __ldg(cr + 2 * (size_t)e.obj + 1);
        ++np;
        if (cv::sphere_may_hit(rk, a.x, a.y, a.z, a.w, b.x)) {
            exact_pair(sc, ray, e.obj, __float_as_uint(b.w), h, ktype);
            ++ne;
            if (ANYHIT) {
                if (h.t < 1.f) break;
            } else if (h.id >= 0) {
                cutoff = h.t * rk.dlen * 1.0001f + 1e-4f;  // hit point distance from the apex: origin == apex
            }
        }
    }
    if (n_pairs) *n_pairs += np;
    if (n_exact) *n_exact += ne;
    return h;
}

}  // namespace rt
