// rt_device.cuh -- device-side building blocks of the B200 render path.
//
// Two kinds of arithmetic live here and are kept strictly apart:
//
//  * ex::  "exact" functions.  They decide everything that is observable in the
//    output (hit / miss, t, the winning object, hit point, normal, reflection
//    vector, Phong terms).  They follow the reference kernels operation by
//    operation -- IEEE binary32, round to nearest, NO contraction, sums in the
//    order the kernels write them (SAR = shade_and_reflect_kernel.cl) -- using
//    the __f*_rn intrinsics, which nvcc never fuses into FMAs.  Hit ids and t are
//    therefore bit-identical to the reference arithmetic on every ray.
//
//  * cv::  "conservative" functions.  They only ever answer "this ray certainly
//    does not produce a hit on this object / inside this node" and are free to use
//    FMAs and approximate arithmetic, because every bound they test against has been
//    inflated on the host (rt_scene.cpp) by more than the rounding error of both the
//    filter and the reference's own arithmetic.  A pair that survives a cv:: filter
//    is handed to ex::; a pair that is rejected would have been a miss in ex:: too.
//
// Nothing here touches the CPU oracle; the product path has no CPU fallback.
#pragma once

#ifndef RT_HOST_EMULATION
#include <cuda_runtime.h>
#endif
#include <stdint.h>

#include "rt_layout.h"

#define RT_MAX_FLOAT 3.402823466e+38F /* SAR:31 */
#define RT_BVH_STACK 64

namespace rt {

// ---------------------------------------------------------------- scene in HBM
// Structure of arrays, one array per access pattern (DESIGN.md "Data layout"):
//   cull   : 32 B / object, read by every ray x object pair of the brute-force filter
//   minv   : 64 B / object, read only by pairs that survive the filter (exact test)
//   mv,mat : 64 + 48 B / object, read once per accepted hit (surface + shading)
struct Light {  // 64 B
    float4 ambient, diffuse, specular, position;
};

struct Scene {
    const float4* __restrict__ minv;   // M x 4 columns of mvInverse
    const float4* __restrict__ mv;     // M x 4 columns of mv
    const float4* __restrict__ mat;    // M x 3: (ambient, absorption) (diffuse, shininess) (specular, 0)
    const CullRec* __restrict__ cull;  // M
    const BvhNode* __restrict__ nodes;
    const uint32_t* __restrict__ leaf_obj;  // object ids in leaf order
    const CullRec* __restrict__ leaf_cull;  // cull records in leaf order
    const Light* __restrict__ lights;
    uint32_t n_objs, n_lights, n_nodes, root_is_leaf;
    int32_t root;  // child-style reference to the root (leaf if negative)
    float filter_h;        // smallest h = 1 - g of the scene (brute-force filter folds it into per-ray constants)
    float root_a, root_f;  // root_f: the scene-wide largest noise coefficient (informational)
};

struct Ray {
    float4 s, d;  // start (w carried, ~1), direction (w carried, 0)
};

struct Hit {
    float t;
    int32_t id;  // -1: none
};

// ------------------------------------------------------------------ exact math
namespace ex {

__device__ __forceinline__ float mul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float add(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float sub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float div(float a, float b) { return __fdiv_rn(a, b); }
__device__ __forceinline__ float sqrt_(float a) { return __fsqrt_rn(a); }

struct V3 {
    float x, y, z;
};
__device__ __forceinline__ V3 mk3(float x, float y, float z) {
    V3 r = {x, y, z};
    return r;
}
__device__ __forceinline__ V3 xyz(float4 v) { return mk3(v.x, v.y, v.z); }
__device__ __forceinline__ V3 vadd(V3 a, V3 b) { return mk3(add(a.x, b.x), add(a.y, b.y), add(a.z, b.z)); }
__device__ __forceinline__ V3 vsub(V3 a, V3 b) { return mk3(sub(a.x, b.x), sub(a.y, b.y), sub(a.z, b.z)); }
__device__ __forceinline__ V3 vneg(V3 a) { return mk3(-a.x, -a.y, -a.z); }
__device__ __forceinline__ V3 vmul(V3 a, V3 b) { return mk3(mul(a.x, b.x), mul(a.y, b.y), mul(a.z, b.z)); }
__device__ __forceinline__ V3 vscale(V3 a, float s) { return mk3(mul(a.x, s), mul(a.y, s), mul(a.z, s)); }
__device__ __forceinline__ V3 svec(float s, V3 a) { return mk3(mul(s, a.x), mul(s, a.y), mul(s, a.z)); }
// dot(): a.x*b.x + a.y*b.y + a.z*b.z summed left to right (cl_shim.hpp contract)
__device__ __forceinline__ float dot(V3 a, V3 b) { return add(add(mul(a.x, b.x), mul(a.y, b.y)), mul(a.z, b.z)); }
// normalize(v) = v / sqrt(dot(v,v)), correctly rounded sqrt and divisions
__device__ __forceinline__ V3 normalize(V3 v) {
    const float len = sqrt_(dot(v, v));
    return mk3(div(v.x, len), div(v.y, len), div(v.z, len));
}
// SAR:68-70 reflect: incident - 2.f * dot(incident, normal) * normal
__device__ __forceinline__ V3 reflect(V3 incident, V3 normal) {
    const float k = mul(2.f, dot(incident, normal));
    return vsub(incident, svec(k, normal));
}

// SAR:60-65 transform: column-major 4x4 times float4, each row summed left to right.
__device__ __forceinline__ float4 xform(const float4 c0, const float4 c1, const float4 c2, const float4 c3, const float4 v) {
    float4 o;
    o.x = add(add(add(mul(c0.x, v.x), mul(c1.x, v.y)), mul(c2.x, v.z)), mul(c3.x, v.w));
    o.y = add(add(add(mul(c0.y, v.x), mul(c1.y, v.y)), mul(c2.y, v.z)), mul(c3.y, v.w));
    o.z = add(add(add(mul(c0.z, v.x), mul(c1.z, v.y)), mul(c2.z, v.z)), mul(c3.z, v.w));
    o.w = add(add(add(mul(c0.w, v.x), mul(c1.w, v.y)), mul(c2.w, v.z)), mul(c3.w, v.w));
    return o;
}
__device__ __forceinline__ float4 xform(const float4* __restrict__ m, const float4 v) {
    return xform(__ldg(m + 0), __ldg(m + 1), __ldg(m + 2), __ldg(m + 3), v);
}

// SAR:33-58 intersectsWidthBoxSide: one slab of the unit cube.
__device__ __forceinline__ bool box_slab(float& tMin, float& tMax, float start, float dir) {
    float t1 = sub(-0.5f, start);
    float t2 = sub(0.5f, start);
    if (dir == 0.f) {
        if (copysignf(t1, t2) == t1) return false;
        tMin = -RT_MAX_FLOAT;
        tMax = RT_MAX_FLOAT;
        return true;
    }
    t1 = div(t1, dir);
    t2 = div(t2, dir);
    if (dir < 0.f) {
        tMin = fminf(t1, t2);
        tMax = fmaxf(t1, t2);
    } else {
        tMin = t1;
        tMax = t2;
    }
    return true;
}

// The per-object body of raycast(), SAR:75-168, up to (not including) the comparison
// against hit->time.  s, d: the ray already transformed by mvInverse (SAR:78-79).
// Returns true and t if the object produces a candidate.
__device__ __forceinline__ bool candidate(uint32_t type, const float4 s, const float4 d, float& t) {
    if (type == 0u) {  // sphere, SAR:82-106
        const float A = add(add(mul(d.x, d.x), mul(d.y, d.y)), mul(d.z, d.z));
        const float B = mul(2.f, add(add(mul(d.x, s.x), mul(d.y, s.y)), mul(d.z, s.z)));
        const float C = sub(add(add(mul(s.x, s.x), mul(s.y, s.y)), mul(s.z, s.z)), 1.f);
        const float radical = sub(mul(B, B), mul(mul(4.f, A), C));
        if (radical < 0.f) return false;
        const float root = sqrt_(radical);
        const float twoA = mul(2.f, A);
        const float t1 = div(sub(-B, root), twoA);
        const float t2 = div(add(-B, root), twoA);
        const float tMin = (t1 >= 0.f && t2 >= 0.f) ? fminf(t1, t2) : fmaxf(t1, t2);
        if (tMin < 0.f) return false;
        t = tMin;
        return true;
    }
    if (type == 1u) {  // box, SAR:123-144
        float txMin, txMax, tyMin, tyMax, tzMin, tzMax;
        if (!box_slab(txMin, txMax, s.x, d.x)) return false;
        if (!box_slab(tyMin, tyMax, s.y, d.y)) return false;
        if (!box_slab(tzMin, tzMax, s.z, d.z)) return false;
        const float tMin = fmaxf(fmaxf(txMin, tyMin), tzMin);
        const float tMax = fminf(fminf(txMax, tyMax), tzMax);
        if (tMax < tMin) return false;
        const float tHit = (tMin >= 0.f && tMax >= 0.f) ? fminf(tMin, tMax) : fmaxf(tMin, tMax);
        if (tHit < 0.f) return false;
        t = tHit;
        return true;
    }
    if (type == 2u) {  // EXTENSION: unit triangle x >= -0.5, y >= -0.5, x + y <= 0 in the plane z = 0 (oracle_rt.c, type 2)
        if (d.z == 0.f) return false;
        const float tHit = div(sub(0.f, s.z), d.z);
        if (tHit < 0.f) return false;
        const float px = add(s.x, mul(tHit, d.x));
        const float py = add(s.y, mul(tHit, d.y));
        if (px < -0.5f || py < -0.5f || add(px, py) > 0.f) return false;
        t = tHit;
        return true;
    }
    return false;  // any other type value: the kernels' switch has no default
}

// Does candidate (tc, c) replace the current best (tk, k)?  Order-independent form of
// the sequential rules "sphere skipped only if hit.time < t" (SAR:108) and "box skipped
// if hit.time <= t" (SAR:147; the triangle extension follows the box rule): among objects tied at the minimum t the winner is the
// highest-index sphere if any sphere is tied, else the lowest-index box.
// A candidate with t >= MAX_FLOAT (an overflowed t, +inf) never becomes a hit in the reference either: the sphere rule
// skips it (MAX < t), the box rule skips it (MAX <= t), and t == MAX_FLOAT reads as "miss" at SAR:173.  A NaN t (a ray whose
// object-space direction is zero or not finite) is OUTSIDE the bit-exact domain: the reference's sequential loop then
// depends on the object order in a way no order-independent rule reproduces (every later candidate overwrites a NaN);
// here such a candidate is dropped.
__device__ __forceinline__ bool replaces(float tc, int32_t c, uint32_t type_c, float tk, int32_t k, uint32_t type_k) {
    if (!(tc < RT_MAX_FLOAT)) return false;
    if (k < 0) return true;
    if (tc < tk) return true;
    if (tc > tk) return false;
    if (type_c == 0u) return (type_k != 0u) || (c > k);
    return (type_k != 0u) && (c < k);
}

struct Surface {
    float4 P;  // hit.intersection (w carried)      SAR:113,161
    V3 N;      // hit.normal                          SAR:118,165
};

// What raycast() stores when it accepts object `id` at time t (SAR:110-119 / 149-166),
// recomputed from (id, t): the same operations on the same inputs give the same bits.
__device__ __forceinline__ Surface surface(const Scene& sc, int32_t id, float t, const Ray& ray) {
    const float4* mi = sc.minv + 4 * (size_t)id;
    const float4 s = xform(mi, ray.s);
    const float4 d = xform(mi, ray.d);
    float4 po;
    po.x = add(s.x, mul(t, d.x));
    po.y = add(s.y, mul(t, d.y));
    po.z = add(s.z, mul(t, d.z));
    po.w = add(s.w, mul(t, d.w));
    const float4* mv = sc.mv + 4 * (size_t)id;
    const float4 c0 = __ldg(mv + 0), c1 = __ldg(mv + 1), c2 = __ldg(mv + 2), c3 = __ldg(mv + 3);
    Surface sf;
    sf.P = xform(c0, c1, c2, c3, po);
    float4 no;
    const uint32_t type = __ldg(&sc.cull[id].type);
    if (type == 0u) {
        no = make_float4(po.x, po.y, po.z, 0.f);
    } else if (type == 2u) {  // EXTENSION: unit triangle, plane normal
        no = make_float4(0.f, 0.f, 1.f, 0.f);
    } else {
        no = make_float4(0.f, 0.f, 0.f, 0.f);
        if (po.x > 0.4998f) no.x = add(no.x, 1.f); else if (po.x < -0.4998f) no.x = sub(no.x, 1.f);
        if (po.y > 0.4998f) no.y = add(no.y, 1.f); else if (po.y < -0.4998f) no.y = sub(no.y, 1.f);
        if (po.z > 0.4998f) no.z = add(no.z, 1.f); else if (po.z < -0.4998f) no.z = sub(no.z, 1.f);
    }
    const float4 n = xform(c0, c1, c2, c3, no);  // SAR:117: mv, not the inverse transpose
    sf.N = normalize(mk3(n.x, n.y, n.z));
    return sf;
}

}  // namespace ex

// ---------------------------------------------------------- conservative filters
namespace cv {

// Per-ray constants of the filters.
struct RayK {
    float ox, oy, oz;     // origin
    float ux, uy, uz;     // unit direction
    float idx, idy, idz;  // 1/d (clamped, never inf/NaN)
    float aix, aiy, aiz;  // |1/d|
    float onorm;          // |o|
    float dlen;           // |d|: t_view = t * dlen
};

__device__ __forceinline__ float safe_inv(float d) {
    const float big = 1e30f;
    if (fabsf(d) < 1e-30f) return copysignf(big, d);
    return __fdiv_rn(1.f, d);
}

__device__ __forceinline__ RayK prepare(const Ray& r) {
    RayK k;
    k.ox = r.s.x; k.oy = r.s.y; k.oz = r.s.z;
    const float l2 = r.d.x * r.d.x + r.d.y * r.d.y + r.d.z * r.d.z;
    const float l = __fsqrt_rn(l2);
    k.dlen = l;
    k.ux = __fdiv_rn(r.d.x, l); k.uy = __fdiv_rn(r.d.y, l); k.uz = __fdiv_rn(r.d.z, l);
    k.idx = safe_inv(r.d.x); k.idy = safe_inv(r.d.y); k.idz = safe_inv(r.d.z);
    k.aix = fabsf(k.idx); k.aiy = fabsf(k.idy); k.aiz = fabsf(k.idz);
    k.onorm = __fsqrt_rn(k.ox * k.ox + k.oy * k.oy + k.oz * k.oz);
    return k;
}

// Bounding-sphere filter: true = the pair may hit, run the exact test.
// Squared distance from the centre to the ray's line, p^2 = |oc|^2 - (oc.u)^2, against the
// padded radius plus the cone term g*|oc|^2 (rt_scene.cpp explains g).  Written so that a
// NaN anywhere lets the pair through.
__device__ __forceinline__ bool sphere_may_hit(const RayK& k, float cx, float cy, float cz, float r2, float h) {
    const float x = cx - k.ox, y = cy - k.oy, z = cz - k.oz;
    const float oc2 = fmaf(z, z, fmaf(y, y, x * x));
    const float dt = fmaf(z, k.uz, fmaf(y, k.uy, x * k.ux));
    const float lhs = fmaf(-dt, dt, oc2 * h);
    return !(lhs > r2);
}

// The same test without normalising the direction (no square root, no division per ray): with dd = |d|^2 both sides are
// scaled by dd,  h |oc|^2 dd - (oc.d)^2 > r2 dd.  Used where a ray meets only a handful of objects (the reference's own
// scenes: 1-9 objects), so that per-ray set-up does not outweigh the tests.  The extra roundings (two products per object)
// are inside the 24 u of slack g carries for the filter's own arithmetic (rt_scene.cpp).
__device__ __forceinline__ bool sphere_may_hit_raw(float ox, float oy, float oz, float dx, float dy, float dz, float dd, float cx, float cy, float cz,
                                                   float r2, float h) {
    const float x = cx - ox, y = cy - oy, z = cz - oz;
    const float oc2 = fmaf(z, z, fmaf(y, y, x * x));
    const float dt = fmaf(z, dz, fmaf(y, dy, x * dx));
    const float lhs = fmaf(-dt, dt, oc2 * (h * dd));
    return !(lhs > r2 * dd);
}

// Slab test of one child box against [0, tlimit].  The box (centre c, half extent e) has to be inflated by
// f * d_far, d_far = distance from the ray origin to the farthest point of the box: every object centre of the subtree
// lies in the box, so |oc| <= d_far and f * d_far bounds the noise term of each object (rt_scene.cpp).  d_far is
// bounded by its 1-norm  sum_k (|c_k - o_k| + e_k)  -- no square root -- whose static part f * sum(e) the scene builder
// has already added to the stored half extent, together with the pad for the rounding of (c - o); the relative
// rounding of the slab arithmetic itself (<= 4e-7 of d_far) is covered by RT_SLAB_REL, also already inside f.
// 3 FADD + 2 FADD + FMUL (inflation), 3 FMUL + 3 FADD + 6 FFMA (slabs), 2 x 3-input min/max.
// Returns the conservative entry parameter through tnear.
#define RT_SLAB_REL 2e-6f
__device__ __forceinline__ bool box_may_hit(float ox, float oy, float oz, float idx, float idy, float idz, float cx, float cy,
                                            float cz, float ex, float ey, float ez, float f, float tlimit, float& tnear) {
    const float x = cx - ox, y = cy - oy, z = cz - oz;
    const float infl = f * (fabsf(x) + fabsf(y) + fabsf(z));
    const float tx = x * idx, ty = y * idy, tz = z * idz;
    const float Ex = ex + infl, Ey = ey + infl, Ez = ez + infl;
    const float ax = fabsf(idx), ay = fabsf(idy), az = fabsf(idz);
    const float tn = fmaxf(fmaxf(fmaf(-Ex, ax, tx), fmaf(-Ey, ay, ty)), fmaf(-Ez, az, tz));
    const float tf = fminf(fminf(fmaf(Ex, ax, tx), fmaf(Ey, ay, ty)), fmaf(Ez, az, tz));
    tnear = tn;
    return !(tn > fminf(tf, tlimit)) && !(tf < 0.f);
}

}  // namespace cv

// ------------------------------------------------------------------ nearest hit
// All three searches return the same (t, id): the exact test decides, the filters only
// skip pairs that the exact test would reject.

// Reference order, exact arithmetic on every pair (SAR:75-168 as written).
__device__ __forceinline__ Hit trace_brute(const Scene& sc, const Ray& ray, unsigned long long* n_exact) {
    Hit h;
    h.t = RT_MAX_FLOAT;
    h.id = -1;
    uint32_t ktype = 0;
    for (uint32_t k = 0; k < sc.n_objs; ++k) {
        const float4* mi = sc.minv + 4 * (size_t)k;
        const float4 c0 = __ldg(mi + 0), c1 = __ldg(mi + 1), c2 = __ldg(mi + 2), c3 = __ldg(mi + 3);
        const uint32_t type = __ldg(&sc.cull[k].type);
        const float4 s = ex::xform(c0, c1, c2, c3, ray.s);
        const float4 d = ex::xform(c0, c1, c2, c3, ray.d);
        float t;
        if (ex::candidate(type, s, d, t) && ex::replaces(t, (int32_t)k, type, h.t, h.id, ktype)) {
            h.t = t;
            h.id = (int32_t)k;
            ktype = type;
        }
    }
    if (n_exact) *n_exact += sc.n_objs;
    return h;
}

__device__ __forceinline__ void exact_pair(const Scene& sc, const Ray& ray, uint32_t k, uint32_t type, Hit& h, uint32_t& ktype) {
    const float4* mi = sc.minv + 4 * (size_t)k;
    const float4 c0 = __ldg(mi + 0), c1 = __ldg(mi + 1), c2 = __ldg(mi + 2), c3 = __ldg(mi + 3);
    const float4 s = ex::xform(c0, c1, c2, c3, ray.s);
    const float4 d = ex::xform(c0, c1, c2, c3, ray.d);
    float t;
    if (ex::candidate(type, s, d, t) && ex::replaces(t, (int32_t)k, type, h.t, h.id, ktype)) {
        h.t = t;
        h.id = (int32_t)k;
        ktype = type;
    }
}

// Every pair goes through the bounding-sphere filter, survivors through the exact test.
__device__ __forceinline__ Hit trace_cull(const Scene& sc, const Ray& ray, unsigned long long* n_exact) {
    Hit h;
    h.t = RT_MAX_FLOAT;
    h.id = -1;
    uint32_t ktype = 0;
    const float dd = fmaf(ray.d.z, ray.d.z, fmaf(ray.d.y, ray.d.y, ray.d.x * ray.d.x));
    const float4* cr = reinterpret_cast<const float4*>(sc.cull);
    unsigned long long ne = 0;
    for (uint32_t k = 0; k < sc.n_objs; ++k) {
        const float4 a = __ldg(cr + 2 * (size_t)k);
        const float4 b = __ldg(cr + 2 * (size_t)k + 1);
        if (cv::sphere_may_hit_raw(ray.s.x, ray.s.y, ray.s.z, ray.d.x, ray.d.y, ray.d.z, dd, a.x, a.y, a.z, a.w, b.x)) {
            exact_pair(sc, ray, k, __float_as_uint(b.w), h, ktype);
            ++ne;
        }
    }
    if (n_exact) *n_exact += ne;
    return h;
}

// BVH: conservative traversal, leaf objects through the sphere filter and the exact test.
// ANYHIT (shadow rays): only "is there an exact t < tcap" matters -- the reference looks at
// nothing but "nearest t in [0,1)" (SAR:229) -- so nodes beyond tcap are pruned and the walk
// stops at the first such t.  The returned t is then "some t < tcap" or ">= tcap / MAX".
template <bool ANYHIT>
__device__ __forceinline__ Hit trace_bvh(const Scene& sc, const Ray& ray, float tcap, unsigned long long* n_exact,
                                         unsigned long long* n_nodes) {
    Hit h;
    h.t = RT_MAX_FLOAT;
    h.id = -1;
    uint32_t ktype = 0;
    if (sc.n_objs == 0) return h;
    const cv::RayK rk = cv::prepare(ray);
    int32_t stack[RT_BVH_STACK];
    int sp = 0;
    int32_t cur = sc.root;
    unsigned long long ne = 0, nn = 0;
    const float4* lc = reinterpret_cast<const float4*>(sc.leaf_cull);
    for (;;) {
        if (cur >= 0) {
            const float4* np = reinterpret_cast<const float4*>(sc.nodes + cur);
            const float4 q0 = __ldg(np + 0), q1 = __ldg(np + 1), q2 = __ldg(np + 2), q3 = __ldg(np + 3);
            ++nn;
            const int32_t c0 = __float_as_int(q3.x), c1 = __float_as_int(q3.y);
            const float tlimit = ANYHIT ? fminf(h.t, tcap) : h.t;
            float tn0, tn1;
            // child 0: ctr (q0.x q0.y q0.z) ext (q0.w q1.x q1.y); child 1: ctr (q1.z q1.w q2.x) ext (q2.y q2.z q2.w)
            const bool h0 = cv::box_may_hit(rk.ox, rk.oy, rk.oz, rk.idx, rk.idy, rk.idz, q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q3.z,
                                            tlimit, tn0);
            const bool h1 = cv::box_may_hit(rk.ox, rk.oy, rk.oz, rk.idx, rk.idy, rk.idz, q1.z, q1.w, q2.x, q2.y, q2.z, q2.w, q3.w,
                                            tlimit, tn1);
            if (h0 && h1) {
                const bool first0 = tn0 <= tn1;
                cur = first0 ? c0 : c1;
                if (sp < RT_BVH_STACK) stack[sp++] = first0 ? c1 : c0;
                continue;
            }
            if (h0) { cur = c0; continue; }
            if (h1) { cur = c1; continue; }
        } else {
            const uint32_t code = (uint32_t)(~cur);
            const uint32_t first = code >> 4, cnt = (code & 15u) + 1u;
            for (uint32_t i = 0; i < cnt; ++i) {
                const float4 a = __ldg(lc + 2 * (size_t)(first + i));
                const float4 b = __ldg(lc + 2 * (size_t)(first + i) + 1);
                if (cv::sphere_may_hit(rk, a.x, a.y, a.z, a.w, b.x)) {
                    exact_pair(sc, ray, __ldg(sc.leaf_obj + first + i), __float_as_uint(b.w), h, ktype);
                    ++ne;
                }
            }
            if (ANYHIT && h.t < tcap) break;
        }
        if (sp == 0) break;
        cur = stack[--sp];
    }
    if (n_exact) *n_exact += ne;
    if (n_nodes) *n_nodes += nn;
    return h;
}

// ---------------------------------------------------------------------- shading
struct Material {
    ex::V3 ambient, diffuse, specular;
    float absorption, shininess;
};

__device__ __forceinline__ Material load_material(const Scene& sc, int32_t id) {
    const float4* m = sc.mat + 3 * (size_t)id;
    const float4 a = __ldg(m + 0), d = __ldg(m + 1), s = __ldg(m + 2);
    Material mt;
    mt.ambient = ex::mk3(a.x, a.y, a.z);
    mt.diffuse = ex::mk3(d.x, d.y, d.z);
    mt.specular = ex::mk3(s.x, s.y, s.z);
    mt.absorption = a.w;
    mt.shininess = d.w;
    return mt;
}

// SAR:195-209: the shadow ray of one light for a hit at P.
__device__ __forceinline__ Ray shadow_ray(const Light& light, const float4 P, ex::V3& lightVec) {
    const ex::V3 fPosition = ex::xyz(P);
    if (light.position.w != 0.f)
        lightVec = ex::vsub(ex::xyz(light.position), fPosition);  // SAR:196
    else
        lightVec = ex::vneg(ex::xyz(light.position));  // SAR:198
    const ex::V3 nd = ex::normalize(lightVec);
    Ray r;
    // rayToLight.start = (float4)(fPosition, 1.f) += 0.01f * (float4)(normalize(dir), 0)
    r.s.x = ex::add(fPosition.x, ex::mul(0.01f, nd.x));
    r.s.y = ex::add(fPosition.y, ex::mul(0.01f, nd.y));
    r.s.z = ex::add(fPosition.z, ex::mul(0.01f, nd.z));
    r.s.w = ex::add(1.f, ex::mul(0.01f, 0.f));
    r.d = make_float4(lightVec.x, lightVec.y, lightVec.z, 0.f);
    return r;
}

struct Phong {
    ex::V3 ambient, diffuse, specular;
};

// SAR:211-237: Phong terms of one light once the shadow ray's nearest t is known.
// ph.specular is carried across lights by the caller (stale when N.L <= 0, SAR:231-232).
// Returns whether this light assigned (or zeroed) ph.specular.
__device__ __forceinline__ bool light_terms(const Light& light, const Material& mat, const float4 P, const ex::V3 N,
                                            ex::V3 lightVec, float shadow_t, Phong& ph) {
    const ex::V3 fPosition = ex::xyz(P);
    lightVec = ex::normalize(lightVec);                // SAR:211
    const ex::V3 normalView = ex::normalize(N);        // SAR:213-214
    const float nDotL = ex::dot(normalView, lightVec);  // SAR:215
    const ex::V3 viewVec = ex::normalize(ex::vneg(fPosition));  // SAR:217-218: camera at the origin
    const ex::V3 reflectVec = ex::normalize(ex::reflect(ex::vneg(lightVec), normalView));  // SAR:220-221
    float rDotV = ex::dot(reflectVec, viewVec);
    rDotV = fmaxf(rDotV, 0.0f);
    ph.ambient = ex::vmul(mat.ambient, ex::xyz(light.ambient));  // SAR:226
    if (shadow_t >= 1.f || shadow_t < 0.f) {                      // SAR:229
        ph.diffuse = ex::vscale(ex::vmul(mat.diffuse, ex::xyz(light.diffuse)), fmaxf(nDotL, 0.f));
        if (nDotL > 0.f) {
            ph.specular = ex::vscale(ex::vmul(mat.specular, ex::xyz(light.specular)), powf(rDotV, fmaxf(mat.shininess, 1.f)));
            return true;
        }
        return false;
    }
    ph.diffuse = ex::mk3(0.f, 0.f, 0.f);
    ph.specular = ex::mk3(0.f, 0.f, 0.f);
    return true;
}

// SAR:261-263 / 275-277: the reflection ray leaving a hit.
__device__ __forceinline__ Ray reflection_ray(const float4 P, const ex::V3 R) {
    const ex::V3 nr = ex::normalize(R);
    Ray r;
    r.d = make_float4(R.x, R.y, R.z, 0.f);
    r.s.x = ex::add(P.x, ex::mul(nr.x, 0.001f));
    r.s.y = ex::add(P.y, ex::mul(nr.y, 0.001f));
    r.s.z = ex::add(P.z, ex::mul(nr.z, 0.001f));
    r.s.w = ex::add(P.w, ex::mul(0.f, 0.001f));
    return r;
}

// OpenCL-Raytracer.cpp:18-26,68-72: primary ray of pixel (ii, jj).
__device__ __forceinline__ Ray camera_ray(uint32_t ii, uint32_t jj, float half_w, float half_h, float height, float zdir) {
    Ray r;
    r.s = make_float4(0.f, 0.f, 0.f, 1.f);
    // pos = (ii, height - jj) as floats; dir = (pos.x - halfWidth, pos.y - halfHeight, zdir)
    const float px = (float)ii;
    const float py = (float)((int)height - (int)jj);
    r.d = make_float4(ex::sub(px, half_w), ex::sub(py, half_h), zdir, 0.f);
    return r;
}

}  // namespace rt
