// rt_wave.cuh -- the wavefront form of the per-pixel path.
//
// The reference runs the whole pixel in one work-item (shade_and_reflect_kernel.cl:244-285).  Here every pixel
// is a *path* with exactly one ray in flight; a generation of the pipeline is
//     trace(queue A)  ->  advance(queue A -> queue B)
// where `advance` consumes the hit of a path's ray, moves the path's state machine one step (this file) and
// emits at most one new ray, compacted into queue B with a warp ballot.  Paths whose primary ray missed, whose
// absorption saturated or whose bounce budget ran out drop out of the queue, so every warp of the next trace
// is full again.  The bounce loop never returns to the host: the queues, their counters and the path state
// stay in HBM.  Output-identical to render_pixel() (rt_pixel.cuh) with skip_dead = 1.
#pragma once

#include "rt_device.cuh"
#include "rt_kernels.cuh"
#include "rt_pixel.cuh"

namespace rt {

// Path state, structure of arrays indexed by path id, 80 B + 4 B per path.
struct PathArrays {
    float4* acc;    // absorbColor.xyz, absorptionPercent            (mode SHADE: accumulated colour, -)
    float4* refl;   // reflectColor.xyz, bounces as uint bits        (mode SHADE: carried specular, -)
    float4* P;      // hit.intersection (w carried)
    float4* Nm;     // hit.normal.xyz, object id as int bits
    float4* Rc;     // hit.reflection.xyz, ctrl as uint bits
    uint32_t* out;  // framebuffer index of the path
};

// A generation's queue is ONE array filled from both ends: tasks for the BVH / filter / brute-force kernels ("class B":
// reflection rays, shadow rays of lights without grids, everything when the grids are off) occupy [0, nB), tasks for the
// perspective-grid kernel ("class A": camera rays, shadow rays of point lights; rt_pgrid.cuh) occupy (cap - 1 - j),
// j in [0, nA).  Counters: count[2 g] = nB, count[2 g + 1] = nA of generation g.
struct TaskQueue {
    float4* s;       // ray start
    float4* d;       // ray direction
    uint32_t* path;  // path id | light code << 27 | RT_TASK_ANYHIT
};
#define RT_TASK_ANYHIT 0x80000000u
#define RT_TASK_LIGHT_SHIFT 27
#define RT_TASK_NO_GRID 15u         // light code of a shadow ray whose light has no grids (and of every non-shadow ray)
#define RT_TASK_PATH_MASK 0x07ffffffu

// ctrl word: phase (2 bits) | last-light-lit flag | light index of the shadow ray in flight
#define RT_PH_PRIMARY 0u
#define RT_PH_SHADOW_PRIMARY 1u  // shading the primary hit
#define RT_PH_SHADOW_BOUNCE 2u   // shading a reflection hit
#define RT_PH_REFLECT 3u
#define RT_CTRL_LASTLIT 4u
#define RT_CTRL_LI_SHIFT 8

struct WaveParams {
    uint32_t max_bounces;
    int mode;
    uint32_t grid_lights;  // bit li: shadow rays of light li go to the perspective-grid kernel (0: grids off)
    uint32_t qcap;         // capacity of one task queue of this launch sequence (class A tasks sit at qcap - 1 - j)
};

struct StepOut {
    bool emit;    // a new ray goes into the next queue
    bool anyhit;  // ... and it is a shadow ray
    uint32_t light;  // ... of this light
    Ray ray;
    bool write;   // the path is finished: px goes to the framebuffer
    float4 px;
};

namespace wave {

__device__ __forceinline__ uint32_t f2u(float f) { return __float_as_uint(f); }
__device__ __forceinline__ float u2f(uint32_t u) { return __int_as_float((int)u); }

__device__ __forceinline__ void emit_shadow(const Scene& sc, uint32_t li, const float4 P, StepOut& o) {
    ex::V3 lv;
    o.ray = shadow_ray(sc.lights[li], P, lv);
    o.emit = true;
    o.anyhit = true;
    o.light = li;
}

// SAR:281-284: the post-loop rule and the pixel store.
__device__ __forceinline__ void finish(const ex::V3 absorbColor, float ap, const ex::V3 reflectColor, uint32_t bounces, StepOut& o) {
    ex::V3 c = absorbColor;
    if (bounces == 0u && ap < 1.f) c = ex::vadd(c, ex::svec(ex::sub(1.f, ap), reflectColor));
    o.write = true;
    o.px = make_float4(c.x, c.y, c.z, 1.f);
}

// The head of one iteration of SAR:268: `bounces-- > 0`, then (skip_dead) the absorption test, then the ray.
template <bool COUNT>
__device__ __forceinline__ void try_reflect(const PathArrays& pa, uint32_t p, const ex::V3 absorbColor, float ap,
                                            const ex::V3 reflectColor, uint32_t bounces, const float4 P, const ex::V3 R,
                                            StepOut& o, LocalCounters& lc) {
    const bool go = bounces-- > 0u;  // uint wrap on purpose (SAR:258,268)
    if (go && COUNT) lc.ref_reflect++;
    if (!go || !(ap <= 0.999f)) {
        finish(absorbColor, ap, reflectColor, bounces, o);
        return;
    }
    pa.acc[p] = make_float4(absorbColor.x, absorbColor.y, absorbColor.z, ap);
    pa.refl[p] = make_float4(reflectColor.x, reflectColor.y, reflectColor.z, u2f(bounces));
    float4 rc = make_float4(R.x, R.y, R.z, u2f(RT_PH_REFLECT));
    pa.Rc[p] = rc;
    o.ray = reflection_ray(P, R);
    o.emit = true;
    o.anyhit = false;
    if (COUNT) lc.traced_reflect++;
}

// A shade() call has produced `color` (SAR:255 for the primary hit, SAR:269 inside the loop).
template <bool COUNT>
__device__ __forceinline__ void shade_done(const Scene& sc, const WaveParams& wp, const PathArrays& pa, uint32_t p, uint32_t phase,
                                           const ex::V3 color, const float4 P, const ex::V3 R, int32_t obj, StepOut& o,
                                           LocalCounters& lc) {
    const float absorption = __ldg(&sc.mat[3 * (size_t)obj].w);
    ex::V3 absorbColor, reflectColor;
    float ap;
    uint32_t bounces;
    if (phase == RT_PH_SHADOW_PRIMARY) {
        absorbColor = ex::svec(absorption, color);  // SAR:255
        ap = absorption;                            // SAR:256
        bounces = wp.max_bounces;                   // SAR:258
        reflectColor = ex::mk3(0.f, 0.f, 0.f);
    } else {
        const float4 a = pa.acc[p], r = pa.refl[p];
        absorbColor = ex::mk3(a.x, a.y, a.z);
        ap = a.w;
        bounces = f2u(r.w);
        reflectColor = color;                                             // SAR:269
        const float ra = ex::mul(ex::sub(1.f, ap), absorption);          // SAR:270
        absorbColor = ex::vadd(absorbColor, ex::svec(ra, reflectColor));  // SAR:271
        ap = ex::add(ap, ra);                                             // SAR:272
    }
    try_reflect<COUNT>(pa, p, absorbColor, ap, reflectColor, bounces, P, R, o, lc);
}

// A hit has been accepted (primary or reflection): start shade() for it.
template <bool COUNT>
__device__ __forceinline__ void begin_shade(const Scene& sc, const WaveParams& wp, const PathArrays& pa, uint32_t p, uint32_t phase,
                                            const float4 P, const ex::V3 N, const ex::V3 R, int32_t obj, StepOut& o, LocalCounters& lc) {
    const uint32_t L = sc.n_lights;
    if (COUNT) lc.ref_shadow += L;
    if (L == 0u) {  // shade() returns (0,0,0)
        if (wp.mode == 1) {
            o.write = true;
            o.px = make_float4(0.f, 0.f, 0.f, 1.f);
            return;
        }
        shade_done<COUNT>(sc, wp, pa, p, phase, ex::mk3(0.f, 0.f, 0.f), P, R, obj, o, lc);
        return;
    }
    // SAR mode walks the lights backwards (only the rays that can reach the output, see shade_sar());
    // SHADE mode (shade_kernel.cl) accumulates over all lights in order.
    const uint32_t li = wp.mode == 1 ? 0u : L - 1u;
    pa.P[p] = P;
    pa.Nm[p] = make_float4(N.x, N.y, N.z, __int_as_float(obj));
    pa.Rc[p] = make_float4(R.x, R.y, R.z, u2f(phase | (li << RT_CTRL_LI_SHIFT)));
    if (wp.mode == 1) {
        pa.acc[p] = make_float4(0.f, 0.f, 0.f, 0.f);
        pa.refl[p] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    emit_shadow(sc, li, P, o);
    if (COUNT) lc.traced_shadow++;
}

// One step of path p: `ray` is the ray that was traced for it, `res` its result.
template <bool COUNT>
__device__ __forceinline__ void step(const Scene& sc, const WaveParams& wp, const PathArrays& pa, uint32_t p, bool first,
                                     const Ray& ray, const Hit res, StepOut& o, LocalCounters& lc) {
    o.emit = o.write = o.anyhit = false;
    o.light = RT_TASK_NO_GRID;
    uint32_t ctrl = RT_PH_PRIMARY;
    float4 rc = make_float4(0.f, 0.f, 0.f, 0.f);
    if (!first) {
        rc = pa.Rc[p];
        ctrl = f2u(rc.w);
    }
    const uint32_t phase = ctrl & 3u;

    if (phase == RT_PH_PRIMARY || phase == RT_PH_REFLECT) {
        if (phase == RT_PH_PRIMARY && COUNT) lc.ref_primary++;
        if (res.id < 0) {
            if (phase == RT_PH_PRIMARY) {  // SAR:251: the work-item returns, the pixel keeps {0,0,0,1}
                o.write = true;
                o.px = make_float4(0.f, 0.f, 0.f, 1.f);
            } else {  // the loop condition of SAR:268 fails on raycast()
                const float4 a = pa.acc[p], r = pa.refl[p];
                finish(ex::mk3(a.x, a.y, a.z), a.w, ex::mk3(r.x, r.y, r.z), f2u(r.w), o);
            }
            return;
        }
        if (phase == RT_PH_PRIMARY) {
            if (COUNT) lc.hit_pixels++;
            if (wp.mode == 0) {  // hittest_kernel.cl:149
                o.write = true;
                o.px = make_float4(res.t, res.t, res.t, 1.f);
                return;
            }
        }
        const ex::Surface sf = ex::surface(sc, res.id, res.t, ray);
        const ex::V3 R = ex::reflect(ex::xyz(ray.d), sf.N);  // SAR:175
        begin_shade<COUNT>(sc, wp, pa, p, phase == RT_PH_PRIMARY ? RT_PH_SHADOW_PRIMARY : RT_PH_SHADOW_BOUNCE, sf.P, sf.N, R,
                           res.id, o, lc);
        return;
    }

    // ---- a shadow ray came back
    const uint32_t li = ctrl >> RT_CTRL_LI_SHIFT;
    const uint32_t L = sc.n_lights;
    const float4 P = pa.P[p];
    const float4 nm = pa.Nm[p];
    const ex::V3 N = ex::mk3(nm.x, nm.y, nm.z);
    const int32_t obj = __float_as_int(nm.w);
    const Material mat = load_material(sc, obj);
    const float shadow_t = res.id >= 0 ? res.t : RT_MAX_FLOAT;
    const Light light = sc.lights[li];
    ex::V3 lv;
    if (light.position.w != 0.f)
        lv = ex::vsub(ex::xyz(light.position), ex::xyz(P));
    else
        lv = ex::vneg(ex::xyz(light.position));
    const ex::V3 R = ex::mk3(rc.x, rc.y, rc.z);

    if (wp.mode == 1) {  // shade_kernel.cl:203-255, lights in order, everything accumulates
        const float4 a = pa.acc[p], sp = pa.refl[p];
        Phong ph;
        ph.ambient = ph.diffuse = ex::mk3(0.f, 0.f, 0.f);
        ph.specular = ex::mk3(sp.x, sp.y, sp.z);
        light_terms(light, mat, P, N, lv, shadow_t, ph);
        const ex::V3 acc = ex::vadd(ex::vadd(ex::vadd(ex::mk3(a.x, a.y, a.z), ph.ambient), ph.diffuse), ph.specular);
        if (li + 1u < L) {
            pa.acc[p] = make_float4(acc.x, acc.y, acc.z, 0.f);
            pa.refl[p] = make_float4(ph.specular.x, ph.specular.y, ph.specular.z, 0.f);
            pa.Rc[p] = make_float4(rc.x, rc.y, rc.z, u2f(phase | ((li + 1u) << RT_CTRL_LI_SHIFT)));
            emit_shadow(sc, li + 1u, P, o);
            if (COUNT) lc.traced_shadow++;
        } else {
            o.write = true;
            o.px = make_float4(acc.x, acc.y, acc.z, 1.f);
        }
        return;
    }

    Phong ph;
    ph.ambient = ph.diffuse = ph.specular = ex::mk3(0.f, 0.f, 0.f);
    const bool touched = light_terms(light, mat, P, N, lv, shadow_t, ph);
    const bool lit = shadow_t >= 1.f || shadow_t < 0.f;
    if (li == L - 1u && lit) ctrl |= RT_CTRL_LASTLIT;
    if (!touched && li > 0u) {  // specular still undecided: the previous light's shadow ray is live
        ctrl = (ctrl & 0xffu) | ((li - 1u) << RT_CTRL_LI_SHIFT);
        pa.Rc[p] = make_float4(rc.x, rc.y, rc.z, u2f(ctrl));
        emit_shadow(sc, li - 1u, P, o);
        if (COUNT) lc.traced_shadow++;
        return;
    }
    // fColor = ambient + diffuse + specular of the LAST light (SAR:238), specular as decided above
    ex::V3 amb = ph.ambient, dif = ph.diffuse;
    const ex::V3 spec = touched ? ph.specular : ex::mk3(0.f, 0.f, 0.f);
    if (li != L - 1u) {  // recompute the last light's ambient / diffuse from its recorded shadow state
        const Light last = sc.lights[L - 1u];
        ex::V3 lv2;
        if (last.position.w != 0.f)
            lv2 = ex::vsub(ex::xyz(last.position), ex::xyz(P));
        else
            lv2 = ex::vneg(ex::xyz(last.position));
        Phong p2;
        p2.ambient = p2.diffuse = p2.specular = ex::mk3(0.f, 0.f, 0.f);
        light_terms(last, mat, P, N, lv2, (ctrl & RT_CTRL_LASTLIT) ? RT_MAX_FLOAT : 0.5f, p2);
        amb = p2.ambient;
        dif = p2.diffuse;
    }
    const ex::V3 color = ex::vadd(ex::vadd(amb, dif), spec);
    shade_done<COUNT>(sc, wp, pa, p, phase, color, P, R, obj, o, lc);
}

}  // namespace wave
}  // namespace rt
