// rt_pixel.cuh -- the per-pixel path (state machine of shade_and_reflect_kernel.cl:244-285 and of the two
// earlier stages shade_kernel.cl:180-258 / hittest_kernel.cl:54-150) as device functions, shared by the
// megakernel and by the wavefront stages.
#pragma once

#include "rt_device.cuh"
#include "rt_kernels.cuh"

namespace rt {

struct LocalCounters {
    unsigned long long ref_primary = 0, ref_shadow = 0, ref_reflect = 0, traced_shadow = 0, traced_reflect = 0;
    unsigned long long pair_tests = 0, exact_tests = 0, node_visits = 0, hit_pixels = 0;
};

template <int ACCEL, bool COUNT>
struct Tracer {
    const Scene& sc;
    LocalCounters& lc;
    __device__ __forceinline__ Tracer(const Scene& s, LocalCounters& l) : sc(s), lc(l) {}

    __device__ __forceinline__ Hit nearest(const Ray& r) {
        unsigned long long ne = 0, nn = 0;
        Hit h;
        if (ACCEL == 0) {
            h = trace_brute(sc, r, COUNT ? &ne : nullptr);
            if (COUNT) lc.pair_tests += sc.n_objs;
        } else if (ACCEL == 1) {
            h = trace_cull(sc, r, COUNT ? &ne : nullptr);
            if (COUNT) lc.pair_tests += sc.n_objs;
        } else {
            h = trace_bvh<false>(sc, r, RT_MAX_FLOAT, COUNT ? &ne : nullptr, COUNT ? &nn : nullptr);
        }
        if (COUNT) {
            lc.exact_tests += ne;
            lc.node_visits += nn;
        }
        return h;
    }
    // nearest t if it is < 1, otherwise any value >= 1 (SAR:229 only distinguishes these)
    __device__ __forceinline__ float shadow(const Ray& r) {
        if (ACCEL == 2) {
            unsigned long long ne = 0, nn = 0;
            const Hit h = trace_bvh<true>(sc, r, 1.f, COUNT ? &ne : nullptr, COUNT ? &nn : nullptr);
            if (COUNT) {
                lc.exact_tests += ne;
                lc.node_visits += nn;
            }
            return h.t;
        }
        return nearest(r).t;
    }
};

// SAR:184-242 shade().  fColor = ambient + diffuse + specular is ASSIGNED per light (SAR:238): only the last
// light's ambient and diffuse survive, and `specular` is whatever the last light that touched it left there
// (assigned when lit and N.L > 0, zeroed when shadowed, otherwise stale from the previous light, SAR:229-237).
// With skip_dead the lights are therefore walked backwards and only the shadow rays that can reach the output
// are traced; without it every light's shadow ray is traced as the reference does.  Same result either way.
template <class TR>
__device__ __forceinline__ ex::V3 shade_sar(TR& tr, const Scene& sc, const Material& mat, const ex::Surface& sf, int skip_dead,
                                            LocalCounters& lc, bool count) {
    const uint32_t L = sc.n_lights;
    if (count) lc.ref_shadow += L;
    if (L == 0) return ex::mk3(0.f, 0.f, 0.f);
    Phong ph;
    ph.ambient = ph.diffuse = ph.specular = ex::mk3(0.f, 0.f, 0.f);
    if (!skip_dead) {
        for (uint32_t li = 0; li < L; ++li) {
            const Light light = sc.lights[li];
            ex::V3 lv;
            const Ray sr = shadow_ray(light, sf.P, lv);
            const float st = tr.shadow(sr);
            if (count) lc.traced_shadow++;
            light_terms(light, mat, sf.P, sf.N, lv, st, ph);
        }
        return ex::vadd(ex::vadd(ph.ambient, ph.diffuse), ph.specular);
    }
    // last light: ambient, diffuse, and possibly specular
    ex::V3 amb = ex::mk3(0.f, 0.f, 0.f), dif = amb, spec = amb;
    bool need_spec = true;
    for (int li = (int)L - 1; li >= 0 && need_spec; --li) {
        const Light light = sc.lights[li];
        ex::V3 lv;
        const Ray sr = shadow_ray(light, sf.P, lv);
        const float st = tr.shadow(sr);
        if (count) lc.traced_shadow++;
        Phong p2;
        p2.ambient = p2.diffuse = p2.specular = ex::mk3(0.f, 0.f, 0.f);
        const bool touched = light_terms(light, mat, sf.P, sf.N, lv, st, p2);
        if (li == (int)L - 1) {
            amb = p2.ambient;
            dif = p2.diffuse;
        }
        if (touched) {  // this light assigned or zeroed specular: earlier lights cannot reach the output
            spec = p2.specular;
            need_spec = false;
        }
    }
    return ex::vadd(ex::vadd(amb, dif), spec);
}

// shade_kernel.cl:203-255: lights accumulate (absorbColor += ambient + diffuse + specular, :252).
template <class TR>
__device__ __forceinline__ ex::V3 shade_accumulate(TR& tr, const Scene& sc, const Material& mat, const ex::Surface& sf,
                                                   LocalCounters& lc, bool count) {
    ex::V3 acc = ex::mk3(0.f, 0.f, 0.f);
    Phong ph;
    ph.ambient = ph.diffuse = ph.specular = ex::mk3(0.f, 0.f, 0.f);
    for (uint32_t li = 0; li < sc.n_lights; ++li) {
        const Light light = sc.lights[li];
        ex::V3 lv;
        const Ray sr = shadow_ray(light, sf.P, lv);
        const float st = tr.shadow(sr);
        if (count) {
            lc.ref_shadow++;
            lc.traced_shadow++;
        }
        light_terms(light, mat, sf.P, sf.N, lv, st, ph);
        acc = ex::vadd(ex::vadd(ex::vadd(acc, ph.ambient), ph.diffuse), ph.specular);
    }
    return acc;
}

// One pixel / one caller ray: everything the reference's work-item does, plus the optional AOVs.
template <int ACCEL, bool COUNT>
__device__ __forceinline__ void render_pixel(const Scene& sc, const Frame& fr, const Ray& ray, const uint64_t oi,
                                             const bool aov_only, LocalCounters& lc) {
    Tracer<ACCEL, COUNT> tr(sc, lc);
    if (COUNT) lc.ref_primary++;
    const Hit hit = tr.nearest(ray);
    if (fr.hit_id) fr.hit_id[oi] = hit.id;
    if (fr.hit_t) fr.hit_t[oi] = hit.id >= 0 ? hit.t : RT_MAX_FLOAT;
    float4 px = make_float4(0.f, 0.f, 0.f, 1.f);  // OpenCLRaytracer.cpp:32: a missed pixel keeps {0,0,0,1}
    if (hit.id >= 0 && !aov_only) {
        if (COUNT) lc.hit_pixels++;
        if (fr.mode == 0) {  // hittest_kernel.cl:149
            px = make_float4(hit.t, hit.t, hit.t, 1.f);
        } else {
            ex::Surface sf = ex::surface(sc, hit.id, hit.t, ray);
            Material mat = load_material(sc, hit.id);
            if (fr.mode == 1) {  // shade_kernel.cl:257
                const ex::V3 c = shade_accumulate(tr, sc, mat, sf, lc, COUNT);
                px = make_float4(c.x, c.y, c.z, 1.f);
            } else {
                // SAR:253-284
                ex::V3 reflectColor = ex::mk3(0.f, 0.f, 0.f);
                ex::V3 absorbColor = ex::svec(mat.absorption, shade_sar(tr, sc, mat, sf, fr.skip_dead, lc, COUNT));
                float absorptionPercent = mat.absorption;
                uint32_t bounces = fr.max_bounces;  // SAR:258: uint, wraps below zero
                Ray rr = reflection_ray(sf.P, ex::reflect(ex::xyz(ray.d), sf.N));
                for (;;) {
                    if (!(bounces-- > 0u)) break;
                    if (COUNT) lc.ref_reflect++;
                    // SAR:268 traces the ray before testing the absorption; when the test fails the
                    // ray's result is never looked at, so skip_dead does not trace it.
                    const bool alive = absorptionPercent <= 0.999f;
                    if (!alive && fr.skip_dead) break;
                    const Hit rh = tr.nearest(rr);
                    if (COUNT) lc.traced_reflect++;
                    if (rh.id < 0) break;
                    if (!alive) break;
                    sf = ex::surface(sc, rh.id, rh.t, rr);
                    mat = load_material(sc, rh.id);
                    reflectColor = shade_sar(tr, sc, mat, sf, fr.skip_dead, lc, COUNT);
                    const float reflectedAbsorbtion = ex::mul(ex::sub(1.f, absorptionPercent), mat.absorption);
                    absorbColor = ex::vadd(absorbColor, ex::svec(reflectedAbsorbtion, reflectColor));
                    absorptionPercent = ex::add(absorptionPercent, reflectedAbsorbtion);
                    rr = reflection_ray(sf.P, ex::reflect(ex::xyz(rr.d), sf.N));
                }
                // SAR:281-282: fires only when the loop ended on the last permitted bounce
                if (bounces == 0u && absorptionPercent < 1.f)
                    absorbColor = ex::vadd(absorbColor, ex::svec(ex::sub(1.f, absorptionPercent), reflectColor));
                px = make_float4(absorbColor.x, absorbColor.y, absorbColor.z, 1.f);
            }
        }
    }
    if (!aov_only) fr.out[oi] = px;  // one 128-bit store per pixel
}

}  // namespace rt
