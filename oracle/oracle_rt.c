/* TEST INFRASTRUCTURE -- not part of the product path.
 *
 * oracle_rt.c: plain-C CPU restatement of the reference's per-pixel render path,
 * written from the kernel sources' behaviour, each function citing the lines it
 * follows (SAR = shade_and_reflect_kernel.cl, SHD = shade_kernel.cl,
 * HT = hittest_kernel.cl under the reference).
 *
 * Why it exists next to oracle/_ref (the reference kernels themselves compiled as
 * C++): the reference outputs only colours.  This restatement additionally returns
 * the winning object index and t of the primary ray and the ray counts by class,
 * which the parity criteria need.  It is PINNED: tests/test_oracle.py asserts that
 * its colours are bit-identical to oracle/_ref on every scene (when _ref is present)
 * and that it reproduces the committed golden vectors (tests/golden/).
 *
 * Arithmetic contract (mirrored by cl_shim.hpp and by the CUDA path): IEEE
 * binary32, no contraction (build with -ffp-contract=off), sums left to right as
 * written in the kernels, correctly rounded sqrt and division,
 * normalize(v) = v / sqrt(dot(v,v)), pow = powf of the C library.
 */
#include <math.h>
#include <pthread.h>
#include <stdatomic.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>

#include "../include/rt_b200.h"

#define OR_MAX_FLOAT 3.402823466e+38F /* SAR:31 */

typedef struct { float x, y, z; } v3;
typedef struct { float x, y, z, w; } v4;

typedef struct {
    int id;      /* winning object index, -1 = none                  */
    float time;  /* SAR:17                                            */
    v4 P;        /* hit.intersection (float4, w carried)   SAR:14     */
    v3 N;        /* hit.normal                              SAR:15     */
    v3 R;        /* hit.reflection                          SAR:16     */
    const rt_material* mat; /* hit.mat = obj->mat          SAR:119,166 */
} hit_rec;

typedef struct {
    uint64_t rays_primary, rays_shadow, rays_reflect, hit_pixels;
} oracle_stats;

static inline v3 v3sub(v3 a, v3 b) { v3 r = {a.x - b.x, a.y - b.y, a.z - b.z}; return r; }
static inline v3 v3add(v3 a, v3 b) { v3 r = {a.x + b.x, a.y + b.y, a.z + b.z}; return r; }
static inline v3 v3neg(v3 a) { v3 r = {-a.x, -a.y, -a.z}; return r; }
static inline v3 v3scale(v3 a, float s) { v3 r = {a.x * s, a.y * s, a.z * s}; return r; }
static inline v3 sv3(float s, v3 a) { v3 r = {s * a.x, s * a.y, s * a.z}; return r; }
static inline float v3dot(v3 a, v3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static inline v3 v3normalize(v3 v) {
    const float len = sqrtf(v3dot(v, v));
    v3 r = {v.x / len, v.y / len, v.z / len};
    return r;
}
/* SAR:179-182 componentWiseMultiply */
static inline v3 v3mul(v3 a, v3 b) { v3 r = {a.x * b.x, a.y * b.y, a.z * b.z}; return r; }
static inline v3 arr3(const float* p) { v3 r = {p[0], p[1], p[2]}; return r; }

/* SAR:60-65 transform: column-major float16 times float4, summed left to right. */
static inline v4 xform(const float* m, v4 v) {
    v4 o;
    o.x = m[0] * v.x + m[4] * v.y + m[8] * v.z + m[12] * v.w;
    o.y = m[1] * v.x + m[5] * v.y + m[9] * v.z + m[13] * v.w;
    o.z = m[2] * v.x + m[6] * v.y + m[10] * v.z + m[14] * v.w;
    o.w = m[3] * v.x + m[7] * v.y + m[11] * v.z + m[15] * v.w;
    return o;
}

/* SAR:68-70 reflect */
static inline v3 reflect3(v3 incident, v3 normal) {
    const float k = 2.f * v3dot(incident, normal);
    return v3sub(incident, sv3(k, normal));
}

/* SAR:33-58 intersectsWidthBoxSide: one slab of the unit cube. */
static int box_slab(float* tMin, float* tMax, float start, float dir) {
    float t1 = (-0.5f - start);
    float t2 = (0.5f - start);
    if (dir == 0) {
        if (copysignf(t1, t2) == t1) return 0; /* both bounds on the same side: origin outside the slab */
        *tMin = -OR_MAX_FLOAT;
        *tMax = OR_MAX_FLOAT;
        return 1;
    }
    t1 /= dir;
    t2 /= dir;
    if (dir < 0) {
        *tMin = fminf(t1, t2);
        *tMax = fmaxf(t1, t2);
    } else {
        *tMin = t1;
        *tMax = t2;
    }
    return 1;
}

/* SAR:72-177 raycast (identical loop in SHD:66-168 and, t only, HT:66-147).
 * Brute force over all objects in index order, no early out.  Tie rules: a sphere
 * is skipped only if hit.time < t (SAR:108, equal t replaces); a box is skipped if
 * hit.time <= t (SAR:147).  hit->time must be OR_MAX_FLOAT on entry. */
static int raycast(const rt_object* objs, uint32_t count, v4 rstart, v4 rdir, hit_rec* hit, int want_reflection) {
    for (uint32_t k = 0; k < count; ++k) {
        const rt_object* obj = &objs[k];
        const v4 s = xform(obj->mvInverse, rstart); /* SAR:78 */
        const v4 d = xform(obj->mvInverse, rdir);   /* SAR:79 */
        if (obj->type == 0) {                       /* sphere, SAR:82-121 */
            const float A = d.x * d.x + d.y * d.y + d.z * d.z;
            const float B = 2.f * (d.x * s.x + d.y * s.y + d.z * s.z);
            const float Cc = s.x * s.x + s.y * s.y + s.z * s.z - 1.f;
            const float radical = B * B - 4.f * A * Cc;
            if (radical < 0) continue;
            const float root = sqrtf(radical);
            const float t1 = (-B - root) / (2.f * A);
            const float t2 = (-B + root) / (2.f * A);
            const float tMin = (t1 >= 0 && t2 >= 0) ? fminf(t1, t2) : fmaxf(t1, t2);
            if (tMin < 0) continue;
            if (hit->time < tMin) continue;
            hit->time = tMin;
            hit->id = (int)k;
            v4 po = {s.x + tMin * d.x, s.y + tMin * d.y, s.z + tMin * d.z, s.w + tMin * d.w}; /* SAR:112 */
            hit->P = xform(obj->mv, po);
            v4 no = po;
            no.w = 0.f;
            const v4 n = xform(obj->mv, no); /* SAR:117: mv, not the inverse transpose */
            v3 n3 = {n.x, n.y, n.z};
            hit->N = v3normalize(n3);
            hit->mat = &obj->mat;
        } else if (obj->type == 1) { /* box, SAR:123-168 */
            float txMin, txMax, tyMin, tyMax, tzMin, tzMax;
            if (!box_slab(&txMin, &txMax, s.x, d.x)) continue;
            if (!box_slab(&tyMin, &tyMax, s.y, d.y)) continue;
            if (!box_slab(&tzMin, &tzMax, s.z, d.z)) continue;
            const float tMin = fmaxf(fmaxf(txMin, tyMin), tzMin);
            const float tMax = fminf(fminf(txMax, tyMax), tzMax);
            if (tMax < tMin) continue;
            const float tHit = (tMin >= 0 && tMax >= 0) ? fminf(tMin, tMax) : fmaxf(tMin, tMax);
            if (tHit < 0) continue;
            if (hit->time <= tHit) continue;
            v4 po = {s.x + tHit * d.x, s.y + tHit * d.y, s.z + tHit * d.z, s.w + tHit * d.w}; /* SAR:149 */
            v4 no = {0.f, 0.f, 0.f, 0.f};
            if (po.x > 0.4998f) no.x += 1.f; else if (po.x < -0.4998f) no.x -= 1.f;
            if (po.y > 0.4998f) no.y += 1.f; else if (po.y < -0.4998f) no.y -= 1.f;
            if (po.z > 0.4998f) no.z += 1.f; else if (po.z < -0.4998f) no.z -= 1.f;
            hit->time = tHit;
            hit->id = (int)k;
            hit->P = xform(obj->mv, po);
            const v4 n = xform(obj->mv, no);
            v3 n3 = {n.x, n.y, n.z};
            hit->N = v3normalize(n3);
            hit->mat = &obj->mat;
        } else if (obj->type == 2) {
            /* EXTENSION (no reference code): canonical unit triangle, the half of the
             * unit square [-0.5,0.5]^2 of the plane z = 0 below its diagonal: x >= -0.5, y >= -0.5, x + y <= 0.
             * Written in the style of the box case: object-space ray, IEEE divide, "skip if hit.time <= t" (lowest
             * index wins ties among non-spheres), normal = normalize(mv * (0,0,1,0)), two-sided. */
            if (d.z == 0) continue;
            const float tHit = (0.f - s.z) / d.z;
            if (tHit < 0) continue;
            const float px = s.x + tHit * d.x;
            const float py = s.y + tHit * d.y;
            if (px < -0.5f || py < -0.5f || px + py > 0.f) continue;
            if (hit->time <= tHit) continue;
            v4 po = {px, py, s.z + tHit * d.z, s.w + tHit * d.w};
            v4 no = {0.f, 0.f, 1.f, 0.f};
            hit->time = tHit;
            hit->id = (int)k;
            hit->P = xform(obj->mv, po);
            const v4 n = xform(obj->mv, no);
            v3 n3 = {n.x, n.y, n.z};
            hit->N = v3normalize(n3);
            hit->mat = &obj->mat;
        }
        /* any other type value: the kernels' switch has no default, the object is skipped */
    }
    if (hit->time == OR_MAX_FLOAT) return 0; /* SAR:173 */
    if (want_reflection) {
        v3 d3 = {rdir.x, rdir.y, rdir.z};
        hit->R = reflect3(d3, hit->N); /* SAR:175, un-normalised incident direction */
    }
    return 1;
}

typedef struct { v3 ambient, diffuse, specular; } phong_terms;

/* One iteration of the light loop, SAR:193-237 / SHD:207-251: shadow ray + Phong terms.
 * `terms->specular` is carried across lights by the caller: it is re-assigned only when
 * the light is visible and N.L > 0, and zeroed when shadowed (SAR:229-237); `diffuse`
 * and `ambient` are assigned every iteration. */
static void light_terms(const rt_object* objs, uint32_t M, const rt_light* light, const hit_rec* hit,
                        phong_terms* terms, oracle_stats* st) {
    const v3 fPosition = {hit->P.x, hit->P.y, hit->P.z};
    v3 lightVec;
    if (light->position[3] != 0)
        lightVec = v3sub(arr3(light->position), fPosition); /* SAR:196 */
    else
        lightVec = v3neg(arr3(light->position)); /* SAR:198 */

    /* SAR:201-209: shadow ray, un-normalised direction, 0.01 skin along the unit direction */
    v4 start = {fPosition.x, fPosition.y, fPosition.z, 1.f};
    const v4 dir = {lightVec.x, lightVec.y, lightVec.z, 0.f};
    const v3 nd = v3normalize(lightVec);
    start.x += 0.01f * nd.x;
    start.y += 0.01f * nd.y;
    start.z += 0.01f * nd.z;
    start.w += 0.01f * 0.f;
    hit_rec sh;
    sh.time = OR_MAX_FLOAT;
    sh.id = -1;
    sh.mat = 0;
    raycast(objs, M, start, dir, &sh, 0);
    st->rays_shadow++;

    lightVec = v3normalize(lightVec);                 /* SAR:211 */
    const v3 normalView = v3normalize(hit->N);        /* SAR:213-214 */
    const float nDotL = v3dot(normalView, lightVec);  /* SAR:215 */
    const v3 viewVec = v3normalize(v3neg(fPosition)); /* SAR:217-218: camera at the origin */
    const v3 reflectVec = v3normalize(reflect3(v3neg(lightVec), normalView)); /* SAR:220-221 */
    float rDotV = v3dot(reflectVec, viewVec);
    rDotV = fmaxf(rDotV, 0.0f);

    terms->ambient = v3mul(arr3(hit->mat->ambient), arr3(light->ambient)); /* SAR:226 */
    if (sh.time >= 1.f || sh.time < 0) {                                   /* SAR:229 */
        terms->diffuse = v3scale(v3mul(arr3(hit->mat->diffuse), arr3(light->diffuse)), fmaxf(nDotL, 0.f));
        if (nDotL > 0)
            terms->specular = v3scale(v3mul(arr3(hit->mat->specular), arr3(light->specular)),
                                      powf(rDotV, fmaxf(hit->mat->shininess, 1.f)));
    } else {
        terms->diffuse.x = terms->diffuse.y = terms->diffuse.z = 0.f;
        terms->specular.x = terms->specular.y = terms->specular.z = 0.f;
    }
}

/* SAR:184-242 shade(): fColor = ambient + diffuse + specular is ASSIGNED per light (SAR:238),
 * so with several lights only the last one's result survives (all shadow rays are still traced). */
static v3 shade_sar(const rt_object* objs, uint32_t M, const rt_light* lights, uint32_t L, const hit_rec* hit,
                    oracle_stats* st) {
    v3 fColor = {0.f, 0.f, 0.f};
    phong_terms t;
    memset(&t, 0, sizeof t);
    for (uint32_t li = 0; li < L; ++li) {
        light_terms(objs, M, &lights[li], hit, &t, st);
        fColor = v3add(v3add(t.ambient, t.diffuse), t.specular);
    }
    return fColor;
}

/* SHD:203-255: absorbColor = absorbColor + ambient + diffuse + specular (SHD:252), lights accumulate. */
static v3 shade_accumulate(const rt_object* objs, uint32_t M, const rt_light* lights, uint32_t L, const hit_rec* hit,
                           oracle_stats* st) {
    v3 acc = {0.f, 0.f, 0.f};
    phong_terms t;
    memset(&t, 0, sizeof t);
    for (uint32_t li = 0; li < L; ++li) {
        light_terms(objs, M, &lights[li], hit, &t, st);
        acc = v3add(v3add(v3add(acc, t.ambient), t.diffuse), t.specular);
    }
    return acc;
}

static inline void reflection_ray(const hit_rec* h, v4* start, v4* dir) {
    /* SAR:261-263 / 275-277: start = intersection (float4, w kept) + normalize(reflection) * 0.001 */
    const v3 nr = v3normalize(h->R);
    dir->x = h->R.x; dir->y = h->R.y; dir->z = h->R.z; dir->w = 0.f;
    start->x = h->P.x + nr.x * 0.001f;
    start->y = h->P.y + nr.y * 0.001f;
    start->z = h->P.z + nr.z * 0.001f;
    start->w = h->P.w + 0.f * 0.001f;
}

/* One work-item.  mode 2: SAR:244-285; mode 1: SHD:180-258; mode 0: HT:54-150.
 * Returns 1 if the pixel was written (primary ray hit). */
static int render_pixel(int mode, uint32_t max_bounces, const rt_object* objs, uint32_t M, const rt_light* lights,
                        uint32_t L, const rt_ray* ray, float* rgba, int32_t* hit_id, float* hit_t, oracle_stats* st) {
    const v4 rs = {ray->start[0], ray->start[1], ray->start[2], ray->start[3]};
    const v4 rd = {ray->direction[0], ray->direction[1], ray->direction[2], ray->direction[3]};
    hit_rec hit;
    hit.time = OR_MAX_FLOAT;
    hit.id = -1;
    hit.mat = 0;
    st->rays_primary++;
    const int got = raycast(objs, M, rs, rd, &hit, mode == 2);
    if (hit_id) *hit_id = got ? hit.id : -1;
    if (hit_t) *hit_t = got ? hit.time : OR_MAX_FLOAT;
    if (!got) return 0; /* SAR:251: return without writing */
    st->hit_pixels++;

    if (mode == 0) { /* HT:149 */
        rgba[0] = rgba[1] = rgba[2] = hit.time;
        return 1;
    }
    if (mode == 1) { /* SHD:257 */
        const v3 c = shade_accumulate(objs, M, lights, L, &hit, st);
        rgba[0] = c.x; rgba[1] = c.y; rgba[2] = c.z;
        return 1;
    }

    v3 reflectColor = {0.f, 0.f, 0.f};
    v3 absorbColor = sv3(hit.mat->absorption, shade_sar(objs, M, lights, L, &hit, st)); /* SAR:255 */
    float absorptionPercent = hit.mat->absorption;
    uint32_t bounces = max_bounces; /* SAR:258: uint, wraps */

    v4 rstart, rdir;
    reflection_ray(&hit, &rstart, &rdir);
    hit_rec rh;
    rh.time = OR_MAX_FLOAT;
    rh.id = -1;
    rh.mat = 0;

    /* SAR:268: the reflection ray is traced BEFORE the absorption test (short-circuit order) */
    for (;;) {
        if (!(bounces-- > 0)) break;
        st->rays_reflect++;
        if (!raycast(objs, M, rstart, rdir, &rh, 1)) break;
        if (!(absorptionPercent <= 0.999f)) break;
        reflectColor = shade_sar(objs, M, lights, L, &rh, st);
        const float reflectedAbsorbtion = (1.f - absorptionPercent) * rh.mat->absorption;
        absorbColor = v3add(absorbColor, sv3(reflectedAbsorbtion, reflectColor));
        absorptionPercent += reflectedAbsorbtion;
        reflection_ray(&rh, &rstart, &rdir);
        rh.time = OR_MAX_FLOAT;
    }
    /* SAR:281-282: fires only when the loop ended on the last permitted bounce */
    if (bounces == 0 && absorptionPercent < 1.f)
        absorbColor = v3add(absorbColor, sv3(1.f - absorptionPercent, reflectColor));

    rgba[0] = absorbColor.x; rgba[1] = absorbColor.y; rgba[2] = absorbColor.z;
    return 1;
}

/* ---- threaded drivers ------------------------------------------------------ */
typedef struct {
    int mode;
    uint32_t max_bounces;
    const rt_object* objs;
    uint32_t M;
    const rt_light* lights;
    uint32_t L;
    const rt_ray* rays;
    uint64_t N;
    float* rgba;
    int32_t* hit_id;
    float* hit_t;
    atomic_ullong* next;
    oracle_stats st;
    int trace_only;
    uint64_t chunk;
} job_t;

static void* worker(void* arg) {
    job_t* j = (job_t*)arg;
    const uint64_t chunk = j->chunk ? j->chunk : 256;
    for (;;) {
        const uint64_t b = atomic_fetch_add(j->next, chunk);
        if (b >= j->N) break;
        const uint64_t e = (b + chunk < j->N) ? b + chunk : j->N;
        for (uint64_t i = b; i < e; ++i) {
            if (j->trace_only) {
                const rt_ray* r = &j->rays[i];
                const v4 rs = {r->start[0], r->start[1], r->start[2], r->start[3]};
                const v4 rd = {r->direction[0], r->direction[1], r->direction[2], r->direction[3]};
                hit_rec h;
                h.time = OR_MAX_FLOAT;
                h.id = -1;
                h.mat = 0;
                const int got = raycast(j->objs, j->M, rs, rd, &h, 0);
                j->hit_id[i] = got ? h.id : -1;
                j->hit_t[i] = got ? h.time : OR_MAX_FLOAT;
            } else {
                render_pixel(j->mode, j->max_bounces, j->objs, j->M, j->lights, j->L, &j->rays[i], j->rgba + 4 * i,
                             j->hit_id ? j->hit_id + i : 0, j->hit_t ? j->hit_t + i : 0, &j->st);
            }
        }
    }
    return 0;
}

static int run_jobs(job_t* proto, int nthreads, oracle_stats* total) {
    if (nthreads <= 0) {
        nthreads = (int)sysconf(_SC_NPROCESSORS_ONLN);
    }
    if (nthreads < 1) nthreads = 1;
    if (nthreads > 256) nthreads = 256;
    atomic_ullong next;
    atomic_init(&next, 0);
    job_t* jobs = (job_t*)calloc((size_t)nthreads, sizeof(job_t));
    pthread_t* th = (pthread_t*)calloc((size_t)nthreads, sizeof(pthread_t));
    if (!jobs || !th) return -1;
    /* every thread gets work even on a few thousand expensive rays */
    uint64_t chunk = proto->N / ((uint64_t)nthreads * 8u);
    chunk = chunk < 1 ? 1 : (chunk > 256 ? 256 : chunk);
    for (int t = 0; t < nthreads; ++t) {
        jobs[t] = *proto;
        jobs[t].chunk = chunk;
        jobs[t].next = &next;
        memset(&jobs[t].st, 0, sizeof(oracle_stats));
    }
    if (nthreads == 1) {
        worker(&jobs[0]);
    } else {
        for (int t = 0; t < nthreads; ++t) pthread_create(&th[t], 0, worker, &jobs[t]);
        for (int t = 0; t < nthreads; ++t) pthread_join(th[t], 0);
    }
    if (total) {
        memset(total, 0, sizeof *total);
        for (int t = 0; t < nthreads; ++t) {
            total->rays_primary += jobs[t].st.rays_primary;
            total->rays_shadow += jobs[t].st.rays_shadow;
            total->rays_reflect += jobs[t].st.rays_reflect;
            total->hit_pixels += jobs[t].st.hit_pixels;
        }
    }
    free(jobs);
    free(th);
    return 0;
}

/* mode as rt_mode.  rgba: N float4, pre-filled here with {0,0,0,1} (OpenCLRaytracer.cpp:32);
 * hit_id / hit_t (optional): primary-ray winner and t (-1 / MAX_FLOAT on miss). */
int oracle_render(int mode, uint32_t max_bounces, const rt_object* objs, uint32_t n_objs, const rt_light* lights,
                  uint32_t n_lights, const rt_ray* rays, uint64_t n_rays, float* rgba, int32_t* hit_id, float* hit_t,
                  oracle_stats* stats, int nthreads) {
    if (mode < 0 || mode > 2 || !rgba) return -1;
    for (uint64_t i = 0; i < n_rays; ++i) {
        rgba[4 * i + 0] = rgba[4 * i + 1] = rgba[4 * i + 2] = 0.f;
        rgba[4 * i + 3] = 1.f;
    }
    job_t j;
    memset(&j, 0, sizeof j);
    j.mode = mode;
    j.max_bounces = max_bounces;
    j.objs = objs;
    j.M = n_objs;
    j.lights = lights;
    j.L = n_lights;
    j.rays = rays;
    j.N = n_rays;
    j.rgba = rgba;
    j.hit_id = hit_id;
    j.hit_t = hit_t;
    return run_jobs(&j, nthreads, stats);
}

/* Nearest hit of arbitrary rays (the raycast() of SAR:72-177 alone). */
int oracle_trace(const rt_object* objs, uint32_t n_objs, const rt_ray* rays, uint64_t n_rays, int32_t* hit_id,
                 float* hit_t, int nthreads) {
    if (!hit_id || !hit_t) return -1;
    job_t j;
    memset(&j, 0, sizeof j);
    j.objs = objs;
    j.M = n_objs;
    j.rays = rays;
    j.N = n_rays;
    j.hit_id = hit_id;
    j.hit_t = hit_t;
    j.trace_only = 1;
    return run_jobs(&j, nthreads, 0);
}
