// TEST INFRASTRUCTURE -- not part of the product path.
//
// Host wrapper around the reference's own OpenCL kernel sources compiled as C++
// (see cl_shim.hpp and Makefile).  SAR_INC / SHADE_INC / HITTEST_INC are the
// sed-rewritten temporary copies of
//   /root/reference/shade_and_reflect_kernel.cl   (live kernel, SAR:244-285)
//   /root/reference/shade_kernel.cl               (dead stage 2, :180-258)
//   /root/reference/hittest_kernel.cl             (dead stage 1, :54-150)
// Nothing of the reference is copied into the repository: the sources are read
// where they lie and only the built library lands in oracle/_ref/.
//
// Exposes what OpenCLRaytracer::Render does on the device side
// (OpenCLRaytracer.cpp:80-105): one work-item per ray, result buffer pre-filled
// with {0,0,0,1} (OpenCLRaytracer.cpp:32), a work-item that misses writes nothing.
#include <atomic>
#include <cstring>
#include <thread>
#include <vector>

#include "../include/rt_b200.h"
#include "cl_shim.hpp"

thread_local long cl_shim_global_id = 0;

namespace sar {
#include SAR_INC
}
namespace shd {
#include SHADE_INC
}
namespace hts {
#include HITTEST_INC
}

namespace {

template <typename Obj, typename Lgt, typename RayT>
struct Scene {
    std::vector<Obj> objs;
    std::vector<Lgt> lights;
};

inline void set3(float3& d, const float* s) { d = float3{s[0], s[1], s[2]}; }
inline void set16(float16& d, const float* m) {
    d = float16{m[0], m[1], m[2], m[3], m[4], m[5], m[6], m[7], m[8], m[9], m[10], m[11], m[12], m[13], m[14], m[15]};
}
inline void set16t(float16& d, const float* m) {  // transpose (ObjectData.cpp:8)
    d = float16{m[0], m[4], m[8], m[12], m[1], m[5], m[9], m[13], m[2], m[6], m[10], m[14], m[3], m[7], m[11], m[15]};
}

// cl_ObjectData(const ObjectData&) OpenCLRaytracer.cpp:133-138
template <typename Obj>
void marshal_objs(const rt_object* in, uint32_t n, std::vector<Obj>& out) {
    out.resize(n);
    for (uint32_t i = 0; i < n; ++i) {
        Obj& o = out[i];
        std::memset(&o, 0, sizeof(Obj));
        set3(o.mat.ambient, in[i].mat.ambient);
        set3(o.mat.diffuse, in[i].mat.diffuse);
        set3(o.mat.specular, in[i].mat.specular);
        o.mat.absorption = in[i].mat.absorption;
        o.mat.reflection = in[i].mat.reflection;
        o.mat.transparency = in[i].mat.transparency;
        o.mat.shininess = in[i].mat.shininess;
        set16(o.mv, in[i].mv);
        set16(o.mvInverse, in[i].mvInverse);
        set16t(o.mvInverseTranspose, in[i].mvInverse);
        o.type = in[i].type;
    }
}

// cl_Light(const Light&) OpenCLRaytracer.cpp:141-146
template <typename Lgt>
void marshal_lights(const rt_light* in, uint32_t n, std::vector<Lgt>& out) {
    out.resize(n);
    for (uint32_t i = 0; i < n; ++i) {
        set3(out[i].ambient, in[i].ambient);
        set3(out[i].diffuse, in[i].diffuse);
        set3(out[i].specular, in[i].specular);
        out[i].position = float4{in[i].position[0], in[i].position[1], in[i].position[2], in[i].position[3]};
    }
}

template <typename F>
void parallel_for(uint64_t n, int nthreads, F&& body) {
    if (nthreads <= 0) nthreads = (int)std::thread::hardware_concurrency();
    if (nthreads < 1) nthreads = 1;
    // dynamic chunks small enough that every thread gets work even on a few thousand expensive rays (a fixed chunk of 256
    // left at most 9 threads busy on the bench's 48 x 48 sample of the 104 096-object scene), at most 256 for cheap scenes
    uint64_t chunk = n / ((uint64_t)nthreads * 8u);
    chunk = chunk < 1 ? 1 : (chunk > 256 ? 256 : chunk);
    std::atomic<uint64_t> next{0};
    auto worker = [&]() {
        for (;;) {
            const uint64_t b = next.fetch_add(chunk);
            if (b >= n) return;
            const uint64_t e = (b + chunk < n) ? b + chunk : n;
            for (uint64_t i = b; i < e; ++i) body(i);
        }
    };
    if (nthreads == 1) {
        worker();
        return;
    }
    std::vector<std::thread> pool;
    for (int t = 0; t < nthreads; ++t) pool.emplace_back(worker);
    for (auto& t : pool) t.join();
}

}  // namespace

extern "C" {

int sar_ref_sizeof_object(void) { return (int)sizeof(sar::ObjectData); }

// kernel: 0 hittest, 1 shade, 2 shade_and_reflect (same numbering as rt_mode).
// out_rgba: n_rays float4; hittest writes (t,t,t,1) for hit rays.
int sar_ref_render(int kernel, uint32_t max_bounces, const rt_object* objs, uint32_t n_objs, const rt_light* lights,
                   uint32_t n_lights, const rt_ray* rays, uint64_t n_rays, float* out_rgba, int nthreads) {
    static_assert(sizeof(sar::Ray) == sizeof(rt_ray), "ray layout");
    static_assert(sizeof(sar::ObjectData) == 320, "OpenCL ObjectData is 320 bytes (OpenCLRaytracer.hpp:41-50)");
    static_assert(sizeof(sar::Light) == 64, "OpenCL Light is 64 bytes");
    static_assert(sizeof(float3) == 16, "float3 stride");

    for (uint64_t i = 0; i < n_rays; ++i) {  // OpenCLRaytracer.cpp:32
        out_rgba[4 * i + 0] = 0.f;
        out_rgba[4 * i + 1] = 0.f;
        out_rgba[4 * i + 2] = 0.f;
        out_rgba[4 * i + 3] = 1.f;
    }
    float3* pixels = reinterpret_cast<float3*>(out_rgba);

    if (kernel == 2) {
        std::vector<sar::ObjectData> o;
        std::vector<sar::Light> l;
        marshal_objs(objs, n_objs, o);
        marshal_lights(lights, n_lights, l);
        const sar::Ray* r = reinterpret_cast<const sar::Ray*>(rays);
        parallel_for(n_rays, nthreads, [&](uint64_t i) {
            cl_shim_global_id = (long)i;
            sar::shade_and_reflect(max_bounces, n_objs, o.data(), n_lights, l.data(), r, pixels);
        });
    } else if (kernel == 1) {
        std::vector<shd::ObjectData> o;
        std::vector<shd::Light> l;
        marshal_objs(objs, n_objs, o);
        marshal_lights(lights, n_lights, l);
        const shd::Ray* r = reinterpret_cast<const shd::Ray*>(rays);
        parallel_for(n_rays, nthreads, [&](uint64_t i) {
            cl_shim_global_id = (long)i;
            shd::shade(max_bounces, n_objs, o.data(), n_lights, l.data(), r, pixels);
        });
    } else if (kernel == 0) {
        std::vector<hts::ObjectData> o;
        marshal_objs(objs, n_objs, o);
        const hts::Ray* r = reinterpret_cast<const hts::Ray*>(rays);
        std::vector<float> hits(n_rays, -1.0f);
        parallel_for(n_rays, nthreads, [&](uint64_t i) {
            cl_shim_global_id = (long)i;
            hts::hittest(n_objs, o.data(), r, hits.data());
        });
        for (uint64_t i = 0; i < n_rays; ++i)
            if (hits[i] >= 0.f) out_rgba[4 * i + 0] = out_rgba[4 * i + 1] = out_rgba[4 * i + 2] = hits[i];
    } else {
        return -1;
    }
    // the float3 store leaves .w of the 16-byte slot implementation-defined; normalise to 1
    for (uint64_t i = 0; i < n_rays; ++i) out_rgba[4 * i + 3] = 1.f;
    return 0;
}

}  // extern "C"
