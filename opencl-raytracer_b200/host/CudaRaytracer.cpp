// CudaRaytracer.cpp -- see CudaRaytracer.hpp.  Host C++ only; everything CUDA is behind include/rt_b200.h.
#include "CudaRaytracer.hpp"

#include <cstring>
#include <stdexcept>
#include <string>
#include <thread>

#include <glm/gtc/type_ptr.hpp>

namespace {

void check(int rc, const char* what) {
    if (rc != RT_OK) throw std::runtime_error(std::string("CudaRaytracer: ") + what + ": " + rt_last_error());
}

// cl_ObjectData(const ObjectData&) / cl_Material(const Material&), OpenCLRaytracer.cpp:121-138: matrices are copied
// column-major exactly as glm::value_ptr yields them; mvInverseTranspose is not read by any kernel.
rt_object marshal(const ObjectData& o) {
    rt_object r;
    std::memset(&r, 0, sizeof r);
    std::memcpy(r.mat.ambient, glm::value_ptr(o.mat.ambient), 3 * sizeof(float));
    std::memcpy(r.mat.diffuse, glm::value_ptr(o.mat.diffuse), 3 * sizeof(float));
    std::memcpy(r.mat.specular, glm::value_ptr(o.mat.specular), 3 * sizeof(float));
    r.mat.absorption = o.mat.absorption;
    r.mat.reflection = o.mat.reflection;
    r.mat.transparency = o.mat.transparency;
    r.mat.shininess = o.mat.shininess;
    std::memcpy(r.mv, glm::value_ptr(o.mv), 16 * sizeof(float));
    std::memcpy(r.mvInverse, glm::value_ptr(o.mvInverse), 16 * sizeof(float));
    r.type = static_cast<uint32_t>(o.type);
    return r;
}

// cl_Light(const Light&), OpenCLRaytracer.cpp:141-146
rt_light marshal(const Light& l) {
    rt_light r;
    std::memcpy(r.ambient, glm::value_ptr(l.ambient), 3 * sizeof(float));
    std::memcpy(r.diffuse, glm::value_ptr(l.diffuse), 3 * sizeof(float));
    std::memcpy(r.specular, glm::value_ptr(l.specular), 3 * sizeof(float));
    std::memcpy(r.position, glm::value_ptr(l.lightPosition), 4 * sizeof(float));
    return r;
}

}  // namespace

CudaRaytracer::CudaRaytracer(const std::vector<ObjectData>& objects, const std::vector<Light>& lights,
                             const std::vector<Ray3D>& rays, const unsigned int MAX_BOUNCES)
    : IRaytracer(objects, lights, rays), MAX_BOUNCES(MAX_BOUNCES) {
    static_assert(sizeof(Ray3D) == sizeof(rt_ray), "Ray3D is two vec4 (Ray3D.hpp:5-10) == cl_Ray (OpenCLRaytracer.hpp:25-30)");
    static_assert(sizeof(cl_float4) == 4 * sizeof(float), "cl_float4 layout");
    check(rt_create(&ctx_, -1), "rt_create");
    try {
        std::vector<rt_object>& objArr = objArr_;
        objArr.reserve(objects.size());
        for (const ObjectData& o : objects) objArr.push_back(marshal(o));
        std::vector<rt_light>& lightArr = lightArr_;
        lightArr.reserve(lights.size());
        for (const Light& l : lights) lightArr.push_back(marshal(l));
        check(rt_upload_scene(ctx_, objArr.data(), (uint32_t)objArr.size(), lightArr.data(), (uint32_t)lightArr.size()),
              "rt_upload_scene");
        check(rt_set_rays(ctx_, reinterpret_cast<const rt_ray*>(rays.data()), rays.size()), "rt_set_rays");
        if (rt_get_camera(ctx_, &camW_, &camH_, &camZ_) != RT_OK) camW_ = camH_ = 0;  // not the pinhole grid: the rays were uploaded
        void* p = nullptr;
        check(rt_host_alloc(&p, rays.size() * sizeof(cl_float4)), "rt_host_alloc");
        pixelDataArr = static_cast<cl_float4*>(p);
        for (size_t i = 0; i < rays.size(); ++i) {  // OpenCLRaytracer.cpp:32
            pixelDataArr[i].s[0] = pixelDataArr[i].s[1] = pixelDataArr[i].s[2] = 0.f;
            pixelDataArr[i].s[3] = 1.f;
        }
    } catch (...) {
        if (pixelDataArr) rt_host_free(pixelDataArr);
        rt_destroy(ctx_);
        throw;
    }
}

CudaRaytracer::~CudaRaytracer() {
    for (rt_ctx* c : extra_) rt_destroy(c);
    if (pixelDataArr) rt_host_free(pixelDataArr);
    if (ctx_) rt_destroy(ctx_);
}

cl_float4* CudaRaytracer::Render() {
    float* out = reinterpret_cast<float*>(pixelDataArr);
    if (extra_.empty()) {
        check(rt_render(ctx_, MAX_BOUNCES, mode_, accel_, out), "rt_render");
        return pixelDataArr;
    }
    // one host thread per device: each rt_render traces its own row-blocks and copies them into the shared pinned frame
    std::vector<int> rc(extra_.size() + 1, RT_OK);
    std::vector<std::string> msg(extra_.size() + 1);
    std::vector<std::thread> pool;
    auto run = [&](size_t i) {
        rt_ctx* c = i == 0 ? ctx_ : extra_[i - 1];
        rc[i] = rt_render(c, MAX_BOUNCES, mode_, accel_, out);
        if (rc[i] != RT_OK) msg[i] = rt_last_error();  // the message is thread-local
    };
    for (size_t i = 1; i <= extra_.size(); ++i) pool.emplace_back(run, i);
    run(0);
    for (std::thread& t : pool) t.join();
    for (size_t i = 0; i < rc.size(); ++i)
        if (rc[i] != RT_OK) throw std::runtime_error("CudaRaytracer: rt_render on device " + std::to_string(i) + ": " + msg[i]);
    return pixelDataArr;
}

void CudaRaytracer::SetOption(rt_option option, long long value) {
    check(rt_set_option(ctx_, option, value), "rt_set_option");
    for (rt_ctx* c : extra_) check(rt_set_option(c, option, value), "rt_set_option");
}

unsigned CudaRaytracer::UseDevices(unsigned n) {
    if (camW_ == 0) throw std::runtime_error("CudaRaytracer: UseDevices needs the pinhole camera (UseCameraIfPinhole) first");
    int avail = 0;
    check(rt_device_count(&avail), "rt_device_count");
    if (n < 1) n = 1;
    if ((int)n > avail) n = (unsigned)avail;
    for (rt_ctx* c : extra_) rt_destroy(c);
    extra_.clear();
    for (unsigned d = 1; d < n; ++d) {
        rt_ctx* c = nullptr;
        check(rt_create(&c, (int)d), "rt_create");
        extra_.push_back(c);
        check(rt_upload_scene(c, objArr_.data(), (uint32_t)objArr_.size(), lightArr_.data(), (uint32_t)lightArr_.size()), "rt_upload_scene");
        check(rt_set_camera(c, camW_, camH_, camZ_), "rt_set_camera");
        check(rt_set_shard(c, 16, d, n), "rt_set_shard");
    }
    check(rt_set_shard(ctx_, n > 1 ? 16 : 1, 0, n), "rt_set_shard");
    return n;
}

void CudaRaytracer::ReuploadRays() {
    if (!extra_.empty()) UseDevices(1);
    check(rt_set_rays(ctx_, reinterpret_cast<const rt_ray*>(rays.data()), rays.size()), "rt_set_rays");
    if (rt_get_camera(ctx_, &camW_, &camH_, &camZ_) != RT_OK) camW_ = camH_ = 0;
}

bool CudaRaytracer::UseCameraIfPinhole(unsigned width, unsigned height) {
    if ((size_t)width * height != rays.size() || rays.empty()) return false;
    const float zdir = rays[0].direction.z;
    const float halfW = (float)width / 2.0f, halfH = (float)height / 2.0f;
    for (unsigned jj = 0; jj < height; ++jj)
        for (unsigned ii = 0; ii < width; ++ii) {
            const Ray3D& r = rays[(size_t)jj * width + ii];
            const float px = (float)ii, py = (float)((int)height - (int)jj);
            if (r.start.x != 0.f || r.start.y != 0.f || r.start.z != 0.f || r.start.w != 1.f || r.direction.w != 0.f ||
                r.direction.x != px - halfW || r.direction.y != py - halfH || r.direction.z != zdir)
                return false;
        }
    check(rt_set_camera(ctx_, width, height, zdir), "rt_set_camera");
    camW_ = width;
    camH_ = height;
    camZ_ = zdir;
    return true;
}

unsigned long long CudaRaytracer::QuantizeRgb8(std::vector<unsigned char>& rgb) const {
    rgb.resize(3 * rays.size());
    uint64_t neg = 0;
    // one device: its framebuffer holds the frame; several: the pieces are on different devices, the host frame is quantised
    const float* host = extra_.empty() ? nullptr : reinterpret_cast<const float*>(pixelDataArr);
    check(rt_quantize_rgb8(ctx_, host, rays.size(), rgb.data(), &neg), "rt_quantize_rgb8");
    return neg;
}

rt_stats CudaRaytracer::Stats() const {
    rt_stats st;
    check(rt_get_stats(ctx_, &st), "rt_get_stats");
    for (rt_ctx* c : extra_) {
        rt_stats s;
        check(rt_get_stats(c, &s), "rt_get_stats");
        st.rays_primary += s.rays_primary; st.rays_shadow += s.rays_shadow; st.rays_reflect += s.rays_reflect;
        st.traced_shadow += s.traced_shadow; st.traced_reflect += s.traced_reflect;
        st.pair_tests += s.pair_tests; st.exact_tests += s.exact_tests; st.node_visits += s.node_visits;
        st.pixels += s.pixels; st.hit_pixels += s.hit_pixels;
        st.kernel_launches += s.kernel_launches;
        if (s.kernel_ms > st.kernel_ms) st.kernel_ms = s.kernel_ms;
        if (s.total_ms > st.total_ms) st.total_ms = s.total_ms;
    }
    return st;
}
