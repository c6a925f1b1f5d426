// rt_cli -- headless driver that replaces the reference's main() (OpenCL-Raytracer.cpp:28-104): same scene loading
// (SceneLoader, unchanged), same ray generation loop (:68-72), the tracer swapped for CudaRaytracer, the GLFW window
// (OpenGLView, a graphics API: unavailable on this box) swapped for a PPM file written by the reference's own
// PPMExporter::ExportP3 -- or by a device-quantised bulk writer that is byte-identical to it (--fast-ppm).
// The flags are the ones the reference leaves as TODOs (OpenCL-Raytracer.cpp:29,34).
#include <glm/glm.hpp>
#include <glm/gtc/matrix_transform.hpp>

#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <string>
#include <vector>

#include "CudaRaytracer.hpp"
#include "Light.hpp"
#include "ObjectData.hpp"
#include "PPMExporter.hpp"
#include "Ray3D.hpp"
#include "SceneLoader.hpp"

// OpenCL-Raytracer.cpp:18-26, verbatim semantics: aspect is computed and unused there too.
static Ray3D screenSpaceToViewSpace(float width, float height, glm::vec2 pos, float angle) {
    const float halfWidth = width / 2.0f;
    const float halfHeight = height / 2.0f;
    return Ray3D(glm::vec3(0, 0, 0), glm::vec3(pos.x - halfWidth, pos.y - halfHeight, -(halfHeight / tan(angle))));
}

static void usage() {
    std::puts("usage: rt_cli --scene file.txt [--width 2560] [--height 1440] [--fov 60] [--bounces 30]\n"
              "              [--mode hittest|shade|sar] [--accel auto|brute|filter|bvh] [--frames 1]\n"
              "              [--gpus 1] [--out render.ppm] [--fast-ppm] [--upload-rays]");
}

int main(int argc, char** argv) {
    int width = 2560, height = 1440;  // OpenCL-Raytracer.cpp:31
    float fovDeg = 60.f;              // :32
    unsigned bounces = 30;            // :75
    int frames = 1, gpus = 1;
    std::string scene, out = "render.ppm";  // :35
    rt_mode mode = RT_MODE_SHADE_AND_REFLECT;
    rt_accel accel = RT_ACCEL_AUTO;
    bool fastPpm = false, uploadRays = false;
    for (int i = 1; i < argc; ++i) {
        const std::string a = argv[i];
        auto next = [&]() -> const char* {
            if (i + 1 >= argc) {
                usage();
                std::exit(2);
            }
            return argv[++i];
        };
        if (a == "--scene") scene = next();
        else if (a == "--width") width = std::atoi(next());
        else if (a == "--height") height = std::atoi(next());
        else if (a == "--fov") fovDeg = (float)std::atof(next());
        else if (a == "--bounces") bounces = (unsigned)std::atoi(next());
        else if (a == "--frames") frames = std::atoi(next());
        else if (a == "--gpus") gpus = std::atoi(next());
        else if (a == "--out") out = next();
        else if (a == "--fast-ppm") fastPpm = true;
        else if (a == "--upload-rays") uploadRays = true;
        else if (a == "--mode") {
            const std::string m = next();
            mode = m == "hittest" ? RT_MODE_HITTEST : m == "shade" ? RT_MODE_SHADE : RT_MODE_SHADE_AND_REFLECT;
        } else if (a == "--accel") {
            const std::string m = next();
            accel = m == "brute" ? RT_ACCEL_BRUTE : m == "filter" ? RT_ACCEL_CULL : m == "bvh" ? RT_ACCEL_BVH : RT_ACCEL_AUTO;
        } else if (scene.empty() && a[0] != '-') {
            scene = a;  // the reference takes the scene as argv[1] (OpenCL-Raytracer.cpp:47)
        } else {
            usage();
            return 2;
        }
    }
    if (scene.empty() || width <= 0 || height <= 0) {
        usage();
        return 2;
    }

    std::vector<ObjectData> objects;
    std::vector<Light> lights;
    try {
        SceneLoader loader;
        loader.Load(scene, objects, lights);
    } catch (const std::exception& err) {
        std::cout << err.what() << std::endl;
        return 1;
    }
    std::cout << "Scene file loaded without any errors.\n";

    float fov = glm::radians(fovDeg);
    fov *= 0.5f;
    std::vector<Ray3D> rays;
    rays.reserve((size_t)width * height);
    for (int jj = 0; jj < height; ++jj)
        for (int ii = 0; ii < width; ++ii)
            rays.emplace_back(screenSpaceToViewSpace((float)width, (float)height, glm::vec2(ii, height - jj), fov));

    try {
        CudaRaytracer tracer(objects, lights, rays, bounces);
        tracer.SetMode(mode);
        tracer.SetAccel(accel);
        if (!uploadRays && tracer.UseCameraIfPinhole((unsigned)width, (unsigned)height)) {
            std::cout << "Primary rays are the pinhole grid: generated on the device.\n";
            if (gpus > 1) std::cout << "Frame sharded over " << tracer.UseDevices((unsigned)gpus) << " GPU(s).\n";
        }
        tracer.SetOption(RT_OPT_COUNT_RAYS, 1);
        IRaytracer* raytracer = &tracer;
        cl_float4* pixelData = nullptr;
        for (int f = 0; f < frames; ++f) {
            const auto t0 = std::chrono::high_resolution_clock::now();
            pixelData = raytracer->Render();
            const auto t1 = std::chrono::high_resolution_clock::now();
            const rt_stats st = tracer.Stats();
            const double ms = std::chrono::duration<double, std::milli>(t1 - t0).count();
            const unsigned long long traced = st.rays_primary + st.traced_shadow + st.traced_reflect;
            std::printf("Frame finished in %.3f ms (kernels %.3f ms, %u launches); rays traced %llu (reference %llu): %.1f Mrays/s\n",
                        ms, st.kernel_ms, st.kernel_launches, traced, (unsigned long long)(st.rays_primary + st.rays_shadow + st.rays_reflect),
                        traced / (ms * 1e-3) / 1e6);
        }
        if (!out.empty() && out != "none") {
            const size_t n = (size_t)width * height;
            if (fastPpm) {
                // PPMExporter.cpp:19-21 quantisation on the device, one bulk write instead of a flush per pixel
                std::vector<int32_t> rgb(3 * n);
                rt_ctx* q = nullptr;
                if (rt_create(&q, -1) != RT_OK || rt_quantize_rgb8(q, reinterpret_cast<const float*>(pixelData), n, rgb.data()) != RT_OK)
                    throw std::runtime_error(rt_last_error());
                rt_destroy(q);
                std::string text = "P3\n" + std::to_string(width) + " " + std::to_string(height) + "\n255\n";
                text.reserve(text.size() + n * 12);
                char buf[48];
                for (size_t i = 0; i < n; ++i) {
                    const int len = std::snprintf(buf, sizeof buf, "%d %d %d\n", rgb[3 * i], rgb[3 * i + 1], rgb[3 * i + 2]);
                    text.append(buf, (size_t)len);
                }
                FILE* fp = std::fopen(out.c_str(), "wb");
                if (!fp) throw std::runtime_error("cannot open " + out);
                std::fwrite(text.data(), 1, text.size(), fp);
                std::fclose(fp);
            } else {
                // float4 -> 3 floats per pixel adapter (PPMExporter.cpp:19-21 reads 3, Render() returns 4)
                std::vector<float> rgb(3 * n);
                for (size_t i = 0; i < n; ++i) {
                    rgb[3 * i] = pixelData[i].s[0];
                    rgb[3 * i + 1] = pixelData[i].s[1];
                    rgb[3 * i + 2] = pixelData[i].s[2];
                }
                PPMExporter::ExportP3(out, (size_t)width, (size_t)height, rgb);
            }
        }
    } catch (const std::exception& err) {
        std::cout << err.what() << std::endl;
        return 1;
    }
    return 0;
}
