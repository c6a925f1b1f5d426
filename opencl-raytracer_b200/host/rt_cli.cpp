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
#include <algorithm>
#include <string>
#include <thread>
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
              "              [--mode hittest|shade|sar] [--accel auto|brute|filter|bvh|grid] [--frames 1]\n"
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
            accel = m == "brute" ? RT_ACCEL_BRUTE : m == "filter" ? RT_ACCEL_CULL : m == "bvh" ? RT_ACCEL_BVH : m == "grid" ? RT_ACCEL_GRID : RT_ACCEL_AUTO;
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
        CudaRaytracer tracer(objects, lights, rays, bounces);  // <- the one line that differs from the reference's main()
        tracer.SetMode(mode);
        tracer.SetAccel(accel);
        if (uploadRays) {  // force the generic path: the ray vector is uploaded and traced as arbitrary rays
            tracer.SetOption(RT_OPT_DETECT_PINHOLE, 0);
            tracer.ReuploadRays();
        }
        if (tracer.IsPinhole()) {
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
            std::vector<unsigned char> rgb8;
            if (fastPpm && tracer.QuantizeRgb8(rgb8) != 0) {
                std::cout << "Negative colour channels: falling back to PPMExporter::ExportP3.\n";
                fastPpm = false;
            }
            if (fastPpm) {
                // PPMExporter.cpp:19-21 quantisation ran on the device (3 bytes per pixel came back); the text is formatted by all
                // host threads, a row range each, and written in one piece instead of a flush per pixel (PPMExporter.cpp:21)
                static char lut[256][4];
                static int lutLen[256];
                for (int v = 0; v < 256; ++v) lutLen[v] = std::snprintf(lut[v], 4, "%d", v);
                const unsigned nt = std::max(1u, std::min(std::thread::hardware_concurrency(), 32u));
                std::vector<std::string> part(nt);
                std::vector<std::thread> pool;
                for (unsigned t = 0; t < nt; ++t)
                    pool.emplace_back([&, t] {
                        const size_t b = n * t / nt, e = n * (t + 1) / nt;
                        std::string& txt = part[t];
                        txt.reserve((e - b) * 12);
                        for (size_t i = b; i < e; ++i) {
                            const unsigned char* px = &rgb8[3 * i];
                            txt.append(lut[px[0]], (size_t)lutLen[px[0]]);
                            txt.push_back(' ');
                            txt.append(lut[px[1]], (size_t)lutLen[px[1]]);
                            txt.push_back(' ');
                            txt.append(lut[px[2]], (size_t)lutLen[px[2]]);
                            txt.push_back('\n');
                        }
                    });
                for (std::thread& th : pool) th.join();
                FILE* fp = std::fopen(out.c_str(), "wb");
                if (!fp) throw std::runtime_error("cannot open " + out);
                const std::string head = "P3\n" + std::to_string(width) + " " + std::to_string(height) + "\n255\n";
                std::fwrite(head.data(), 1, head.size(), fp);
                for (const std::string& txt : part) std::fwrite(txt.data(), 1, txt.size(), fp);
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
