// CudaRaytracer -- drop-in backend behind the reference's IRaytracer interface (IRaytracer.hpp:10-21),
// next to OpenCLRaytracer (OpenCLRaytracer.hpp:24-89).  Same constructor signature as
// OpenCLRaytracer.hpp:61, same ownership rules as OpenCLRaytracer.cpp:13-105:
//   * objects / lights / rays are borrowed references (the caller keeps them alive);
//   * the scene and the rays are uploaded once, in the constructor (OpenCLRaytracer.cpp:70-73);
//   * Render() launches, reads the frame back synchronously and returns a pointer to a buffer this object
//     owns: rays.size() x cl_float4, index = ray index, xyz = unclamped linear RGB, w = 1; stable across
//     calls, overwritten by each call, never freed by the caller.
// Differences: the buffer is pinned host memory and is released in the destructor (the reference leaks it,
// OpenCLRaytracer.cpp:76-78); errors surface as std::runtime_error carrying rt_last_error() instead of
// boost::compute::opencl_error; there is no run-time kernel source and no CPU fallback.
//
// Reused unchanged from the reference: IRaytracer.hpp, ObjectData.hpp, Material.hpp, Light.hpp, Ray3D.hpp,
// HitRecord.hpp (put the reference directory on the include path).  This class calls nothing but the
// extern "C" layer of include/rt_b200.h.
#pragma once

#include <string>
#include <vector>

#include "IRaytracer.hpp"
#include "rt_b200.h"

class CudaRaytracer : public IRaytracer {
public:
    CudaRaytracer(const std::vector<ObjectData>& objects, const std::vector<Light>& lights, const std::vector<Ray3D>& rays,
                  const unsigned int MAX_BOUNCES);
    ~CudaRaytracer();
    CudaRaytracer(const CudaRaytracer&) = delete;
    CudaRaytracer& operator=(const CudaRaytracer&) = delete;

    // Inherited via IRaytracer
    cl_float4* Render() override;

    // Optional knobs the reference leaves as TODOs (OpenCL-Raytracer.cpp:29,34); defaults reproduce it.
    void SetMode(rt_mode mode) { mode_ = mode; }        // default RT_MODE_SHADE_AND_REFLECT (the live kernel)
    void SetAccel(rt_accel accel) { accel_ = accel; }   // default RT_ACCEL_AUTO
    void SetOption(rt_option option, long long value);
    // The constructor hands `rays` to rt_set_rays, which recognises the pinhole grid main() builds (OpenCL-Raytracer.cpp:68-72)
    // and then generates the rays on the device (bit-identical; nothing uploaded; camera grid, bands and lanes apply): the
    // one-line integration is the fast path.  IsPinhole() tells whether that happened; UseCameraIfPinhole() is the explicit
    // form (e.g. after SetOption(RT_OPT_DETECT_PINHOLE, 0) was used to force the upload).
    bool IsPinhole() const { return camW_ != 0; }
    bool UseCameraIfPinhole(unsigned width, unsigned height);
    void ReuploadRays();  // hand `rays` to rt_set_rays again (after changing RT_OPT_DETECT_PINHOLE)
    // PPMExporter's quantisation (PPMExporter.cpp:19-21) of the frame of the last Render(), on the device where that frame
    // still is (one device) -- 3 bytes per pixel come back instead of 16.  Returns the number of negative channels (clamped to
    // 0; the exporter would print them as negative numbers).
    unsigned long long QuantizeRgb8(std::vector<unsigned char>& rgb) const;
    // Shard the camera frame over the first n CUDA devices of this process (row-blocks of 16 rows dealt round-robin,
    // SURVEY.md 8(e)): one context and one host thread per device, every device renders its rows and copies them into
    // the one pinned frame Render() returns.  Needs the pinhole camera (IsPinhole()); n = 1 undoes it.  Returns the number of
    // devices actually used (at most rt_device_count()).  The reference has no multi-device path.
    unsigned UseDevices(unsigned n);
    rt_stats Stats() const;  // ray counts summed over the devices, times of the slowest device
    size_t PixelCount() const { return rays.size(); }

private:
    rt_ctx* ctx_ = nullptr;
    std::vector<rt_ctx*> extra_;  // contexts of devices 1 .. n-1 (UseDevices)
    std::vector<rt_object> objArr_;  // marshalled scene, kept for UseDevices
    std::vector<rt_light> lightArr_;
    uint32_t camW_ = 0, camH_ = 0;
    float camZ_ = 0.f;
    cl_float4* pixelDataArr = nullptr;  // pinned, rays.size() entries
    const unsigned int MAX_BOUNCES;
    rt_mode mode_ = RT_MODE_SHADE_AND_REFLECT;
    rt_accel accel_ = RT_ACCEL_AUTO;
};
