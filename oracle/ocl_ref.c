// TEST / BASELINE INFRASTRUCTURE -- not part of the product path.
//
// Runs the reference's OWN OpenCL kernel source, unchanged, on whatever OpenCL platform the box offers
// (SURVEY.md 8(f) N3).  This is the host side of OpenCLRaytracer (OpenCLRaytracer.cpp:13-105: buffers, JIT build
// from source, 7 kernel arguments, 4 uploads, 1-D NDRange with local size 32, blocking read) written against the
// raw OpenCL C API instead of Boost.Compute.  Neither CL/cl.h nor an ICD registry exists in this image, so the
// entry points are declared by hand and resolved at run time:
//   1. dlopen an OpenCL library (OCL_REF_LIB, libOpenCL.so.1, CUDA's copy of it, libnvidia-opencl.so.1);
//   2. find a platform through clGetPlatformIDs or, for a bare vendor ICD, clIcdGetPlatformIDsKHR
//      (clGetExtensionFunctionAddress);
//   3. call everything else through the platform object's ICD dispatch table (cl_khr_icd: every OpenCL object
//      starts with a pointer to it; the slot order below is that of the Khronos cl_icd.h, OpenCL 1.0 block).
// OpenCL is a compute API; nothing graphical is touched.  If no platform enumerates, ocl_ref_probe() says so and
// callers record "unavailable on this box".
#define _GNU_SOURCE
#include <dlfcn.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

typedef int32_t cl_int;
typedef uint32_t cl_uint;
typedef uint64_t cl_ulong;
typedef cl_ulong cl_bitfield;
typedef void* cl_obj;

enum {  // cl_icd.h dispatch slots (OpenCL 1.0)
    S_GetPlatformIDs = 0, S_GetPlatformInfo = 1, S_GetDeviceIDs = 2, S_GetDeviceInfo = 3, S_CreateContext = 4,
    S_ReleaseContext = 7, S_CreateCommandQueue = 9, S_ReleaseCommandQueue = 11, S_CreateBuffer = 14,
    S_ReleaseMemObject = 18, S_CreateProgramWithSource = 26, S_ReleaseProgram = 29, S_BuildProgram = 30,
    S_GetProgramBuildInfo = 33, S_CreateKernel = 34, S_ReleaseKernel = 37, S_SetKernelArg = 38, S_ReleaseEvent = 44,
    S_GetEventProfilingInfo = 45, S_Finish = 47, S_EnqueueReadBuffer = 48, S_EnqueueWriteBuffer = 49,
    S_EnqueueNDRangeKernel = 59
};

typedef cl_int (*fn_GetPlatformIDs)(cl_uint, cl_obj*, cl_uint*);
typedef cl_int (*fn_GetPlatformInfo)(cl_obj, cl_uint, size_t, void*, size_t*);
typedef cl_int (*fn_GetDeviceIDs)(cl_obj, cl_bitfield, cl_uint, cl_obj*, cl_uint*);
typedef cl_int (*fn_GetDeviceInfo)(cl_obj, cl_uint, size_t, void*, size_t*);
typedef cl_obj (*fn_CreateContext)(const intptr_t*, cl_uint, const cl_obj*, void*, void*, cl_int*);
typedef cl_int (*fn_Release)(cl_obj);
typedef cl_obj (*fn_CreateCommandQueue)(cl_obj, cl_obj, cl_bitfield, cl_int*);
typedef cl_obj (*fn_CreateBuffer)(cl_obj, cl_bitfield, size_t, void*, cl_int*);
typedef cl_obj (*fn_CreateProgramWithSource)(cl_obj, cl_uint, const char**, const size_t*, cl_int*);
typedef cl_int (*fn_BuildProgram)(cl_obj, cl_uint, const cl_obj*, const char*, void*, void*);
typedef cl_int (*fn_GetProgramBuildInfo)(cl_obj, cl_obj, cl_uint, size_t, void*, size_t*);
typedef cl_obj (*fn_CreateKernel)(cl_obj, const char*, cl_int*);
typedef cl_int (*fn_SetKernelArg)(cl_obj, cl_uint, size_t, const void*);
typedef cl_int (*fn_GetEventProfilingInfo)(cl_obj, cl_uint, size_t, void*, size_t*);
typedef cl_int (*fn_Finish)(cl_obj);
typedef cl_int (*fn_EnqueueRW)(cl_obj, cl_obj, cl_uint, size_t, size_t, void*, cl_uint, const cl_obj*, cl_obj*);
typedef cl_int (*fn_EnqueueNDRange)(cl_obj, cl_obj, cl_uint, const size_t*, const size_t*, const size_t*, cl_uint, const cl_obj*, cl_obj*);
typedef void* (*fn_GetExtensionFunctionAddress)(const char*);

#define CL_DEVICE_TYPE_GPU 4ull
#define CL_MEM_READ_ONLY 4ull
#define CL_MEM_WRITE_ONLY 2ull
#define CL_QUEUE_PROFILING_ENABLE 2ull
#define CL_PLATFORM_VERSION 0x0901
#define CL_PLATFORM_NAME 0x0902
#define CL_DEVICE_NAME 0x102B
#define CL_PROGRAM_BUILD_LOG 0x1183
#define CL_PROFILING_COMMAND_START 0x1282
#define CL_PROFILING_COMMAND_END 0x1283

static struct {
    int tried, ok;
    void* lib;
    cl_obj platform, device;
    void** disp;
    char info[512];
} G;

#define D(slot, type) ((type)G.disp[slot])

static double now_ms(void) {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6;
}

static int try_library(const char* name, char* why, size_t cap) {
    void* lib = dlopen(name, RTLD_NOW | RTLD_LOCAL);
    if (!lib) {
        snprintf(why + strlen(why), cap - strlen(why), "[%s: %s] ", name, "dlopen failed");
        return 0;
    }
    cl_obj plats[8];
    cl_uint n = 0;
    cl_int rc = -1;
    fn_GetPlatformIDs gp = (fn_GetPlatformIDs)dlsym(lib, "clGetPlatformIDs");
    if (gp) rc = gp(8, plats, &n);
    if (rc != 0 || n == 0) {
        fn_GetExtensionFunctionAddress ge = (fn_GetExtensionFunctionAddress)dlsym(lib, "clGetExtensionFunctionAddress");
        fn_GetPlatformIDs icd = ge ? (fn_GetPlatformIDs)ge("clIcdGetPlatformIDsKHR") : NULL;
        n = 0;
        if (icd) rc = icd(8, plats, &n);
    }
    if (rc != 0 || n == 0) {
        snprintf(why + strlen(why), cap - strlen(why), "[%s: no platform, rc %d] ", name, (int)rc);
        dlclose(lib);
        return 0;
    }
    for (cl_uint p = 0; p < n; ++p) {
        void** disp = *(void***)plats[p];
        if (!disp) continue;
        cl_obj dev = NULL;
        cl_uint nd = 0;
        if (((fn_GetDeviceIDs)disp[S_GetDeviceIDs])(plats[p], CL_DEVICE_TYPE_GPU, 1, &dev, &nd) == 0 && nd > 0) {
            G.lib = lib;
            G.platform = plats[p];
            G.device = dev;
            G.disp = disp;
            char pn[128] = "", pv[128] = "", dn[128] = "";
            D(S_GetPlatformInfo, fn_GetPlatformInfo)(G.platform, CL_PLATFORM_NAME, sizeof pn, pn, NULL);
            D(S_GetPlatformInfo, fn_GetPlatformInfo)(G.platform, CL_PLATFORM_VERSION, sizeof pv, pv, NULL);
            D(S_GetDeviceInfo, fn_GetDeviceInfo)(G.device, CL_DEVICE_NAME, sizeof dn, dn, NULL);
            snprintf(G.info, sizeof G.info, "%s | %s | %s | via %s", pn, pv, dn, name);
            return 1;
        }
    }
    snprintf(why + strlen(why), cap - strlen(why), "[%s: platform without a GPU device] ", name);
    dlclose(lib);
    return 0;
}

static void init_once(void) {
    if (G.tried) return;
    G.tried = 1;
    char why[400] = "";
    // let an ICD loader find the vendor library without an /etc/OpenCL/vendors registry
    setenv("OCL_ICD_FILENAMES", "libnvidia-opencl.so.1", 0);
    const char* cands[6] = {getenv("OCL_REF_LIB"), "libOpenCL.so.1", "/usr/local/cuda/lib64/libOpenCL.so.1",
                            "libnvidia-opencl.so.1", "libnvidia-opencl.so", NULL};
    for (int i = 0; i < 5; ++i)
        if (cands[i] && *cands[i] && try_library(cands[i], why, sizeof why)) {
            G.ok = 1;
            return;
        }
    snprintf(G.info, sizeof G.info, "no OpenCL GPU platform: %s", why);
}

// 0: a GPU platform is available (info = platform | version | device | library); 1: unavailable (info = why)
int ocl_ref_probe(char* info, int cap) {
    init_once();
    if (info && cap > 0) snprintf(info, (size_t)cap, "%s", G.info);
    return G.ok ? 0 : 1;
}

// OpenCL device records (OpenCLRaytracer.hpp:25-58; shade_and_reflect_kernel.cl:1-29)
typedef struct {
    float ambient[4], diffuse[4], specular[4];
    float absorption, reflection, transparency, shininess;
} ocl_material;  // 64 B
typedef struct {
    ocl_material mat;
    float mv[16], mvInverse[16], mvInverseTranspose[16];
    uint32_t type;
    uint8_t spacer[60];
} ocl_object;  // 320 B
typedef struct {
    float ambient[4], diffuse[4], specular[4], position[4];
} ocl_light;  // 64 B

#include "../include/rt_b200.h"

static void marshal(const rt_object* in, uint32_t n, ocl_object* out) {  // OpenCLRaytracer.cpp:108-138
    memset(out, 0, (size_t)n * sizeof *out);
    for (uint32_t i = 0; i < n; ++i) {
        for (int k = 0; k < 3; ++k) {
            out[i].mat.ambient[k] = in[i].mat.ambient[k];
            out[i].mat.diffuse[k] = in[i].mat.diffuse[k];
            out[i].mat.specular[k] = in[i].mat.specular[k];
        }
        out[i].mat.absorption = in[i].mat.absorption;
        out[i].mat.reflection = in[i].mat.reflection;
        out[i].mat.transparency = in[i].mat.transparency;
        out[i].mat.shininess = in[i].mat.shininess;
        memcpy(out[i].mv, in[i].mv, 64);
        memcpy(out[i].mvInverse, in[i].mvInverse, 64);
        for (int r = 0; r < 4; ++r)
            for (int c = 0; c < 4; ++c) out[i].mvInverseTranspose[4 * c + r] = in[i].mvInverse[4 * r + c];
        out[i].type = in[i].type;
    }
}

#define CHECK(cond, ...)                               \
    do {                                               \
        if (!(cond)) {                                 \
            if (log && logcap > 0) snprintf(log, (size_t)logcap, __VA_ARGS__); \
            rc_out = -1;                               \
            goto done;                                 \
        }                                              \
    } while (0)

// Constructor + `frames` x Render() of OpenCLRaytracer for kernel `kernel_name` of the source file `cl_path`.
// out_rgba: n_rays float4 (pre-filled {0,0,0,1} like OpenCLRaytracer.cpp:32; a missing work-item writes nothing).
// kernel_ms: mean device time of the NDRange (event profiling); render_ms: mean wall time of launch + blocking read;
// build_ms: JIT build.  n_rays is padded up to a multiple of the reference's local size 32 with copies of ray 0.
int ocl_ref_render(const char* cl_path, const char* kernel_name, uint32_t max_bounces, const rt_object* objs, uint32_t n_objs,
                   const rt_light* lights, uint32_t n_lights, const rt_ray* rays, uint64_t n_rays, int frames, float* out_rgba,
                   double* kernel_ms, double* render_ms, double* build_ms, char* log, int logcap) {
    init_once();
    int rc_out = 0;
    cl_int e = 0;
    cl_obj ctx = NULL, q = NULL, prog = NULL, ker = NULL, b_obj = NULL, b_light = NULL, b_ray = NULL, b_pix = NULL;
    ocl_object* o = NULL;
    ocl_light* l = NULL;
    rt_ray* r = NULL;
    float* pix = NULL;
    char* src = NULL;
    if (log && logcap > 0) log[0] = 0;
    CHECK(G.ok, "%s", G.info);
    CHECK(n_rays > 0 && frames > 0, "nothing to do");
    {
        FILE* f = fopen(cl_path, "rb");
        CHECK(f != NULL, "cannot open %s", cl_path);
        fseek(f, 0, SEEK_END);
        long sz = ftell(f);
        fseek(f, 0, SEEK_SET);
        src = (char*)malloc((size_t)sz + 1);
        size_t got = fread(src, 1, (size_t)sz, f);
        fclose(f);
        src[got] = 0;
    }
    const uint64_t n_pad = (n_rays + 31u) & ~31ull;
    o = (ocl_object*)malloc((size_t)(n_objs ? n_objs : 1) * sizeof *o);
    l = (ocl_light*)calloc(n_lights ? n_lights : 1, sizeof *l);
    r = (rt_ray*)malloc(n_pad * sizeof *r);
    pix = (float*)malloc(n_pad * 16);
    marshal(objs, n_objs, o);
    for (uint32_t i = 0; i < n_lights; ++i) {
        memcpy(l[i].ambient, lights[i].ambient, 12);
        memcpy(l[i].diffuse, lights[i].diffuse, 12);
        memcpy(l[i].specular, lights[i].specular, 12);
        memcpy(l[i].position, lights[i].position, 16);
    }
    memcpy(r, rays, n_rays * sizeof *r);
    for (uint64_t i = n_rays; i < n_pad; ++i) r[i] = rays[0];
    for (uint64_t i = 0; i < n_pad; ++i) {
        pix[4 * i] = pix[4 * i + 1] = pix[4 * i + 2] = 0.f;
        pix[4 * i + 3] = 1.f;
    }

    ctx = D(S_CreateContext, fn_CreateContext)(NULL, 1, &G.device, NULL, NULL, &e);
    CHECK(e == 0 && ctx, "clCreateContext: %d", (int)e);
    q = D(S_CreateCommandQueue, fn_CreateCommandQueue)(ctx, G.device, CL_QUEUE_PROFILING_ENABLE, &e);
    CHECK(e == 0 && q, "clCreateCommandQueue: %d", (int)e);
    const size_t sz_obj = (size_t)(n_objs ? n_objs : 1) * sizeof *o, sz_light = (size_t)(n_lights ? n_lights : 1) * sizeof *l;
    b_obj = D(S_CreateBuffer, fn_CreateBuffer)(ctx, CL_MEM_READ_ONLY, sz_obj, NULL, &e);
    CHECK(e == 0, "clCreateBuffer objs: %d", (int)e);
    b_light = D(S_CreateBuffer, fn_CreateBuffer)(ctx, CL_MEM_READ_ONLY, sz_light, NULL, &e);
    CHECK(e == 0, "clCreateBuffer lights: %d", (int)e);
    b_ray = D(S_CreateBuffer, fn_CreateBuffer)(ctx, CL_MEM_READ_ONLY, n_pad * sizeof *r, NULL, &e);
    CHECK(e == 0, "clCreateBuffer rays: %d", (int)e);
    b_pix = D(S_CreateBuffer, fn_CreateBuffer)(ctx, CL_MEM_WRITE_ONLY, n_pad * 16, NULL, &e);
    CHECK(e == 0, "clCreateBuffer pixels: %d", (int)e);

    {
        const char* srcs[1] = {src};
        const double t0 = now_ms();
        prog = D(S_CreateProgramWithSource, fn_CreateProgramWithSource)(ctx, 1, srcs, NULL, &e);
        CHECK(e == 0 && prog, "clCreateProgramWithSource: %d", (int)e);
        e = D(S_BuildProgram, fn_BuildProgram)(prog, 1, &G.device, NULL, NULL, NULL);  // program.build(): no options
        if (e != 0) {
            char bl[1500] = "";
            D(S_GetProgramBuildInfo, fn_GetProgramBuildInfo)(prog, G.device, CL_PROGRAM_BUILD_LOG, sizeof bl - 1, bl, NULL);
            CHECK(0, "clBuildProgram: %d: %s", (int)e, bl);
        }
        if (build_ms) *build_ms = now_ms() - t0;
    }
    ker = D(S_CreateKernel, fn_CreateKernel)(prog, kernel_name, &e);
    CHECK(e == 0 && ker, "clCreateKernel(%s): %d", kernel_name, (int)e);
    {
        fn_SetKernelArg sa = D(S_SetKernelArg, fn_SetKernelArg);
        const cl_uint oc = n_objs, lc = n_lights, mb = max_bounces;
        // OpenCLRaytracer.cpp:61-68 (shade_kernel.cl's `shade` takes the same seven arguments)
        e = sa(ker, 0, sizeof mb, &mb) | sa(ker, 1, sizeof oc, &oc) | sa(ker, 2, sizeof(cl_obj), &b_obj) | sa(ker, 3, sizeof lc, &lc) |
            sa(ker, 4, sizeof(cl_obj), &b_light) | sa(ker, 5, sizeof(cl_obj), &b_ray) | sa(ker, 6, sizeof(cl_obj), &b_pix);
        CHECK(e == 0, "clSetKernelArg: %d", (int)e);
    }
    {
        fn_EnqueueRW wr = D(S_EnqueueWriteBuffer, fn_EnqueueRW);
        e = wr(q, b_obj, 1, 0, sz_obj, o, 0, NULL, NULL);
        if (e == 0) e = wr(q, b_light, 1, 0, sz_light, l, 0, NULL, NULL);
        if (e == 0) e = wr(q, b_ray, 1, 0, n_pad * sizeof *r, r, 0, NULL, NULL);
        if (e == 0) e = wr(q, b_pix, 1, 0, n_pad * 16, pix, 0, NULL, NULL);
        CHECK(e == 0, "clEnqueueWriteBuffer: %d", (int)e);
    }
    {
        double k_sum = 0, r_sum = 0;
        for (int f = 0; f < frames; ++f) {  // OpenCLRaytracer::Render(), OpenCLRaytracer.cpp:80-105
            const size_t global = (size_t)n_pad, local = 32;
            cl_obj ev = NULL;
            const double t0 = now_ms();
            e = D(S_EnqueueNDRangeKernel, fn_EnqueueNDRange)(q, ker, 1, NULL, &global, &local, 0, NULL, &ev);
            CHECK(e == 0, "clEnqueueNDRangeKernel: %d", (int)e);
            e = D(S_EnqueueReadBuffer, fn_EnqueueRW)(q, b_pix, 1, 0, n_pad * 16, pix, 0, NULL, NULL);
            CHECK(e == 0, "clEnqueueReadBuffer: %d", (int)e);
            r_sum += now_ms() - t0;
            cl_ulong a = 0, b = 0;
            D(S_GetEventProfilingInfo, fn_GetEventProfilingInfo)(ev, CL_PROFILING_COMMAND_START, sizeof a, &a, NULL);
            D(S_GetEventProfilingInfo, fn_GetEventProfilingInfo)(ev, CL_PROFILING_COMMAND_END, sizeof b, &b, NULL);
            k_sum += (double)(b - a) * 1e-6;
            D(S_ReleaseEvent, fn_Release)(ev);
        }
        if (kernel_ms) *kernel_ms = k_sum / frames;
        if (render_ms) *render_ms = r_sum / frames;
    }
    memcpy(out_rgba, pix, n_rays * 16);
    for (uint64_t i = 0; i < n_rays; ++i) out_rgba[4 * i + 3] = 1.f;  // .w of a float3 store is implementation-defined
done:
    if (ker) D(S_ReleaseKernel, fn_Release)(ker);
    if (prog) D(S_ReleaseProgram, fn_Release)(prog);
    if (b_obj) D(S_ReleaseMemObject, fn_Release)(b_obj);
    if (b_light) D(S_ReleaseMemObject, fn_Release)(b_light);
    if (b_ray) D(S_ReleaseMemObject, fn_Release)(b_ray);
    if (b_pix) D(S_ReleaseMemObject, fn_Release)(b_pix);
    if (q) D(S_ReleaseCommandQueue, fn_Release)(q);
    if (ctx) D(S_ReleaseContext, fn_Release)(ctx);
    free(o);
    free(l);
    free(r);
    free(pix);
    free(src);
    return rc_out;
}
