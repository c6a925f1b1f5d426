/* rt_b200.h -- C ABI of the B200-native per-pixel render path.
 *
 * This is the drop-in boundary for the reference's tracer backend: everything
 * `OpenCLRaytracer` does against Boost.Compute/OpenCL (OpenCLRaytracer.cpp:13-105)
 * is reached through these entry points instead.  Plain pointers and sizes only;
 * no C++/torch types.  The C++ class `CudaRaytracer : public IRaytracer`
 * (opencl-raytracer_b200/host/CudaRaytracer.hpp) and the ctypes binding
 * (opencl-raytracer_b200/pyrt) are thin callers of this header.
 *
 * All functions return RT_OK (0) or a negative rt_status; rt_last_error() gives
 * the message of the last failure on the calling thread.  There is NO CPU
 * fallback: without a CUDA device of compute capability 10.x rt_create fails.
 *
 * Reference interface each entry point replaces (file:line under the reference):
 *   rt_create / rt_destroy   device, context, queue    OpenCLRaytracer.cpp:36-44, 76-78
 *   rt_upload_scene          objs/lights marshalling   OpenCLRaytracer.cpp:16-24, 47-48, 70-71, 108-146
 *   rt_set_rays              ray buffer upload         OpenCLRaytracer.cpp:26-33, 49, 72
 *   rt_set_camera            host ray generation       OpenCL-Raytracer.cpp:18-26, 68-72
 *   rt_render                Render(): launch + read   OpenCLRaytracer.cpp:80-105
 *   rt_render_aov / rt_trace (no equivalent: hit ids / t are not exported by the reference;
 *                             added so that parity of the nearest-hit search can be tested)
 *   rt_trace_static          (no equivalent: rt_trace without device allocation)
 *   rt_quantize_rgb8         PPM quantisation          PPMExporter.cpp:19-21
 *   rt_plan_sequences / rt_plan_bands / rt_host_build_bvh   (no equivalent: host-only views of the render plan and the scene
 *                             builder, for tests on machines without a GPU)
 */
#ifndef RT_B200_H
#define RT_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RT_B200_ABI_VERSION 2

typedef struct rt_ctx rt_ctx;

typedef enum rt_status {
    RT_OK = 0,
    RT_ERR_INVALID = -1,     /* bad argument / call order                       */
    RT_ERR_NO_DEVICE = -2,   /* no usable sm_100 device (no CPU fallback)       */
    RT_ERR_CUDA = -3,        /* CUDA runtime error, see rt_last_error()         */
    RT_ERR_UNSUPPORTED = -4, /* scene outside the supported domain (non-affine) */
    RT_ERR_NOMEM = -5
} rt_status;

/* Kernel semantics selected per render (three stages of one path). */
typedef enum rt_mode {
    RT_MODE_HITTEST = 0,          /* hittest_kernel.cl:54-150  -> t in .x, (t,t,t,1) */
    RT_MODE_SHADE = 1,            /* shade_kernel.cl:180-258   -> lights accumulate  */
    RT_MODE_SHADE_AND_REFLECT = 2 /* shade_and_reflect_kernel.cl:244-285 (the live kernel) */
} rt_mode;

/* Nearest-hit search strategy.  All of them return bit-identical hits. */
typedef enum rt_accel {
    RT_ACCEL_AUTO = -1,
    RT_ACCEL_BRUTE = 0, /* every ray x every object, the reference's arithmetic on every pair   */
    RT_ACCEL_CULL = 1,  /* every ray x every object, conservative bounding-sphere reject first  */
    RT_ACCEL_BVH = 2,   /* bounding volume hierarchy, conservative traversal, exact leaf tests  */
    RT_ACCEL_GRID = 3   /* wavefront pipeline: rays through a common point -- camera rays (all start at the view-space origin,
                         * OpenCL-Raytracer.cpp:68-72) and shadow rays of point lights (their line passes through the light,
                         * shade_and_reflect_kernel.cl:196-209) -- look up ONE cell of a perspective grid around that point and
                         * walk its sorted object list; every other ray uses the BVH.  On the megakernel it is the BVH. */
} rt_accel;

/* Material.hpp:6-11 / cl_Material OpenCLRaytracer.hpp:32-39 (tightly packed here). */
typedef struct rt_material {
    float ambient[3];
    float diffuse[3];
    float specular[3];
    float absorption, reflection, transparency, shininess;
} rt_material;

/* ObjectData.hpp:20-22 / cl_ObjectData OpenCLRaytracer.hpp:41-50.  Matrices are
 * column-major exactly as glm::value_ptr yields them (m[12..14] = translation).
 * mvInverseTranspose is not read by any kernel of the path and is not passed. */
typedef struct rt_object {
    rt_material mat;
    float mv[16];
    float mvInverse[16];
    uint32_t type; /* 0 = unit sphere, 1 = unit cube [-0.5,0.5]^3 (ObjectData.hpp:9-12);
                    * 2 = EXTENSION, unit triangle x >= -0.5, y >= -0.5, x + y <= 0 in the plane z = 0 (the reference has no
                    * triangle primitive; static_cast<ObjectData::PrimativeType>(2) encodes it without a header change);
                    * any other value is never hit, as in the reference kernels' switch without default */
} rt_object;

/* Light.hpp:4-16 / cl_Light OpenCLRaytracer.hpp:52-58. position is view space. */
typedef struct rt_light {
    float ambient[3];
    float diffuse[3];
    float specular[3];
    float position[4];
} rt_light;

/* Ray3D.hpp:5-10 / cl_Ray OpenCLRaytracer.hpp:25-30. */
typedef struct rt_ray {
    float start[4];
    float direction[4];
} rt_ray;

typedef struct rt_stats {
    /* rays the reference kernel traces for the rendered pixels (SURVEY.md 8(d) formula) ...  */
    uint64_t rays_primary;
    uint64_t rays_shadow;
    uint64_t rays_reflect;
    /* ... and the shadow / reflection rays this path really traced: with RT_OPT_SKIP_DEAD_RAYS rays whose
     * result cannot reach the output (SAR:238 last-light-wins, SAR:268 absorption test) are not traced.
     * Throughput is always computed from rays_primary + traced_shadow + traced_reflect. */
    uint64_t traced_shadow;
    uint64_t traced_reflect;
    uint64_t pair_tests;     /* ray x object pairs examined by the brute-force searches      */
    uint64_t exact_tests;    /* pairs that ran the reference's exact arithmetic              */
    uint64_t node_visits;    /* BVH inner nodes visited                                      */
    uint64_t pixels;         /* pixels rendered by this context (its shard)                  */
    uint64_t hit_pixels;
    float kernel_ms;         /* device time from the start of the last render to the end of its last kernel (CUDA events;
                                rt_render: copies of finished bands run underneath)           */
    float total_ms;          /* kernels + device->host copy of the last render               */
    uint32_t kernel_launches;
    uint32_t accel_used;     /* rt_accel actually used                                       */
    uint32_t pipeline_used;  /* 0 = megakernel, 1 = wavefront                                */
    uint32_t grid_walk_used; /* RT_ACCEL_GRID: 1 = the rays without a common apex walked the uniform grid (node_visits then counts
                              * cells), 0 = they walked the BVH */
} rt_stats;

/* Integer options for rt_set_option. */
typedef enum rt_option {
    RT_OPT_PIPELINE = 1,   /* -1 auto, 0 megakernel, 1 wavefront */
    RT_OPT_COUNT_RAYS = 2, /* 1: fill the ray/test counters of rt_stats (small atomics cost) */
    RT_OPT_BVH_LEAF = 3,   /* max objects per BVH leaf (1..16) */
    RT_OPT_SKIP_DEAD_RAYS = 4, /* 1 (default): do not trace rays whose result is dead; 0: trace every ray the reference traces */
    RT_OPT_BVH_BUILDER = 6,    /* 0 (default): binned SAH on the host in rt_upload_scene; 1: Morton-order radix tree built on the
                                  device (scenes that change every frame: ~1 ms instead of ~150 ms per 10^5 objects, a looser tree) */
    RT_OPT_LANES = 7,      /* frames that stay on the device (rt_render_device): 2 (default) = the camera launches of the wavefront
                              pipeline run as two halves on two streams, each filling the thin generation tails of the other;
                              1: one launch sequence */
    RT_OPT_TAIL_GEN = 12,  /* wavefront pipeline: from this generation on (1 = after the primary rays, 1 + lights = after the first
                              shading, ...) ONE kernel finishes every path that is still alive instead of a trace/advance pair per
                              generation -- removes the latency-bound thin tail of small frames.  -1 (default) auto, 0 off */
    RT_OPT_DETECT_PINHOLE = 11, /* 1 (default): rt_set_rays recognises the pinhole grid main() builds (OpenCL-Raytracer.cpp:68-72) and
                                   switches to rt_set_camera (device ray generation, no 32 B/ray upload); 0: upload whatever arrives */
    RT_OPT_ZERO_COPY = 9,  /* rt_render into page-locked, device-addressable host memory (rt_host_alloc / rt_host_register / cudaHostAlloc):
                              the kernels store finished pixels straight into the caller's frame instead of a framebuffer in HBM that is
                              copied afterwards.  -1 (default): on the wavefront pipeline, 0: never, 1: on both pipelines */
    RT_OPT_GRID3_DENSITY = 10, /* RT_ACCEL_GRID: the rays without a common apex (reflection rays, ...) walk a uniform 3-D grid with this
                                  many cells per 100 objects (default 400) instead of the BVH; 0 = BVH.  Scenes whose lists come out
                                  long (clustered / very uneven object sizes) keep the BVH regardless. */
    RT_OPT_GRID_RES = 8,   /* cells per side of each of the six perspective grids around a light (default 512, 32..4096) */
    RT_OPT_READBACK_BANDS = 5  /* rt_render of a camera frame (or shard of one) on the wavefront pipeline: the frame is cut into n launch
                                  sequences of rows / row-blocks, each its own readback band, dealt round-robin to three streams;
                                  a finished band is copied under the kernels of the others.  Sizes shrink geometrically (x 0.7)
                                  where the copy is faster than the kernels, and are equal where it is not (ranks sharing the
                                  host's ingest bandwidth) -- decided from the copy and kernel times of the previous frame.
                                  0 (default) = auto (6-8 sequences, fewer for small shards), n = n sequences,
                                  1 = render everything, then copy it */
} rt_option;

const char* rt_last_error(void);
int rt_abi_version(void);

/* Number of CUDA devices this library can use (compute capability 10.x). */
int rt_device_count(int* count);
/* device < 0 selects the current CUDA device. */
int rt_create(rt_ctx** out, int device);
int rt_destroy(rt_ctx* ctx);
int rt_set_option(rt_ctx* ctx, int option, int64_t value);

/* Repack the AoS scene into the device SoA (and build the BVH lazily). */
int rt_upload_scene(rt_ctx* ctx, const rt_object* objs, uint32_t n_objs, const rt_light* lights, uint32_t n_lights);

/* Rays: either an explicit host array (any rays, index = output index; the pinhole grid is recognised, RT_OPT_DETECT_PINHOLE) ... */
int rt_set_rays(rt_ctx* ctx, const rt_ray* host_rays, uint64_t n_rays);
/* ... or the reference's pinhole grid generated on the device, bit-identical to
 * OpenCL-Raytracer.cpp:18-26,68-72: start (0,0,0,1), dir (ii - W/2, (H - jj) - H/2, zdir, 0),
 * zdir = -(H/2)/tan(fov/2) computed by the caller in float. */
int rt_set_camera(rt_ctx* ctx, uint32_t width, uint32_t height, float zdir);
/* The camera in use (set by rt_set_camera, or recognised by rt_set_rays); RT_ERR_INVALID when the context renders caller rays. */
int rt_get_camera(rt_ctx* ctx, uint32_t* width, uint32_t* height, float* zdir);
/* Restrict this context to a shard of the camera frame: row-blocks of
 * `block_rows` rows dealt round-robin, this context owning blocks b with
 * b % n_shards == shard.  (1 block, shard 0 of 1) = whole frame. */
int rt_set_shard(rt_ctx* ctx, uint32_t block_rows, uint32_t shard, uint32_t n_shards);

/* Render into a HOST buffer of n_rays float4 (xyz = unclamped colour, w = 1;
 * pixels whose primary ray misses are (0,0,0,1), OpenCLRaytracer.cpp:32).
 * With a shard set, only this shard's rows are written.  Synchronous. */
int rt_render(rt_ctx* ctx, uint32_t max_bounces, int mode, int accel, float* host_out);
/* Same, result left in DEVICE memory (dev_out may be a peer-mapped pointer of
 * another GPU: stores go over NVLink).  dev_out == NULL renders into the
 * context's own framebuffer.  Returns after the kernels finished. */
int rt_render_device(rt_ctx* ctx, uint32_t max_bounces, int mode, int accel, float* dev_out);
/* Device framebuffer of the context (n_rays float4), valid until rays/camera change. */
int rt_framebuffer(rt_ctx* ctx, float** dev_ptr, uint64_t* n_pixels);

/* Copy the context's whole device framebuffer (all n_rays float4, including rows other ranks stored into
 * it through rt_ipc_open) to host memory.  Synchronous. */
int rt_readback(rt_ctx* ctx, float* host_out);

/* Primary-ray nearest hit as AOVs: object index (-1 = miss) and t, host arrays of n_rays. */
int rt_render_aov(rt_ctx* ctx, int accel, int32_t* host_hit_id, float* host_hit_t);
/* Nearest hit of arbitrary host rays against the uploaded scene. */
int rt_trace(rt_ctx* ctx, const rt_ray* host_rays, uint64_t n_rays, int accel, int32_t* host_hit_id, float* host_hit_t);

/* rt_trace for callers that may not allocate device memory (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc).  No counterpart in the reference.  Same kernel and
 * same results as rt_trace, but the call runs no cudaMalloc / cudaFree or other allocating routine: the rays and the results are
 * staged in a fixed `static __device__` array inside the library, RT_TRACE_STATIC_BYTES (10 MiB) that every GPU which
 * loads the library holds until the process exits.
 *   - n_rays > RT_TRACE_STATIC_MAX_RAYS is rejected with RT_ERR_UNSUPPORTED (split the batch, or use rt_trace).
 *   - The scene must be resident and prepared for the rays' origins (rt_upload_scene done; origins no farther from the
 *     world origin than the scene's own extent, or than the rays of an earlier rt_trace / rt_set_rays): otherwise
 *     RT_ERR_UNSUPPORTED instead of a re-preparation, which would allocate.
 *   - Every context and every call on a GPU shares the one array, which keeps the last call's contents (each call overwrites
 *     all it reads back).  The call returns after its results reached the host, so calls in sequence are safe; calls that
 *     overlap on the same GPU -- other host threads, other contexts -- race and must be serialised by the caller.
 *   - n_rays == 0 only resolves the array's device address for the context's GPU (cudaGetSymbolAddress, once per device),
 *     so that a caller can do this set-up outside the region where even that is unwelcome.
 *   - No rt_stats are produced. */
#define RT_TRACE_STATIC_MAX_RAYS 262144u
#define RT_TRACE_STATIC_BYTES (RT_TRACE_STATIC_MAX_RAYS * 40ull)
int rt_trace_static(rt_ctx* ctx, const rt_ray* host_rays, uint64_t n_rays, int accel, int32_t* host_hit_id, float* host_hit_t);

int rt_get_stats(rt_ctx* ctx, rt_stats* out);

/* Device version of PPMExporter.cpp:19-21 quantisation, min(255, (int)floorf(c * 255.f)) per channel, packed rgb8
 * (3 bytes per pixel).  host_rgba == NULL: the frame the last render left in the context's DEVICE framebuffer (nothing is
 * uploaded, n_pixels is ignored); otherwise a host float4 frame of n_pixels.  The exporter prints a negative channel as a
 * negative number; bytes cannot, so negative channels are clamped to 0 and counted in *n_negative (may be NULL): a caller
 * that needs the exporter's exact text falls back to it when the count is not zero. */
int rt_quantize_rgb8(rt_ctx* ctx, const float* host_rgba, uint64_t n_pixels, uint8_t* host_rgb_out, uint64_t* n_negative);

/* Multi-GPU, one process per GPU: the gather root exports its device framebuffer as a
 * 64-byte CUDA IPC handle; every other rank opens it and passes the mapped pointer to rt_render_device,
 * so its kernels store their shard's pixels straight into the root's HBM over NVLink (no separate
 * collective).  The reference has no multi-device path (OpenCLRaytracer.cpp:36-44 picks one device). */
int rt_ipc_export_framebuffer(rt_ctx* ctx, void* handle_64_bytes);
int rt_ipc_open(rt_ctx* ctx, const void* handle_64_bytes, void** dev_ptr);
int rt_ipc_close(rt_ctx* ctx, void* dev_ptr);

/* Pinned host memory for render targets (cudaHostAlloc); rt_render accepts any host pointer,
 * a pinned one makes the readback a single DMA. */
int rt_host_alloc(void** out, uint64_t bytes);
int rt_host_free(void* p);
/* Page-lock an existing host range (e.g. a POSIX shared-memory segment that every rank of a multi-GPU job maps, so that
 * each GPU copies its own rows of the frame straight into the one host frame over its own PCIe link). */
int rt_host_register(void* p, uint64_t bytes);
int rt_host_unregister(void* p);

/* Host-only, no GPU needed: the scene builder that rt_upload_scene runs (conservative per-object bounds + binned-SAH tree,
 * what OpenCLRaytracer.cpp:13-33 marshals minus the tree, which the reference does not have), with the result copied out --
 * for tests of the builder on a machine without a GPU and for timing it.  nodes_out: n_nodes x 64 bytes, per node
 * float ctr0[3] ext0[3] ctr1[3] ext1[3]; int32 c0 c1 (>= 0 inner node, < 0 leaf: ~c = first << 4 | count - 1); float f0 f1.
 * leaf_obj_out: object indices in leaf order.  Either may be NULL; info always reports the sizes.  threads: 0 = all
 * host threads (at most 16); the tree does not depend on it. */
typedef struct rt_bvh_info {
    uint32_t n_nodes, n_leaf_objs, depth;
    int32_t root;         /* child-style reference */
    float omax, scene_radius;
    double build_ms;
} rt_bvh_info;
int rt_host_build_bvh(const rt_object* objs, uint32_t n_objs, const rt_light* lights, uint32_t n_lights, float omax_user, int leaf_size,
                      int threads, rt_bvh_info* info, void* nodes_out, uint64_t nodes_capacity_bytes, uint32_t* leaf_obj_out,
                      uint64_t leaf_capacity);

/* Host-only: how rt_render cuts a camera frame (n_shards == 1: its rows; else the row-blocks shard owns) into n_bands readback
 * bands shrinking by `ratio` (the two-lane band plan, RT_BAND_MODE=1 RT_STAGGER_K=0; the default plan is described at
 * RT_OPT_READBACK_BANDS).  bounds[0 .. n_bands]: band b covers units bounds[b] .. bounds[b+1]. */
int rt_plan_bands(uint32_t height, uint32_t block_rows, uint32_t shard, uint32_t n_shards, int n_bands, double ratio, uint32_t* bounds);

/* Host-only: the default readback plan of rt_render (RT_OPT_READBACK_BANDS): the rows of a width x height frame (n_shards == 1) or
 * the row-blocks `shard` owns are cut into launch sequences, each its own readback band.  n_sequences 0 = auto, q 0 = size ratio
 * of consecutive sequences from kernel_copy_ratio (what rt_render measures per frame: 2.5 on one GPU, ~1.2 when 8 ranks share
 * the host).  bounds must hold 17 entries; *n_out sequences, sequence b covers units bounds[b] .. bounds[b+1].  For tests. */
int rt_plan_sequences(uint32_t width, uint32_t height, uint32_t block_rows, uint32_t shard, uint32_t n_shards, int n_sequences, double q,
                      double kernel_copy_ratio, uint32_t* bounds, int* n_out);

/* Register-only FFMA loop: measured FP32 peak of this device in TFLOP/s (roofline denominator). */
int rt_measure_fp32_peak(rt_ctx* ctx, float* tflops, float* sm_mhz_hint);
/* The two loops of the last rt_measure_fp32_peak separately: scalar FFMA, packed FFMA2 (its result is their maximum). */
int rt_fp32_peak_detail(rt_ctx* ctx, float* scalar_tflops, float* packed_tflops);

#ifdef __cplusplus
}
#endif

#endif /* RT_B200_H */
