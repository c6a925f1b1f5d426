/* Stand-in for <CL/cl.h>: the reference's IRaytracer.hpp:4,13 needs only the
 * cl_float4 host type for the Render() return value. OpenCL itself is not used
 * by the CUDA path. Layout follows the Khronos header: 16-byte aligned union
 * with .s[4] and .x/.y/.z/.w views; cl_float3 is an alias of cl_float4. */
#ifndef RT_B200_COMPAT_CL_H
#define RT_B200_COMPAT_CL_H

#include <stdint.h>

typedef float cl_float;
typedef uint32_t cl_uint;
typedef int32_t cl_int;
typedef uint64_t cl_ulong;

typedef union {
    cl_float __attribute__((aligned(16))) s[4];
    struct { cl_float x, y, z, w; };
    struct { cl_float s0, s1, s2, s3; };
} cl_float4;

typedef cl_float4 cl_float3;

typedef union {
    cl_float __attribute__((aligned(64))) s[16];
    struct { cl_float s0, s1, s2, s3, s4, s5, s6, s7, s8, s9, sa, sb, sc, sd, se, sf; };
} cl_float16;

#endif
