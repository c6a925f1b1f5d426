// rt_layout.h -- byte layouts shared by the host scene builder (rt_scene.cpp) and the
// device code (rt_device.cuh).  Plain structs, no CUDA types.
#pragma once
#include <stdint.h>

namespace rt {

struct CullRec {       // 32 B, two 128-bit loads
    float cx, cy, cz;  // view-space centre of the bounding sphere
    float r2;          // padded squared radius
    float h;           // 1 - g (cone term): reject iff h*|oc|^2 - (oc.u)^2 > r2
    float cc;          // |c|^2
    float e0;          // h*cc - r2: common-origin (o = 0) form, reject iff e0 > (c.u)^2
    uint32_t type;     // 0 sphere, 1 box, 2 triangle (extension), other: never hit (the kernels' switch has no default)
};

struct BvhNode {  // 64 B: both children's boxes as centre + half extent (the slab test needs no min/max then)
    float ctr0[3], ext0[3];
    float ctr1[3], ext1[3];
    int32_t c0, c1;  // >= 0: inner node index; < 0: leaf, ~c = (first << 4) | (count - 1)
    float f0, f1;    // noise coefficient of each child: its box is inflated by f * (distance origin -> farthest box point)
};                   // (the static pad for the rounding of the slab arithmetic is already inside ext)

}  // namespace rt
