// rt_scene.hpp -- host-side repacking of the AoS scene (rt_object / rt_light, the layout of
// cl_ObjectData / cl_Light in OpenCLRaytracer.hpp:41-58) into the device SoA, the conservative
// bounds of every object and the BVH over them.
#pragma once

#include <stdint.h>

#include <string>
#include <vector>

#include "../../include/rt_b200.h"
#include "rt_layout.h"

namespace rt {

struct F4 {
    float x, y, z, w;
};

// Input of the device-side tree builder (rt_lbvh.cu): one record per live object, bounds already padded and rounded
// outwards to binary32.
struct LbvhLeaf {  // 32 B
    float c[3];     // centre
    float half[3];  // padded half extents
    float f;        // noise coefficient (box inflation per unit distance)
    uint32_t obj;   // object index
};

struct HostScene {
    std::vector<F4> minv;  // 4 per object (columns of mvInverse)
    std::vector<F4> mv;    // 4 per object
    std::vector<F4> mat;   // 3 per object
    std::vector<CullRec> cull;
    std::vector<F4> aabb;  // per object: half extents of its box INCLUDING the noise inflation f * (omax + |c|) a ray from anywhere
                           // in the scene can need (the uniform grid enters objects by this box; w unused); zeros = never hit
    std::vector<BvhNode> nodes;
    std::vector<uint32_t> leaf_obj;
    std::vector<CullRec> leaf_cull;
    std::vector<F4> lights;  // 4 per light
    uint32_t n_objs = 0, n_lights = 0;
    int32_t root = 0;      // child-style reference
    float omax = 0.f;      // bound on |ray origin| the pads were computed for
    float scene_radius = 0.f;
    uint32_t bvh_depth = 0;
    float h_min = 1.f;     // smallest filter coefficient h = 1 - g over the live objects
    float f_global = 0.f;  // largest noise coefficient of the scene (per-child values are in the nodes)
    // leaf_size == 0 ("no host tree"): the records for the device builder and the bounds of their centres
    std::vector<LbvhLeaf> lbvh_leaves;
    float lbvh_lo[3] = {0.f, 0.f, 0.f}, lbvh_hi[3] = {0.f, 0.f, 0.f};
};

// Returns RT_OK or RT_ERR_UNSUPPORTED / RT_ERR_INVALID with a message in err.
// omax_user: largest |origin| of caller-supplied rays (0 for the pinhole camera).
// leaf_size: objects per BVH leaf of the host SAH builder (1..16); 0 = build no tree, fill lbvh_leaves instead.
// threads: host threads for the per-object bounds and the subtrees of the SAH build (0 = all, at most 16); the result does
// not depend on it.
int build_scene(const rt_object* objs, uint32_t n_objs, const rt_light* lights, uint32_t n_lights, float omax_user,
                int leaf_size, HostScene& out, std::string& err, int threads = 0);

}  // namespace rt
