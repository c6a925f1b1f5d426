// Stand-in for <glm/gtc/type_ptr.hpp> (Material.hpp:4, OpenCLRaytracer.cpp:115).
#pragma once
#include "../glm.hpp"

namespace glm {
inline const float* value_ptr(const mat4& m) { return &m[0].x; }
inline float* value_ptr(mat4& m) { return &m[0].x; }
inline const float* value_ptr(const vec4& v) { return &v.x; }
inline const float* value_ptr(const vec3& v) { return &v.x; }
}  // namespace glm
