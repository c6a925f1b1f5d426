// Stand-in for <glm/gtx/string_cast.hpp> (OpenCL-Raytracer.cpp:3; debugging aid only).
#pragma once
#include <cstdio>
#include <string>
#include "../glm.hpp"

namespace glm {
inline std::string to_string(const vec3& v) {
    char b[96];
    std::snprintf(b, sizeof b, "vec3(%f, %f, %f)", v.x, v.y, v.z);
    return b;
}
inline std::string to_string(const vec4& v) {
    char b[128];
    std::snprintf(b, sizeof b, "vec4(%f, %f, %f, %f)", v.x, v.y, v.z, v.w);
    return b;
}
inline std::string to_string(const mat4& m) {
    return "mat4x4(" + to_string(m[0]) + ", " + to_string(m[1]) + ", " + to_string(m[2]) + ", " + to_string(m[3]) + ")";
}
}  // namespace glm
