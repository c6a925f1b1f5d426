// Stand-in for <glm/gtc/matrix_transform.hpp> (GLM 0.9.9 formulas, right-handed
// default), used by SceneLoader.cpp:215,270,282,294 and OpenCL-Raytracer.cpp:2.
// See glm/glm.hpp in this directory for scope and the parity note.
#pragma once
#include "../glm.hpp"

namespace glm {

// translate(m, v): Result[3] = m[0]*v[0] + m[1]*v[1] + m[2]*v[2] + m[3]
inline mat4 translate(const mat4& m, const vec3& v) {
    mat4 r(m);
    r[3] = m[0] * v[0] + m[1] * v[1] + m[2] * v[2] + m[3];
    return r;
}

// scale(m, v): columns scaled, translation column kept
inline mat4 scale(const mat4& m, const vec3& v) {
    mat4 r(0.0f);
    r[0] = m[0] * v[0];
    r[1] = m[1] * v[1];
    r[2] = m[2] * v[2];
    r[3] = m[3];
    return r;
}

// rotate(m, angle, axis): Rodrigues matrix with temp = (1 - c) * axis
inline mat4 rotate(const mat4& m, float angle, const vec3& v) {
    const float a = angle;
    const float c = std::cos(a);
    const float s = std::sin(a);

    const vec3 axis(normalize(v));
    const vec3 temp((1.0f - c) * axis);

    mat4 Rotate(0.0f);
    Rotate[0][0] = c + temp[0] * axis[0];
    Rotate[0][1] = temp[0] * axis[1] + s * axis[2];
    Rotate[0][2] = temp[0] * axis[2] - s * axis[1];

    Rotate[1][0] = temp[1] * axis[0] - s * axis[2];
    Rotate[1][1] = c + temp[1] * axis[1];
    Rotate[1][2] = temp[1] * axis[2] + s * axis[0];

    Rotate[2][0] = temp[2] * axis[0] + s * axis[1];
    Rotate[2][1] = temp[2] * axis[1] - s * axis[0];
    Rotate[2][2] = c + temp[2] * axis[2];

    mat4 r(0.0f);
    r[0] = m[0] * Rotate[0][0] + m[1] * Rotate[0][1] + m[2] * Rotate[0][2];
    r[1] = m[0] * Rotate[1][0] + m[1] * Rotate[1][1] + m[2] * Rotate[1][2];
    r[2] = m[0] * Rotate[2][0] + m[1] * Rotate[2][1] + m[2] * Rotate[2][2];
    r[3] = m[3];
    return r;
}

// lookAtRH(eye, center, up)
inline mat4 lookAt(const vec3& eye, const vec3& center, const vec3& up) {
    const vec3 f(normalize(center - eye));
    const vec3 s(normalize(cross(f, up)));
    const vec3 u(cross(s, f));

    mat4 r(1.0f);
    r[0][0] = s.x;
    r[1][0] = s.y;
    r[2][0] = s.z;
    r[0][1] = u.x;
    r[1][1] = u.y;
    r[2][1] = u.z;
    r[0][2] = -f.x;
    r[1][2] = -f.y;
    r[2][2] = -f.z;
    r[3][0] = -dot(s, eye);
    r[3][1] = -dot(u, eye);
    r[3][2] = dot(f, eye);
    return r;
}

}  // namespace glm
