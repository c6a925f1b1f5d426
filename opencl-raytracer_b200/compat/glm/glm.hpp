// Minimal stand-in for the subset of GLM that the reference's host headers and
// sources use (IRaytracer.hpp:3, Material.hpp:3-4, ObjectData.cpp:7-8,
// SceneLoader.cpp:215-294, Light.hpp:13, OpenCL-Raytracer.cpp:18-32).
//
// GLM is a third-party dependency of the reference (vcpkg.json:8, unversioned)
// and is NOT present under /root/reference or in this image. This header
// restates the published GLM 0.9.9 algorithms with the same operation order so
// that host-side matrices come out as close to a real GLM build as can be done
// from the published formulas.  It is used ONLY when the real package is absent
// (put the real glm earlier on the include path to override it).
//
// Parity note: "unpinned at the glm boundary" -- no reference test pins any
// matrix. The CUDA path and the oracle both consume the same host ObjectData,
// so CUDA-vs-oracle parity does not depend on this file.
#pragma once

#include <cmath>
#include <cstddef>
#include <cstdint>

namespace glm {

struct vec2 {
    float x, y;
    vec2() : x(0), y(0) {}
    template <typename A, typename B>
    vec2(A a, B b) : x(static_cast<float>(a)), y(static_cast<float>(b)) {}
    explicit vec2(float s) : x(s), y(s) {}
    float& operator[](int i) { return (&x)[i]; }
    const float& operator[](int i) const { return (&x)[i]; }
};

struct vec4;

struct vec3 {
    float x, y, z;
    vec3() : x(0), y(0), z(0) {}
    template <typename A, typename B, typename C>
    vec3(A a, B b, C c) : x(static_cast<float>(a)), y(static_cast<float>(b)), z(static_cast<float>(c)) {}
    explicit vec3(float s) : x(s), y(s), z(s) {}
    // GLM_EXPLICIT is empty unless GLM_FORCE_EXPLICIT_CTOR is defined, so the
    // vec4 -> vec3 truncation is an implicit conversion (ObjectData.cpp:55,131).
    vec3(const vec4& v);
    float& operator[](int i) { return (&x)[i]; }
    const float& operator[](int i) const { return (&x)[i]; }
};

struct vec4 {
    float x, y, z, w;
    vec4() : x(0), y(0), z(0), w(0) {}
    template <typename A, typename B, typename C, typename D>
    vec4(A a, B b, C c, D d)
        : x(static_cast<float>(a)), y(static_cast<float>(b)), z(static_cast<float>(c)), w(static_cast<float>(d)) {}
    explicit vec4(float s) : x(s), y(s), z(s), w(s) {}
    template <typename D>
    vec4(const vec3& v, D d) : x(v.x), y(v.y), z(v.z), w(static_cast<float>(d)) {}
    float& operator[](int i) { return (&x)[i]; }
    const float& operator[](int i) const { return (&x)[i]; }
};

inline vec3::vec3(const vec4& v) : x(v.x), y(v.y), z(v.z) {}

// ---- vec3 arithmetic -------------------------------------------------------
inline vec3 operator+(const vec3& a, const vec3& b) { return vec3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline vec3 operator-(const vec3& a, const vec3& b) { return vec3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline vec3 operator-(const vec3& a) { return vec3(-a.x, -a.y, -a.z); }
inline vec3 operator*(const vec3& a, const vec3& b) { return vec3(a.x * b.x, a.y * b.y, a.z * b.z); }
inline vec3 operator*(const vec3& a, float s) { return vec3(a.x * s, a.y * s, a.z * s); }
inline vec3 operator*(float s, const vec3& a) { return vec3(s * a.x, s * a.y, s * a.z); }
inline vec3 operator/(const vec3& a, float s) { return vec3(a.x / s, a.y / s, a.z / s); }
inline bool operator==(const vec3& a, const vec3& b) { return a.x == b.x && a.y == b.y && a.z == b.z; }

// ---- vec4 arithmetic -------------------------------------------------------
inline vec4 operator+(const vec4& a, const vec4& b) { return vec4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
inline vec4 operator-(const vec4& a, const vec4& b) { return vec4(a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w); }
inline vec4 operator-(const vec4& a) { return vec4(-a.x, -a.y, -a.z, -a.w); }
inline vec4 operator*(const vec4& a, const vec4& b) { return vec4(a.x * b.x, a.y * b.y, a.z * b.z, a.w * b.w); }
inline vec4 operator*(const vec4& a, float s) { return vec4(a.x * s, a.y * s, a.z * s, a.w * s); }
inline vec4 operator*(float s, const vec4& a) { return vec4(s * a.x, s * a.y, s * a.z, s * a.w); }
inline vec4 operator/(const vec4& a, float s) { return vec4(a.x / s, a.y / s, a.z / s, a.w / s); }
inline bool operator==(const vec4& a, const vec4& b) { return a.x == b.x && a.y == b.y && a.z == b.z && a.w == b.w; }

// ---- scalar helpers (func_common / func_trigonometric) ---------------------
inline float min(float a, float b) { return (b < a) ? b : a; }
inline float max(float a, float b) { return (a < b) ? b : a; }
inline float sign(float x) { return static_cast<float>((0.0f < x) - (x < 0.0f)); }
inline float radians(float degrees) { return degrees * static_cast<float>(0.01745329251994329576923690768489); }
inline float degrees(float radians) { return radians * static_cast<float>(57.295779513082320876798154814105); }
inline float inversesqrt(float x) { return 1.0f / std::sqrt(x); }

// ---- geometric -------------------------------------------------------------
// compute_dot<vec3>: tmp = a*b; tmp.x + tmp.y + tmp.z
inline float dot(const vec3& a, const vec3& b) {
    vec3 t = a * b;
    return t.x + t.y + t.z;
}
// compute_dot<vec4>: (tmp.x + tmp.y) + (tmp.z + tmp.w)
inline float dot(const vec4& a, const vec4& b) {
    vec4 t = a * b;
    return (t.x + t.y) + (t.z + t.w);
}
inline vec3 cross(const vec3& a, const vec3& b) {
    return vec3(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y);
}
inline float length(const vec3& v) { return std::sqrt(dot(v, v)); }
inline float length(const vec4& v) { return std::sqrt(dot(v, v)); }
// compute_normalize: v * inversesqrt(dot(v, v))
inline vec3 normalize(const vec3& v) { return v * inversesqrt(dot(v, v)); }
inline vec4 normalize(const vec4& v) { return v * inversesqrt(dot(v, v)); }

// ---- mat4 (column-major: m[c] is column c) ----------------------------------
struct mat4 {
    vec4 col[4];
    mat4() : mat4(1.0f) {}
    explicit mat4(float s) {
        col[0] = vec4(s, 0, 0, 0);
        col[1] = vec4(0, s, 0, 0);
        col[2] = vec4(0, 0, s, 0);
        col[3] = vec4(0, 0, 0, s);
    }
    mat4(const vec4& c0, const vec4& c1, const vec4& c2, const vec4& c3) {
        col[0] = c0; col[1] = c1; col[2] = c2; col[3] = c3;
    }
    vec4& operator[](int i) { return col[i]; }
    const vec4& operator[](int i) const { return col[i]; }
    mat4& operator*=(const mat4& m);
    mat4& operator*=(float s) {
        col[0] = col[0] * s; col[1] = col[1] * s; col[2] = col[2] * s; col[3] = col[3] * s;
        return *this;
    }
};

// mat4x4 * vec4: (m0*x + m1*y) + (m2*z + m3*w)
inline vec4 operator*(const mat4& m, const vec4& v) {
    const vec4 Mul0 = m[0] * vec4(v.x);
    const vec4 Mul1 = m[1] * vec4(v.y);
    const vec4 Add0 = Mul0 + Mul1;
    const vec4 Mul2 = m[2] * vec4(v.z);
    const vec4 Mul3 = m[3] * vec4(v.w);
    const vec4 Add1 = Mul2 + Mul3;
    return Add0 + Add1;
}

// mat4x4 * mat4x4: column i = A0*B[i][0] + A1*B[i][1] + A2*B[i][2] + A3*B[i][3], left to right
inline mat4 operator*(const mat4& a, const mat4& b) {
    mat4 r(0.0f);
    for (int i = 0; i < 4; ++i) r[i] = a[0] * b[i][0] + a[1] * b[i][1] + a[2] * b[i][2] + a[3] * b[i][3];
    return r;
}
inline mat4 operator*(const mat4& m, float s) { return mat4(m[0] * s, m[1] * s, m[2] * s, m[3] * s); }
inline mat4& mat4::operator*=(const mat4& m) { return (*this = *this * m); }
inline bool operator==(const mat4& a, const mat4& b) { return a[0] == b[0] && a[1] == b[1] && a[2] == b[2] && a[3] == b[3]; }

inline mat4 transpose(const mat4& m) {
    mat4 r(0.0f);
    for (int c = 0; c < 4; ++c)
        for (int k = 0; k < 4; ++k) r[c][k] = m[k][c];
    return r;
}

// compute_inverse<4,4>: cofactor expansion with the 18 2x2 sub-determinants.
inline mat4 inverse(const mat4& m) {
    const float Coef00 = m[2][2] * m[3][3] - m[3][2] * m[2][3];
    const float Coef02 = m[1][2] * m[3][3] - m[3][2] * m[1][3];
    const float Coef03 = m[1][2] * m[2][3] - m[2][2] * m[1][3];

    const float Coef04 = m[2][1] * m[3][3] - m[3][1] * m[2][3];
    const float Coef06 = m[1][1] * m[3][3] - m[3][1] * m[1][3];
    const float Coef07 = m[1][1] * m[2][3] - m[2][1] * m[1][3];

    const float Coef08 = m[2][1] * m[3][2] - m[3][1] * m[2][2];
    const float Coef10 = m[1][1] * m[3][2] - m[3][1] * m[1][2];
    const float Coef11 = m[1][1] * m[2][2] - m[2][1] * m[1][2];

    const float Coef12 = m[2][0] * m[3][3] - m[3][0] * m[2][3];
    const float Coef14 = m[1][0] * m[3][3] - m[3][0] * m[1][3];
    const float Coef15 = m[1][0] * m[2][3] - m[2][0] * m[1][3];

    const float Coef16 = m[2][0] * m[3][2] - m[3][0] * m[2][2];
    const float Coef18 = m[1][0] * m[3][2] - m[3][0] * m[1][2];
    const float Coef19 = m[1][0] * m[2][2] - m[2][0] * m[1][2];

    const float Coef20 = m[2][0] * m[3][1] - m[3][0] * m[2][1];
    const float Coef22 = m[1][0] * m[3][1] - m[3][0] * m[1][1];
    const float Coef23 = m[1][0] * m[2][1] - m[2][0] * m[1][1];

    const vec4 Fac0(Coef00, Coef00, Coef02, Coef03);
    const vec4 Fac1(Coef04, Coef04, Coef06, Coef07);
    const vec4 Fac2(Coef08, Coef08, Coef10, Coef11);
    const vec4 Fac3(Coef12, Coef12, Coef14, Coef15);
    const vec4 Fac4(Coef16, Coef16, Coef18, Coef19);
    const vec4 Fac5(Coef20, Coef20, Coef22, Coef23);

    const vec4 Vec0(m[1][0], m[0][0], m[0][0], m[0][0]);
    const vec4 Vec1(m[1][1], m[0][1], m[0][1], m[0][1]);
    const vec4 Vec2(m[1][2], m[0][2], m[0][2], m[0][2]);
    const vec4 Vec3(m[1][3], m[0][3], m[0][3], m[0][3]);

    const vec4 Inv0(Vec1 * Fac0 - Vec2 * Fac1 + Vec3 * Fac2);
    const vec4 Inv1(Vec0 * Fac0 - Vec2 * Fac3 + Vec3 * Fac4);
    const vec4 Inv2(Vec0 * Fac1 - Vec1 * Fac3 + Vec3 * Fac5);
    const vec4 Inv3(Vec0 * Fac2 - Vec1 * Fac4 + Vec2 * Fac5);

    const vec4 SignA(+1, -1, +1, -1);
    const vec4 SignB(-1, +1, -1, +1);
    const mat4 Inverse(Inv0 * SignA, Inv1 * SignB, Inv2 * SignA, Inv3 * SignB);

    const vec4 Row0(Inverse[0][0], Inverse[1][0], Inverse[2][0], Inverse[3][0]);
    const vec4 Dot0(m[0] * Row0);
    const float Dot1 = (Dot0.x + Dot0.y) + (Dot0.z + Dot0.w);

    const float OneOverDeterminant = 1.0f / Dot1;
    return Inverse * OneOverDeterminant;
}

}  // namespace glm
