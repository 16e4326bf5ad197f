// Test infrastructure (oracle/_ref build only) -- NOT product code.
//
// Minimal stand-in for the un-vendored third-party dependency `glm` (reference CMakeLists.txt:45
// `find_package(glm REQUIRED)`, version unpinned, absent from /root/reference and from this image).
// The reference calls exactly: glm::mat4x4(1.0f), translate, rotate, scale (scene.h:11-16),
// inverse (scene.h:22), mat[c][r] (types.h:391-399), and reinterprets vec3/vec4 as float3
// (use_cuda.h:42-59).  This file restates glm's published scalar algorithms for those calls
// (column-major, post-multiplying transforms, angle in radians; `inverse` by cofactor expansion with
// the Coef/Fac/Vec formulation of glm/detail/func_matrix.inl).  No reference test pins glm's
// arithmetic => parity is UNPINNED at this boundary.
#ifndef PTB200_ORACLE_GLM_SHIM_HPP
#define PTB200_ORACLE_GLM_SHIM_HPP
#include <cmath>

namespace glm {

struct vec3 {
    float x, y, z;
    vec3() : x(0), y(0), z(0) {}
    vec3(float a, float b, float c) : x(a), y(b), z(c) {}
    float& operator[](int i) { return (&x)[i]; }
    const float& operator[](int i) const { return (&x)[i]; }
};

struct vec4 {
    float x, y, z, w;
    vec4() : x(0), y(0), z(0), w(0) {}
    vec4(float a, float b, float c, float d) : x(a), y(b), z(c), w(d) {}
    float& operator[](int i) { return (&x)[i]; }
    const float& operator[](int i) const { return (&x)[i]; }
};

inline vec4 operator*(const vec4& a, float s) { return vec4(a.x * s, a.y * s, a.z * s, a.w * s); }
inline vec4 operator*(const vec4& a, const vec4& b) { return vec4(a.x * b.x, a.y * b.y, a.z * b.z, a.w * b.w); }
inline vec4 operator+(const vec4& a, const vec4& b) { return vec4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
inline vec4 operator-(const vec4& a, const vec4& b) { return vec4(a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w); }

struct mat4x4 {
    vec4 c[4];
    mat4x4() {}
    explicit mat4x4(float d) {
        c[0] = vec4(d, 0, 0, 0); c[1] = vec4(0, d, 0, 0); c[2] = vec4(0, 0, d, 0); c[3] = vec4(0, 0, 0, d);
    }
    vec4& operator[](int i) { return c[i]; }
    const vec4& operator[](int i) const { return c[i]; }
};
typedef mat4x4 mat4;

inline mat4x4 translate(const mat4x4& m, const vec3& v) {
    mat4x4 r = m;
    r[3] = m[0] * v[0] + m[1] * v[1] + m[2] * v[2] + m[3];
    return r;
}

inline mat4x4 rotate(const mat4x4& m, float angle, const vec3& v) {
    const float a = angle;
    const float c = std::cos(a);
    const float s = std::sin(a);
    const float il = 1.0f / std::sqrt(v.x * v.x + v.y * v.y + v.z * v.z);
    const vec3 axis(v.x * il, v.y * il, v.z * il);
    const vec3 temp((1.0f - c) * axis.x, (1.0f - c) * axis.y, (1.0f - c) * axis.z);
    float R[3][3];
    R[0][0] = c + temp[0] * axis[0];
    R[0][1] = temp[0] * axis[1] + s * axis[2];
    R[0][2] = temp[0] * axis[2] - s * axis[1];
    R[1][0] = temp[1] * axis[0] - s * axis[2];
    R[1][1] = c + temp[1] * axis[1];
    R[1][2] = temp[1] * axis[2] + s * axis[0];
    R[2][0] = temp[2] * axis[0] + s * axis[1];
    R[2][1] = temp[2] * axis[1] - s * axis[0];
    R[2][2] = c + temp[2] * axis[2];
    mat4x4 r;
    r[0] = m[0] * R[0][0] + m[1] * R[0][1] + m[2] * R[0][2];
    r[1] = m[0] * R[1][0] + m[1] * R[1][1] + m[2] * R[1][2];
    r[2] = m[0] * R[2][0] + m[1] * R[2][1] + m[2] * R[2][2];
    r[3] = m[3];
    return r;
}

inline mat4x4 scale(const mat4x4& m, const vec3& v) {
    mat4x4 r;
    r[0] = m[0] * v[0];
    r[1] = m[1] * v[1];
    r[2] = m[2] * v[2];
    r[3] = m[3];
    return r;
}

inline mat4x4 inverse(const mat4x4& m) {
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
    mat4x4 Inverse;
    Inverse[0] = Inv0 * SignA;
    Inverse[1] = Inv1 * SignB;
    Inverse[2] = Inv2 * SignA;
    Inverse[3] = Inv3 * SignB;

    const vec4 Row0(Inverse[0][0], Inverse[1][0], Inverse[2][0], Inverse[3][0]);
    const vec4 Dot0(m[0] * Row0);
    const float Dot1 = (Dot0.x + Dot0.y) + (Dot0.z + Dot0.w);
    const float OneOverDeterminant = 1.0f / Dot1;

    mat4x4 r;
    for (int i = 0; i < 4; i++) r[i] = Inverse[i] * OneOverDeterminant;
    return r;
}

}  // namespace glm
#endif
