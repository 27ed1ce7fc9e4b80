// mcpt_math.h — tiny Eigen-free linear algebra for the host-side mirror of the
// reference's scene/BVH interface.
//
// The reference uses Eigen through easycppogl/gl_eigen.h (GLVec3/GLVec4/GLMat4,
// namespace Transfo, easycppogl/gl_eigen.cpp:30-121). Eigen is not available in
// this image, so the few operations the scene path needs are provided here with
// the same conventions: column-major 4x4 float storage (Eigen default, which is
// also what glUniformMatrix4fv / GLSL mat4 expect), angles of rotateX/Y/Z in
// DEGREES, M(row, col) element access.
#pragma once
#include <cmath>
#include <cstring>

namespace mcpt {

struct Vec3 {
    float x = 0, y = 0, z = 0;
    Vec3() = default;
    Vec3(float a, float b, float c) : x(a), y(b), z(c) {}
    float& operator[](int i) { return (&x)[i]; }
    float operator[](int i) const { return (&x)[i]; }
};
inline Vec3 operator+(Vec3 a, Vec3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline Vec3 operator-(Vec3 a, Vec3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline Vec3 operator*(Vec3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
inline Vec3 operator*(float s, Vec3 a) { return a * s; }
inline Vec3 operator/(Vec3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }
inline float dot(Vec3 a, Vec3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline Vec3 cross(Vec3 a, Vec3 b) {
    return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
inline float norm(Vec3 a) { return std::sqrt(dot(a, a)); }

struct Vec4 {
    float x = 0, y = 0, z = 0, w = 0;
    Vec4() = default;
    Vec4(float a, float b, float c, float d) : x(a), y(b), z(c), w(d) {}
    float& operator[](int i) { return (&x)[i]; }
    float operator[](int i) const { return (&x)[i]; }
};
inline Vec4 operator*(Vec4 a, float s) { return {a.x * s, a.y * s, a.z * s, a.w * s}; }
inline Vec4 operator*(float s, Vec4 a) { return a * s; }
inline Vec4 operator+(Vec4 a, Vec4 b) { return {a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w}; }

// 4x4 float matrix, column-major: m[4*col + row].
struct Mat4 {
    float m[16];
    Mat4() { set_identity(); }
    void set_identity() {
        std::memset(m, 0, sizeof m);
        m[0] = m[5] = m[10] = m[15] = 1.0f;
    }
    float& operator()(int r, int c) { return m[4 * c + r]; }
    float operator()(int r, int c) const { return m[4 * c + r]; }
    const float* data() const { return m; }
    float* data() { return m; }

    Mat4 inverse() const;
};

inline Mat4 operator*(const Mat4& A, const Mat4& B) {
    Mat4 C;
    for (int c = 0; c < 4; ++c)
        for (int r = 0; r < 4; ++r) {
            float s = 0.0f;
            for (int k = 0; k < 4; ++k) s += A(r, k) * B(k, c);
            C(r, c) = s;
        }
    return C;
}

inline Vec4 operator*(const Mat4& A, const Vec4& v) {
    Vec4 r;
    for (int i = 0; i < 4; ++i) r[i] = A(i, 0) * v.x + A(i, 1) * v.y + A(i, 2) * v.z + A(i, 3) * v.w;
    return r;
}

// General 4x4 inverse by cofactor expansion, evaluated in double and rounded
// once to float (Eigen's float path rounds at every step; the tables this feeds
// are *inputs* of the traced path, so either is a valid PrimData).
inline Mat4 Mat4::inverse() const {
    double a[16], inv[16];
    for (int i = 0; i < 16; ++i) a[i] = m[i];
    inv[0] = a[5] * a[10] * a[15] - a[5] * a[11] * a[14] - a[9] * a[6] * a[15] + a[9] * a[7] * a[14] + a[13] * a[6] * a[11] - a[13] * a[7] * a[10];
    inv[4] = -a[4] * a[10] * a[15] + a[4] * a[11] * a[14] + a[8] * a[6] * a[15] - a[8] * a[7] * a[14] - a[12] * a[6] * a[11] + a[12] * a[7] * a[10];
    inv[8] = a[4] * a[9] * a[15] - a[4] * a[11] * a[13] - a[8] * a[5] * a[15] + a[8] * a[7] * a[13] + a[12] * a[5] * a[11] - a[12] * a[7] * a[9];
    inv[12] = -a[4] * a[9] * a[14] + a[4] * a[10] * a[13] + a[8] * a[5] * a[14] - a[8] * a[6] * a[13] - a[12] * a[5] * a[10] + a[12] * a[6] * a[9];
    inv[1] = -a[1] * a[10] * a[15] + a[1] * a[11] * a[14] + a[9] * a[2] * a[15] - a[9] * a[3] * a[14] - a[13] * a[2] * a[11] + a[13] * a[3] * a[10];
    inv[5] = a[0] * a[10] * a[15] - a[0] * a[11] * a[14] - a[8] * a[2] * a[15] + a[8] * a[3] * a[14] + a[12] * a[2] * a[11] - a[12] * a[3] * a[10];
    inv[9] = -a[0] * a[9] * a[15] + a[0] * a[11] * a[13] + a[8] * a[1] * a[15] - a[8] * a[3] * a[13] - a[12] * a[1] * a[11] + a[12] * a[3] * a[9];
    inv[13] = a[0] * a[9] * a[14] - a[0] * a[10] * a[13] - a[8] * a[1] * a[14] + a[8] * a[2] * a[13] + a[12] * a[1] * a[10] - a[12] * a[2] * a[9];
    inv[2] = a[1] * a[6] * a[15] - a[1] * a[7] * a[14] - a[5] * a[2] * a[15] + a[5] * a[3] * a[14] + a[13] * a[2] * a[7] - a[13] * a[3] * a[6];
    inv[6] = -a[0] * a[6] * a[15] + a[0] * a[7] * a[14] + a[4] * a[2] * a[15] - a[4] * a[3] * a[14] - a[12] * a[2] * a[7] + a[12] * a[3] * a[6];
    inv[10] = a[0] * a[5] * a[15] - a[0] * a[7] * a[13] - a[4] * a[1] * a[15] + a[4] * a[3] * a[13] + a[12] * a[1] * a[7] - a[12] * a[3] * a[5];
    inv[14] = -a[0] * a[5] * a[14] + a[0] * a[6] * a[13] + a[4] * a[1] * a[14] - a[4] * a[2] * a[13] - a[12] * a[1] * a[6] + a[12] * a[2] * a[5];
    inv[3] = -a[1] * a[6] * a[11] + a[1] * a[7] * a[10] + a[5] * a[2] * a[11] - a[5] * a[3] * a[10] - a[9] * a[2] * a[7] + a[9] * a[3] * a[6];
    inv[7] = a[0] * a[6] * a[11] - a[0] * a[7] * a[10] - a[4] * a[2] * a[11] + a[4] * a[3] * a[10] + a[8] * a[2] * a[7] - a[8] * a[3] * a[6];
    inv[11] = -a[0] * a[5] * a[11] + a[0] * a[7] * a[9] + a[4] * a[1] * a[11] - a[4] * a[3] * a[9] - a[8] * a[1] * a[7] + a[8] * a[3] * a[5];
    inv[15] = a[0] * a[5] * a[10] - a[0] * a[6] * a[9] - a[4] * a[1] * a[10] + a[4] * a[2] * a[9] + a[8] * a[1] * a[6] - a[8] * a[2] * a[5];
    double det = a[0] * inv[0] + a[1] * inv[4] + a[2] * inv[8] + a[3] * inv[12];
    Mat4 R;
    double id = 1.0 / det;
    for (int i = 0; i < 16; ++i) R.m[i] = float(inv[i] * id);
    return R;
}

// namespace Transfo of easycppogl/gl_eigen.cpp:30-121 (same names, degrees).
namespace Transfo {
inline Mat4 translate(float x, float y, float z) {
    Mat4 T;
    T(0, 3) = x;
    T(1, 3) = y;
    T(2, 3) = z;
    return T;
}
inline Mat4 translate(Vec3 v) { return translate(v.x, v.y, v.z); }
inline Mat4 scale(float sx, float sy, float sz) {
    Mat4 S;
    S(0, 0) = sx;
    S(1, 1) = sy;
    S(2, 2) = sz;
    return S;
}
inline Mat4 scale(float s) { return scale(s, s, s); }
inline Mat4 scale(Vec3 v) { return scale(v.x, v.y, v.z); }
inline void sincos_deg(float deg, float& s, float& c) {
    float rad = float(M_PI / 180) * deg;
    s = std::sin(rad);
    c = std::cos(rad);
}
inline Mat4 rotateX(float deg) {
    float s, c;
    sincos_deg(deg, s, c);
    Mat4 R;
    R(1, 1) = c;
    R(1, 2) = -s;
    R(2, 1) = s;
    R(2, 2) = c;
    return R;
}
inline Mat4 rotateY(float deg) {
    float s, c;
    sincos_deg(deg, s, c);
    Mat4 R;
    R(0, 0) = c;
    R(0, 2) = s;
    R(2, 0) = -s;
    R(2, 2) = c;
    return R;
}
inline Mat4 rotateZ(float deg) {
    float s, c;
    sincos_deg(deg, s, c);
    Mat4 R;
    R(0, 0) = c;
    R(0, 1) = -s;
    R(1, 0) = s;
    R(1, 1) = c;
    return R;
}
// point transform (w = 1), gl_eigen.h:91-94
inline Vec3 apply(const Mat4& T, Vec3 p) {
    Vec4 r = T * Vec4(p.x, p.y, p.z, 1.0f);
    return {r.x, r.y, r.z};
}
}  // namespace Transfo

}  // namespace mcpt
