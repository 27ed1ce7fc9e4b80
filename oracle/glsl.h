// oracle/glsl.h — TEST INFRASTRUCTURE (CPU oracle). Not part of the product.
//
// GLSL 4.30 vector types and built-ins, written out with ONE fixed evaluation
// order so that the oracle is a deterministic function on any IEEE-754 machine:
// every operator is a separately rounded binary32 operation (compile with
// -ffp-contract=off), sums are evaluated left to right, division and sqrt are
// the IEEE ones, normalize(v) = v * (1 / sqrt(dot(v,v))). GLSL leaves these
// choices to the driver; the oracle pins them.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>

namespace glsl {

struct vec2 {
    float x, y;
};
struct vec3 {
    float x, y, z;
    float& operator[](int i) { return (&x)[i]; }
    float operator[](int i) const { return (&x)[i]; }
};
struct vec4 {
    float x, y, z, w;
    vec3 xyz() const { return {x, y, z}; }
};
struct uvec3 {
    uint32_t x, y, z;
};
struct ivec3 {
    int x, y, z;
};

inline vec3 V3(float a) { return {a, a, a}; }
inline vec3 V3(float a, float b, float c) { return {a, b, c}; }
inline vec4 V4(vec3 v, float w) { return {v.x, v.y, v.z, w}; }

inline vec3 operator+(vec3 a, vec3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline vec3 operator-(vec3 a, vec3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline vec3 operator-(vec3 a) { return {-a.x, -a.y, -a.z}; }
inline vec3 operator*(vec3 a, vec3 b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }
inline vec3 operator*(vec3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
inline vec3 operator*(float s, vec3 a) { return {s * a.x, s * a.y, s * a.z}; }
inline vec3 operator/(vec3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }
inline vec3 operator/(vec3 a, vec3 b) { return {a.x / b.x, a.y / b.y, a.z / b.z}; }
inline vec2 operator+(vec2 a, vec2 b) { return {a.x + b.x, a.y + b.y}; }
inline vec2 operator*(float s, vec2 a) { return {s * a.x, s * a.y}; }

inline float dot(vec3 a, vec3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline float dot(vec2 a, vec2 b) { return a.x * b.x + a.y * b.y; }
inline float dot(vec4 a, vec4 b) { return a.x * b.x + a.y * b.y + a.z * b.z + a.w * b.w; }
inline vec3 cross(vec3 a, vec3 b) { return {a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y}; }
inline float length(vec3 a) { return sqrtf(dot(a, a)); }
inline float length(vec2 a) { return sqrtf(dot(a, a)); }
inline float distance(vec3 a, vec3 b) { return length(a - b); }
inline vec3 normalize(vec3 a) {
    float inv = 1.0f / sqrtf(dot(a, a));
    return a * inv;
}
inline vec4 normalize(vec4 a) {
    float inv = 1.0f / sqrtf(dot(a, a));
    return {a.x * inv, a.y * inv, a.z * inv, a.w * inv};
}
inline float mix(float a, float b, float t) { return a * (1.0f - t) + b * t; }
inline vec3 mix(vec3 a, vec3 b, float t) { return a * (1.0f - t) + b * t; }
inline float clamp(float x, float lo, float hi) { return fminf(fmaxf(x, lo), hi); }
inline float max(float a, float b) { return fmaxf(a, b); }
inline float min(float a, float b) { return (b < a) ? b : a; }
inline vec3 abs(vec3 a) { return {fabsf(a.x), fabsf(a.y), fabsf(a.z)}; }
inline float abs(float a) { return fabsf(a); }
inline vec3 reflect(vec3 I, vec3 N) { return I - N * (2.0f * dot(N, I)); }
inline vec3 refract(vec3 I, vec3 N, float eta) {
    float d = dot(N, I);
    float k = 1.0f - eta * eta * (1.0f - d * d);
    if (k < 0.0f) return V3(0.0f);
    return I * eta - N * (eta * d + sqrtf(k));
}
inline uint32_t floatBitsToUint(float f) {
    uint32_t u;
    std::memcpy(&u, &f, 4);
    return u;
}
inline float uintBitsToFloat(uint32_t u) {
    float f;
    std::memcpy(&f, &u, 4);
    return f;
}

// column-major 4x4: c[col] is a vec4
struct mat4 {
    vec4 c[4];
};
inline vec4 operator*(const mat4& m, vec4 v) {
    return {m.c[0].x * v.x + m.c[1].x * v.y + m.c[2].x * v.z + m.c[3].x * v.w,
            m.c[0].y * v.x + m.c[1].y * v.y + m.c[2].y * v.z + m.c[3].y * v.w,
            m.c[0].z * v.x + m.c[1].z * v.y + m.c[2].z * v.z + m.c[3].z * v.w,
            m.c[0].w * v.x + m.c[1].w * v.y + m.c[2].w * v.z + m.c[3].w * v.w};
}

}  // namespace glsl
