// vecmath.cuh — float3 arithmetic with the evaluation order of the GLSL built-ins
// the reference shaders use. This translation unit is compiled with -fmad=false:
// every product and sum below is a separately rounded IEEE binary32 operation,
// evaluated left to right, with IEEE division and square root. That is what makes
// hit records reproducible bit for bit on any IEEE machine (tests compare them
// with a CPU restatement of the shaders). Culling code that only needs to be
// conservative calls __fmaf_rn explicitly.
#pragma once
#include <cuda_runtime.h>
#include <math_constants.h>

namespace mcpt {

#define MCPT_DEV __device__ __forceinline__

struct f3 {
    float x, y, z;
};

MCPT_DEV f3 mk3(float x, float y, float z) { return f3{x, y, z}; }
MCPT_DEV f3 operator+(f3 a, f3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
MCPT_DEV f3 operator-(f3 a, f3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
MCPT_DEV f3 operator-(f3 a) { return {-a.x, -a.y, -a.z}; }
MCPT_DEV f3 operator*(f3 a, f3 b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }
MCPT_DEV f3 operator*(f3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
MCPT_DEV f3 operator*(float s, f3 a) { return {s * a.x, s * a.y, s * a.z}; }
MCPT_DEV f3 operator/(f3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }
MCPT_DEV float dot3(f3 a, f3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
MCPT_DEV float dot2(float ax, float ay, float bx, float by) { return ax * bx + ay * by; }
MCPT_DEV f3 cross3(f3 a, f3 b) {
    return {a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y};
}
MCPT_DEV float length3(f3 a) { return sqrtf(dot3(a, a)); }
MCPT_DEV float distance3(f3 a, f3 b) { return length3(a - b); }
// normalize(v) = v * (1 / sqrt(dot(v,v)))
MCPT_DEV f3 normalize3(f3 a) {
    float inv = 1.0f / sqrtf(dot3(a, a));
    return a * inv;
}
MCPT_DEV f3 mix3(f3 a, f3 b, float t) { return a * (1.0f - t) + b * t; }
MCPT_DEV float mixf(float a, float b, float t) { return a * (1.0f - t) + b * t; }
MCPT_DEV f3 reflect3(f3 I, f3 N) { return I - N * (2.0f * dot3(N, I)); }
// GLSL refract: k = 1 - eta^2 (1 - (N.I)^2); k < 0 -> vec3(0)
MCPT_DEV f3 refract3(f3 I, f3 N, float eta) {
    float d = dot3(N, I);
    float k = 1.0f - eta * eta * (1.0f - d * d);
    if (k < 0.0f) return mk3(0.0f, 0.0f, 0.0f);
    return I * eta - N * (eta * d + sqrtf(k));
}

// Affine part of a column-major mat4, stored as three rows so that
// (M * vec4(v, w)).xyz is three dot products: row r = (M[0][r], M[1][r], M[2][r], M[3][r]).
struct Aff {
    float4 r0, r1, r2;
};
// (M * vec4(p, 1)).xyz
MCPT_DEV f3 xform_point(const Aff& m, f3 p) {
    return {m.r0.x * p.x + m.r0.y * p.y + m.r0.z * p.z + m.r0.w,
            m.r1.x * p.x + m.r1.y * p.y + m.r1.z * p.z + m.r1.w,
            m.r2.x * p.x + m.r2.y * p.y + m.r2.z * p.z + m.r2.w};
}
// (M * vec4(v, 0)).xyz — the "+ M[3][r]*0" term is kept: it is what the shader evaluates
MCPT_DEV f3 xform_vector(const Aff& m, f3 v) {
    return {m.r0.x * v.x + m.r0.y * v.y + m.r0.z * v.z + m.r0.w * 0.0f,
            m.r1.x * v.x + m.r1.y * v.y + m.r1.z * v.z + m.r1.w * 0.0f,
            m.r2.x * v.x + m.r2.y * v.y + m.r2.z * v.z + m.r2.w * 0.0f};
}

MCPT_DEV float comp(f3 v, int i) { return i == 0 ? v.x : (i == 1 ? v.y : v.z); }

}  // namespace mcpt
