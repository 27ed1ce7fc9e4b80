// oracle/ref_shim/glsl_compat.h — TEST INFRASTRUCTURE. Not part of the product.
//
// GLSL 4.30 types and built-ins as C++, so that the reference's shader TEXT
// (shaders/raytracer_func.frag + tp/*.frag + shaders/main.frag, shaders/raytracer.vert),
// after the mechanical rewriting of oracle/build_ref.py (qualifiers dropped, `out T x` ->
// `T& x`, float literals suffixed, swizzles turned into calls), compiles as the body of a C++
// class and runs on the host. No graphics API is involved.
//
// GLSL leaves evaluation order and rounding of the built-ins to the driver; this header fixes
// one choice — every operator a separately rounded binary32 operation (build with
// -ffp-contract=off), sums left to right, IEEE sqrt and divide, the formulae of the GLSL 4.30
// specification section 8 — the same choice as oracle/glsl.h, which the hand-written oracle uses.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>

namespace glslc {

typedef uint32_t uint;

struct vec2 {
    float x, y;
    vec2() : x(0), y(0) {}
    explicit vec2(float a) : x(a), y(a) {}
    vec2(float a, float b) : x(a), y(b) {}
    float& operator[](int i) { return (&x)[i]; }
    float operator[](int i) const { return (&x)[i]; }
};

struct vec3 {
    union {
        struct {
            float x, y, z;
        };
        struct {
            float r, g, b;
        };
    };
    vec3() : x(0), y(0), z(0) {}
    explicit vec3(float a) : x(a), y(a), z(a) {}
    vec3(float a, float b, float c) : x(a), y(b), z(c) {}
    vec3(vec2 v, float c) : x(v.x), y(v.y), z(c) {}
    float& operator[](int i) { return (&x)[i]; }
    float operator[](int i) const { return (&x)[i]; }
    vec2 xy() const { return vec2(x, y); }
    vec3 xyz() const { return *this; }
    vec3 rgb() const { return *this; }
};

struct vec4 {
    union {
        struct {
            float x, y, z, w;
        };
        struct {
            float r, g, b, a;
        };
    };
    vec4() : x(0), y(0), z(0), w(0) {}
    explicit vec4(float s) : x(s), y(s), z(s), w(s) {}
    vec4(float a_, float b_, float c_, float d_) : x(a_), y(b_), z(c_), w(d_) {}
    vec4(vec3 v, float d_) : x(v.x), y(v.y), z(v.z), w(d_) {}
    vec4(vec2 v, float c_, float d_) : x(v.x), y(v.y), z(c_), w(d_) {}
    float& operator[](int i) { return (&x)[i]; }
    float operator[](int i) const { return (&x)[i]; }
    vec2 xy() const { return vec2(x, y); }
    vec3 xyz() const { return vec3(x, y, z); }
    vec3 rgb() const { return vec3(x, y, z); }
};

struct ivec2 {
    int x, y;
    ivec2() : x(0), y(0) {}
    ivec2(int a, int b) : x(a), y(b) {}
};
struct ivec3 {
    int x, y, z;
    ivec3() : x(0), y(0), z(0) {}
    ivec3(int a, int b, int c) : x(a), y(b), z(c) {}
    ivec3 xyz() const { return *this; }
    ivec3 rgb() const { return *this; }
};
struct ivec4 {
    int x, y, z, w;
    ivec4() : x(0), y(0), z(0), w(0) {}
    ivec4(int a, int b, int c, int d) : x(a), y(b), z(c), w(d) {}
    ivec3 xyz() const { return ivec3(x, y, z); }
    ivec3 rgb() const { return ivec3(x, y, z); }
};
struct uvec3 {
    uint x, y, z;
    uvec3() : x(0), y(0), z(0) {}
    uvec3(uint a, uint b, uint c) : x(a), y(b), z(c) {}
};
struct bvec3 {
    bool x, y, z;
};

// ---- vec3
inline vec3 operator+(vec3 a, vec3 b) { return vec3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline vec3 operator-(vec3 a, vec3 b) { return vec3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline vec3 operator-(vec3 a) { return vec3(-a.x, -a.y, -a.z); }
inline vec3 operator*(vec3 a, vec3 b) { return vec3(a.x * b.x, a.y * b.y, a.z * b.z); }
inline vec3 operator*(vec3 a, float s) { return vec3(a.x * s, a.y * s, a.z * s); }
inline vec3 operator*(float s, vec3 a) { return vec3(s * a.x, s * a.y, s * a.z); }
inline vec3 operator/(vec3 a, float s) { return vec3(a.x / s, a.y / s, a.z / s); }
inline vec3 operator/(vec3 a, vec3 b) { return vec3(a.x / b.x, a.y / b.y, a.z / b.z); }
inline vec3 operator+(vec3 a, float s) { return vec3(a.x + s, a.y + s, a.z + s); }
inline vec3 operator-(vec3 a, float s) { return vec3(a.x - s, a.y - s, a.z - s); }
inline vec3& operator+=(vec3& a, vec3 b) { return a = a + b; }
inline vec3& operator-=(vec3& a, vec3 b) { return a = a - b; }
inline vec3& operator*=(vec3& a, float s) { return a = a * s; }
inline vec3& operator*=(vec3& a, vec3 b) { return a = a * b; }
// ---- vec2
inline vec2 operator+(vec2 a, vec2 b) { return vec2(a.x + b.x, a.y + b.y); }
inline vec2 operator-(vec2 a, vec2 b) { return vec2(a.x - b.x, a.y - b.y); }
inline vec2 operator*(float s, vec2 a) { return vec2(s * a.x, s * a.y); }
inline vec2 operator*(vec2 a, float s) { return vec2(a.x * s, a.y * s); }
inline vec2 operator-(vec2 a, float s) { return vec2(a.x - s, a.y - s); }
inline vec2 operator+(vec2 a, float s) { return vec2(a.x + s, a.y + s); }
inline vec2 operator/(vec2 a, float s) { return vec2(a.x / s, a.y / s); }
// ---- vec4
inline vec4 operator+(vec4 a, vec4 b) { return vec4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
inline vec4 operator-(vec4 a, vec4 b) { return vec4(a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w); }
inline vec4 operator*(vec4 a, float s) { return vec4(a.x * s, a.y * s, a.z * s, a.w * s); }
inline vec4 operator*(float s, vec4 a) { return vec4(s * a.x, s * a.y, s * a.z, s * a.w); }
inline vec4 operator/(vec4 a, float s) { return vec4(a.x / s, a.y / s, a.z / s, a.w / s); }
// ---- uvec3 (the unused hash / pcg3d helpers of raytracer_func.frag:63-90 must still compile)
inline uvec3 operator*(uvec3 a, uint s) { return uvec3(a.x * s, a.y * s, a.z * s); }
inline uvec3 operator+(uvec3 a, uint s) { return uvec3(a.x + s, a.y + s, a.z + s); }
inline uvec3 operator+(uvec3 a, uvec3 b) { return uvec3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline uvec3 operator>>(uvec3 a, uint s) { return uvec3(a.x >> s, a.y >> s, a.z >> s); }
inline uvec3 operator^(uvec3 a, uvec3 b) { return uvec3(a.x ^ b.x, a.y ^ b.y, a.z ^ b.z); }
inline uvec3& operator^=(uvec3& a, uvec3 b) { return a = a ^ b; }
inline uvec3& operator+=(uvec3& a, uvec3 b) { return a = a + b; }

// ---- scalar built-ins (named like GLSL; found before ::sqrt etc. by ordinary lookup)
inline float sqrt(float x) { return ::sqrtf(x); }
inline float sin(float x) { return ::sinf(x); }
inline float cos(float x) { return ::cosf(x); }
inline float acos(float x) { return ::acosf(x); }
inline float asin(float x) { return ::asinf(x); }
inline float log(float x) { return ::logf(x); }
inline float pow(float x, float y) { return ::powf(x, y); }
inline float abs(float x) { return ::fabsf(x); }
inline int abs(int x) { return x < 0 ? -x : x; }
inline float fract(float x) { return x - ::floorf(x); }
inline float max(float a, float b) { return ::fmaxf(a, b); }
inline float min(float a, float b) { return (b < a) ? b : a; }
inline float clamp(float x, float lo, float hi) { return ::fminf(::fmaxf(x, lo), hi); }
inline float mix(float a, float b, float t) { return a * (1.0f - t) + b * t; }
inline vec3 mix(vec3 a, vec3 b, float t) { return a * (1.0f - t) + b * t; }
inline vec3 abs(vec3 a) { return vec3(::fabsf(a.x), ::fabsf(a.y), ::fabsf(a.z)); }

inline float dot(vec2 a, vec2 b) { return a.x * b.x + a.y * b.y; }
inline float dot(vec3 a, vec3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline float dot(vec4 a, vec4 b) { return a.x * b.x + a.y * b.y + a.z * b.z + a.w * b.w; }
inline vec3 cross(vec3 a, vec3 b) { return vec3(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y); }
inline float length(vec2 a) { return ::sqrtf(dot(a, a)); }
inline float length(vec3 a) { return ::sqrtf(dot(a, a)); }
inline float distance(vec3 a, vec3 b) { return length(a - b); }
inline vec3 normalize(vec3 a) {
    float inv = 1.0f / ::sqrtf(dot(a, a));
    return a * inv;
}
inline vec4 normalize(vec4 a) {
    float inv = 1.0f / ::sqrtf(dot(a, a));
    return a * inv;
}
inline vec3 reflect(vec3 I, vec3 N) { return I - N * (2.0f * dot(N, I)); }
inline vec3 refract(vec3 I, vec3 N, float eta) {
    float d = dot(N, I);
    float k = 1.0f - eta * eta * (1.0f - d * d);
    if (k < 0.0f) return vec3(0.0f);
    return I * eta - N * (eta * d + ::sqrtf(k));
}
inline bvec3 lessThan(vec3 a, vec3 b) { return bvec3{a.x < b.x, a.y < b.y, a.z < b.z}; }
inline bool all(bvec3 b) { return b.x && b.y && b.z; }
inline bool any(bvec3 b) { return b.x || b.y || b.z; }

inline uint floatBitsToUint(float f) {
    uint u;
    std::memcpy(&u, &f, 4);
    return u;
}
inline uvec3 floatBitsToUint(vec3 v) { return uvec3(floatBitsToUint(v.x), floatBitsToUint(v.y), floatBitsToUint(v.z)); }
inline float uintBitsToFloat(uint u) {
    float f;
    std::memcpy(&f, &u, 4);
    return f;
}

// ---- matrices: column-major, constructed from columns
struct mat3 {
    vec3 c[3];
    mat3() {}
    mat3(vec3 a, vec3 b, vec3 d) {
        c[0] = a;
        c[1] = b;
        c[2] = d;
    }
};
inline vec3 operator*(const mat3& m, vec3 v) {
    return vec3(m.c[0].x * v.x + m.c[1].x * v.y + m.c[2].x * v.z, m.c[0].y * v.x + m.c[1].y * v.y + m.c[2].y * v.z,
                m.c[0].z * v.x + m.c[1].z * v.y + m.c[2].z * v.z);
}
struct mat4 {
    vec4 c[4];
    mat4() {}
    mat4(vec4 a, vec4 b, vec4 d, vec4 e) {
        c[0] = a;
        c[1] = b;
        c[2] = d;
        c[3] = e;
    }
};
inline vec4 operator*(const mat4& m, vec4 v) {
    return vec4(m.c[0].x * v.x + m.c[1].x * v.y + m.c[2].x * v.z + m.c[3].x * v.w,
                m.c[0].y * v.x + m.c[1].y * v.y + m.c[2].y * v.z + m.c[3].y * v.w,
                m.c[0].z * v.x + m.c[1].z * v.y + m.c[2].z * v.z + m.c[3].z * v.w,
                m.c[0].w * v.x + m.c[1].w * v.y + m.c[2].w * v.z + m.c[3].w * v.w);
}

// ---- samplers: texelFetch over host arrays (NEAREST, lod 0). Missing components read as
// (0, 0, 0, 1) like an OpenGL texture of that internal format.
struct sampler2D {
    const float* data = nullptr;
    int width = 0, height = 0, comps = 0;
};
struct isampler2D {
    const int32_t* data = nullptr;
    int width = 0, height = 0, comps = 0;
};
// Out-of-range fetches return zeros (robust-buffer-access behaviour) instead of reading wild.
inline vec4 texelFetch(const sampler2D& s, ivec2 p, int /*lod*/) {
    if (p.x < 0 || p.y < 0 || p.x >= s.width || p.y >= s.height) return vec4(0.0f);
    const float* t = s.data + (size_t(p.y) * s.width + p.x) * s.comps;
    vec4 v(0.0f, 0.0f, 0.0f, 1.0f);
    for (int k = 0; k < s.comps; ++k) v[k] = t[k];
    return v;
}
inline ivec4 texelFetch(const isampler2D& s, ivec2 p, int /*lod*/) {
    if (p.x < 0 || p.y < 0 || p.x >= s.width || p.y >= s.height) return ivec4(0, 0, 0, 0);
    const int32_t* t = s.data + (size_t(p.y) * s.width + p.x) * s.comps;
    ivec4 v(0, 0, 0, 1);
    if (s.comps > 0) v.x = t[0];
    if (s.comps > 1) v.y = t[1];
    if (s.comps > 2) v.z = t[2];
    if (s.comps > 3) v.w = t[3];
    return v;
}

}  // namespace glslc
