// oracle/ref_shader.cpp — TEST INFRASTRUCTURE (CPU oracle). See ref_shader.h.
// Compile with -ffp-contract=off.
#include "ref_shader.h"

#include <cmath>

namespace oracle {

static const float FLT_MAX_ = 3.402823e38f;
static const float PI = 2.0f * acosf(0.0f);
static const float EPSILON = 1e-10f;
static const float BIAS = 1e-2f;
static const int PRIM_WIDTH = 16;
static const int HALF_TEX_WIDTH = TEX_WIDTH / 2;
static const int PRIMS_PER_LINE = TEX_WIDTH / PRIM_WIDTH;

enum { CODE_MESH = 0, CODE_SPHERE = 1, CODE_CUBE = 2, CODE_CYLINDER = 3, CODE_CONE = 4, CODE_ORIENTED_QUAD = 5 };

// ---- texelFetch emulation
static inline vec4 fetch4(const Tex& t, int x, int y) {
    const float* p = &t.f[(size_t(y) * t.width + x) * 4];
    return {p[0], p[1], p[2], p[3]};
}
static inline vec3 fetch3(const Tex& t, int x, int y) {
    const float* p = &t.f[(size_t(y) * t.width + x) * 3];
    return {p[0], p[1], p[2]};
}
static inline ivec3 fetch3i(const Tex& t, int x, int y) {
    const int32_t* p = &t.i[(size_t(y) * t.width + x) * 3];
    return {p[0], p[1], p[2]};
}

// ---- RNG, raytracer_func.frag:90-124
static uint32_t xxhash32(uvec3 p) {
    const uint32_t PRIME32_2 = 2246822519U, PRIME32_3 = 3266489917U;
    const uint32_t PRIME32_4 = 668265263U, PRIME32_5 = 374761393U;
    uint32_t h32 = p.z + PRIME32_5 + p.x * PRIME32_3;
    h32 = PRIME32_4 * ((h32 << 17) | (h32 >> (32 - 17)));
    h32 += p.y * PRIME32_3;
    h32 = PRIME32_4 * ((h32 << 17) | (h32 >> (32 - 17)));
    h32 = PRIME32_2 * (h32 ^ (h32 >> 15));
    h32 = PRIME32_3 * (h32 ^ (h32 >> 13));
    return h32 ^ (h32 >> 16);
}

void Shader::srand() {
    vec3 w = V3(screen_tc.x, +float(U.numero_pass) * 3.14f + U.date, screen_tc.y) * float(1 + U.numero_pass) * 1.125f;
    random_seed = {floatBitsToUint(w.x), floatBitsToUint(w.y), floatBitsToUint(w.z)};
}

float Shader::random_float() {
    uint32_t m = xxhash32(random_seed);
    const uint32_t ieeeMantissa = 0x007FFFFFu;
    const uint32_t ieeeOne = 0x3F800000u;
    m &= ieeeMantissa;
    m |= ieeeOne;
    float f = uintBitsToFloat(m);
    random_seed.x += 11u;
    random_seed.y += 43u;
    random_seed.z += 67u;
    return f - 1.0f;
}

vec3 Shader::random_vec3() {
    float a = random_float(), b = random_float(), c = random_float();
    return V3(a, b, c);
}

// ---- table readers, raytracer_func.frag:138-255
ivec3 Shader::read_mesh_info(int mesh_line) { return fetch3i(T.tex_tri, 0, mesh_line); }

void Shader::read_triangle_pos(int mesh_line, int tri, vec3& Pa, vec3& Pb, vec3& Pc) {
    tri++;  // skip info
    ivec3 indices = fetch3i(T.tex_tri, tri % TEX_WIDTH, mesh_line + tri / TEX_WIDTH);
    Pa = fetch3(T.tex_p, indices.x % TEX_WIDTH, indices.x / TEX_WIDTH);
    Pb = fetch3(T.tex_p, indices.y % TEX_WIDTH, indices.y / TEX_WIDTH);
    Pc = fetch3(T.tex_p, indices.z % TEX_WIDTH, indices.z / TEX_WIDTH);
}

void Shader::read_triangle_PN(int mesh_line, int tri, vec3& Pa, vec3& Na, vec3& Pb, vec3& Nb, vec3& Pc, vec3& Nc) {
    tri++;
    ivec3 indices = fetch3i(T.tex_tri, tri % TEX_WIDTH, mesh_line + tri / TEX_WIDTH);
    Pa = fetch3(T.tex_p, indices.x % TEX_WIDTH, indices.x / TEX_WIDTH);
    Na = fetch3(T.tex_n, indices.x % TEX_WIDTH, indices.x / TEX_WIDTH);
    Pb = fetch3(T.tex_p, indices.y % TEX_WIDTH, indices.y / TEX_WIDTH);
    Nb = fetch3(T.tex_n, indices.y % TEX_WIDTH, indices.y / TEX_WIDTH);
    Pc = fetch3(T.tex_p, indices.z % TEX_WIDTH, indices.z / TEX_WIDTH);
    Nc = fetch3(T.tex_n, indices.z % TEX_WIDTH, indices.z / TEX_WIDTH);
}

int Shader::read_primitive_type(int i, int& mesh_line) {
    int l = i / PRIMS_PER_LINE;
    int c = (i % PRIMS_PER_LINE) * PRIM_WIDTH;
    vec4 texel = fetch4(T.tex_prim, c + 12, l);
    mesh_line = int(texel.y);
    return int(texel.x);
}

mat4 Shader::read_mat(int i, int first) {
    int l = i / PRIMS_PER_LINE;
    int c = (i % PRIMS_PER_LINE) * PRIM_WIDTH + first;
    mat4 m;
    for (int k = 0; k < 4; ++k) m.c[k] = fetch4(T.tex_prim, c + k, l);
    return m;
}

vec4 Shader::read_color(int i) {
    int l = i / PRIMS_PER_LINE;
    int c = (i % PRIMS_PER_LINE) * PRIM_WIDTH;
    return fetch4(T.tex_prim, c + 13, l);
}

vec4 Shader::read_material(int i) {
    int l = i / PRIMS_PER_LINE;
    int c = (i % PRIMS_PER_LINE) * PRIM_WIDTH;
    return fetch4(T.tex_prim, c + 14, l);
}

int Shader::read_bvh_ind(int bl, int i) { return T.tex_ind.i[size_t(i / TEX_WIDTH + bl) * TEX_WIDTH + i % TEX_WIDTH]; }

void Shader::read_bvh_bb(int bl, int i, vec3& bbmin, vec3& bbmax) {
    int l = i / HALF_TEX_WIDTH + bl;
    int c = (i % HALF_TEX_WIDTH) * 2;
    bbmin = fetch3(T.tex_bb, c, l);
    bbmax = fetch3(T.tex_bb, c + 1, l);
}

// ---- AABB tests, raytracer_func.frag:273-352
static inline bool inside_unit(vec3 Oi) { return fabsf(Oi.x) < 1.0f && fabsf(Oi.y) < 1.0f && fabsf(Oi.z) < 1.0f; }

// nearest entry parameter over the six faces of the unit cube (shared loop of
// intersect_bv / intersect_bvm / Cube_intersect); returns FLT_MAX if none
static inline float unit_cube_entry(vec3 Oi, vec3 Di, int& cl) {
    float al = FLT_MAX_;
    for (int c = 0; c < 6; ++c) {
        int c0 = c / 2;
        if (fabsf(Di[c0]) > EPSILON) {
            int c1 = (c0 + 1) % 3;
            int c2 = (c0 + 2) % 3;
            float cd = -1.0f + 2.0f * float(c % 2);
            float a = (cd - Oi[c0]) / Di[c0];
            if ((a > EPSILON) && (fabsf(Oi[c1] + a * Di[c1]) <= 1.0f) && (fabsf(Oi[c2] + a * Di[c2]) <= 1.0f))
                if (a < al) {
                    al = a;
                    cl = c;
                }
        }
    }
    return al;
}

bool Shader::intersect_bvm(int bl, int i, vec3 O, vec3 D, vec3 Ol, const mat4& trf) {
    ++cnt.box_tests;
    vec3 bbmin, bbmax;
    read_bvh_bb(bl, i, bbmin, bbmax);
    vec3 center = (bbmin + bbmax) / 2.0f;
    vec3 width = 0.5f * (bbmax - bbmin);
    vec3 Oi = (O - center) / width;
    vec3 Di = D / width;
    if (inside_unit(Oi)) return true;
    int cl = 0;
    float al = unit_cube_entry(Oi, Di, cl);
    if (al < FLT_MAX_) {
        vec3 Pl = Oi + al * Di;
        vec3 Pg = (trf * V4(Pl * width + center, 1.0f)).xyz();
        float dist = distance(Ol, Pg);
        return (dist <= closest_intersection.dist);
    }
    return false;
}

bool Shader::intersect_bv(int bl, int i, vec3 O, vec3 D, vec3 Ol) {
    ++cnt.box_tests;
    vec3 bbmin, bbmax;
    read_bvh_bb(bl, i, bbmin, bbmax);
    vec3 center = (bbmin + bbmax) / 2.0f;
    vec3 width = 0.5f * (bbmax - bbmin);
    vec3 Oi = (O - center) / width;
    vec3 Di = D / width;
    if (inside_unit(Oi)) return true;
    int cl = 0;
    float al = unit_cube_entry(Oi, Di, cl);
    if (al < FLT_MAX_) {
        vec3 Pg = (al * Di + Oi) * width + center;
        float dist = distance(Ol, Pg);
        return (dist <= closest_intersection.dist);
    }
    return false;
}

// common tail of every intersector: Pg = trf*Pl, dist = |Ol-Pg|, keep if closer
void Shader::candidate(int index, int shape, int dir, vec3 Pl, vec3 Ol) {
    mat4 trf = read_transfo(index);
    vec3 Pg = (trf * V4(Pl, 1.0f)).xyz();
    float dist = distance(Ol, Pg);
    if (dist < closest_intersection.dist) {
        closest_intersection.pl = Pl;
        closest_intersection.pg = Pg;
        closest_intersection.dist = dist;
        closest_intersection.dir = dir;
        closest_intersection.index = index;
        closest_intersection.shape = shape;
    }
}

// raytracer_func.frag:354-396
void Shader::Triangle_intersect(int mesh_line, int tri_index, int bb_index, vec3 O, vec3 D, vec3 Ol, const mat4& trf) {
    ++cnt.tri_tests;
    vec3 vA, vB, vC;
    read_triangle_pos(mesh_line, tri_index, vA, vB, vC);
    vec3 edge1 = vB - vA;
    vec3 edge2 = vC - vA;
    vec3 h = cross(D, edge2);
    float det = dot(edge1, h);
    if (fabsf(det) < EPSILON) return;
    float invdet = 1.0f / det;
    vec3 s = O - vA;
    float u = dot(s, h) * invdet;
    if (u < 0.0f || u > 1.0f) return;
    vec3 q = cross(s, edge1);
    float v = dot(D, q) * invdet;
    if (v < 0.0f || (u + v) > 1.0f) return;
    float a = dot(edge2, q) * invdet;
    if (a > EPSILON) {
        vec3 Pl = O + a * D;
        vec3 Pg = (trf * V4(Pl, 1.0f)).xyz();
        float dist = distance(Ol, Pg);
        if (dist < closest_intersection.dist) {
            closest_intersection.pl = Pl;
            closest_intersection.pg = Pg;
            closest_intersection.dist = dist;
            closest_intersection.tri_index = tri_index;
            closest_intersection.index = bb_index;
            closest_intersection.mesh_line = mesh_line;
            closest_intersection.shape = CODE_MESH;
            closest_intersection.dir = 0;
        }
    }
}

// raytracer_func.frag:398-441
void Shader::Sphere_intersect(int index, vec3 O, vec3 D, vec3 Ol) {
    float OO = dot(O, O);
    float OD = dot(O, D);
    float D2 = dot(D, D);
    float delta4 = OD * OD - D2 * (OO - 1.0f);
    if (delta4 > 0) {
        float a = -(OD + sqrtf(delta4)) / D2;
        if (a > EPSILON) candidate(index, CODE_SPHERE, 0, O + a * D, Ol);
        a = -(OD - sqrtf(delta4)) / D2;
        if (a > EPSILON) candidate(index, CODE_SPHERE, 0, O + a * D, Ol);
    }
}

// raytracer_func.frag:443-470
void Shader::OrientedQuad_intersect(int index, vec3 O, vec3 D, vec3 Ol) {
    if (D.z > -EPSILON) return;
    float a = -O.z / D.z;
    vec3 Pl = O + a * D;
    if (fabsf(Pl.x) > 1 || fabsf(Pl.y) > 1) return;
    candidate(index, CODE_ORIENTED_QUAD, 0, Pl, Ol);
}

// raytracer_func.frag:472-512
void Shader::Cube_intersect(int index, vec3 O, vec3 D, vec3 Ol) {
    int cl = 0;
    float al = unit_cube_entry(O, D, cl);
    if (al < FLT_MAX_) candidate(index, CODE_CUBE, cl, O + al * D, Ol);
}

// raytracer_func.frag:515-577
void Shader::Cylinder_intersect(int index, vec3 O, vec3 D, vec3 Ol) {
    int cl = -1;
    float al = FLT_MAX_;
    if (fabsf(D.z) > EPSILON) {
        float a = (-1.0f - O.z) / D.z;
        if (a > EPSILON) {
            vec2 R = vec2{O.x, O.y} + a * vec2{D.x, D.y};
            if ((dot(R, R) < 1.0f) && (a < al)) {
                cl = 0;
                al = a;
            }
        }
        a = (1.0f - O.z) / D.z;
        if (a > EPSILON) {
            vec2 R = vec2{O.x, O.y} + a * vec2{D.x, D.y};
            if ((dot(R, R) < 1.0f) && (a < al)) {
                cl = 1;
                al = a;
            }
        }
    }
    vec2 Oxy{O.x, O.y}, Dxy{D.x, D.y};
    float O2 = dot(Oxy, Oxy);
    float OD = dot(Oxy, Dxy);
    float D2 = dot(Dxy, Dxy);
    float delta4 = OD * OD - D2 * (O2 - 1);
    if (delta4 > 0.0f) {
        float a = -(OD + sqrtf(delta4)) / D2;
        if ((a > EPSILON) && (a < al)) {
            float z = O.z + a * D.z;
            if (fabsf(z) < 1.0f) {
                cl = 2;
                al = a;
            }
        }
    }
    if (al < FLT_MAX_) candidate(index, CODE_CYLINDER, cl, O + al * D, Ol);
}

// raytracer_func.frag:579-640
void Shader::Cone_intersect(int index, vec3 O, vec3 D, vec3 Ol) {
    int cl = -1;
    float tl = FLT_MAX_;
    if (fabsf(D.z) > EPSILON) {
        float t0 = (-1.0f - O.z) / D.z;
        if (t0 > EPSILON) {
            vec2 R = vec2{O.x, O.y} + t0 * vec2{D.x, D.y};
            if ((dot(R, R) < 1.0f) && (t0 < tl)) {
                cl = 0;
                tl = t0;
            }
        }
    }
    vec3 co = O;
    co.z -= 1.0f;
    float a = D.z * D.z - 0.8f;
    float b = 2.0f * (D.z * co.z - dot(D, co) * 0.8f);
    float c = co.z * co.z - dot(co, co) * 0.8f;
    float det = b * b - 4.0f * a * c;
    if (det > 0.0f) {
        det = sqrtf(det);
        float t1 = (-b - det) / (2.0f * a);
        if (fabsf(O.z + t1 * D.z) > 1) t1 = FLT_MAX_;
        float t2 = (-b + det) / (2.0f * a);
        if (fabsf(O.z + t2 * D.z) > 1) t2 = FLT_MAX_;
        float t = glsl::min(t1, t2);
        if (t < tl) {
            cl = 2;
            tl = t;
        }
    }
    if (tl < FLT_MAX_) candidate(index, CODE_CONE, cl, O + tl * D, Ol);
}

}  // namespace oracle
