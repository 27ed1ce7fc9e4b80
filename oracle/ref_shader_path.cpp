// oracle/ref_shader_path.cpp — TEST INFRASTRUCTURE (CPU oracle). See ref_shader.h.
// Second half: tree walks, geometry info, and the tp/ bounce shaders.
// Compile with -ffp-contract=off.
#include <cmath>

#include "ref_shader.h"

namespace oracle {

static const float FLT_MAX_ = 3.402823e38f;
static const float PI = 2.0f * acosf(0.0f);
static const float BIAS = 1e-2f;

// Sensitivity probe (tests only): move the result of every transcendental call of the bounce sampler by
// g_ulp_perturb ulps. GLSL leaves sin / cos / log / pow to the driver with errors of several ulps, so two
// conforming implementations of the reference differ by this much; the tests use it to measure how far fixed-seed
// images of a scene may drift apart for that reason alone (chaotic scenes: tests/test_oracle_golden.py).
int g_ulp_perturb = 0;
static inline float ulps(float v) {
    for (int k = 0; k < g_ulp_perturb; ++k) v = nextafterf(v, INFINITY);
    for (int k = 0; k > g_ulp_perturb; --k) v = nextafterf(v, -INFINITY);
    return v;
}
enum { CODE_MESH = 0, CODE_SPHERE = 1, CODE_CUBE = 2, CODE_CYLINDER = 3, CODE_CONE = 4, CODE_ORIENTED_QUAD = 5 };

// raytracer_func.frag:642-678
void Shader::Mesh_intersect(int mesh_line, int bb_index, vec3 O, vec3 D, vec3 Ol) {
    int stack_ind[29];
    int head = 1;
    stack_ind[0] = 0;
    ivec3 mesh_info = read_mesh_info(mesh_line);
    int max_line_depth = mesh_info.z;
    mat4 trfm = read_mesh_transfo(bb_index);
    while (head > 0) {
        head--;
        int i = stack_ind[head];
        if (i >= max_line_depth) {
            int p = read_bvh_ind(mesh_info.y, i - max_line_depth);
            if (p >= 0) {
                Triangle_intersect(mesh_line, p, bb_index, O, D, Ol, trfm);
                if (hit_only && closest_intersection.shape >= 0) return;
            }
        } else {
            int j = 2 * i + 1;
            if (intersect_bvm(mesh_info.x, j, O, D, Ol, trfm)) stack_ind[head++] = j;
            j++;
            if (intersect_bvm(mesh_info.x, j, O, D, Ol, trfm)) stack_ind[head++] = j;
        }
    }
}

// raytracer_func.frag:681-705
void Shader::intersect_prim(int i, vec3 O, vec3 D) {
    ++cnt.prim_tests;
    int mesh_line = -1;
    int prim_t = read_primitive_type(i, mesh_line);
    if (prim_t < 0) return;
    mat4 inv = read_inv_transfo(i);
    vec3 Oi = (inv * V4(O, 1.0f)).xyz();
    vec3 Di = normalize(inv * V4(D, 0.0f)).xyz();
    switch (prim_t) {
        case CODE_MESH: Mesh_intersect(mesh_line, i, Oi, Di, O); break;
        case CODE_SPHERE: Sphere_intersect(i, Oi, Di, O); break;
        case CODE_CUBE: Cube_intersect(i, Oi, Di, O); break;
        case CODE_CYLINDER: Cylinder_intersect(i, Oi, Di, O); break;
        case CODE_CONE: Cone_intersect(i, Oi, Di, O); break;
        case CODE_ORIENTED_QUAD: OrientedQuad_intersect(i, Oi, Di, O); break;
    }
}

// raytracer_func.frag:708-716
void Shader::reset_inter() {
    closest_intersection.index = -1;
    closest_intersection.dist = FLT_MAX_;
    closest_intersection.tri_index = -1;
    closest_intersection.shape = -1;
    closest_intersection.dir = -1;
}

// raytracer_func.frag:734-769
void Shader::intersect_bvh(vec3 O, vec3 D) {
    ++cnt.traversals;
    int stack_ind[29];
    int head = 1;
    stack_ind[0] = 0;
    reset_inter();
    int max_line_depth = int(powf(2.0f, float(U.bvh_depth)) - 1.0f);
    while (head > 0) {
        head--;
        int i = stack_ind[head];
        if (i >= max_line_depth) {
            int p = read_bvh_ind(0, i - max_line_depth);
            if (p >= 0) {
                intersect_prim(p, O, D);
                if (hit_only && closest_intersection.shape >= 0) return;
            }
        } else {
            int j = 2 * i + 1;
            if (intersect_bv(0, j, O, D, O)) stack_ind[head++] = j;
            j++;
            if (intersect_bv(0, j, O, D, O)) stack_ind[head++] = j;
        }
    }
}

void Shader::just_hit_bvh(vec3 O, vec3 D) {
    hit_only = true;
    intersect_bvh(O, D);
}

void Shader::traverse_all_bvh(vec3 O, vec3 D) {
    hit_only = false;
    intersect_bvh(O, D);
}

// raytracer_func.frag:783-897. N and Pg are GLSL `out` parameters: a shape that
// writes nothing (shape < 0, cone dir outside {0,1,2}) leaves the caller's
// variables as they were.
void Shader::intersection_info(vec3& N, vec3& Pg) {
    const sInter& ci = closest_intersection;
    switch (ci.shape) {
        case CODE_MESH: {
            vec3 A, B, C, nA, nB, nC;
            read_triangle_PN(ci.mesh_line, ci.tri_index, A, nA, B, nB, C, nC);
            mat4 transfo = read_mesh_transfo(ci.index);
            Pg = ci.pg;
            if (U.flat_face)
                N = normalize((transfo * V4(ci.pl + cross(B - A, C - A), 1.0f)).xyz() - Pg);
            else {
                vec3 PA = A - ci.pl;
                vec3 PB = B - ci.pl;
                vec3 PC = C - ci.pl;
                float tA = length(cross(PB, PC));
                float tB = length(cross(PA, PC));
                float tC = length(cross(PA, PB));
                vec3 No = nA * tA + nB * tB + nC * tC;
                N = normalize((transfo * V4(ci.pl + No, 1.0f)).xyz() - Pg);
            }
            break;
        }
        case CODE_SPHERE: {
            Pg = ci.pg;
            vec3 Pn = 2.0f * ci.pl;
            mat4 transfo = read_transfo(ci.index);
            N = normalize((transfo * V4(Pn, 1.0f)).xyz() - Pg);
            break;
        }
        case CODE_CUBE: {
            Pg = ci.pg;
            vec3 No = V3(0.0f);
            No[ci.dir / 2] = (ci.dir % 2 != 0) ? 1.0f : -1.0f;
            mat4 transfo = read_transfo(ci.index);
            N = normalize((transfo * V4(ci.pl + No, 1.0f)).xyz() - Pg);
            break;
        }
        case CODE_CYLINDER: {
            Pg = ci.pg;
            vec3 No = V3(0.0f);
            if (ci.dir < 2)
                No.z = (ci.dir % 2 != 0) ? 1.0f : -1.0f;
            else
                No = V3(ci.pl.x, ci.pl.y, 0.0f);
            mat4 transfo = read_transfo(ci.index);
            N = normalize((transfo * V4(ci.pl + No, 1.0f)).xyz() - Pg);
            break;
        }
        case CODE_CONE: {
            Pg = ci.pg;
            mat4 transfo = read_transfo(ci.index);
            switch (ci.dir) {
                case 0: N = normalize((transfo * vec4{ci.pl.x, ci.pl.y, ci.pl.z - 1.0f, 1.0f}).xyz() - Pg); break;
                case 1: N = V3(0, 0, 0); break;
                case 2: {
                    vec3 No = V3(ci.pl.x, ci.pl.y, length(vec2{ci.pl.x, ci.pl.y}) / 2.0f);
                    N = normalize((transfo * V4(ci.pl + No, 1.0f)).xyz() - Pg);
                    break;
                }
            }
            break;
        }
        case CODE_ORIENTED_QUAD: {
            Pg = ci.pg;
            mat4 transfo = read_transfo(ci.index);
            vec3 No = V3(0, 0, 1);
            N = normalize((transfo * V4(ci.pl + No, 1.0f)).xyz() - Pg);
            break;
        }
    }
}

vec4 Shader::intersection_color_info() { return read_color(closest_intersection.index); }
vec4 Shader::intersection_mat_info() { return read_material(closest_intersection.index); }

// ---- tp/montecarlo.frag:49-70
vec3 Shader::sample_hemisphere(float roughness) {
    float alpha = roughness * roughness;
    float beta = 2.0f * PI * random_float();
    float tanTheta2 = -alpha * alpha * ulps(logf(1.0f - random_float()));
    float cosTheta = 1.0f / sqrtf(1.0f + tanTheta2);
    float sinTheta = sqrtf(glsl::max(0.0f, 1.0f - cosTheta * cosTheta));
    vec3 localDir = V3(ulps(cosf(beta)) * sinTheta, ulps(sinf(beta)) * sinTheta, cosTheta);
    return normalize(localDir);
}

// tp/montecarlo.frag:72-89
vec3 Shader::random_ray(vec3 D, float roughness) {
    vec3 W = normalize(V3(D.x, D.y + 5.0f, D.z + 3.0f));
    vec3 Uv = normalize(cross(D, W));
    vec3 V = normalize(cross(D, Uv));
    vec3 s = sample_hemisphere(roughness);
    // mat3 M = mat3(U,V,D); M * s
    vec3 r = V3(Uv.x * s.x + V.x * s.y + D.x * s.z, Uv.y * s.x + V.y * s.y + D.y * s.z, Uv.z * s.x + V.z * s.y + D.z * s.z);
    return normalize(r);
}

// tp/montecarlo.frag:91-98
float Shader::rSchlick(vec3 I, vec3 N) {
    float r0 = (U.refract_ind - 1.0f) / (U.refract_ind + 1.0f);
    r0 *= r0;
    float x = 1.0f - dot(N, I);
    return clamp(r0 + (1.0f - r0) * x * x * x * x * x, 0.0f, 1.0f);
}

// tp/montecarlo.frag:100-179. The GLSL stack (:5-47) never holds more than one
// entry (one empile per iteration, one depile at the top), so stack_full() is
// never true; it is kept as a one-slot stack here.
vec3 Shader::random_path(vec3 Dir, vec3 Orig) {
    vec3 D = Dir;
    vec3 O = Orig;
    vec3 total = V3(0.0f);
    struct {
        vec3 O, D, Attenu;
    } stack[9];
    int stHead = -1;
    auto empile = [&](vec3 o, vec3 d, vec3 a) {
        ++stHead;
        stack[stHead].O = o;
        stack[stHead].D = d;
        stack[stHead].Attenu = a;
    };
    empile(O, D, V3(0.8f));
    for (int i = 0; (i < U.NB_BOUNCES) && !(stHead < 0); ++i) {
        vec3 attenu;
        O = stack[stHead].O;
        D = stack[stHead].D;
        attenu = stack[stHead].Attenu;
        stHead--;
        traverse_all_bvh(O, D);
        if (!hit()) return total + attenu * mix(V3(0.5f, 0.5f, 0.9f), V3(1.0f, 1.0f, 0.8f), glsl::max(0.0f, D.z));
        ++cnt.segments;
        vec3 new_attenu;
        vec3 N = V3(0.0f);  // GLSL leaves these undefined until intersection_info writes them
        vec3 P = V3(0.0f);
        intersection_info(N, P);
        vec4 mat = intersection_mat_info();
        vec4 col = intersection_color_info();
        vec3 colrgb = col.xyz();
        vec3 ray = random_ray(N, 1 - mat.y);
        float rs = rSchlick(D, N);
        vec3 R = reflect(-ray, N);
        vec3 E = normalize(O - P);
        float se = mix(100.0f, 2.0f, mat.y);
        float spec = ulps(powf(glsl::max(0.0f, dot(E, R)), se));
        total = total + (colrgb * 0.1f + attenu * mat.z * (1.0f - mat.x) * col.w);
        if (mat.z <= 0.5f && !(stHead == 9 - 1)) {
            if (mat.x > 0 && col.w == 1) {  // reflection
                new_attenu = colrgb * attenu + attenu * col.w * rs * spec * mix(attenu, colrgb, mat.x);
                vec3 nd = random_ray(reflect(D, N), 1 - mat.x * mat.y);
                empile(P + BIAS * N, nd, new_attenu);
            } else if (col.w < 1 && mat.x == 0) {  // refraction
                new_attenu = colrgb * attenu + attenu * (1.0f - col.w) * (1.0f - rs) * spec * mix(attenu, colrgb, mat.x);
                O = P - BIAS * N;
                D = refract(D, N, U.refract_ind);
                traverse_all_bvh(O, D);
                intersection_info(N, P);
                empile(P + BIAS * N, refract(D, -N, 1 / U.refract_ind), new_attenu);
            } else if (col.w < 1 && mat.x > 0) {
                float r = random_float();
                if (r > 0.5f) {
                    new_attenu = colrgb * attenu + attenu * col.w * rs * spec * mix(attenu, colrgb, mat.x);
                    vec3 nd = random_ray(reflect(D, N), 1 - mat.x * mat.y);
                    empile(P + BIAS * N, nd, new_attenu);
                } else {
                    new_attenu = colrgb * attenu + attenu * (1.0f - col.w) * (1.0f - rs) * spec * mix(attenu, colrgb, mat.x);
                    O = P - BIAS * N;
                    traverse_all_bvh(O, D);
                    intersection_info(N, P);
                    empile(P + BIAS * N, refract(D, -N, 1 / U.refract_ind), new_attenu);
                }
            } else {  // diffuse
                new_attenu = colrgb * attenu + attenu * spec * mix(attenu, colrgb, mat.x);
                empile(P + BIAS * N, ray, new_attenu);
            }
        } else
            return total;
    }
    return V3(0, 0, 0);
}

// raytrace(): tp/montecarlo.frag:182-188, tp/montecarlo_mat.frag:5-28,
// tp/montecarlo_mat_tr.frag:5-28
vec3 Shader::raytrace(vec3 Dir, vec3 Orig, int mode) {
    srand();
    vec3 D = normalize(Dir);
    if (mode == 0) return random_path(D, Orig);
    traverse_all_bvh(Orig, D);
    if (!hit()) return V3(0, 0, 0.2f);
    vec3 N = V3(0.0f), P = V3(0.0f);
    intersection_info(N, P);
    vec4 col = intersection_color_info();
    if (mode == 1) return glsl::abs(N) * random_vec3();
    return col.xyz() * random_float();
}

// raytracer.vert:9-22 at the four strip vertices, then the rasteriser's linear
// interpolation (gl_Position.w = 1 everywhere, so it is plain barycentric) at
// the centre of pixel (x, y). Strip triangles: (v0,v1,v2) and (v2,v1,v3) with
// v = (id%2, id/2); a pixel centre with u+v <= 1 lies in the first.
void setup_fragment(Shader& s, int x, int y, int w, int h) {
    mat4 invV, invPV;
    for (int c = 0; c < 4; ++c) {
        invV.c[c] = {s.U.invV[4 * c], s.U.invV[4 * c + 1], s.U.invV[4 * c + 2], s.U.invV[4 * c + 3]};
        invPV.c[c] = {s.U.invPV[4 * c], s.U.invPV[4 * c + 1], s.U.invPV[4 * c + 2], s.U.invPV[4 * c + 3]};
    }
    vec3 dir[4];
    vec3 ori{0, 0, 0};
    for (int id = 0; id < 4; ++id) {
        vec2 tc{float(id % 2), float(id / 2)};
        vec2 c{2.0f * tc.x - 1.0f, 2.0f * tc.y - 1.0f};
        vec4 P4 = invV * vec4{0, 0, 0, 1};
        vec4 Q4 = invPV * vec4{c.x, c.y, 1, 1};
        ori = P4.xyz();
        dir[id] = normalize(Q4.xyz() / Q4.w - P4.xyz());
    }
    float u = (float(x) + 0.5f) / float(w);
    float v = (float(y) + 0.5f) / float(h);
    vec3 d;
    if (u + v <= 1.0f)
        d = dir[0] * ((1.0f - u) - v) + dir[1] * u + dir[2] * v;
    else
        d = dir[1] * (1.0f - v) + dir[2] * (1.0f - u) + dir[3] * ((u + v) - 1.0f);
    s.Ori_in = ori;
    s.Dir_in = d;
    s.screen_tc = {u, v};
}

}  // namespace oracle
