// oracle/ref_shader.h — TEST INFRASTRUCTURE (CPU oracle). Not part of the product.
//
// Function-by-function C++ restatement of the reference's fragment program:
//   shaders/raytracer_func.frag   RNG :90-124, table readers :138-255, sInter :257-271,
//                                 intersect_bvm/bv :273-352, intersectors :354-640,
//                                 Mesh_intersect :642-678, intersect_prim :681-705,
//                                 intersect_bvh & co :708-781, geometry info :783-907
//   tp/montecarlo.frag            stack :5-47, sample_hemisphere :49-70, random_ray :72-89,
//                                 rSchlick :91-98, random_path :100-179, raytrace :182-188
//   tp/montecarlo_mat.frag, tp/montecarlo_mat_tr.frag   (modes 1 and 2)
//   shaders/raytracer.vert :9-22 + shaders/main.frag :4-7
// Names, control flow and expression order follow the GLSL; the GLSL built-ins are
// the ones of glsl.h. One Shader object = one fragment invocation's global state.
#pragma once
#include "glsl.h"
#include "ref_tables.h"

namespace oracle {

using namespace glsl;

struct Uniforms {
    int nb_prims = 0;       // location 1
    int bvh_depth = 0;      // location 2
    bool flat_face = false; // location 3
    int numero_pass = 0;    // location 4
    float date = 0.0f;      // location 5
    float invV[16];         // location 6
    float invPV[16];        // location 10
    int nb_emissives = 0;   // location 20
    int NB_BOUNCES = 3;     // location 21
    float refract_ind = 1.0f; // location 24
};

struct sInter {
    vec3 pl{0, 0, 0};
    vec3 pg{0, 0, 0};
    float dist = 3.402823e38f;
    int dir = -1;
    int index = -1;
    int tri_index = -1;
    int mesh_line = -1;
    int shape = -1;
};

// counters for the roofline accounting: how much work the
// reference's algorithm does on a given input
struct Counters {
    uint64_t traversals = 0, box_tests = 0, prim_tests = 0, tri_tests = 0, segments = 0;
};

class Shader {
public:
    const Tables& T;
    Uniforms U;
    // fragment inputs
    vec3 Ori_in{0, 0, 0}, Dir_in{0, 0, 0};
    vec2 screen_tc{0, 0};
    // globals of the GLSL program
    sInter closest_intersection;
    bool hit_only = false;
    uvec3 random_seed{0, 0, 0};
    Counters cnt;

    explicit Shader(const Tables& t) : T(t) {}

    // RNG
    void srand();
    float random_float();
    vec3 random_vec3();

    // traversal
    bool hit() const { return closest_intersection.shape >= 0; }
    void reset_inter();
    void intersect_bvh(vec3 O, vec3 D);
    void traverse_all_bvh(vec3 O, vec3 D);
    void just_hit_bvh(vec3 O, vec3 D);
    void intersect_prim(int i, vec3 O, vec3 D);
    void intersection_info(vec3& N, vec3& Pg);
    vec4 intersection_color_info();
    vec4 intersection_mat_info();

    // tp/montecarlo.frag
    vec3 sample_hemisphere(float roughness);
    vec3 random_ray(vec3 D, float roughness);
    float rSchlick(vec3 I, vec3 N);
    vec3 random_path(vec3 Dir, vec3 Orig);
    // mode: 0 montecarlo.frag, 1 montecarlo_mat.frag, 2 montecarlo_mat_tr.frag
    vec3 raytrace(vec3 Dir, vec3 Orig, int mode);

private:
    // table readers
    ivec3 read_mesh_info(int mesh_line);
    void read_triangle_pos(int mesh_line, int tri, vec3& Pa, vec3& Pb, vec3& Pc);
    void read_triangle_PN(int mesh_line, int tri, vec3& Pa, vec3& Na, vec3& Pb, vec3& Nb, vec3& Pc, vec3& Nc);
    int read_primitive_type(int i, int& mesh_line);
    mat4 read_mat(int i, int first_texel);
    mat4 read_transfo(int i) { return read_mat(i, 0); }
    mat4 read_inv_transfo(int i) { return read_mat(i, 4); }
    mat4 read_mesh_transfo(int i) { return read_mat(i, 8); }
    vec4 read_color(int i);
    vec4 read_material(int i);
    int read_bvh_ind(int bl, int i);
    void read_bvh_bb(int bl, int i, vec3& bbmin, vec3& bbmax);

    bool intersect_bvm(int bl, int i, vec3 O, vec3 D, vec3 Ol, const mat4& trf);
    bool intersect_bv(int bl, int i, vec3 O, vec3 D, vec3 Ol);
    void Triangle_intersect(int mesh_line, int tri_index, int bb_index, vec3 O, vec3 D, vec3 Ol, const mat4& trf);
    void Sphere_intersect(int index, vec3 O, vec3 D, vec3 Ol);
    void OrientedQuad_intersect(int index, vec3 O, vec3 D, vec3 Ol);
    void Cube_intersect(int index, vec3 O, vec3 D, vec3 Ol);
    void Cylinder_intersect(int index, vec3 O, vec3 D, vec3 Ol);
    void Cone_intersect(int index, vec3 O, vec3 D, vec3 Ol);
    void Mesh_intersect(int mesh_line, int bb_index, vec3 O, vec3 D, vec3 Ol);
    void candidate(int index, int shape, int dir, vec3 Pl, vec3 Ol);
};

// raytracer.vert:9-22 for the 4 strip vertices + the rasteriser's interpolation
// at pixel (x, y) of a w x h target (pixel centres, two triangles (v0,v1,v2) and
// (v1,v2,v3)); sets Ori_in, Dir_in, screen_tc of `s`.
void setup_fragment(Shader& s, int x, int y, int w, int h);

extern int g_ulp_perturb;  // sensitivity probe, ref_shader_path.cpp

}  // namespace oracle
