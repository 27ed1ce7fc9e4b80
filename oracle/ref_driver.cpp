// oracle/ref_driver.cpp — TEST INFRASTRUCTURE. Not part of the product.
//
// C entry points around the REFERENCE'S OWN code, built into oracle/_ref/libref.so by
// oracle/build_ref.py (this container only; see that script for what is compiled from where):
//   * scene tables: ScenePrimitives / BVH_KDtree / BVH_GPU_Scene compiled unmodified from
//     /root/reference/bvh_gpu, Transfo and Camera from /root/reference/easycppogl, the scene
//     builders cut out of MontecarloGPU/montecarlo.cpp (gen/scenes.inc);
//   * the traced path: the shader text itself, compiled as C++ (gen/programs.inc).
// What is NOT the reference's here, because the reference leaves it to OpenGL: the six textures
// live in host arrays (ref_shim/texture2d.h), the rasteriser's attribute interpolation over the
// two-triangle strip is the definition of SURVEY 7 (pixel centres, barycentric over (v0,v1,v2)
// / (v1,v2,v3)), blending is a float add per pass in pass order (montecarlo.cpp:450-466), and
// `date` is a parameter instead of the wall clock (montecarlo.cpp:398-401,453).
//
// The exported names and signatures are those of oracle/oracle_capi.cpp, so oracle/binding.py
// drives either library and tests can compare the hand restatement with this one call by call.
#include <omp.h>

#include <cstdint>
#include <cstring>
#include <memory>
#include <vector>

// bvh_data_ / mesh_data_ are private in BVH_GPU_Scene; the tests read the trees the reference
// built. Only this translation unit sees the members as public; layout is unaffected.
#include <gl_eigen.h>
#include <mesh.h>
#include <texture2d.h>

#include <iostream>
#include <map>
#include <algorithm>
#include <cmath>
#define private public
#define class struct  // members before the first access specifier are private in a class
#include "gpu_bvh_scene.h"
#undef class
#undef private
#include "scenes.inc"  // palette, RTViewer::scene_*
#include <camera.h>

#include "programs.inc"  // glslc::FragProgram0/1/2, glslc::VertProgram

namespace {

struct Uni {
    int NB_BOUNCES = 3;
    float refract_ind = 1.0f, date = 0.0f;
    bool flat_face = false;
    float invV[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
    float invPV[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
};

struct RefCtx {
    std::unique_ptr<RTViewer> v{new RTViewer()};
    std::vector<SP_Mesh> meshes;
    Uni U;
    bool finalized = false;
};

struct RefHit {  // == OracleHit == mcpt_hit
    int32_t index, tri_index, shape, dir;
    float dist;
    float pl[3], pg[3];
};

GLMat4 mat_of(const float m[16]) {
    GLMat4 M;
    std::memcpy(M.data(), m, 16 * sizeof(float));
    return M;
}

template <class S, class T>
void bind_tex(S& s, const SP_Texture2D& t) {
    if (!t) return;
    s.data = reinterpret_cast<const T*>(t->raw());
    s.width = t->width();
    s.height = t->height();
    s.comps = t->comps();
}

glslc::mat4 glsl_mat(const float m[16]) {
    using glslc::vec4;
    return glslc::mat4(vec4(m[0], m[1], m[2], m[3]), vec4(m[4], m[5], m[6], m[7]), vec4(m[8], m[9], m[10], m[11]),
                       vec4(m[12], m[13], m[14], m[15]));
}

// uniforms 20 / 21 / 24 are declared by tp/montecarlo.frag only (:1-3); the two debug shaders do not have them
template <class P>
auto set_bounce_uniforms(P& p, const Uni& u, int nb_emissives, int) -> decltype(p.refract_ind, void()) {
    p.nb_emissives = nb_emissives;
    p.NB_BOUNCES = u.NB_BOUNCES;
    p.refract_ind = u.refract_ind;
}
template <class P>
void set_bounce_uniforms(P&, const Uni&, int, long) {}

// the state RTViewer::draw_ogl binds before each draw (montecarlo.cpp:430-448)
template <class P>
void bind_program(P& p, const RefCtx* c, int numero_pass) {
    const BVH_GPU_Scene& g = c->v->bvh_gpu_scene_;
    bind_tex<glslc::sampler2D, float>(p.primsData, g.tex_prim_);
    bind_tex<glslc::sampler2D, float>(p.bvh_bb, g.tex_bb_);
    bind_tex<glslc::isampler2D, int32_t>(p.bvh_ind, g.tex_ind_);
    bind_tex<glslc::isampler2D, int32_t>(p.tri_ind, g.tex_tri_);
    bind_tex<glslc::sampler2D, float>(p.tri_pos, g.tex_p_);
    bind_tex<glslc::sampler2D, float>(p.tri_norm, g.tex_n_);
    p.nb_prims = g.nb_prim();
    p.bvh_depth = c->finalized ? g.depth(0) : 0;
    p.flat_face = c->U.flat_face;
    p.numero_pass = numero_pass;
    p.date = c->U.date;
    set_bounce_uniforms(p, c->U, g.nb_emissives(), 0);
}

struct Corner {
    glslc::vec3 ori, dir;
};

// shaders/raytracer.vert run for the four strip vertices
void run_vertex_stage(const RefCtx* c, Corner out[4]) {
    for (int id = 0; id < 4; ++id) {
        glslc::VertProgram vp;
        vp.invV = glsl_mat(c->U.invV);
        vp.invPV = glsl_mat(c->U.invPV);
        vp.gl_VertexID = id;
        vp.main();
        out[id].ori = vp.Ori_in;
        out[id].dir = vp.Dir_in;
    }
}

// rasteriser stand-in: varyings at the centre of pixel (x, y) of a w x h target
template <class P>
void set_fragment_inputs(P& p, const Corner cn[4], int x, int y, int w, int h) {
    using glslc::vec3;
    float u = (float(x) + 0.5f) / float(w);
    float v = (float(y) + 0.5f) / float(h);
    vec3 d;
    if (u + v <= 1.0f)
        d = cn[0].dir * ((1.0f - u) - v) + cn[1].dir * u + cn[2].dir * v;
    else
        d = cn[1].dir * (1.0f - v) + cn[2].dir * (1.0f - u) + cn[3].dir * ((u + v) - 1.0f);
    p.Ori_in = cn[3].ori;
    p.Dir_in = d;
    p.screen_tc = glslc::vec2(u, v);
}

template <class P>
void render_t(const RefCtx* c, int w, int h, int first_pass, int n_passes, float* accum_rgb, int y0, int y1) {
    Corner cn[4];
    run_vertex_stage(c, cn);
    P proto;
    bind_program(proto, c, 0);
#pragma omp parallel for schedule(dynamic, 1)
    for (int y = y0; y < y1; ++y)
        for (int x = 0; x < w; ++x) {
            float* px = accum_rgb + 3 * (size_t(y) * w + x);
            for (int k = 0; k < n_passes; ++k) {
                P p(proto);  // a fragment invocation starts from freshly initialised globals
                p.numero_pass = first_pass + k;
                set_fragment_inputs(p, cn, x, y, w, h);
                p.main();
                px[0] += p.frag_out.x;  // glBlendFunc(GL_ONE, GL_ONE)
                px[1] += p.frag_out.y;
                px[2] += p.frag_out.z;
            }
        }
}

const BVH_KDtree& tree(const RefCtx* c, int which) { return *c->v->bvh_gpu_scene_.bvh_data_[which < 0 ? 0 : 1 + which].bvh_; }

}  // namespace

extern "C" {

RefCtx* oracle_new() { return new RefCtx(); }
void oracle_free(RefCtx* c) { delete c; }
void oracle_clear(RefCtx* c) {
    c->v->bvh_gpu_scene_.clear();  // BVH_GPU_Scene::clear also clears scene_
    c->meshes.clear();
    c->finalized = false;
}

int oracle_add_prim(RefCtx* c, int type, const float trf[16], const float color[4], float shin, float rough, float emi) {
    BVH_GPU_Scene& g = c->v->bvh_gpu_scene_;
    Material mat(GLVec4(color[0], color[1], color[2], color[3]), shin, rough, emi);
    GLMat4 M = mat_of(trf);
    switch (type) {
        case 1: g.add_sphere(M, mat); break;
        case 2: g.add_cube(M, mat); break;
        case 3: g.add_cylinder(M, mat); break;
        case 4: g.add_cone(M, mat); break;
        case 5: g.add_orientedQuad(M, mat); break;
        default: return -1;
    }
    return g.nb_prim() - 1;
}

int oracle_add_mesh(RefCtx* c, const float* pos, int n_verts, const float* nrm, const uint32_t* tri, int n_tris) {
    SP_Mesh m = std::make_shared<Mesh>();
    for (int v = 0; v < n_verts; ++v) {
        m->vertices_.push_back(GLVec3(pos[3 * v], pos[3 * v + 1], pos[3 * v + 2]));
        m->normals_.push_back(GLVec3(nrm[3 * v], nrm[3 * v + 1], nrm[3 * v + 2]));
        m->bb_.add_point(m->vertices_.back());
    }
    m->tri_indices.assign(tri, tri + 3 * size_t(n_tris));
    c->meshes.push_back(m);
    return c->v->bvh_gpu_scene_.add_mesh(m);
}

int oracle_place_mesh(RefCtx* c, int mesh_id, const float trf[16], const float color[4], float shin, float rough, float emi) {
    BVH_GPU_Scene& g = c->v->bvh_gpu_scene_;
    g.place_mesh(mesh_id, mat_of(trf), Material(GLVec4(color[0], color[1], color[2], color[3]), shin, rough, emi));
    return g.nb_prim() - 1;
}

// RTViewer::scene_* in hot-key order (montecarlo.cpp:236-290): Q W E R T Y U I
int oracle_build_repo_scene(RefCtx* c, int id, float light_intensity) {
    oracle_clear(c);
    RTViewer& v = *c->v;
    v.light_intensity_ = light_intensity;
    switch (id) {
        case 1: v.scene_box_diffuse(); break;
        case 2: v.scene_box_balls(); break;
        case 3: v.scene_menger(); break;
        case 4: v.scene_box_no_top(); break;
        case 5: v.scene_materials(); break;
        case 6: v.scene_4boules(); break;
        case 7: v.scene_menger_lights(); break;
        case 8: v.scene_colonnes(); break;
        default: return -1;
    }
    return 0;
}

void oracle_poke_prim(RefCtx* c, int prim, int word, float value) {
    float* rec = reinterpret_cast<float*>(const_cast<PrimData*>(c->v->scene_.prim_data()) + prim);
    rec[word] = value;
    if (c->finalized) {  // keep the uploaded texture in step (rows of 1024 texels inside a 1204-wide texture)
        const BVH_GPU_Scene& g = c->v->bvh_gpu_scene_;
        g.tex_prim_->update(16 * (prim % 64) + word / 4, prim / 64, 1, 1, rec + 4 * (word / 4));
    }
}

void oracle_finalize(RefCtx* c) {
    c->v->bvh_gpu_scene_.finalize();
    c->finalized = true;
}

int oracle_n_prims(const RefCtx* c) { return c->v->bvh_gpu_scene_.nb_prim(); }
int oracle_nb_emissives(const RefCtx* c) { return c->v->bvh_gpu_scene_.nb_emissives(); }
const void* oracle_prim_data(const RefCtx* c) { return c->v->scene_.prim_data(); }
int oracle_n_meshes(const RefCtx* c) { return int(c->meshes.size()); }
int oracle_bvh_depth(const RefCtx* c, int which) { return tree(c, which).depth(); }
const void* oracle_bvh_bb(const RefCtx* c, int which) { return tree(c, which).data_BB(); }
const int* oracle_bvh_ind(const RefCtx* c, int which) { return tree(c, which).data_ind(); }
int oracle_mesh_line(const RefCtx* c, int m) { return c->v->bvh_gpu_scene_.mesh_data_[m].tri_begins_; }
int oracle_mesh_n_tris(const RefCtx* c, int m) { return c->meshes[m]->nb_triangles(); }
int oracle_mesh_n_verts(const RefCtx* c, int m) { return c->meshes[m]->nb_vertices(); }
const uint32_t* oracle_mesh_tri(const RefCtx* c, int m) { return c->meshes[m]->tri_indices.data(); }
const float* oracle_mesh_pos(const RefCtx* c, int m) { return c->meshes[m]->vertices_[0].data(); }
const float* oracle_mesh_nrm(const RefCtx* c, int m) { return c->meshes[m]->normals_[0].data(); }

static const SP_Texture2D& tex(const RefCtx* c, int k) {
    const BVH_GPU_Scene& g = c->v->bvh_gpu_scene_;
    switch (k) {
        case 0: return g.tex_prim_;
        case 1: return g.tex_bb_;
        case 2: return g.tex_ind_;
        case 3: return g.tex_tri_;
        case 4: return g.tex_p_;
        default: return g.tex_n_;
    }
}
int oracle_tex_height(const RefCtx* c, int k) { return tex(c, k) ? tex(c, k)->height() : 0; }
int oracle_tex_width(const RefCtx* c, int k) { return tex(c, k) ? tex(c, k)->width() : 0; }
int oracle_tex_comps(const RefCtx* c, int k) { return tex(c, k) ? tex(c, k)->comps() : 0; }
const void* oracle_tex_data(const RefCtx* c, int k) { return tex(c, k) ? tex(c, k)->raw() : nullptr; }

void oracle_set_uniforms(RefCtx* c, int nb_bounces, float refract_ind, float date, int flat_face, const float invV[16],
                         const float invPV[16]) {
    c->U.NB_BOUNCES = nb_bounces;
    c->U.refract_ind = refract_ind;
    c->U.date = date;
    c->U.flat_face = flat_face != 0;
    if (invV) std::memcpy(c->U.invV, invV, sizeof c->U.invV);
    if (invPV) std::memcpy(c->U.invPV, invPV, sizeof c->U.invPV);
}

int oracle_set_threads(int n) {
    if (n > 0) omp_set_num_threads(n);
    return omp_get_max_threads();
}

// traverse_all_bvh / just_hit_bvh of the shader library on a ray batch
void oracle_trace(const RefCtx* c, int64_t n, const float* o, const float* d, RefHit* out, int any_hit, uint64_t* counters) {
    glslc::FragProgram0 proto;
    bind_program(proto, c, 0);
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; ++i) {
        glslc::FragProgram0 p(proto);
        glslc::vec3 O(o[3 * i], o[3 * i + 1], o[3 * i + 2]), D(d[3 * i], d[3 * i + 1], d[3 * i + 2]);
        if (any_hit)
            p.just_hit_bvh(O, D);
        else
            p.traverse_all_bvh(O, D);
        const auto& ci = p.closest_intersection;
        RefHit& h = out[i];
        h.index = ci.index;
        h.shape = ci.shape;
        h.dist = ci.dist;
        h.tri_index = (ci.shape == 0) ? ci.tri_index : -1;
        h.dir = ci.dir;
        for (int k = 0; k < 3; ++k) {
            h.pl[k] = (ci.shape >= 0) ? ci.pl[k] : 0.0f;
            h.pg[k] = (ci.shape >= 0) ? ci.pg[k] : 0.0f;
        }
    }
    if (counters)
        for (int k = 0; k < 5; ++k) counters[k] = 0;  // the reference's shaders are not instrumented
}

// the draw loop + additive blend, montecarlo.cpp:450-466
void oracle_render(const RefCtx* c, int w, int h, int first_pass, int n_passes, int mode, float* accum_rgb, int y0, int y1,
                   uint64_t* counters) {
    if (y0 < 0) y0 = 0;
    if (y1 < 0 || y1 > h) y1 = h;
    if (mode == 1)
        render_t<glslc::FragProgram1>(c, w, h, first_pass, n_passes, accum_rgb, y0, y1);
    else if (mode == 2)
        render_t<glslc::FragProgram2>(c, w, h, first_pass, n_passes, accum_rgb, y0, y1);
    else
        render_t<glslc::FragProgram0>(c, w, h, first_pass, n_passes, accum_rgb, y0, y1);
    if (counters)
        for (int k = 0; k < 5; ++k) counters[k] = 0;
}

void oracle_rng_stream(float tcx, float tcy, int numero_pass, float date, int n, float* out, uint32_t* seed_out) {
    glslc::FragProgram0 p;
    p.screen_tc = glslc::vec2(tcx, tcy);
    p.numero_pass = numero_pass;
    p.date = date;
    p.srand();
    if (seed_out) {
        seed_out[0] = p.random_seed.x;
        seed_out[1] = p.random_seed.y;
        seed_out[2] = p.random_seed.z;
    }
    for (int i = 0; i < n; ++i) out[i] = p.random_float();
}

void oracle_random_rays(const float D[3], float roughness, float tcx, float tcy, int n, float* out) {
    glslc::FragProgram0 p;
    p.screen_tc = glslc::vec2(tcx, tcy);
    p.numero_pass = 0;
    p.date = 0.25f;
    p.srand();
    for (int i = 0; i < n; ++i) {
        glslc::vec3 r = p.random_ray(glslc::vec3(D[0], D[1], D[2]), roughness);
        out[3 * i] = r.x;
        out[3 * i + 1] = r.y;
        out[3 * i + 2] = r.z;
    }
}

void oracle_primary_ray(const RefCtx* c, int x, int y, int w, int h, float o[3], float d[3]) {
    Corner cn[4];
    run_vertex_stage(c, cn);
    glslc::FragProgram0 p;
    set_fragment_inputs(p, cn, x, y, w, h);
    glslc::vec3 dn = glslc::normalize(p.Dir_in);
    for (int k = 0; k < 3; ++k) {
        o[k] = p.Ori_in[k];
        d[k] = dn[k];
    }
}

// The matrices RTViewer::draw_ogl uploads for an untouched camera: GLViewer's Camera with
// set_scene_center(0), set_scene_radius(145) (montecarlo.cpp:388-389), aspect = w/h
// (gl_viewer.cpp:405), view = modelview * rotateX(-80) (montecarlo.cpp:404-405,439-440).
void oracle_default_camera(int w, int h, float invV[16], float invPV[16]) {
    Camera cam;
    cam.set_pivot_point(Eigen::Vector3d(0.0, 0.0, 0.0));
    cam.set_scene_radius(145.0);
    cam.set_aspect_ratio(double(w) / h);
    const GLMat4 proj = cam.get_projection_matrix();
    GLMat4 view = cam.get_modelview_matrix() * Transfo::rotateX(-80);
    GLMat4 ipv = (proj * view).inverse(), iv = view.inverse();
    std::memcpy(invPV, ipv.data(), 16 * sizeof(float));
    std::memcpy(invV, iv.data(), 16 * sizeof(float));
}

// Transfo products for the tests: ops over {t,s,u,X,Y,Z} consuming args left to right
void oracle_transfo(const char* ops, const float* a, float out[16]) {
    GLMat4 M;
    M.setIdentity();
    for (const char* p = ops; *p; ++p) switch (*p) {
            case 't': M = M * Transfo::translate(a[0], a[1], a[2]); a += 3; break;
            case 's': M = M * Transfo::scale(a[0], a[1], a[2]); a += 3; break;
            case 'u': M = M * Transfo::scale(a[0]); a += 1; break;
            case 'X': M = M * Transfo::rotateX(a[0]); a += 1; break;
            case 'Y': M = M * Transfo::rotateY(a[0]); a += 1; break;
            case 'Z': M = M * Transfo::rotateZ(a[0]); a += 1; break;
        }
    std::memcpy(out, M.data(), 16 * sizeof(float));
}

const char* oracle_kind() { return "reference"; }

}  // extern "C"
