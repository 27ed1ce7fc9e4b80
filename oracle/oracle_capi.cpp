// oracle/oracle_capi.cpp — TEST INFRASTRUCTURE (CPU oracle). Not part of the product.
//
// C entry points (loaded with ctypes by tests/, __graft_entry__.smoke() and the
// cpu_baseline / --impl reference legs of bench.py — nothing else may use them).
// Parity status: the reference ships no tests, golden vectors or fixtures for
// this path and its GLSL cannot be executed on this box (no graphics stack), so
// the restatement is pinned by (a) oracle/_ref, the reference's own .frag/.cpp
// sources compiled as C++ where that build is available, and (b) the committed
// fixtures under tests/golden/ generated from it.
#include <omp.h>

#include <cstdio>
#include <cstring>
#include <vector>

#include "ref_shader.h"

using namespace oracle;

struct OracleCtx {
    Tables T;
    Uniforms U;
    bool finalized = false;
};

// mirror of include/mcpt.h mcpt_hit (kept textually separate on purpose: the
// oracle does not include product headers)
struct OracleHit {
    int32_t index, tri_index, shape, dir;
    float dist;
    float pl[3], pg[3];
};

extern "C" {

OracleCtx* oracle_new() { return new OracleCtx(); }
void oracle_free(OracleCtx* c) { delete c; }
void oracle_clear(OracleCtx* c) {
    c->T.clear();
    c->finalized = false;
}

int oracle_add_prim(OracleCtx* c, int type, const float trf[16], const float color[4], float shin, float rough, float emi) {
    M4 m;
    std::memcpy(m.m, trf, sizeof m.m);
    return c->T.add_prim(type, m, color, shin, rough, emi);
}

int oracle_add_mesh(OracleCtx* c, const float* pos, int n_verts, const float* nrm, const uint32_t* tri, int n_tris) {
    MeshIn m;
    m.pos.assign(pos, pos + 3 * size_t(n_verts));
    m.nrm.assign(nrm, nrm + 3 * size_t(n_verts));
    m.tri.assign(tri, tri + 3 * size_t(n_tris));
    for (int i = 0; i < 3; ++i) {
        m.bb.mn[i] = 3.402823e38f;
        m.bb.mx[i] = -3.402823e38f;
    }
    for (int v = 0; v < n_verts; ++v)
        for (int i = 0; i < 3; ++i) {
            if (pos[3 * v + i] < m.bb.mn[i]) m.bb.mn[i] = pos[3 * v + i];
            if (pos[3 * v + i] > m.bb.mx[i]) m.bb.mx[i] = pos[3 * v + i];
        }
    return c->T.add_mesh(m);
}

int oracle_place_mesh(OracleCtx* c, int mesh_id, const float trf[16], const float color[4], float shin, float rough, float emi) {
    M4 m;
    std::memcpy(m.m, trf, sizeof m.m);
    return c->T.place_mesh(mesh_id, m, color, shin, rough, emi);
}

// overwrite one float of a primitive record before finalize (material overrides
// of BASELINE configs 2 and 5); word = index into the 64-float record
void oracle_poke_prim(OracleCtx* c, int prim, int word, float value) {
    reinterpret_cast<float*>(&c->T.prims[prim])[word] = value;
}

void oracle_finalize(OracleCtx* c) {
    c->T.finalize();
    c->U.nb_prims = int(c->T.prims.size());
    c->U.bvh_depth = c->T.root.depth;
    c->U.nb_emissives = c->T.nb_emissives;
    c->finalized = true;
}

// ---- table access
int oracle_n_prims(const OracleCtx* c) { return int(c->T.prims.size()); }
int oracle_nb_emissives(const OracleCtx* c) { return c->T.nb_emissives; }
const void* oracle_prim_data(const OracleCtx* c) { return c->T.prims.data(); }
int oracle_n_meshes(const OracleCtx* c) { return int(c->T.meshes.size()); }
// which = -1: root tree, else mesh id
static const KdTree& tree(const OracleCtx* c, int which) { return which < 0 ? c->T.root : c->T.mesh_bvh[which]; }
int oracle_bvh_depth(const OracleCtx* c, int which) { return tree(c, which).depth; }
const void* oracle_bvh_bb(const OracleCtx* c, int which) { return tree(c, which).BB_bvh.data(); }
const int* oracle_bvh_ind(const OracleCtx* c, int which) { return tree(c, which).ind_bvh.data(); }
int oracle_mesh_line(const OracleCtx* c, int m) { return c->T.mesh_tri_begins[m]; }
int oracle_mesh_n_tris(const OracleCtx* c, int m) { return int(c->T.meshes[m].tri.size() / 3); }
int oracle_mesh_n_verts(const OracleCtx* c, int m) { return int(c->T.meshes[m].pos.size() / 3); }
const uint32_t* oracle_mesh_tri(const OracleCtx* c, int m) { return c->T.meshes[m].tri.data(); }
const float* oracle_mesh_pos(const OracleCtx* c, int m) { return c->T.meshes[m].pos.data(); }
const float* oracle_mesh_nrm(const OracleCtx* c, int m) { return c->T.meshes[m].nrm.data(); }
// textures: 0 prim, 1 bb, 2 ind, 3 tri, 4 pos, 5 nrm
static const Tex& tex(const OracleCtx* c, int k) {
    switch (k) {
        case 0: return c->T.tex_prim;
        case 1: return c->T.tex_bb;
        case 2: return c->T.tex_ind;
        case 3: return c->T.tex_tri;
        case 4: return c->T.tex_p;
        default: return c->T.tex_n;
    }
}
int oracle_tex_height(const OracleCtx* c, int k) { return tex(c, k).height; }
int oracle_tex_comps(const OracleCtx* c, int k) { return tex(c, k).comps; }
const void* oracle_tex_data(const OracleCtx* c, int k) {
    const Tex& t = tex(c, k);
    return t.f.empty() ? static_cast<const void*>(t.i.data()) : static_cast<const void*>(t.f.data());
}

// ---- uniforms
void oracle_set_uniforms(OracleCtx* c, int nb_bounces, float refract_ind, float date, int flat_face, const float invV[16],
                         const float invPV[16]) {
    c->U.NB_BOUNCES = nb_bounces;
    c->U.refract_ind = refract_ind;
    c->U.date = date;
    c->U.flat_face = flat_face != 0;
    if (invV) std::memcpy(c->U.invV, invV, sizeof c->U.invV);
    if (invPV) std::memcpy(c->U.invPV, invPV, sizeof c->U.invPV);
}

// sensitivity probe: shift every sin / cos / log / pow result of the bounce sampler by n ulps (process-wide; 0 = off)
void oracle_set_ulp_perturb(int n) { oracle::g_ulp_perturb = n; }

int oracle_set_threads(int n) {
    if (n > 0) omp_set_num_threads(n);
    return omp_get_max_threads();
}

// ---- traverse_all_bvh / just_hit_bvh on a ray batch
void oracle_trace(const OracleCtx* c, int64_t n, const float* o, const float* d, OracleHit* out, int any_hit, uint64_t* counters) {
    Counters total;
#pragma omp parallel
    {
        Shader s(c->T);
        s.U = c->U;
#pragma omp for schedule(static)
        for (int64_t i = 0; i < n; ++i) {
            vec3 O{o[3 * i], o[3 * i + 1], o[3 * i + 2]}, D{d[3 * i], d[3 * i + 1], d[3 * i + 2]};
            s.closest_intersection = sInter();
            if (any_hit)
                s.just_hit_bvh(O, D);
            else
                s.traverse_all_bvh(O, D);
            const sInter& ci = s.closest_intersection;
            OracleHit& h = out[i];
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
#pragma omp critical
        {
            total.traversals += s.cnt.traversals;
            total.box_tests += s.cnt.box_tests;
            total.prim_tests += s.cnt.prim_tests;
            total.tri_tests += s.cnt.tri_tests;
        }
    }
    if (counters) {
        counters[0] = total.traversals;
        counters[1] = total.box_tests;
        counters[2] = total.prim_tests;
        counters[3] = total.tri_tests;
        counters[4] = 0;
    }
}

// ---- the draw loop + additive blend (montecarlo.cpp:450-466): for each pass in
// order, accum_rgb[pixel] += raytrace(...). accum_rgb is float[3*w*h], row 0 =
// bottom row (gl_FragCoord), NOT cleared here. x0..x1/y0..y1 restrict the work to
// a window of the target (for bounded CPU baselines); pass -1 for the full frame.
void oracle_render(const OracleCtx* c, int w, int h, int first_pass, int n_passes, int mode, float* accum_rgb, int y0, int y1,
                   uint64_t* counters) {
    if (y0 < 0) y0 = 0;
    if (y1 < 0 || y1 > h) y1 = h;
    Counters total;
#pragma omp parallel
    {
        Shader s(c->T);
        s.U = c->U;
#pragma omp for schedule(dynamic, 1)
        for (int y = y0; y < y1; ++y)
            for (int x = 0; x < w; ++x) {
                setup_fragment(s, x, y, w, h);
                float* px = accum_rgb + 3 * (size_t(y) * w + x);
                for (int k = 0; k < n_passes; ++k) {
                    s.U.numero_pass = first_pass + k;
                    s.closest_intersection = sInter();
                    vec3 rgb = s.raytrace(s.Dir_in, s.Ori_in, mode);
                    px[0] += rgb.x;
                    px[1] += rgb.y;
                    px[2] += rgb.z;
                }
            }
#pragma omp critical
        {
            total.traversals += s.cnt.traversals;
            total.box_tests += s.cnt.box_tests;
            total.prim_tests += s.cnt.prim_tests;
            total.tri_tests += s.cnt.tri_tests;
            total.segments += s.cnt.segments;
        }
    }
    if (counters) {
        counters[0] = total.traversals;
        counters[1] = total.box_tests;
        counters[2] = total.prim_tests;
        counters[3] = total.tri_tests;
        counters[4] = total.segments;
    }
}

// ---- unit-level access for the tests
void oracle_rng_stream(float tcx, float tcy, int numero_pass, float date, int n, float* out, uint32_t* seed_out) {
    Tables t;
    Shader s(t);
    s.screen_tc = {tcx, tcy};
    s.U.numero_pass = numero_pass;
    s.U.date = date;
    s.srand();
    if (seed_out) {
        seed_out[0] = s.random_seed.x;
        seed_out[1] = s.random_seed.y;
        seed_out[2] = s.random_seed.z;
    }
    for (int i = 0; i < n; ++i) out[i] = s.random_float();
}

// n samples of random_ray(D, roughness) from a fixed seed (sampler distribution tests)
void oracle_random_rays(const float D[3], float roughness, float tcx, float tcy, int n, float* out) {
    Tables t;
    Shader s(t);
    s.screen_tc = {tcx, tcy};
    s.U.date = 0.25f;
    s.srand();
    for (int i = 0; i < n; ++i) {
        vec3 r = s.random_ray(vec3{D[0], D[1], D[2]}, roughness);
        out[3 * i] = r.x;
        out[3 * i + 1] = r.y;
        out[3 * i + 2] = r.z;
    }
}

// primary ray of a pixel as the oracle defines it (raytracer.vert + interpolation)
void oracle_primary_ray(const OracleCtx* c, int x, int y, int w, int h, float o[3], float d[3]) {
    Shader s(c->T);
    s.U = c->U;
    setup_fragment(s, x, y, w, h);
    vec3 dn = normalize(s.Dir_in);
    o[0] = s.Ori_in.x;
    o[1] = s.Ori_in.y;
    o[2] = s.Ori_in.z;
    d[0] = dn.x;
    d[1] = dn.y;
    d[2] = dn.z;
}

}  // extern "C"
