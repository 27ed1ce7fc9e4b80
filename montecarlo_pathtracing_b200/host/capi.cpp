// capi.cpp — C wrappers of include/mcpt_host.h over the C++ host mirror.
#include <new>

#include "camera.h"
#include "mcpt_host.h"
#include "scenes.h"

using namespace mcpt;

struct mcpth_scene {
    ScenePrimitives prims;
    BVH_GPU_Scene gpu;
    bool finalized = false;
    mcpth_scene() : gpu(prims) {}
};

static Mat4 to_mat(const float m[16]) {
    Mat4 M;
    std::memcpy(M.data(), m, 16 * sizeof(float));
    return M;
}

extern "C" {

mcpth_scene* mcpth_scene_new(void) { return new (std::nothrow) mcpth_scene(); }
void mcpth_scene_free(mcpth_scene* s) { delete s; }
void mcpth_scene_clear(mcpth_scene* s) {
    if (!s) return;
    s->gpu.clear();
    s->finalized = false;
}

int mcpth_add_prim(mcpth_scene* s, int type, const float trf[16], const float color[4], float shin, float rough, float emi) {
    if (!s || !trf || !color) return -1;
    Material mat(Vec4(color[0], color[1], color[2], color[3]), shin, rough, emi);
    Mat4 T = to_mat(trf);
    s->finalized = false;
    switch (type) {
        case MCPT_PRIM_SPHERE: s->gpu.add_sphere(T, mat); break;
        case MCPT_PRIM_CUBE: s->gpu.add_cube(T, mat); break;
        case MCPT_PRIM_CYLINDER: s->gpu.add_cylinder(T, mat); break;
        case MCPT_PRIM_CONE: s->gpu.add_cone(T, mat); break;
        case MCPT_PRIM_ORIENTED_QUAD: s->gpu.add_orientedQuad(T, mat); break;
        default: return -1;
    }
    return s->gpu.nb_prim() - 1;
}

int mcpth_add_mesh(mcpth_scene* s, const float* pos, int n_verts, const float* nrm, const uint32_t* tri, int n_tris) {
    if (!s || !pos || !tri || n_verts <= 0 || n_tris <= 0) return -1;
    auto m = std::make_shared<Mesh>();
    m->vertices_.resize(n_verts);
    for (int v = 0; v < n_verts; ++v) m->vertices_[v] = Vec3(pos[3 * v], pos[3 * v + 1], pos[3 * v + 2]);
    m->tri_indices.assign(tri, tri + 3 * size_t(n_tris));
    for (uint32_t i : m->tri_indices)
        if (i >= uint32_t(n_verts)) return -1;
    m->compute_bb();
    if (nrm) {
        m->normals_.resize(n_verts);
        for (int v = 0; v < n_verts; ++v) m->normals_[v] = Vec3(nrm[3 * v], nrm[3 * v + 1], nrm[3 * v + 2]);
    } else {
        m->compute_normals();
    }
    s->finalized = false;
    return s->gpu.add_mesh(m);
}

int mcpth_place_mesh(mcpth_scene* s, int mesh_id, const float trf[16], const float color[4], float shin, float rough, float emi) {
    if (!s || mesh_id < 0 || mesh_id >= s->gpu.nb_meshes()) return -1;
    s->gpu.place_mesh(mesh_id, to_mat(trf), Material(Vec4(color[0], color[1], color[2], color[3]), shin, rough, emi));
    s->finalized = false;
    return s->gpu.nb_prim() - 1;
}

int mcpth_build_repo_scene(mcpth_scene* s, int id, float li) {
    if (!s) return MCPT_E_INVALID;
    s->finalized = false;
    return build_repo_scene(s->gpu, id, li) ? MCPT_OK : MCPT_E_INVALID;
}
int mcpth_build_scene6_roughness(mcpth_scene* s, float li, float roughness) {
    if (!s) return MCPT_E_INVALID;
    s->finalized = false;
    build_scene6_roughness(s->gpu, li, roughness);
    return MCPT_OK;
}
int mcpth_build_divergence_stress(mcpth_scene* s, float li) {
    if (!s) return MCPT_E_INVALID;
    s->finalized = false;
    build_divergence_stress(s->gpu, li);
    return MCPT_OK;
}
int mcpth_build_synthetic_meshes(mcpth_scene* s, int n_meshes, int tris_log2, uint32_t seed) {
    if (!s || tris_log2 > 26) return MCPT_E_INVALID;
    s->finalized = false;
    build_synthetic_meshes(s->gpu, n_meshes, tris_log2, seed);
    return MCPT_OK;
}

int mcpth_finalize(mcpth_scene* s) {
    if (!s) return MCPT_E_INVALID;
    s->gpu.finalize();
    s->finalized = true;
    return MCPT_OK;
}

int mcpth_desc(const mcpth_scene* s, mcpt_scene_desc* out) {
    if (!s || !out || !s->finalized) return MCPT_E_INVALID;
    *out = s->gpu.desc();
    return MCPT_OK;
}

int mcpth_upload(const mcpth_scene* s, mcpt_ctx* ctx) {
    if (!s || !ctx || !s->finalized) return MCPT_E_INVALID;
    return s->gpu.upload(ctx);
}

int mcpth_nb_prim(const mcpth_scene* s) { return s ? s->gpu.nb_prim() : 0; }
int mcpth_nb_emissives(const mcpth_scene* s) { return s ? s->gpu.nb_emissives() : 0; }
int mcpth_depth(const mcpth_scene* s, int i) { return (s && s->finalized) ? s->gpu.depth(i) : -1; }
const mcpt_prim_data* mcpth_prim_data(const mcpth_scene* s) { return s ? s->prims.prim_data() : nullptr; }
int mcpth_nb_meshes(const mcpth_scene* s) { return s ? s->gpu.nb_meshes() : 0; }

void mcpth_default_camera(int w, int h, float inv_v[16], float inv_pv[16]) { default_camera(w, h, inv_v, inv_pv); }

// ops: 't' translate(x,y,z) 's' scale(x,y,z) 'u' scale(s) 'X' 'Y' 'Z' rotate(deg);
// the product is taken left to right, e.g. "tZs" = translate * rotateZ * scale.
void mcpth_transfo(const char* ops, const float* a, float out[16]) {
    Mat4 M;
    for (const char* o = ops; o && *o; ++o) {
        switch (*o) {
            case 't': M = M * Transfo::translate(a[0], a[1], a[2]); a += 3; break;
            case 's': M = M * Transfo::scale(a[0], a[1], a[2]); a += 3; break;
            case 'u': M = M * Transfo::scale(a[0]); a += 1; break;
            case 'X': M = M * Transfo::rotateX(a[0]); a += 1; break;
            case 'Y': M = M * Transfo::rotateY(a[0]); a += 1; break;
            case 'Z': M = M * Transfo::rotateZ(a[0]); a += 1; break;
        }
    }
    std::memcpy(out, M.data(), 16 * sizeof(float));
}

}  // extern "C"
