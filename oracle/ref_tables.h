// oracle/ref_tables.h — TEST INFRASTRUCTURE (CPU oracle). Not part of the product.
//
// Restatement of the reference's CPU scene path, ending in the SAME six
// 1024-texel-wide tables the reference uploads as OpenGL textures:
//   bvh_gpu/scene.h:30-49,64-73,128-172 + scene.cpp:18-100   Material, PrimData, areas, prim_bb, sortEmissiveFirst, merge
//   bvh_gpu/bvh.cpp:5-93                                     BVH_KDtree::init / split_step / compute
//   bvh_gpu/gpu_bvh_scene.cpp:16-187                         add_bvh / send_bvh / add_mesh / send_mesh / finalize
//   easycppogl/gl_eigen.cpp:30-121                           Transfo (degrees, column-major)
// The shader restatement (ref_shader.h) reads these tables through texelFetch
// emulation with the reference's row/column arithmetic, so packing is part of
// what the oracle checks.
#pragma once
#include <cstdint>
#include <vector>

namespace oracle {

constexpr int TEX_WIDTH = 1024;

struct M4 {
    float m[16];  // column-major
};
M4 m4_identity();
M4 m4_mul(const M4& a, const M4& b);
M4 m4_inverse(const M4& a);
void m4_apply(const M4& a, const float v[4], float out[4]);

// 64-float (256-byte) record, bvh_gpu/scene.h:64-73
struct PrimRec {
    float transfo[16], inv_transfo[16], inv_mesh_bb_transfo[16];
    float type[4], color[4], mat_info[4], padding[4];
};
static_assert(sizeof(PrimRec) == 256, "PrimData is 256 bytes");

struct Box {
    float mn[3], mx[3];
};

struct KdTree {  // BVH_KDtree
    std::vector<float> centers;  // 3 per item
    std::vector<Box> bbs;
    std::vector<int> ids;
    std::vector<Box> BB_bvh;
    std::vector<int> ind_bvh;
    int depth = 0;
    void compute();
};

struct MeshIn {
    std::vector<float> pos, nrm;  // 3 per vertex
    std::vector<uint32_t> tri;    // 3 per triangle
    Box bb;                       // mesh bounding box (Mesh::BB())
};

// a 2-D texture in host memory; `comps` floats or ints per texel
struct Tex {
    int width = TEX_WIDTH, height = 0, comps = 0;
    std::vector<float> f;
    std::vector<int32_t> i;
};

struct Tables {
    // builder state
    std::vector<PrimRec> prims;
    std::vector<MeshIn> meshes;
    std::vector<KdTree> mesh_bvh;
    std::vector<int> mesh_tri_begins, mesh_vert_begins;  // Mesh_Data::{tri,vert}_begins_
    int next_tri_row = 0, next_vert_row = 0;
    KdTree root;
    int nb_emissives = 0;
    // finalize() output: the six textures
    Tex tex_prim, tex_bb, tex_ind, tex_tri, tex_p, tex_n;
    std::vector<int> bb_begins, ind_begins;  // per BVH (0 = root, then meshes)

    void clear();
    int add_prim(int type, const M4& trf, const float color[4], float shin, float rough, float emi);
    int add_mesh(const MeshIn& m);  // returns mesh id
    int place_mesh(int mesh_id, const M4& trf, const float color[4], float shin, float rough, float emi);
    void finalize();
};

}  // namespace oracle
