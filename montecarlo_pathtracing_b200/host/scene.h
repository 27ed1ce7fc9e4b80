// scene.h — host-side mirror of the reference's scene / BVH interface
// (bvh_gpu/scene.h, bvh_gpu/bvh.h, bvh_gpu/gpu_bvh_scene.h), Eigen- and GL-free.
//
// Same class and method names, same argument meaning, same table contents:
// code written against the reference's BVH_GPU_Scene (add_sphere, add_cube,
// add_cylinder, add_cone, add_orientedQuad, add_mesh, place_mesh, clear,
// finalize, depth, nb_prim, nb_emissives) compiles against this one. Where the
// reference's finalize() creates six OpenGL textures, this one fills a
// mcpt_scene_desc (include/mcpt_scene.h) and hands it to mcpt_upload_scene.
#pragma once
#include <cstdint>
#include <memory>
#include <vector>

#include "mcpt.h"
#include "mcpt_math.h"

namespace mcpt {

// bvh_gpu/scene.h:15-19
struct BB {
    Vec3 min_;
    Vec3 max_;
};
static_assert(sizeof(BB) == sizeof(mcpt_bb), "BB must stay 24 bytes");

BB merge(const BB& a, const BB& b);  // bvh_gpu/scene.cpp:91-100

// bvh_gpu/scene.h:22-28
class Scene {
public:
    virtual ~Scene() = default;
    virtual Vec3 prim_bb(int p, BB& bb) const = 0;
    virtual int nb() const = 0;
    virtual void clear() = 0;
};

// bvh_gpu/scene.h:30-49
class Material {
public:
    Vec4 color_;
    float shininess_;
    float roughness_;
    float emissivity_;
    Material(const Vec4& color, float shin, float rough, float emi)
        : color_(color), shininess_(shin), roughness_(rough), emissivity_(emi) {}
    Material(const Vec4& color, float shin, float rough)
        : color_(color), shininess_(shin), roughness_(rough), emissivity_(0.0f) {}
    explicit Material(const Vec4& color) : color_(color), shininess_(0.0f), roughness_(0.0f), emissivity_(0.0f) {}
    static Material light(const Vec4& col, float emi) { return Material(col, 0.0f, 0.0f, emi); }
};

// The three arrays of easycppogl/mesh.h:85-90 the mesh path consumes, plus the
// bounding box whose matrix() enters PrimData.transfo_ (mesh.h:41-72).
class Mesh {
public:
    std::vector<Vec3> vertices_;
    std::vector<Vec3> normals_;
    std::vector<uint32_t> tri_indices;
    BB bb_;
    int32_t nb_vertices() const { return int32_t(vertices_.size()); }
    int32_t nb_triangles() const { return int32_t(tri_indices.size() / 3); }
    void compute_bb();
    void compute_normals();  // area-weighted vertex normals
    Mat4 bb_matrix() const;  // BoundingBox::matrix(), mesh.h:67-71
};
using SP_Mesh = std::shared_ptr<Mesh>;

// bvh_gpu/scene.h:51-62, scene.cpp:106-131
class SceneMesh : public Scene {
    SP_Mesh mesh_;

public:
    explicit SceneMesh(SP_Mesh m) : mesh_(std::move(m)) {}
    Vec3 prim_bb(int p, BB& bb) const override;
    int nb() const override { return mesh_->nb_triangles(); }
    void clear() override {}
};

using PrimData = mcpt_prim_data;  // bvh_gpu/scene.h:64-73

// bvh_gpu/scene.h:75-185
class ScenePrimitives : public Scene {
    std::vector<PrimData> buffer_;
    int add_prim(int prim, const Mat4& trf, const Material& mat, float area);

public:
    void clear() override;
    const PrimData* prim_data() const { return buffer_.data(); }
    PrimData* prim_data_mut() { return buffer_.data(); }
    int nb() const override { return int(buffer_.size()); }
    int sortEmissiveFirst();
    int type(int i) const { return int(buffer_[i].type[0]); }
    int mesh_line(int i) const { return int(buffer_[i].type[1]); }

    int add_mesh(const Mesh& mesh, int mesh_line, const Mat4& trf, const Material& mat);
    int add_sphere(const Mat4& trf, const Material& mat);
    int add_cube(const Mat4& trf, const Material& mat);
    int add_cylinder(const Mat4& trf, const Material& mat);
    int add_cone(const Mat4& trf, const Material& mat);
    int add_orientedQuad(const Mat4& trf, const Material& mat);

    Vec3 prim_bb(int p, BB& bb) const override;
};

// bvh_gpu/bvh.h:10-60, bvh.cpp:5-93: complete binary tree, median split with the
// axis cycling x,y,z per level, implicit heap layout.
class BVH_KDtree {
    std::vector<Vec3> centers_;
    std::vector<BB> bbs_;
    std::vector<int> ids_;
    std::vector<BB> BB_bvh_;
    std::vector<int> ind_bvh_;
    int depth_ = 0;

public:
    int depth() const { return depth_; }
    int nb_bb() const { return int(BB_bvh_.size()); }
    int nb_indices() const { return int(ind_bvh_.size()); }
    const int* data_ind() const { return ind_bvh_.data(); }
    const BB* data_BB() const { return BB_bvh_.data(); }
    void init(const Scene& s);
    void compute();
    mcpt_bvh_desc desc() const {
        return mcpt_bvh_desc{reinterpret_cast<const mcpt_bb*>(BB_bvh_.data()), ind_bvh_.data(), depth_};
    }
};
using SP_BVH_KDtree = std::shared_ptr<BVH_KDtree>;

// bvh_gpu/gpu_bvh_scene.h:35-121
class BVH_GPU_Scene {
    struct MeshEntry {
        SP_Mesh mesh;
        SP_BVH_KDtree bvh;
        int tri_begins;  // row of the mesh in the reference's tex_tri_ == mesh_line
    };
    std::vector<MeshEntry> meshes_;
    int next_tri_row_ = 0;
    SP_BVH_KDtree root_;
    int nb_emi_ = 0;
    ScenePrimitives& scene_;
    std::vector<mcpt_mesh_desc> mesh_descs_;

public:
    explicit BVH_GPU_Scene(ScenePrimitives& sc) : scene_(sc) {}
    void clear();
    int add_mesh(SP_Mesh m);  // builds the per-mesh BVH, returns the mesh id
    void place_mesh(int i, const Mat4& trf, const Material& mat) {
        scene_.add_mesh(*meshes_[i].mesh, meshes_[i].tri_begins, trf, mat);
    }
    void add_sphere(const Mat4& trf, const Material& mat) { scene_.add_sphere(trf, mat); }
    void add_cube(const Mat4& trf, const Material& mat) { scene_.add_cube(trf, mat); }
    void add_cylinder(const Mat4& trf, const Material& mat) { scene_.add_cylinder(trf, mat); }
    void add_cone(const Mat4& trf, const Material& mat) { scene_.add_cone(trf, mat); }
    void add_orientedQuad(const Mat4& trf, const Material& mat) { scene_.add_orientedQuad(trf, mat); }

    // sort emissives first, build the root BVH, assemble the table descriptor
    void finalize();
    // finalize()'s GPU half: hand the tables to the CUDA library
    int upload(mcpt_ctx* ctx) const;
    mcpt_scene_desc desc() const;

    int depth(int i) const { return i == 0 ? root_->depth() : meshes_[i - 1].bvh->depth(); }
    int nb_prim() const { return scene_.nb(); }
    int nb_bb() const { return 1 + int(meshes_.size()); }
    int nb_emissives() const { return nb_emi_; }
    int nb_meshes() const { return int(meshes_.size()); }
    const ScenePrimitives& primitives() const { return scene_; }
    ScenePrimitives& primitives() { return scene_; }
    const BVH_KDtree& root() const { return *root_; }
};

}  // namespace mcpt
