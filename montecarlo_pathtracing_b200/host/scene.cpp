// scene.cpp — host-side mirror of bvh_gpu/{scene,bvh,gpu_bvh_scene}.cpp.
// Produces the same PrimData records, the same emissive-first order and the same
// implicit-heap BVH arrays as the reference's CPU code; see scene.h.
#include "scene.h"

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <limits>

namespace mcpt {

BB merge(const BB& a, const BB& b) {
    BB r;
    for (int k = 0; k < 3; ++k) {
        r.min_[k] = std::min(a.min_[k], b.min_[k]);
        r.max_[k] = std::max(a.max_[k], b.max_[k]);
    }
    return r;
}

// ----------------------------------------------------------------- Mesh
void Mesh::compute_bb() {
    const float big = std::numeric_limits<float>::max();
    bb_.min_ = Vec3(big, big, big);
    bb_.max_ = Vec3(-big, -big, -big);
    for (const Vec3& p : vertices_)
        for (int k = 0; k < 3; ++k) {
            bb_.min_[k] = std::min(bb_.min_[k], p[k]);
            bb_.max_[k] = std::max(bb_.max_[k], p[k]);
        }
}

void Mesh::compute_normals() {
    normals_.assign(vertices_.size(), Vec3(0, 0, 0));
    for (size_t t = 0; t + 2 < tri_indices.size(); t += 3) {
        uint32_t a = tri_indices[t], b = tri_indices[t + 1], c = tri_indices[t + 2];
        Vec3 n = cross(vertices_[b] - vertices_[a], vertices_[c] - vertices_[a]);
        normals_[a] = normals_[a] + n;
        normals_[b] = normals_[b] + n;
        normals_[c] = normals_[c] + n;
    }
    for (Vec3& n : normals_) {
        float l = norm(n);
        if (l > 0.0f) n = n / l;
    }
}

// BoundingBox::matrix(), easycppogl/mesh.h:67-71: unit cube [-1,1]^3 -> mesh box
Mat4 Mesh::bb_matrix() const {
    Vec3 half = (bb_.max_ - bb_.min_) / 2.0f;
    Vec3 centre = (bb_.min_ + bb_.max_) / 2.0f;
    return Transfo::translate(centre) * Transfo::scale(half);
}

// bvh_gpu/scene.cpp:106-122: triangle box padded by 1e-3 on every side
Vec3 SceneMesh::prim_bb(int p, BB& bb) const {
    const uint32_t* t = &mesh_->tri_indices[3 * size_t(p)];
    const Vec3& P0 = mesh_->vertices_[t[0]];
    const Vec3& P1 = mesh_->vertices_[t[1]];
    const Vec3& P2 = mesh_->vertices_[t[2]];
    for (int k = 0; k < 3; ++k) {
        bb.min_[k] = std::min({P0[k], P1[k], P2[k]});
        bb.max_[k] = std::max({P0[k], P1[k], P2[k]});
    }
    const Vec3 pad(0.001f, 0.001f, 0.001f);
    bb.min_ = bb.min_ - pad;
    bb.max_ = bb.max_ + pad;
    return (bb.min_ + bb.max_) / 2.0f;
}

// ------------------------------------------------------ ScenePrimitives
void ScenePrimitives::clear() {
    buffer_.clear();
    buffer_.shrink_to_fit();
}

static void store(float dst[16], const Mat4& M) { std::memcpy(dst, M.data(), 16 * sizeof(float)); }
static void store4(float dst[4], float a, float b, float c, float d) {
    dst[0] = a;
    dst[1] = b;
    dst[2] = c;
    dst[3] = d;
}

// bvh_gpu/scene.cpp:44-53
int ScenePrimitives::add_prim(int prim, const Mat4& trf, const Material& mat, float area) {
    PrimData pd;
    store(pd.transfo, trf);
    store(pd.inv_transfo, trf.inverse());
    store(pd.mesh_transfo, trf);
    store4(pd.type, float(prim), 0, 0, 0);
    store4(pd.color, mat.color_.x, mat.color_.y, mat.color_.z, mat.color_.w);
    store4(pd.mat_info, mat.shininess_, mat.roughness_, mat.emissivity_, area);
    store4(pd.padding, 0, 0, 0, 0);
    buffer_.push_back(pd);
    return int(buffer_.size()) - 1;
}

// bvh_gpu/scene.cpp:56-67
int ScenePrimitives::add_mesh(const Mesh& mesh, int mesh_line, const Mat4& trf, const Material& mat) {
    PrimData pd;
    store(pd.transfo, trf * mesh.bb_matrix());
    store(pd.inv_transfo, trf.inverse());
    store(pd.mesh_transfo, trf);
    store4(pd.type, 0.0f, float(mesh_line), 0, 0);
    store4(pd.color, mat.color_.x, mat.color_.y, mat.color_.z, mat.color_.w);
    store4(pd.mat_info, mat.shininess_, mat.roughness_, mat.emissivity_, 0.0f);
    store4(pd.padding, 0, 0, 0, 0);
    buffer_.push_back(pd);
    return int(buffer_.size()) - 1;
}

namespace {
// edge vectors of the transformed unit cube from its (-1,-1,zlo) corner
struct CubeEdges {
    Vec3 U, V, W;
};
CubeEdges cube_edges(const Mat4& trf, float zlo) {
    Vec3 o = Transfo::apply(trf, Vec3(-1, -1, zlo));
    return {Transfo::apply(trf, Vec3(1, -1, zlo)) - o, Transfo::apply(trf, Vec3(-1, 1, zlo)) - o,
            Transfo::apply(trf, Vec3(-1, -1, 1)) - o};
}
}  // namespace

// areas as in bvh_gpu/scene.h:128-172
int ScenePrimitives::add_sphere(const Mat4& trf, const Material& mat) {
    float r = norm(Vec3(trf(0, 0), trf(1, 0), trf(2, 0)));  // uniform scale assumed
    float area = float(2.0 * M_PI) * r * r;
    return add_prim(MCPT_PRIM_SPHERE, trf, mat, area);
}
int ScenePrimitives::add_cube(const Mat4& trf, const Material& mat) {
    CubeEdges e = cube_edges(trf, -1);
    float area = 2.0f * (norm(cross(e.U, e.V)) + norm(cross(e.U, e.W)) + norm(cross(e.W, e.V)));
    return add_prim(MCPT_PRIM_CUBE, trf, mat, area);
}
int ScenePrimitives::add_cylinder(const Mat4& trf, const Material& mat) {
    CubeEdges e = cube_edges(trf, -1);
    float area = (dot(e.U, e.U) + dot(e.V, e.V)) / 4.0f * sqrtf(2.0f) * float(M_PI) * norm(e.W);
    return add_prim(MCPT_PRIM_CYLINDER, trf, mat, area);
}
int ScenePrimitives::add_cone(const Mat4& trf, const Material& mat) {
    return add_prim(MCPT_PRIM_CONE, trf, mat, 0.0f);  // area is a TODO in the reference too
}
int ScenePrimitives::add_orientedQuad(const Mat4& trf, const Material& mat) {
    CubeEdges e = cube_edges(trf, 0);
    return add_prim(MCPT_PRIM_ORIENTED_QUAD, trf, mat, norm(cross(e.U, e.V)));
}

// bvh_gpu/scene.cpp:18-42: world box of the unit cube inflated to +-1.005
// (quads: z flattened to +-0.001) pushed through transfo_.
Vec3 ScenePrimitives::prim_bb(int p, BB& bb) const {
    const float hi = std::numeric_limits<float>::max();
    const float lo = std::numeric_limits<float>::lowest();
    bb.min_ = Vec3(hi, hi, hi);
    bb.max_ = Vec3(lo, lo, lo);
    const PrimData& pd = buffer_[p];
    Mat4 T;
    std::memcpy(T.data(), pd.transfo, sizeof pd.transfo);
    for (uint32_t v = 0; v < 8; ++v) {
        float x = float(v & 1) * 2.01f - 1.005f;
        float y = float((v >> 1) & 1) * 2.01f - 1.005f;
        float z = float((v >> 2) & 1) * 2.01f - 1.005f;
        if (pd.type[0] == 5.0f) z /= std::abs(z) * 1000.0f;
        Vec4 B = T * Vec4(x, y, z, 1.0f);
        for (int k = 0; k < 3; ++k) {
            bb.min_[k] = std::min(bb.min_[k], B[k]);
            bb.max_[k] = std::max(bb.max_[k], B[k]);
        }
    }
    return (bb.min_ + bb.max_) / 2.0f;
}

// bvh_gpu/scene.cpp:70-88: move emissive records to the front, keeping the
// reference's exact (swap-based, not stable) order.
int ScenePrimitives::sortEmissiveFirst() {
    size_t n = buffer_.size();
    size_t front = 0;
    while (front < n && buffer_[front].mat_info[2] > 0.0f) ++front;
    for (size_t i = front; i < n; ++i)
        if (buffer_[i].mat_info[2] > 0.0f) {
            std::swap(buffer_[front], buffer_[i]);
            ++front;
        }
    return int(front);
}

// ----------------------------------------------------------- BVH_KDtree
void BVH_KDtree::init(const Scene& s) {
    int n = s.nb();
    centers_.resize(n);
    bbs_.resize(n);
    ids_.resize(n);
    for (int i = 0; i < n; ++i) {
        ids_[i] = i;
        centers_[i] = s.prim_bb(i, bbs_[i]);
    }
}

// bvh_gpu/bvh.cpp:18-93. depth = ceil(log2 n); (depth-1) rounds of median
// partition of every current range (std::nth_element on the box centre along
// axis round%3); the 2^(depth-1) final ranges hold 1 or 2 primitives and fill
// the leaf slots pairwise ((id,-1) or (id0,id1)); interior boxes are unions.
// The reference indexes out of bounds for n < 2 (bvh.cpp:44,59-68); here n < 2
// yields a depth-1 tree with slots (id,-1) / (-1,-1).
void BVH_KDtree::compute() {
    const int n = int(ids_.size());
    depth_ = (n < 2) ? 1 : int(ceilf(log2f(float(n))));
    std::vector<int> cuts{0, n}, next;
    for (int round = 1, axis = 0; round < depth_; ++round, axis = (axis + 1) % 3) {
        next.clear();
        next.push_back(cuts[0]);
        for (size_t r = 1; r < cuts.size(); ++r) {
            int64_t lo = cuts[r - 1], hi = cuts[r], mid = (lo + hi) / 2;
            std::nth_element(ids_.begin() + lo, ids_.begin() + mid, ids_.begin() + hi,
                             [&](int a, int b) { return centers_[a][axis] < centers_[b][axis]; });
            next.push_back(int(mid));
            next.push_back(int(hi));
        }
        cuts.swap(next);
    }

    const int n_leaves = 1 << depth_;
    ind_bvh_.assign(n_leaves, -1);
    BB empty;
    empty.min_ = Vec3(FLT_MAX, FLT_MAX, FLT_MAX);
    empty.max_ = Vec3(-FLT_MAX, -FLT_MAX, -FLT_MAX);
    BB_bvh_.assign(2 * n_leaves - 1, empty);
    const int first_leaf = n_leaves - 1;
    // range r (0-based) owns leaf slots 2r and 2r+1
    for (size_t r = 0; r + 1 < cuts.size(); ++r) {
        int a = cuts[r], cnt = cuts[r + 1] - a;
        int s0 = 2 * int(r), s1 = s0 + 1;
        if (cnt == 1) {
            int id = ids_[a];
            ind_bvh_[s0] = id;
            ind_bvh_[s1] = -1;
            BB_bvh_[first_leaf + s0] = bbs_[id];
            BB_bvh_[first_leaf + s1] = bbs_[id];
        } else if (cnt >= 2) {
            int id0 = ids_[a], id1 = ids_[a + 1];
            ind_bvh_[s0] = id0;
            ind_bvh_[s1] = id1;
            BB_bvh_[first_leaf + s0] = bbs_[id0];
            BB_bvh_[first_leaf + s1] = bbs_[id1];
        }
    }
    for (int i = first_leaf - 1; i >= 0; --i) BB_bvh_[i] = merge(BB_bvh_[2 * i + 2], BB_bvh_[2 * i + 1]);
}

// -------------------------------------------------------- BVH_GPU_Scene
void BVH_GPU_Scene::clear() {
    meshes_.clear();
    mesh_descs_.clear();
    next_tri_row_ = 0;
    root_.reset();
    nb_emi_ = 0;
    scene_.clear();
}

// bvh_gpu/gpu_bvh_scene.cpp:51-74. mesh_line of a mesh is its first row in the
// reference's 1024-wide triangle table: rows advance by ceil(nb_triangles/1024).
int BVH_GPU_Scene::add_mesh(SP_Mesh m) {
    SceneMesh sm(m);
    auto bvh = std::make_shared<BVH_KDtree>();
    bvh->init(sm);
    bvh->compute();
    int nb = m->nb_triangles();
    meshes_.push_back({m, bvh, next_tri_row_});
    next_tri_row_ += nb / 1024 + ((nb % 1024 != 0) ? 1 : 0);
    return int(meshes_.size()) - 1;
}

// bvh_gpu/gpu_bvh_scene.cpp:121-130 (CPU half of finalize)
void BVH_GPU_Scene::finalize() {
    nb_emi_ = scene_.sortEmissiveFirst();
    root_ = std::make_shared<BVH_KDtree>();
    root_->init(scene_);
    root_->compute();
    mesh_descs_.clear();
    for (const MeshEntry& e : meshes_) {
        mcpt_mesh_desc d;
        d.bvh = e.bvh->desc();
        d.tri_idx = e.mesh->tri_indices.data();
        d.pos = &e.mesh->vertices_[0].x;
        d.nrm = &e.mesh->normals_[0].x;
        d.n_tris = uint32_t(e.mesh->nb_triangles());
        d.n_verts = uint32_t(e.mesh->nb_vertices());
        d.mesh_line = e.tri_begins;
        mesh_descs_.push_back(d);
    }
}

mcpt_scene_desc BVH_GPU_Scene::desc() const {
    mcpt_scene_desc d;
    d.prim_data = scene_.prim_data();
    d.n_prims = scene_.nb();
    d.nb_emissives = nb_emi_;
    d.root = root_ ? root_->desc() : mcpt_bvh_desc{nullptr, nullptr, 0};
    d.meshes = mesh_descs_.data();
    d.n_meshes = int(mesh_descs_.size());
    return d;
}

// bvh_gpu/gpu_bvh_scene.cpp:143-175 (GPU half of finalize)
int BVH_GPU_Scene::upload(mcpt_ctx* ctx) const {
    mcpt_scene_desc d = desc();
    return mcpt_upload_scene(ctx, &d);
}

}  // namespace mcpt
