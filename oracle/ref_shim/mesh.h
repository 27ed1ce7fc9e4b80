// oracle/ref_shim/mesh.h — TEST INFRASTRUCTURE. Not part of the product.
//
// Stand-in for easycppogl/mesh.h, which cannot be compiled here (it pulls vao.h / ebo.h -> GL
// and assimp). Only what bvh_gpu/*.cpp touches is declared, with the same names and meaning:
// BoundingBox (mesh.h:41-75, mesh.cpp:93-113) and the three arrays + accessors of Mesh
// (mesh.h:85-141).
#pragma once
#include <gl_eigen.h>

#include <algorithm>
#include <cstdint>
#include <memory>
#include <vector>

#include "texture2d.h"  // GLuint

namespace EZCOGL {

class BoundingBox {
protected:
    GLVec3 min_;
    GLVec3 max_;
    bool initialized_;

public:
    BoundingBox() : initialized_(false) {}
    void add_point(const GLVec3& P) {
        if (!initialized_) {
            min_ = P;
            max_ = P;
            initialized_ = true;
        } else
            for (int i = 0; i < 3; ++i) {
                min_[i] = std::min(min_[i], P[i]);
                max_[i] = std::max(max_[i], P[i]);
            }
    }
    GLVec3 center() const { return (min_ + max_) / 2.0; }
    float radius() const {
        GLVec3 dv = max_ - min_;
        return dv.norm() / 2.0f;
    }
    GLMat4 matrix() const {
        GLVec3 sc = (max_ - min_) / 2.0f;
        return Transfo::translate(center()) * Transfo::scale(sc);
    }
    const GLVec3& min() const { return min_; }
    const GLVec3& max() const { return max_; }
};

class Mesh;
using SP_Mesh = std::shared_ptr<Mesh>;

class Mesh {
public:
    std::vector<GLVec3> vertices_;
    std::vector<GLVec3> normals_;
    std::vector<GLuint> tri_indices;
    BoundingBox bb_;

    Mesh() {}
    Mesh(const Mesh&) = delete;
    int32_t nb_vertices() const { return int32_t(vertices_.size()); }
    int32_t nb_triangles() const { return int32_t(tri_indices.size() / 3); }
    const BoundingBox& BB() const { return bb_; }
    std::vector<GLuint>::const_iterator triangle_index(int t) const { return tri_indices.begin() + 3 * t; }
    const GLVec3& triangle_vertex0(std::vector<GLuint>::const_iterator t) const { return vertices_[*t]; }
    const GLVec3& triangle_vertex1(std::vector<GLuint>::const_iterator t) const { return vertices_[*(t + 1)]; }
    const GLVec3& triangle_vertex2(std::vector<GLuint>::const_iterator t) const { return vertices_[*(t + 2)]; }
};

}  // namespace EZCOGL
