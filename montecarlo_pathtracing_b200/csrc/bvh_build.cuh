// bvh_build.cuh — GPU construction of a mesh tree at upload time (bvh_build.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <vector>

#include "layout.cuh"

namespace mcpt {

// One LBVH build (bvh_build.cu). The tree is fitted either to the mesh-local vertices (fwd == null: the tree of a mesh,
// walked in the mesh frame) or to the world positions fwd * pos of ONE placement of the mesh (the subtree of that
// instance in the unified world-space tree); the DevTri slots always hold the local vertices, because the exact
// triangle test runs in the mesh frame (intersect.cuh).
struct LbvhBuild {
    const float* d_pos = nullptr;          // 3 * n_verts, mesh-local
    const uint32_t* d_idx = nullptr;       // 3 * n_tris
    int n_tris = 0, n_verts = 0;
    const int32_t* d_slot_of_tri = nullptr;  // the reference's leaf slot of every triangle (tie-break rank; -1: not in the
                                             // reference tree, left out); null: no reference tree, rank = triangle index
    const Aff* fwd = nullptr;              // host pointer: mesh -> world of the instance, or null
    DevNode2* d_nodes = nullptr;           // max(n_leaves - 1, 1) two-children nodes; null: kept in the workspace only
    DevNode4* d_nodes4 = nullptr;          // as many four-children nodes (indexed like d_nodes, sparse)
    DevTri* d_tris = nullptr;              // kMeshLeafTris * n_leaves slots
    int n_leaves = 0;                      // ceil(n_tris / kMeshLeafTris) >= 1
    int node_base = 0, leaf_base = 0;      // added to the node / leaf tokens of d_nodes4 (position inside the unified arrays)
    int prim = -1;                         // primitive id of the instance, stored in DevTri.c.w
};

// Enqueues the build on `stream` (no host synchronisation, no host pass over the mesh: the vertex box is reduced on
// the device). The scratch arrays live in a caller-provided workspace of at least lbvh_workspace_bytes(), which
// builds on the same stream may share. `launches` is advanced by the kernels launched.
size_t lbvh_workspace_bytes(int n_tris, int n_leaves, int n_verts);
cudaError_t build_mesh_lbvh(cudaStream_t stream, const LbvhBuild& b, void* workspace, size_t workspace_bytes,
                            unsigned long long* launches);


// upload-time checks and tables that used to be host loops over the mesh:
// *d_flag |= 1 if a vertex index is >= n_verts
void check_indices(cudaStream_t stream, const uint32_t* d_idx, size_t n, uint32_t n_verts, int* d_flag);
// d_slot_of_tri[t] = leaf slot of triangle t in the uploaded reference tree (d_ind: its 2^depth leaf slots), -1 if absent;
// *d_flag |= 2 if a slot holds an index outside [-1, n_tris)
cudaError_t slots_from_leaves(cudaStream_t stream, const int32_t* d_ind, int n_slots, int n_tris, int32_t* d_slot_of_tri, int* d_flag);

}  // namespace mcpt
