// bvh_build.cuh — GPU construction of a mesh tree at upload time (bvh_build.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <vector>

#include "layout.cuh"

namespace mcpt {

// Enqueues the LBVH build of one mesh on `stream` (no host synchronisation: the scratch arrays live
// in a caller-provided workspace of at least lbvh_workspace_bytes(), which builds on the same stream
// may share). d_nodes must hold max(n_leaves - 1, 1) DevNode2 (and d_nodes4 as many DevNode4), d_tris kMeshLeafTris * n_leaves
// DevTri, n_leaves = ceil(n_tris / kMeshLeafTris) >= 1. d_slot_of_tri[t] is the reference's leaf slot
// of triangle t (-1: the reference tree does not hold it, it is left out). lo/hi: box of the mesh
// vertices. `launches` is advanced by the kernels launched.
size_t lbvh_workspace_bytes(int n_tris, int n_leaves);
cudaError_t build_mesh_lbvh(cudaStream_t stream, const float* d_pos, const uint32_t* d_idx, int n_tris,
                            const int32_t* d_slot_of_tri, const float lo[3], const float hi[3], DevNode2* d_nodes,
                            DevNode4* d_nodes4, DevTri* d_tris, int n_leaves, void* workspace, size_t workspace_bytes,
                            unsigned long long* launches);

}  // namespace mcpt
