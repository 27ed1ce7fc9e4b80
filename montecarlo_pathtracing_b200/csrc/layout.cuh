// layout.cuh — device-side data layout of the scene and of the wavefront state.
//
// The reference keeps its tables in six 1024-wide OpenGL textures
// (bvh_gpu/gpu_bvh_scene.cpp:121-187) and pays 5-9 dependent texel fetches per
// candidate primitive. Here the same information is re-packed once at upload:
//
//  * BVH nodes, two forms. DevNode, 32 bytes = two float4 {min.xyz, aux0} {max.xyz, aux1}, in the
//    reference's implicit heap order (node k has children 2k, 2k+1, 1-based): no child pointers, and
//    a walk needs no stack in memory -- a 32-bit trail of "sibling still pending" bits plus the heap
//    index is the whole state; leaf-level nodes of the top-level tree carry the primitive id in aux0
//    (flat list of small scenes, one-ray-per-lane tree walk, mcpt_trace_rays). DevNode2, 64 bytes =
//    both child boxes + two child tokens: the mesh trees built on the GPU at upload (bvh_build.cu)
//    and the top-level tree re-packed into the same form, walked with one explicit stack by the
//    persistent warps. Boxes are inflated by a few ulps so the (FMA, non-exact) slab test stays
//    conservative.
//  * primitive records: 128 bytes = inverse transform (3 rows), forward transform
//    (3 rows), {type, mesh id, leaf slot}, instead of the 256-byte PrimData.
//  * triangles: 48 bytes per LEAF SLOT (three float4: A,B,C with the triangle id
//    in A.w, the reference's leaf slot in B.w for the GPU-built trees), stored in leaf order so
//    that the leaf level needs neither the index table nor the three scattered vertex fetches of
//    read_triangle_pos (shaders/raytracer_func.frag:144-152).
//  * shading tables: colour + material float4 per primitive; per-mesh vertex
//    positions/normals/indices for the smooth-normal reconstruction.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "vecmath.cuh"

namespace mcpt {

struct DevNode {
    float4 lo;  // min.xyz, aux0
    float4 hi;  // max.xyz, aux1
};

struct DevPrim {
    Aff inv;      // world -> local            (PrimData.inv_transfo)
    Aff fwd;      // local -> world: primitives PrimData.transfo, meshes PrimData.mesh_transfo
    int32_t type; // MCPT_PRIM_*
    int32_t mesh; // mesh id for type 0, else -1
    int32_t slot; // leaf slot of this primitive in the top-level tree (visit rank)
    int32_t pad;
    float4 pad2;
};
static_assert(sizeof(DevPrim) == 128, "DevPrim must be 128 bytes");

struct DevTri {
    float4 a;  // A.xyz, w = triangle id as int bits (-1: empty slot)
    float4 b;  // B.xyz
    float4 c;  // C.xyz
};

// node of a GPU-built mesh tree (bvh_build.cu): the boxes of both children and their tokens
// (>= 0: node index, < 0: ~leaf; leaf l owns kMeshLeafTris consecutive DevTri slots, whose b.w holds the
// reference's leaf slot of the triangle = its visit rank)
// Triangles per leaf of the GPU-built mesh trees. Measured on the 1 M-triangle workload: 1 -> 720,
// 2 -> 535, 3 -> 487, 4 -> 456, 8 -> 379 Msamples/s: the exact (IEEE divide, no FMA) triangle test is the
// expensive step, a box pair is cheap.
#ifndef MCPT_LEAF_TRIS
#define MCPT_LEAF_TRIS 1
#endif
constexpr int kMeshLeafTris = MCPT_LEAF_TRIS;

struct DevNode2 {
    float4 lo0;  // child 0 min.xyz, w = token of child 0
    float4 hi0;  // child 0 max.xyz, w = token of child 1
    float4 lo1;  // child 1 min.xyz
    float4 hi1;  // child 1 max.xyz
};

// Four-children node of the persistent walk (MCPT_WIDE): every second level of a two-children tree is
// folded into its parent, so a ray takes half as many dependent node fetches, loop iterations and
// stack operations for the same boxes. 128 bytes = one L2 line, boxes stored per component so that one
// float4 feeds the same slab plane of all four children. Child j: token tok[j] (>= 0 node index,
// < 0 ~leaf, kTokEmpty: no child).
struct DevNode4 {
    float4 lox, loy, loz;  // min of children 0..3
    float4 hix, hiy, hiz;  // max of children 0..3
    int4 tok;
    int4 pad;
};
static_assert(sizeof(DevNode4) == 128, "DevNode4 must be 128 bytes");

struct DevMesh {
    const DevNode* nodes;   // 2^depth - 1 interior nodes, heap order   (reference tree; null when rebuilt)
    const DevTri* tris;     // 2^depth leaf slots                       (reference tree; null when rebuilt)
    const DevNode2* nodes2; // GPU-built tree (MCPT_OPT_MESH_BVH), null: walk the reference tree
    const DevTri* tris2;    // its kMeshLeafTris * n_leaves2 triangle slots
    const DevNode4* nodes4; // the same tree with every second level folded away (indexed like nodes2, sparse)
    const uint32_t* idx;    // 3 * n_tris vertex ids (shading)
    const float* pos;       // 3 * n_verts
    const float* nrm;       // 3 * n_verts
    int32_t depth;
    int32_t n_tris;
    int32_t smem_off;       // offset (in nodes) of this mesh's staged top levels, -1 if none
    int32_t smem_nodes;     // number of staged nodes (2^L - 1)
    int32_t n_leaves2;
    int32_t pad;
    int64_t pad2;
};

struct DevScene {
    const DevNode* top_nodes;  // 2^(depth+1) - 1 nodes, heap order
    const DevPrim* prims;
    const float4* color;       // PrimData.color_
    const float4* mat;         // PrimData.mat_info
    const DevMesh* meshes;
    const float* leaf_bb;      // 6 floats per primitive: its un-inflated top-level leaf box
    const DevNode* flat;       // n_flat leaf boxes (aux0 = primitive id) when the scene is small, else null
    int32_t n_flat;            // > 0 selects the warp-coherent flat walk
    int32_t top_depth;
    int32_t n_prims;
    int32_t n_meshes;
    int32_t persistent;        // 1: tree scenes are traced by persistent warps (MCPT_OPT_PERSISTENT)
    int32_t refill_lanes;      // persistent warps: idle lanes that trigger a refill (MCPT_OPT_REFILL_LANES)
    const DevNode2* top2;      // the top-level tree in two-children node format (persistent warps), root = node 0
    const DevNode4* top4;      // ... and in four-children format (MCPT_WIDE), root = node 0
    // shared-memory staging plan (in units of DevNode / DevPrim)
    int32_t smem_top_nodes;    // first N top-level nodes staged
    int32_t smem_prims;        // first N primitive records staged
    int32_t smem_mesh_nodes;   // total staged mesh nodes (sum over meshes)
    int32_t smem_bytes;
};

// closest-hit registers of one ray (the live part of struct sInter,
// raytracer_func.frag:257-267; pg is recomputed from pl where needed)
struct Hit {
    float dist;
    f3 pl;
    int32_t prim;
    int32_t sub;   // triangle id for meshes, face id `dir` otherwise
    int32_t rank;  // visit rank inside the winning primitive's mesh (tie-break only)
};

// uniforms of the traced path
struct DevParams {
    float inv_v[16];
    float inv_pv[16];
    int32_t width, height;
    int32_t nb_bounces;
    float refract_ind;
    float date;
    int32_t flat_face;
    int32_t mode;
    uint32_t batch;    // generation of the batch being traced (path tags, see Wavefront::attenu)
};

// SoA wavefront state, one entry per path slot
struct Wavefront {
    float4* ray_o;     // origin.xyz
    float4* ray_d;     // direction.xyz, w = number of RNG draws the path has consumed (uint bits)
    float4* hit_a;     // pl.xyz, dist
    int2* hit_b;       // prim, sub
    float4* attenu;    // attenuation.xyz, w = path tag (uint bits): 2*batch: alive, 2*batch + 1: ended, the value
                       // random_path returns is `total`; anything else: never shaded in this batch
    float4* total;     // radiance gathered so far .xyz
    float4* keep_n;    // N of the entry hit (interior rays)
    float4* keep_p;    // P of the entry hit (interior rays)
};

}  // namespace mcpt
