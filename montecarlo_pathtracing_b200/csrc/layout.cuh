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
//    persistent warps; DevNode4, 128 bytes = four child boxes + tokens: the same trees with every second
//    level folded away, walked on scenes with meshes. Boxes are inflated by a few ulps so the (FMA,
//    non-exact) slab test stays conservative.
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
    float inv_smin; // 1 / the smallest column norm of fwd: world noise -> local units (sphere_test)
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
// The two planes of one axis are neighbours (float4 index 2 * axis and 2 * axis + 1), so a ray that keeps the index of
// the plane it ENTERS through per axis (0 / 1, 2 / 3, 4 / 5, by the sign of its direction) loads entry and exit planes
// directly and needs no per-axis min / max (walk_node_u).
struct DevNode4 {
    float4 lox, hix;  // min / max x of children 0..3
    float4 loy, hiy;
    float4 loz, hiz;
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

// One light of next-event estimation (MCPT_OPT_NEE): an emissive oriented quad among the first nb_emissives records
// (the reference sorts emissive primitives first, bvh_gpu/scene.cpp:70-88, and stores an area figure in mat_info.a,
// bvh_gpu/scene.h:128-172, for a light sampler it never finished, shaders/raytracer_func.frag:908-956).
struct DevLight {
    float4 c;      // centre (world), w = 1 / true area of the quad
    float4 ex;     // half edge along local x (world), w = selection probability of this light
    float4 ey;     // half edge along local y (world), w = cumulative selection probability up to and including it
    float4 n;      // unit normal (world), w = primitive id (int bits)
    float4 emit;   // colour * 0.1 (what a hit adds to `total` besides the emission), w = emissivity * (1 - shininess) * alpha
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
    int32_t flat_pool;         // 1: after the first bounce the flat walk pools its exact tests per warp and primitive type
    int32_t flat_max_type;     // the largest number of list entries that hold primitives of one type (sizes the item list)
    int32_t top_depth;
    int32_t n_prims;
    int32_t n_meshes;
    int32_t persistent;        // 1: tree scenes are traced by persistent warps (MCPT_OPT_PERSISTENT)
    int32_t refill_lanes;      // persistent warps: idle lanes that trigger a refill (MCPT_OPT_REFILL_LANES)
    int32_t refill_first;      // ... on the identity queue (camera rays: the 32 rays of a warp are neighbours in the image)
    const DevNode2* top2;      // the top-level tree in two-children node format (persistent warps), root = node 0
    const DevNode4* top4;      // ... and in four-children format (MCPT_WIDE), root = node 0
    // The unified world-space tree (MCPT_OPT_UNIFIED, scenes with meshes): the top-level tree and, below each placement of
    // a mesh, a subtree fitted to the WORLD positions of that instance's triangles, all in one array of four-children
    // nodes (root = node 0). A ray is walked in one frame: no mesh entry / exit, one kind of node, one kind of leaf step.
    // Leaf token ~L: L < u_n_tris is triangle slot L of u_tris (local vertices; c.w = primitive id of the instance),
    // else analytic primitive L - u_n_tris. null: the two-level walk.
    const DevNode4* u_nodes;
    const DevTri* u_tris;
    int32_t u_n_tris;
    // The exact tests run in primitive-local frames reached through the uploaded inverse transforms, which are not the
    // exact inverses of the forward ones: a hit the exact test accepts lies off the world ray by up to
    // |fwd * inv - 1| * (|O| + distance). Every box of the unified tree is therefore padded PER RAY by
    // u_pad_rel * (|O|_1 + u_radius) + u_pad_abs (both bounds computed at upload from the matrices themselves).
    float u_pad_rel, u_pad_abs, u_radius;
    // Pooled leaf tests of the unified walk (MCPT_OPT_LEAF_POOL, k_traverse_unified_pooled): u_pool > 0 = the number of queued
    // (ray, leaf) items at which a warp runs a round of exact tests. A candidate is ranked by a 64-bit key: distance bits,
    // then (u_slot_max - leaf slot of the primitive) << u_key_shift | (u_rank_max - (rank + 1)) << 1 | candidate number --
    // offer()'s order (intersect.cuh). 0: the upload found that the three fields do not fit 32 bits, or the option is off.
    int32_t u_pool, u_key_shift, u_slot_max, u_rank_max;
    const DevLight* lights;    // next-event estimation: the quads that are sampled directly (null: none)
    int32_t n_lights;
    int32_t u_smem_nodes;      // the first N nodes of u_nodes are staged in shared memory (the top part is stored in
                               // breadth-first order, so these are the top levels of the tree); 0: none
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

// Wavefront state: one 128-byte record per path slot, four 32-byte sectors, each moved by ONE 256-bit
// load or store (LDG / STG.E.ENL2.256 on sm_100a) so that every access of a stage is a whole sector:
//   sector 0  ray     origin.xyz | direction.xyz, w = number of RNG draws the path has consumed (uint bits)
//                     read by k_traverse, written by k_raygen / k_shade / the exit of interior rays
//   sector 1  hit     pl.xyz, prim | sub, D.xyz (direction of the ray that hit)   written by k_traverse, read by k_shade
//   sector 2  rad     attenuation.xyz, w = path tag | radiance gathered so far (`total`), w = RNG draws consumed
//                     tag (uint bits): 2*batch: alive, 2*batch + 1: ended, the value random_path returns is
//                     `total`; anything else: never shaded in this batch. k_shade, k_accumulate
//   sector 3  keep    N | P of the entry hit (interior rays): written by k_shade, read where they exit
// After the first bounce paths are reached through index queues. With one array per field (the layout this
// replaces) every 16-byte access of a scattered path moved a 32-byte sector of which half belonged to a
// neighbour, and a path's state sat in seven DRAM pages: k_shade 11.6 -> 10.1 ms per C2 step with records.
constexpr size_t kPathRecordBytes = 128;
struct alignas(32) PathSector {
    float4 a, b;
};
#ifndef MCPT_PATH_INTERLEAVED
#define MCPT_PATH_INTERLEAVED 0  // 1: the four sectors of a path side by side (one 128-byte record per path)
#endif
struct Wavefront {
    char* rec;        // 128 bytes per slot
    size_t plane;     // bytes between the sector arrays (0 when interleaved)
    __device__ __forceinline__ PathSector* sec(size_t p, int s) const {
#if MCPT_PATH_INTERLEAVED
        return reinterpret_cast<PathSector*>(rec + p * kPathRecordBytes + 32 * s);
#else
        return reinterpret_cast<PathSector*>(rec + plane * s + p * 32);
#endif
    }
    __device__ __forceinline__ void load_ray(size_t p, float4& o, float4& d) const {
        const PathSector v = *sec(p, 0);
        o = v.a;
        d = v.b;
    }
    __device__ __forceinline__ float4 load_ray_d(size_t p) const { return sec(p, 0)->b; }
    __device__ __forceinline__ void store_ray(size_t p, float4 o, float4 d) const {
        PathSector v;
        v.a = o;
        v.b = d;
        *sec(p, 0) = v;
    }
    // hit sector: pl.xyz, prim | sub, D.xyz -- the direction of the ray that hit rides along, so that
    // k_shade does not read the ray sector at all (the hit distance is not needed downstream)
    __device__ __forceinline__ void load_hit(size_t p, f3& pl, int& prim, int& sub, f3& D) const {
        const PathSector v = *sec(p, 1);
        pl = mk3(v.a.x, v.a.y, v.a.z);
        prim = __float_as_int(v.a.w);
        sub = __float_as_int(v.b.x);
        D = mk3(v.b.y, v.b.z, v.b.w);
    }
    __device__ __forceinline__ void store_hit(size_t p, f3 pl, int prim, int sub, f3 D) const {
        PathSector v;
        v.a = make_float4(pl.x, pl.y, pl.z, __int_as_float(prim));
        v.b = make_float4(__int_as_float(sub), D.x, D.y, D.z);
        *sec(p, 1) = v;
    }
    __device__ __forceinline__ void load_rad(size_t p, float4& attenu, float4& total) const {
        const PathSector v = *sec(p, 2);
        attenu = v.a;
        total = v.b;
    }
    __device__ __forceinline__ void store_rad(size_t p, float4 attenu, float4 total) const {
        PathSector v;
        v.a = attenu;
        v.b = total;
        *sec(p, 2) = v;
    }
    __device__ __forceinline__ void load_keep(size_t p, float4& n, float4& pt) const {
        const PathSector v = *sec(p, 3);
        n = v.a;
        pt = v.b;
    }
    __device__ __forceinline__ void store_keep(size_t p, float4 n, float4 pt) const {
        PathSector v;
        v.a = n;
        v.b = pt;
        *sec(p, 3) = v;
    }
};

}  // namespace mcpt
