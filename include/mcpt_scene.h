/* mcpt_scene.h — POD data contract between the reference's host-side scene/BVH
 * structures and the B200 path-tracing library.
 *
 * Every struct here is a bit-compatible mirror of a structure the reference
 * keeps on the CPU and uploads as an OpenGL texture; citations are relative to
 * the reference repository root.
 *
 *   mcpt_prim_data  <->  struct PrimData            bvh_gpu/scene.h:64-73
 *   mcpt_bb         <->  struct BB                  bvh_gpu/scene.h:15-19
 *   mcpt_bvh_desc   <->  BVH_KDtree::{BB_bvh_,ind_bvh_,depth_}   bvh_gpu/bvh.h:17-19
 *   mcpt_mesh_desc  <->  Mesh::{vertices_,normals_,tri_indices}  easycppogl/mesh.h:85-90
 *                        + the per-mesh BVH registered by BVH_GPU_Scene::add_mesh
 *                        (bvh_gpu/gpu_bvh_scene.cpp:51-74)
 *   mcpt_scene_desc <->  what BVH_GPU_Scene::finalize() uploads
 *                        (bvh_gpu/gpu_bvh_scene.cpp:121-187)
 *
 * All matrices are 4x4 column-major floats (Eigen default storage == GLSL mat4).
 */
#ifndef MCPT_SCENE_H_
#define MCPT_SCENE_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* primitive type codes, shaders/raytracer_func.frag:38-43 */
enum {
    MCPT_PRIM_MESH = 0,
    MCPT_PRIM_SPHERE = 1,
    MCPT_PRIM_CUBE = 2,
    MCPT_PRIM_CYLINDER = 3,
    MCPT_PRIM_CONE = 4,
    MCPT_PRIM_ORIENTED_QUAD = 5
};

/* 256-byte primitive record = 16 RGBA32F texels of tex_prim_
 * (readers: shaders/raytracer_func.frag:171-233). */
typedef struct mcpt_prim_data {
    float transfo[16];      /* texels 0-3 : unit shape -> world (meshes: trf * bb.matrix()) */
    float inv_transfo[16];  /* texels 4-7 : world -> local                                  */
    float mesh_transfo[16]; /* texels 8-11: prims: trf ; meshes: placement trf (local->world) */
    float type[4];          /* texel 12   : x = type code, y = mesh_line                    */
    float color[4];         /* texel 13   : RGBA, a = opacity                               */
    float mat_info[4];      /* texel 14   : r shininess, g roughness, b emissivity, a area  */
    float padding[4];       /* texel 15                                                     */
} mcpt_prim_data;

/* axis-aligned box, 24 bytes = 2 RGB32F texels of tex_bb_ */
typedef struct mcpt_bb {
    float min[3];
    float max[3];
} mcpt_bb;

/* Complete binary tree in implicit heap order (children of i: 2i+1, 2i+2).
 * bb has 2^(depth+1)-1 entries, ind has 2^depth entries (-1 = empty leaf slot).
 * Built by BVH_KDtree::compute, bvh_gpu/bvh.cpp:34-93. */
typedef struct mcpt_bvh_desc {
    const mcpt_bb* bb;
    const int32_t* ind;
    int32_t depth;
} mcpt_bvh_desc;

/* One triangle mesh with its own BVH in mesh-local space. tri_idx holds plain
 * vertex indices local to this mesh (the reference pre-offsets them by
 * 1024*vert_row for its texture addressing, gpu_bvh_scene.cpp:92-100; that
 * offset is a texture-layout detail and is NOT applied here). mesh_line is the
 * value stored in PrimData.type.y by place_mesh (= tri_begins_ row), used to
 * bind a PrimData record of type 0 to this mesh. */
typedef struct mcpt_mesh_desc {
    mcpt_bvh_desc bvh;
    const uint32_t* tri_idx; /* 3 * n_tris */
    const float* pos;        /* 3 * n_verts */
    const float* nrm;        /* 3 * n_verts */
    uint32_t n_tris;
    uint32_t n_verts;
    int32_t mesh_line;
} mcpt_mesh_desc;

/* Everything finalize() sends to the GPU. All pointers are borrowed host
 * pointers and are copied before mcpt_upload_scene returns. */
typedef struct mcpt_scene_desc {
    const mcpt_prim_data* prim_data;
    int32_t n_prims;
    int32_t nb_emissives;
    mcpt_bvh_desc root;
    const mcpt_mesh_desc* meshes;
    int32_t n_meshes;
} mcpt_scene_desc;

#ifdef __cplusplus
}
#endif
#endif /* MCPT_SCENE_H_ */
