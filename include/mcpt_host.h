/* mcpt_host.h — C view of the host-side mirror of the reference's scene API.
 *
 * The C++ classes in montecarlo_pathtracing_b200/host/ (ScenePrimitives,
 * BVH_KDtree, BVH_GPU_Scene, Camera, RTViewer) mirror bvh_gpu/scene.h,
 * bvh_gpu/bvh.h, bvh_gpu/gpu_bvh_scene.h, easycppogl/camera.h and the RT branch
 * of MontecarloGPU/montecarlo.cpp. C++ callers use those classes directly;
 * these entry points expose them to C / ctypes (tests, bench.py).
 */
#ifndef MCPT_HOST_H_
#define MCPT_HOST_H_

#include "mcpt.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct mcpth_scene mcpth_scene; /* ScenePrimitives + BVH_GPU_Scene */

mcpth_scene* mcpth_scene_new(void);
void mcpth_scene_free(mcpth_scene* s);
void mcpth_scene_clear(mcpth_scene* s); /* BVH_GPU_Scene::clear, gpu_bvh_scene.cpp:76-85 */

/* BVH_GPU_Scene::add_sphere/add_cube/add_cylinder/add_cone/add_orientedQuad
 * (gpu_bvh_scene.h:86-109); type = MCPT_PRIM_* (1..5); trf column-major.
 * Returns the record index, or -1 for a bad type. */
int mcpth_add_prim(mcpth_scene* s, int type, const float trf[16], const float color[4], float shininess, float roughness,
                   float emissivity);
/* BVH_GPU_Scene::add_mesh (gpu_bvh_scene.cpp:51-74); nrm may be NULL (area-weighted
 * normals are computed). Returns the mesh id. */
int mcpth_add_mesh(mcpth_scene* s, const float* pos, int n_verts, const float* nrm, const uint32_t* tri, int n_tris);
/* BVH_GPU_Scene::place_mesh (gpu_bvh_scene.h:81-84) */
int mcpth_place_mesh(mcpth_scene* s, int mesh_id, const float trf[16], const float color[4], float shininess,
                     float roughness, float emissivity);

/* RTViewer::scene_* (montecarlo.cpp:629-795), ids 1..8 in hot-key order. */
int mcpth_build_repo_scene(mcpth_scene* s, int id, float light_intensity);
int mcpth_build_scene6_roughness(mcpth_scene* s, float light_intensity, float roughness);
int mcpth_build_divergence_stress(mcpth_scene* s, float light_intensity);
int mcpth_build_synthetic_meshes(mcpth_scene* s, int n_meshes, int tris_log2, uint32_t seed);

/* BVH_GPU_Scene::finalize, CPU half (gpu_bvh_scene.cpp:121-141) */
int mcpth_finalize(mcpth_scene* s);
/* table view; pointers stay valid until the scene is modified or freed */
int mcpth_desc(const mcpth_scene* s, mcpt_scene_desc* out);
/* BVH_GPU_Scene::finalize, GPU half (gpu_bvh_scene.cpp:143-175) */
int mcpth_upload(const mcpth_scene* s, mcpt_ctx* ctx);

int mcpth_nb_prim(const mcpth_scene* s);
int mcpth_nb_emissives(const mcpth_scene* s);
int mcpth_depth(const mcpth_scene* s, int i); /* 0 = root, i>0 = mesh i-1 */
const mcpt_prim_data* mcpth_prim_data(const mcpth_scene* s);
int mcpth_nb_meshes(const mcpth_scene* s);

/* matrices RTViewer::draw_ogl uploads for the untouched default camera
 * (montecarlo.cpp:388-389,404-405,439-440) */
void mcpth_default_camera(int width, int height, float inv_v[16], float inv_pv[16]);
/* Transfo::translate/scale/rotateX/Y/Z products, for building matrices from C */
void mcpth_transfo(const char* ops, const float* args, float out[16]);

#ifdef __cplusplus
}
#endif
#endif
