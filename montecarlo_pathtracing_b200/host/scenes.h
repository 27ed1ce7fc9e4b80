// scenes.h — the reference's eight procedural scenes (RTViewer::scene_*,
// MontecarloGPU/montecarlo.cpp:629-795, menger :143-179) re-expressed through the
// mirrored BVH_GPU_Scene::add_* calls, plus the synthetic scenes BASELINE.json
// names (divergence stress, N-triangle procedural meshes).
#pragma once
#include "scene.h"

namespace mcpt {

// Scene numbers follow the reference's hot keys Q,W,E,R,T,Y,U,I
// (montecarlo.cpp:236-290) and the captures/N-*.png naming:
// 1 box_diffuse, 2 box_balls, 3 menger, 4 box_no_top, 5 materials, 6 4boules,
// 7 menger_lights, 8 colonnes. Returns false for an unknown id.
// The scene is appended to `s` (call s.clear() first, then s.finalize()).
bool build_repo_scene(BVH_GPU_Scene& s, int id, float light_intensity);

// BASELINE config 2: scene 6 with roughness forced to `roughness` on every
// non-emissive primitive.
void build_scene6_roughness(BVH_GPU_Scene& s, float light_intensity, float roughness);

// BASELINE config 5 (divergence stress): scene_materials layout + scene 6's quad
// light; every non-light primitive is transparent+reflective
// (color.a = 0.3, shininess 0.8, roughness 0.1).
void build_divergence_stress(BVH_GPU_Scene& s, float light_intensity);

// BASELINE configs 3/4: n_meshes displaced spheres of 2^tris_log2 triangles each
// (deterministic hash noise, seed), materials cycling diffuse / reflective /
// transparent / transparent+reflective, over a 9000^2 floor quad, lit by one
// quad light of emissivity 12.
void build_synthetic_meshes(BVH_GPU_Scene& s, int n_meshes, int tris_log2, uint32_t seed);

// one displaced lat-long sphere with exactly 2^tris_log2 triangles (tris_log2 >= 3)
SP_Mesh make_displaced_sphere(int tris_log2, uint32_t seed, float radius, float amplitude);

}  // namespace mcpt
