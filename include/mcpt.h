/* mcpt.h — C ABI of the B200-native Monte Carlo path-tracing core.
 *
 * This library replaces ONE hot path of ksaintmarc/Montecarlo-PathTracing: the
 * per-pixel path tracer the reference runs as an OpenGL fragment program
 * (shaders/raytracer_func.frag + tp/montecarlo.frag + shaders/main.frag), the
 * additive progressive accumulation (MontecarloGPU/montecarlo.cpp:450-467) and
 * the resolve pass (montecarlo.cpp:59-70,470-476 / shaders/average.frag).
 *
 * The reference has no FFI; the seam is two host call sites, and every entry
 * point below names the reference code it stands in for:
 *   (1) BVH_GPU_Scene::finalize()            bvh_gpu/gpu_bvh_scene.cpp:121-187
 *   (2) RTViewer::draw_ogl(), RT branch      MontecarloGPU/montecarlo.cpp:408-477
 *       RTViewer::resize_ogl()               MontecarloGPU/montecarlo.cpp:616-626
 *
 * Conventions: plain C types only; every function returns 0 (MCPT_OK) or a
 * negative MCPT_E_* code and never throws or exits; mcpt_last_error(ctx) gives
 * a human-readable message for the last failure on that context. A context is
 * owned by one caller thread at a time (the reference's "GL thread" rule);
 * several contexts (one per GPU) may be driven from several threads/processes.
 * There is NO CPU fallback: without a CUDA device mcpt_create fails.
 */
#ifndef MCPT_H_
#define MCPT_H_

#include <stddef.h>
#include <stdint.h>
#include "mcpt_scene.h"

#ifdef __cplusplus
extern "C" {
#endif

#define MCPT_OK 0
#define MCPT_E_INVALID -1   /* bad argument / bad state                    */
#define MCPT_E_CUDA -2      /* a CUDA runtime call failed                  */
#define MCPT_E_NODEVICE -3  /* no CUDA device / device index out of range  */
#define MCPT_E_NOMEM -4     /* host or device allocation failed            */
#define MCPT_E_NCCL -5      /* NCCL unavailable or a collective failed     */
#define MCPT_E_UNSUPPORTED -6

#define MCPT_ABI_VERSION 1

typedef struct mcpt_ctx mcpt_ctx;

/* Viewer-facing parameter block = the uniforms RTViewer::draw_ogl sets
 * (montecarlo.cpp:441-448,453; declared at raytracer_func.frag:23-27 and
 * tp/montecarlo.frag:1-3). numero_pass is an argument of mcpt_render. */
typedef struct mcpt_params {
    int32_t nb_bounces;   /* uniform 21 NB_BOUNCES   (UI 0..9, any >=0 accepted) */
    float refract_ind;    /* uniform 24 refract_ind  (UI 1.0..2.5)              */
    float date;           /* uniform 5  date: seconds since previous frame      */
    int32_t flat_face;    /* uniform 3  flat_face (never set by the host => 0)  */
    int32_t mode;         /* tp shader: 0 montecarlo.frag, 1 montecarlo_mat.frag,
                             2 montecarlo_mat_tr.frag (keys O/P, montecarlo.cpp:292-304) */
    int32_t nb_prims;     /* uniform 1  (uploaded, unused by montecarlo.frag)   */
    int32_t nb_emissives; /* uniform 20 (uploaded, unused by montecarlo.frag)   */
} mcpt_params;

/* closest-hit record = struct sInter, raytracer_func.frag:257-267 */
typedef struct mcpt_hit {
    int32_t index;     /* primitive index, -1 on miss                        */
    int32_t tri_index; /* triangle index inside the mesh, -1 if not a mesh   */
    int32_t shape;     /* type code of the hit primitive, -1 on miss         */
    int32_t dir;       /* face id (cube 0..5, cylinder/cone 0/1/2), else 0   */
    float dist;        /* world-space distance |O - pg| (3.402823e38f on miss, the shader's FLT_MAX)    */
    float pl[3];       /* hit point in primitive-local space                 */
    float pg[3];       /* hit point in world space                           */
} mcpt_hit;

/* destination kinds for mcpt_resolve */
enum {
    MCPT_DST_HOST_RGB32F = 0,   /* float[3*w*h], unclamped (what fs_frag computes)      */
    MCPT_DST_HOST_RGBA8 = 1,    /* uint8[4*w*h], clamped to [0,1], no gamma (back buffer) */
    MCPT_DST_DEVICE_RGB32F = 2, /* device pointer, float[3*w*h]                          */
    MCPT_DST_DEVICE_RGBA8 = 3   /* device pointer, uint8[4*w*h]                          */
};

/* tuning knobs for mcpt_set_option */
enum {
    MCPT_OPT_PASSES_PER_BATCH = 0, /* passes traced together in one wavefront (default auto) */
    MCPT_OPT_SMEM_STAGING = 1,     /* shared-memory staging of scene data in the traversal kernels: 0 = none, 1 = the leaf
                                      boxes of small scenes (all of them, once per ray octant), the first 64 primitive
                                      records, and the top levels of the trees the one-ray-per-lane walk reads (default),
                                      2 = also the top levels (256 nodes) of the unified tree (measured slower: those
                                      nodes already sit in L1). Set before mcpt_upload_scene. */
    MCPT_OPT_FAST_MATH = 2,        /* 1 = MUFU sin/cos/log/pow in the bounce sampler         */
    MCPT_OPT_MESH_BVH = 3,         /* 1 = rebuild every mesh tree on the GPU at upload (LBVH, one triangle
                                      per leaf; default 1), 0 = walk the uploaded median-split trees;
                                      set before mcpt_upload_scene. Hit results are the same.   */
    MCPT_OPT_PROFILE_STAGES = 4,   /* 1 = bracket every stage launch with CUDA events         */
    MCPT_OPT_COUNT_WORK = 5,       /* 1 = traversal kernels count boxes / records / triangles */
    MCPT_OPT_FLAT_WALK = 6,        /* 1 = warp-coherent flat walk for scenes of <= 16 primitives (default 1);
                                      set before mcpt_upload_scene */
    MCPT_OPT_TRAVERSE_OCC = 7,     /* resident 256-thread CTAs per SM the traversal kernel is register-capped
                                      for: 2 (<=128 regs), 3 (<=80), 4 (<=64), unified walk also 5 (<=48), 6 (<=40);
                                      0 = auto (default) */
    MCPT_OPT_PERSISTENT = 8,       /* 1 = scenes walked through the tree (> 16 primitives, or meshes) are traced by
                                      persistent warps whose idle lanes fetch new rays (default), 0 = one ray per
                                      lane per tile. Set before mcpt_upload_scene. Same results. */
    MCPT_OPT_REFILL_LANES = 9,     /* persistent warps: idle lanes of a warp that trigger a refill from the queue,
                                      1..32; 0 = auto (8 with meshes, 12 on other unified trees, 32 else; camera rays 32).
                                      Set before mcpt_upload_scene. */
    MCPT_OPT_TOP_SAH = 10,         /* 1 = the persistent warps walk a top-level tree built at upload by a full-sweep
                                      SAH over the primitives' leaf boxes (default), 0 = the topology of the uploaded
                                      median-split tree. Set before mcpt_upload_scene. Same results. */
    MCPT_OPT_TRACE_WALKER = 11,    /* mcpt_trace_rays (closest hit): 0 = through the production traversal kernel of
                                      mcpt_render with whatever walker the scene selected (default), 1 = the
                                      one-ray-per-lane walk over the uploaded trees in the reference's visit order.
                                      any_hit always takes the latter. Same records. */
    MCPT_OPT_UNIFIED = 12,         /* scenes with meshes, and trees of more than 256 leaves: 1 = one world-space tree of
                                      four-children nodes over the top level and every mesh instance (instance subtrees
                                      built on the GPU at upload), walked in a single frame (default; needs
                                      MCPT_OPT_MESH_BVH and MCPT_OPT_PERSISTENT), 0 = two-level walk (top-level tree,
                                      then the mesh's own tree in mesh space). Set before mcpt_upload_scene. Same
                                      records. mcpt_mesh_desc.bvh may be empty (bb == ind == NULL) whenever
                                      MCPT_OPT_MESH_BVH is 1: no host tree is needed, triangles hit at exactly the
                                      same distance are then ranked by index instead of by the reference's leaf slot. */
    MCPT_OPT_NEE = 13,             /* 1 = next-event estimation: every vertex whose next ray is drawn from a cone also
                                      samples one point on one emissive quad (the emissive-first records with
                                      emissivity > 0.5, picked by mat_info.a) and delivers the light's contribution
                                      through a shadow ray, balance-weighted against the bounce ray. A NEW estimator
                                      (the reference's light sampler is commented out, raytracer_func.frag:908-956):
                                      same expectation per pixel -- it converges to the same image -- with less
                                      noise at equal spp where lights are small; the bounce rays are unchanged.
                                      Default 0 (the reference's estimator, bit-level parity). */
    MCPT_OPT_FLAT_POOL = 14,       /* scenes of <= 16 primitives, bounces after the first: 1 = a warp pools the exact
                                      tests of 64 rays per primitive type and runs them 32 at a time, 2 = ... and
                                      copies the next 64 rays into shared memory with cp.async while it does so
                                      (default; falls back to 1 where four CTAs of it would not fit one SM),
                                      0 = the warp-coherent walk of the first bounce everywhere. Set before
                                      mcpt_upload_scene. Same records. */
    MCPT_OPT_LEAF_POOL = 15,       /* scenes traced through the unified world-space tree (meshes, large trees): N > 0 =
                                      a lane that reaches a leaf queues a (ray, leaf) item and walks on; the warp
                                      runs the exact tests of the queue 32 at a time once N items (1..32) wait, the
                                      closest hit of a ray ranked by a 64-bit key in shared memory
                                      (k_traverse_unified_pooled). 0 = every lane tests its own leaf (default:
                                      measured 3 .. 6 % faster, csrc/kernels.cu). Set before mcpt_upload_scene. Same
                                      records. */
    MCPT_OPT_FUSED_RAYGEN = 16     /* 1 (default) = the first traversal of a batch computes the camera ray of every
                                      path it takes (raytracer.vert's arithmetic) and stores it; 0 = a kernel of its
                                      own writes all camera rays first. Same rays, same images. */
};

/* wavefront stages, in launch order inside one bounce */
enum {
    MCPT_STAGE_RAYGEN = 0,
    MCPT_STAGE_TRAVERSE = 1,
    MCPT_STAGE_SHADE = 2,
    MCPT_STAGE_BVH_BUILD = 3,      /* upload-time mesh tree build (MCPT_OPT_MESH_BVH) */
    MCPT_STAGE_TRAVERSE_INTERIOR = 4,
    MCPT_STAGE_SHADE_EXIT = 5,     /* scenes with meshes only; elsewhere the interior traversal builds the exit rays */
    MCPT_STAGE_ACCUMULATE = 6,
    MCPT_STAGE_RESOLVE = 7,
    MCPT_STAGE_COUNT = 8
};

/* per-stage device time (CUDA events on the context's stream) and launch counts
 * gathered while MCPT_OPT_PROFILE_STAGES is on; reading them drains the record */
typedef struct mcpt_stage_times {
    double ms[MCPT_STAGE_COUNT];
    uint64_t launches[MCPT_STAGE_COUNT];
} mcpt_stage_times;

/* counters filled by the wavefront since the last mcpt_clear / mcpt_reset_stats */
typedef struct mcpt_stats {
    uint64_t paths;      /* camera paths started (pixels x passes)              */
    uint64_t rays;       /* traverse_all_bvh calls (primary + bounce + interior) */
    uint64_t segments;   /* shaded hits                                         */
    uint64_t launches;   /* kernels launched by this library                    */
    uint64_t boxes;      /* boxes tested      (only with MCPT_OPT_COUNT_WORK)   */
    uint64_t prim_tests; /* records tested    (only with MCPT_OPT_COUNT_WORK)   */
    uint64_t tri_tests;  /* triangles tested  (only with MCPT_OPT_COUNT_WORK)   */
} mcpt_stats;

int mcpt_abi_version(void);
const char* mcpt_last_error(const mcpt_ctx* ctx);
const char* mcpt_error_string(int code);

/* GL context + program creation, montecarlo.cpp:342-348. device = CUDA ordinal. */
int mcpt_create(int device, mcpt_ctx** out);
int mcpt_destroy(mcpt_ctx* ctx);

/* Launch on an existing cudaStream_t (e.g. torch's current stream) instead of
 * the context's own stream; pass NULL to go back to the private stream. */
int mcpt_set_stream(mcpt_ctx* ctx, void* cuda_stream);
int mcpt_synchronize(mcpt_ctx* ctx);
int mcpt_set_option(mcpt_ctx* ctx, int option, int value);

/* finalize() uploads, gpu_bvh_scene.cpp:143-175: tables are copied to the
 * device and re-packed into the library's compact node / record layout. */
int mcpt_upload_scene(mcpt_ctx* ctx, const mcpt_scene_desc* desc);

/* resize_ogl -> fbo->resize, montecarlo.cpp:616-626: reallocates the
 * accumulation buffer (zeroed) and resets the pass counter. */
int mcpt_resize(mcpt_ctx* ctx, int width, int height);

/* uniforms 6 and 10, montecarlo.cpp:439-440: column-major view.inverse() and
 * (proj*view).inverse(), consumed as in shaders/raytracer.vert:9-22. */
int mcpt_set_camera(mcpt_ctx* ctx, const float inv_v[16], const float inv_pv[16]);

int mcpt_set_params(mcpt_ctx* ctx, const mcpt_params* p);

/* glClear of the FBO when not locked, montecarlo.cpp:423-424. */
int mcpt_clear(mcpt_ctx* ctx);

/* The draw loop, montecarlo.cpp:454-466: pass k of this call uses
 * numero_pass = first_pass + k and is blended additively (ONE,ONE) into the
 * accumulation buffer. Asynchronous on the context's stream. */
int mcpt_render(mcpt_ctx* ctx, int first_pass, int n_passes);

/* The prg_fs draw, montecarlo.cpp:470-476 (frag = accum / nb), with the
 * optional nearest-neighbour sub-sampling upscale of shaders/average.frag:9
 * (dst is (w*subsampling) x (h*subsampling); subsampling >= 1). Host
 * destinations synchronise; device destinations stay asynchronous. */
int mcpt_resolve(mcpt_ctx* ctx, float divisor, int subsampling, void* dst, int dst_kind);

/* Accumulation buffer access (new; multi-GPU spp split and checkpoint/resume).
 * Layout: float4 per pixel (rgb sum, w unused), row-major, row 0 = bottom row
 * (gl_FragCoord convention). mcpt_bind_accum makes the library accumulate
 * straight into a caller-owned device buffer (e.g. an NCCL/torch send buffer) of
 * at least w*h*16 bytes. mcpt_resize ENDS the binding: the size changed, the context
 * goes back to a zeroed buffer of its own and the caller binds again if it wants to. */
int mcpt_get_accum(mcpt_ctx* ctx, float** dev_ptr, size_t* bytes);
int mcpt_bind_accum(mcpt_ctx* ctx, float* dev_ptr, size_t bytes);
int mcpt_read_accum(mcpt_ctx* ctx, float* host_rgba, size_t bytes);
int mcpt_write_accum(mcpt_ctx* ctx, const float* host_rgba, size_t bytes);

/* One ncclReduce(sum, fp32) of the accumulation buffer to `root` on the
 * context's stream. `nccl_comm` is an ncclComm_t; NCCL symbols are resolved at
 * run time from the libnccl already loaded in the process. */
int mcpt_reduce_nccl(mcpt_ctx* ctx, void* nccl_comm, int root);

/* traverse_all_bvh as a callable, raytracer_func.frag:777-781 (any_hit != 0:
 * just_hit_bvh, :771-775 — only `index >= 0` is meaningful then). Host arrays:
 * o and d are float[3*n]. d need not be unit length. */
int mcpt_trace_rays(mcpt_ctx* ctx, int64_t n, const float* o, const float* d,
                    mcpt_hit* out, int any_hit);
/* same, with device-resident inputs/outputs (float4 o[n], float4 d[n]); asynchronous. */
int mcpt_trace_rays_device(mcpt_ctx* ctx, int64_t n, const float* o4, const float* d4,
                           mcpt_hit* out_dev, int any_hit);

/* The bounce sampler on its own: n consecutive random_ray(D, roughness) draws (tp/montecarlo.frag:49-89) of the RNG
 * stream srand() seeds from screen_tc = (tcx, tcy), numero_pass and date (shaders/raytracer_func.frag:105-110), written
 * to out_host as float[3*n]. This is what the reference's DrawSampling program draws (tp/hsphere.vert:1-41: the same
 * two functions, one point per draw, the picture of the roughness cone in README.md:5-14); here it feeds the sampler
 * parity test and montecarlo_pathtracing_b200.imageio.sampling_plot. Transcendentals as MCPT_OPT_FAST_MATH says. */
int mcpt_sample_directions(mcpt_ctx* ctx, const float D[3], float roughness, float tcx, float tcy, int numero_pass,
                           float date, int64_t n, float* out_host);

/* The same closest-hit search as mcpt_trace_rays_device for callers that cannot allocate device memory (code under
 * CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc). NO COUNTERPART IN THE REFERENCE.
 * The ray / hit sectors the traversal kernel works in live in a fixed block
 * inside the library image (a file-scope __device__ array of MCPT_TRACE_STATIC_MAX_RAYS x 64 bytes = 4 MiB, held by
 * every GPU that loads the library until the process exits); the call performs no cudaMalloc / cudaFree and calls no
 * allocating library routine. Inputs and outputs are the caller's device buffers (float4 o[n], float4 d[n],
 * mcpt_hit out[n]); asynchronous on the context's stream; records are bit-identical to mcpt_trace_rays_device.
 * n > MCPT_TRACE_STATIC_MAX_RAYS is rejected with MCPT_E_INVALID (split the batch). All calls on one device share
 * the block, which keeps the previous call's contents (nothing in it needs clearing: every slot a call reads, it
 * has written first); calls that OVERLAP on the GPU -- other streams, other contexts of the same device -- race on
 * it and must be serialised by the caller. */
#define MCPT_TRACE_STATIC_MAX_RAYS 65536
int mcpt_trace_rays_static(mcpt_ctx* ctx, int64_t n, const float* o4, const float* d4, mcpt_hit* out_dev);

/* Host-only inspection of the top-level tree the persistent warps would walk for `desc` (built exactly as
 * mcpt_upload_scene builds it; no context and no CUDA device needed): sizes, depths, and the invariants the
 * traversal relies on. top_sah as MCPT_OPT_TOP_SAH. Used by the CPU tests. */
typedef struct mcpt_tree_stats {
    int32_t nodes;          /* two-children nodes                                                        */
    int32_t leaves;         /* leaf references (a sliced primitive counts once per slab box)             */
    int32_t prims_reached;  /* distinct primitives that own at least one leaf                            */
    int32_t prims_expected; /* primitives that own a leaf slot of the uploaded tree                      */
    int32_t depth;          /* longest root-to-leaf path in two-children levels (the walk's stack bound) */
    int32_t depth4;         /* ... in four-children levels                                               */
    int32_t leaves4;        /* leaf references reachable through the four-children nodes                 */
    int32_t violations;     /* child boxes that do not contain what is below them + sample points of a
                               primitive's bounding cube that none of its leaf boxes contains            */
    double sah_cost;        /* sum over nodes of child-box half-areas x leaves below / root half-area    */
} mcpt_tree_stats;
int mcpt_inspect_walk_tree(const mcpt_scene_desc* desc, int top_sah, mcpt_tree_stats* out);

int mcpt_get_stats(mcpt_ctx* ctx, mcpt_stats* out);
int mcpt_reset_stats(mcpt_ctx* ctx);
/* synchronises the stream */
int mcpt_get_stage_times(mcpt_ctx* ctx, mcpt_stage_times* out);

#ifdef __cplusplus
}
#endif
#endif /* MCPT_H_ */
