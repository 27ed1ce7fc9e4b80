// kernels.cuh — launchers of the wavefront stages (defined in kernels.cu).
//
// One bounce of the wavefront is the kernel chain
//   traverse (ends the misses, queues the hits) -> shade (appends the survivors to the next
//   queues) -> [traverse of the interior rays, which also builds their exit rays]
// run over device-resident queues whose lengths never leave the GPU; raygen
// opens a batch of passes and accumulate closes it.
#pragma once
#include "layout.cuh"

namespace mcpt {

struct CameraRays {
    float ori[3];     // Ori_in (same for the four corners)
    float dir[4][3];  // Dir_in at strip vertices (0,0) (1,0) (0,1) (1,1)
};

// A queue of path slots, split into one private region of `cap` entries per CTA of the wavefront
// grid (see kernels.cu). items == nullptr as input: the identity over n_identity slots.
// Camera rays made where they are consumed (MCPT_OPT_FUSED_RAYGEN): the first traversal of a batch computes the ray of
// slot p itself (and stores it, so that every later stage finds the ray sector as k_raygen would have left it).
struct RayGen {
    CameraRays cam;
    const float4* dirs;   // the camera direction of every pixel (launch_camera_table: npix entries, L2 resident, made once per
                          // mcpt_render -- the same for every pass); the origin is cam.ori
    int32_t width, height;
    uint32_t npix;
    float inv_npix;
    int32_t on;   // 1: the ray is stored where it is made; 2: it is not stored at all -- nothing reads a primary ray back except the
                  // fake sky of k_accumulate for paths that were never shaded, and that takes the direction from `dirs`
};
struct QueueIn {
    const uint32_t* items;
    const uint32_t* counts;  // entries in each CTA's region
    uint32_t n_identity;
    uint32_t cap;
    RayGen gen;              // on != 0 (identity queues only): slot p's ray is the camera ray of pixel p % npix
};
struct QueueOut {
    uint32_t* items;  // nullptr: no queue
    uint32_t* counts;
    uint32_t cap;
};

struct LaunchCfg {
    int sm_count;
    cudaStream_t stream;
    int traverse_occ = 2;  // resident CTAs per SM the traversal kernel is compiled for (register cap)
    int grid = 1;          // CTAs of every queue-driven kernel of the current batch (fixed: regions are per CTA)
    unsigned long long* launches = nullptr;  // the owning context's launch counter (host side, one caller thread per context)
    void count(unsigned long long n = 1) const {
        if (launches) *launches += n;
    }
};

// the table RayGen::dirs points to (k_raygen's arithmetic per pixel)
void launch_camera_table(const LaunchCfg& lc, const DevParams& prm, const CameraRays& cam, float4* dirs, uint32_t npix);
void launch_raygen(const LaunchCfg& lc, const DevParams& prm, const CameraRays& cam, Wavefront wf, int first_pass,
                   int n_passes, uint32_t npix);
// hitq.items != null: rays that hit something are appended to hitq; the others are already settled
// exit_prm != null: interior rays -- the kernel finishes the transparent case itself (new ray through the exit hit)
void launch_traverse(const LaunchCfg& lc, const DevScene& S, Wavefront wf, const QueueIn& qin, const QueueOut& hitq,
                     const DevParams* exit_prm, unsigned long long* ray_counter, unsigned long long* work);
void launch_shade(const LaunchCfg& lc, const DevScene& S, const DevParams& prm, Wavefront wf, const QueueIn& qin,
                  const QueueOut& next, const QueueOut& interior, bool first_bounce, bool last_bounce, int first_pass,
                  uint32_t npix, bool fast_math, unsigned long long* seg_counter);
// next-event estimation (MCPT_OPT_NEE): `nee` holds three sectors per path slot (shadow ray, its hit record, sums);
// shade_nee also appends the slots with a shadow ray to `shadow`; nee_resolve delivers the visible ones
void launch_shade_nee(const LaunchCfg& lc, const DevScene& S, const DevParams& prm, Wavefront wf, Wavefront nee, const QueueIn& qin,
                      const QueueOut& next, const QueueOut& interior, const QueueOut& shadow, bool first_bounce, bool last_bounce,
                      int first_pass, uint32_t npix, bool fast_math, unsigned long long* seg_counter);
void launch_nee_resolve(const LaunchCfg& lc, Wavefront nee, const QueueIn& qin);
void launch_accumulate_nee(const LaunchCfg& lc, Wavefront wf, Wavefront nee, float4* accum, uint32_t npix, int n_passes, uint32_t batch);
void launch_shade_exit(const LaunchCfg& lc, const DevScene& S, const DevParams& prm, Wavefront wf, const QueueIn& qin);
void launch_accumulate(const LaunchCfg& lc, Wavefront wf, float4* accum, uint32_t npix, int n_passes, uint32_t batch, const float4* cam_dirs);
void launch_resolve(const LaunchCfg& lc, const float4* accum, int w, int h, float divisor, int subsampling, void* dst,
                    bool rgba8);
void launch_trace(const LaunchCfg& lc, const DevScene& S, int64_t n, const float4* o, const float4* d, void* hits_out,
                  bool any_hit);
// closest-hit search of caller-provided rays through the PRODUCTION traversal kernel (k_traverse with whatever walker
// the scene selected): rays are packed into ray sectors of `wf`, traced over the identity queue with every record
// delivered to the hit sector, and unpacked into mcpt_hit records (mcpt_trace_rays, MCPT_OPT_TRACE_WALKER = 0)
void launch_trace_wavefront(const LaunchCfg& lq, const DevScene& S, int64_t n, const float4* o, const float4* d, Wavefront wf,
                            uint32_t cap, void* hits_out);

// n consecutive random_ray(D, roughness) draws of the RNG stream seeded by srand(screen_tc = (tcx, tcy), numero_pass, date)
void launch_sample_dirs(const LaunchCfg& lc, const float D[3], float roughness, float tcx, float tcy, int numero_pass, float date,
                        int64_t n, float* out_dev, bool fast_math);

}  // namespace mcpt
