// kernels.cuh — launchers of the wavefront stages (defined in kernels.cu).
//
// One bounce of the wavefront is the kernel chain
//   traverse (ends the misses, queues the hits) -> shade (appends the survivors to the next
//   queues) -> [traverse -> shade_exit on interior rays]
// run over device-resident queues whose lengths never leave the GPU; raygen
// opens a batch of passes and accumulate closes it.
#pragma once
#include "layout.cuh"

namespace mcpt {

struct CameraRays {
    float ori[3];     // Ori_in (same for the four corners)
    float dir[4][3];  // Dir_in at strip vertices (0,0) (1,0) (0,1) (1,1)
};

// device-side queue bookkeeping of one shade launch
struct QueueSet {
    uint32_t* cur;       // path slots to shade: the hit queue of this bounce (nullptr: identity)
    uint32_t* next;      // survivors
    uint32_t* interior;  // paths that need the interior traversal
    const uint32_t* n_cur;  // device count of `cur` (nullptr: use n_cur_host)
    uint32_t n_cur_host;
    uint32_t* n_next;
    uint32_t* n_interior;
    unsigned long long* seg_counter;  // shaded-hit statistics (may be null)
};

struct LaunchCfg {
    int sm_count;
    cudaStream_t stream;
    int traverse_occ = 2;  // resident CTAs per SM the traversal kernel is compiled for (register cap)
};

void launch_raygen(const LaunchCfg& lc, const DevParams& prm, const CameraRays& cam, Wavefront wf, int first_pass,
                   int n_passes, uint32_t npix);
// hitq != null: rays that leave the scene end in the kernel (sky), the others are appended to hitq
void launch_traverse(const LaunchCfg& lc, const DevScene& S, Wavefront wf, const uint32_t* queue, const uint32_t* n_dev,
                     uint32_t n_host, uint32_t* hitq, uint32_t* n_hit, bool first_bounce, unsigned long long* ray_counter,
                     unsigned long long* work);
void launch_shade(const LaunchCfg& lc, const DevScene& S, const DevParams& prm, Wavefront wf, const QueueSet& q,
                  bool first_bounce, bool last_bounce, int first_pass, uint32_t npix, bool fast_math);
void launch_shade_exit(const LaunchCfg& lc, const DevScene& S, const DevParams& prm, Wavefront wf, const uint32_t* queue,
                       const uint32_t* n_dev);
void launch_accumulate(const LaunchCfg& lc, Wavefront wf, float4* accum, uint32_t npix, int n_passes);
void launch_resolve(const LaunchCfg& lc, const float4* accum, int w, int h, float divisor, int subsampling, void* dst,
                    bool rgba8);
void launch_trace(const LaunchCfg& lc, const DevScene& S, int64_t n, const float4* o, const float4* d, void* hits_out,
                  bool any_hit);
// counts launches issued through the functions above (host side)
unsigned long long launches_issued();
void add_launches(unsigned long long n);

}  // namespace mcpt
