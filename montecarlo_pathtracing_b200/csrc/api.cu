// api.cu — context management and the C ABI of include/mcpt.h.
//
// Owns all device memory of a context: the re-packed scene (layout.cuh), the
// accumulation buffer, and the wavefront state (32-byte sector planes, layout.cuh) sized for a batch of passes.
// mcpt_render() drives the wavefront: for each batch of passes
//   raygen -> { traverse -> shade -> [traverse of the interior rays] } x bounces -> accumulate
// with queue lengths kept on the device.
#include <cuda_runtime.h>
#include <dlfcn.h>

#include <algorithm>
#include <array>
#include <cfloat>
#include <chrono>
#include <climits>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <mutex>
#include <new>
#include <string>
#include <unordered_map>
#include <utility>
#include <vector>

#include "bvh_build.cuh"
#include "kernels.cuh"
#include "mcpt.h"

using namespace mcpt;

struct mcpt_ctx {
    int device = 0;
    int sm_count = 148;
    cudaStream_t own_stream = nullptr;
    cudaStream_t stream = nullptr;
    std::string err;

    // side streams of the upload-time tree builds (created on first use)
    static constexpr int kSideStreams = 4;
    cudaStream_t side[kSideStreams] = {nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t ev_fork = nullptr, ev_join[kSideStreams] = {nullptr, nullptr, nullptr, nullptr};

    // scene
    std::vector<void*> scene_allocs;
    DevScene scene{};
    bool have_scene = false;

    // viewer state
    DevParams prm{};
    CameraRays cam{};
    mcpt_params user_params{};
    bool have_camera = false;

    // accumulation target
    float4* accum = nullptr;
    bool accum_owned = false;
    size_t accum_bytes = 0;
    float4* cam_dirs = nullptr;   // camera direction of every pixel (MCPT_OPT_FUSED_RAYGEN), rebuilt by every mcpt_render
    size_t cam_dirs_px = 0;
    void* resolve_tmp = nullptr;  // device staging for host-destination resolves
    size_t resolve_tmp_bytes = 0;
    char* trace_wf = nullptr;     // ray + hit sectors of mcpt_trace_rays through the production kernel (grow-only)
    uint64_t trace_wf_slots = 0;

    // wavefront
    Wavefront wf{};
    std::vector<void*> wf_allocs;
    std::unordered_map<void*, size_t> alloc_bytes;         // every live block of dev_alloc
    std::vector<std::pair<size_t, void*>> alloc_cache;     // parked blocks (size, pointer)
    uint64_t wf_capacity = 0;  // path slots
    uint32_t batch_gen = 0;    // generation of the last batch traced in these buffers (path tags)
    uint32_t* queue[4] = {nullptr, nullptr, nullptr, nullptr};  // next / hit / interior / shadow (MCPT_OPT_NEE)
    Wavefront nee{};           // next-event estimation: shadow ray, its hit record, sums (three sectors per slot)
    uint64_t nee_capacity = 0;
    uint32_t* counters = nullptr;  // per-CTA region lengths of the three queues
    int counters_len = 0;
    unsigned long long* dev_stats = nullptr;  // [0] rays, [1] segments, [2] boxes, [3] records, [4] triangles

    // options
    int opt_passes_per_batch = 0;  // 0 = auto
    int opt_smem_staging = 1;
    int opt_fast_math = 0;
    int opt_profile = 0;
    int opt_flat_walk = 1;
    int opt_top_sah = 1;
    int opt_mesh_bvh = 1;
    int opt_persistent = 1;
    int opt_refill_lanes = 0;
    int opt_traverse_occ = 0;  // 0 = auto
    int opt_count_work = 0;
    int opt_trace_walker = 0;  // mcpt_trace_rays: 0 = the production kernel (k_traverse), 1 = the reference-order walk (k_trace)
    int opt_fused_raygen = 1;  // the first traversal of a batch makes its camera rays itself (no k_raygen launch)
    int opt_leaf_pool = 0;     // unified walk: > 0 = leaf tests queued per warp, a round of tests once this many items wait
    int opt_flat_pool = 2;     // flat scenes, bounces after the first: exact tests pooled per warp and primitive type
    int opt_nee = 0;           // next-event estimation on the emissive quads (a new estimator, off by default)
    int opt_unified = 1;       // scenes with meshes: one world-space tree (traverse.cuh, UWalk) instead of the two-level walk

    // stage profiling: (stage, start, stop) triples waiting to be read, plus a free list
    struct Span {
        int stage;
        cudaEvent_t a, b;
    };
    std::vector<Span> spans;
    std::vector<cudaEvent_t> free_events;

    uint64_t paths_started = 0;
    unsigned long long launches = 0;  // kernels launched by this context since the last mcpt_reset_stats
};

namespace {

constexpr int kStatWords = 5;

int fail(mcpt_ctx* c, int code, const char* fmt, ...) {
    if (c) {
        char buf[512];
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(buf, sizeof buf, fmt, ap);
        va_end(ap);
        c->err = buf;
    }
    return code;
}

#define CK(ctx, call)                                                                                   \
    do {                                                                                                \
        cudaError_t e_ = (call);                                                                        \
        if (e_ != cudaSuccess)                                                                          \
            return fail(ctx, e_ == cudaErrorMemoryAllocation ? MCPT_E_NOMEM : MCPT_E_CUDA, "%s: %s", #call, \
                        cudaGetErrorString(e_));                                                        \
    } while (0)

// Device allocations go through a small per-context cache: free_all() parks the blocks, dev_alloc()
// takes a parked block of exactly the requested size before asking the driver, trim_cache() returns
// what nobody took. Re-uploading a scene of the same shape (every frame of an animated viewer, every
// step of the end-to-end bench) therefore costs no cudaMalloc / cudaFree at all -- those calls
// synchronise the device and were measured at 0.2 .. 1.3 s spikes per upload of the 16-mesh scene.
template <class T>
int dev_alloc(mcpt_ctx* c, std::vector<void*>& owner, T** out, size_t count) {
    void* p = nullptr;
    size_t bytes = std::max<size_t>(count, 1) * sizeof(T);
    for (size_t i = 0; i < c->alloc_cache.size(); ++i)
        if (c->alloc_cache[i].first == bytes) {
            p = c->alloc_cache[i].second;
            c->alloc_cache[i] = c->alloc_cache.back();
            c->alloc_cache.pop_back();
            break;
        }
    if (!p) {
        cudaError_t e = cudaMalloc(&p, bytes);
        if (e != cudaSuccess) return fail(c, MCPT_E_NOMEM, "cudaMalloc(%zu bytes): %s", bytes, cudaGetErrorString(e));
        c->alloc_bytes[p] = bytes;
    }
    owner.push_back(p);
    *out = static_cast<T*>(p);
    return MCPT_OK;
}

// the caller has made sure the device no longer uses the blocks
void free_all(mcpt_ctx* c, std::vector<void*>& v) {
    for (void* p : v) c->alloc_cache.emplace_back(c->alloc_bytes[p], p);
    v.clear();
}

void trim_cache(mcpt_ctx* c) {
    for (auto& e : c->alloc_cache) {
        cudaFree(e.second);
        c->alloc_bytes.erase(e.second);
    }
    c->alloc_cache.clear();
}

int ensure_side_streams(mcpt_ctx* c) {
    if (c->ev_fork) return MCPT_OK;
    for (int k = 0; k < mcpt_ctx::kSideStreams; ++k) {
        CK(c, cudaStreamCreateWithFlags(&c->side[k], cudaStreamNonBlocking));
        CK(c, cudaEventCreateWithFlags(&c->ev_join[k], cudaEventDisableTiming));
    }
    CK(c, cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming));
    return MCPT_OK;
}

// register cap of the traversal kernel: the flat walk fits 64 registers without spilling (4 CTAs
// of 256 threads per SM); the tree and mesh walks spill below their natural 80 / 103 registers.
LaunchCfg lcfg(const mcpt_ctx* c) {
    int occ = c->opt_traverse_occ;
    // measured: flat walk and persistent warps are best at 4 (latency hiding beats the spills of the
    // 64-register cap: +7..10 % over 3); the one-ray-per-lane tree walk at 3
    if (occ == 0) occ = (c->have_scene && (c->scene.n_flat > 0 || c->scene.persistent)) ? 4 : 3;
    LaunchCfg lc{c->sm_count, c->stream, occ};
    lc.launches = const_cast<unsigned long long*>(&c->launches);
    return lc;
}

// rows of the affine part of a column-major mat4
Aff rows_of(const float m[16]) {
    Aff a;
    a.r0 = make_float4(m[0], m[4], m[8], m[12]);
    a.r1 = make_float4(m[1], m[5], m[9], m[13]);
    a.r2 = make_float4(m[2], m[6], m[10], m[14]);
    return a;
}

// Inflate a box so the FMA slab test (traverse.cuh) never rejects what the exact
// arithmetic would accept. Empty boxes (min > max) stay empty.
DevNode pack_node(const mcpt_bb& b, int aux) {
    DevNode n;
    float lo[3], hi[3];
    for (int k = 0; k < 3; ++k) {
        lo[k] = b.min[k];
        hi[k] = b.max[k];
        if (lo[k] <= hi[k]) {
            float mag = std::max(std::fabs(lo[k]), std::fabs(hi[k]));
            float pad = 1e-5f * std::max(mag, hi[k] - lo[k]) + 1e-7f;
            lo[k] -= pad;
            hi[k] += pad;
        }
    }
    n.lo = make_float4(lo[0], lo[1], lo[2], 0.0f);
    n.hi = make_float4(hi[0], hi[1], hi[2], 0.0f);
    std::memcpy(&n.lo.w, &aux, sizeof aux);
    return n;
}

DevNode empty_node() {
    DevNode n;
    int aux = -1;
    n.lo = make_float4(FLT_MAX, FLT_MAX, FLT_MAX, 0.0f);
    n.hi = make_float4(-FLT_MAX, -FLT_MAX, -FLT_MAX, 0.0f);
    std::memcpy(&n.lo.w, &aux, sizeof aux);
    return n;
}

}  // namespace

extern "C" {

int mcpt_abi_version(void) { return MCPT_ABI_VERSION; }

const char* mcpt_last_error(const mcpt_ctx* ctx) { return ctx ? ctx->err.c_str() : "null context"; }

const char* mcpt_error_string(int code) {
    switch (code) {
        case MCPT_OK: return "ok";
        case MCPT_E_INVALID: return "invalid argument or state";
        case MCPT_E_CUDA: return "CUDA error";
        case MCPT_E_NODEVICE: return "no CUDA device";
        case MCPT_E_NOMEM: return "out of memory";
        case MCPT_E_NCCL: return "NCCL error";
        case MCPT_E_UNSUPPORTED: return "unsupported";
    }
    return "unknown error";
}

int mcpt_create(int device, mcpt_ctx** out) {
    if (!out) return MCPT_E_INVALID;
    *out = nullptr;
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0) return MCPT_E_NODEVICE;
    if (device < 0 || device >= n) return MCPT_E_NODEVICE;
    mcpt_ctx* c = new (std::nothrow) mcpt_ctx();
    if (!c) return MCPT_E_NOMEM;
    c->device = device;
    if (cudaSetDevice(device) != cudaSuccess) {
        delete c;
        return MCPT_E_CUDA;
    }
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) == cudaSuccess) c->sm_count = prop.multiProcessorCount;
    if (cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking) != cudaSuccess) {
        delete c;
        return MCPT_E_CUDA;
    }
    c->stream = c->own_stream;
    void* p = nullptr;
    if (cudaMalloc(&p, kStatWords * sizeof(unsigned long long)) != cudaSuccess) {
        cudaStreamDestroy(c->own_stream);
        delete c;
        return MCPT_E_NOMEM;
    }
    c->dev_stats = static_cast<unsigned long long*>(p);
    cudaMemsetAsync(c->dev_stats, 0, kStatWords * sizeof(unsigned long long), c->stream);
    c->prm.nb_bounces = 3;       // RTViewer defaults, montecarlo.cpp:126-141
    c->prm.refract_ind = 1.0f;
    c->user_params.nb_bounces = 3;
    c->user_params.refract_ind = 1.0f;
    *out = c;
    return MCPT_OK;
}

int mcpt_destroy(mcpt_ctx* c) {
    if (!c) return MCPT_E_INVALID;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    free_all(c, c->scene_allocs);
    free_all(c, c->wf_allocs);
    trim_cache(c);
    if (c->accum_owned && c->accum) cudaFree(c->accum);
    if (c->dev_stats) cudaFree(c->dev_stats);
    if (c->cam_dirs) cudaFree(c->cam_dirs);
    if (c->resolve_tmp) cudaFree(c->resolve_tmp);
    if (c->trace_wf) cudaFree(c->trace_wf);
    for (auto& sp : c->spans) {
        cudaEventDestroy(sp.a);
        cudaEventDestroy(sp.b);
    }
    for (cudaEvent_t e : c->free_events) cudaEventDestroy(e);
    for (int k = 0; k < mcpt_ctx::kSideStreams; ++k) {
        if (c->side[k]) cudaStreamDestroy(c->side[k]);
        if (c->ev_join[k]) cudaEventDestroy(c->ev_join[k]);
    }
    if (c->ev_fork) cudaEventDestroy(c->ev_fork);
    if (c->own_stream) cudaStreamDestroy(c->own_stream);
    delete c;
    return MCPT_OK;
}

int mcpt_set_stream(mcpt_ctx* c, void* s) {
    if (!c) return MCPT_E_INVALID;
    c->stream = s ? static_cast<cudaStream_t>(s) : c->own_stream;
    return MCPT_OK;
}

int mcpt_synchronize(mcpt_ctx* c) {
    if (!c) return MCPT_E_INVALID;
    CK(c, cudaSetDevice(c->device));
    CK(c, cudaStreamSynchronize(c->stream));
    return MCPT_OK;
}

int mcpt_set_option(mcpt_ctx* c, int option, int value) {
    if (!c) return MCPT_E_INVALID;
    switch (option) {
        case MCPT_OPT_PASSES_PER_BATCH:
            if (value < 0) return fail(c, MCPT_E_INVALID, "passes per batch must be >= 0");
            c->opt_passes_per_batch = value;
            return MCPT_OK;
        case MCPT_OPT_SMEM_STAGING:
            c->opt_smem_staging = std::min(std::max(value, 0), 2);
            if (c->have_scene) return fail(c, MCPT_E_INVALID, "set MCPT_OPT_SMEM_STAGING before mcpt_upload_scene");
            return MCPT_OK;
        case MCPT_OPT_FAST_MATH: c->opt_fast_math = value ? 1 : 0; return MCPT_OK;
        case MCPT_OPT_MESH_BVH: c->opt_mesh_bvh = value ? 1 : 0; return MCPT_OK;
        case MCPT_OPT_REFILL_LANES: c->opt_refill_lanes = std::min(std::max(value, 0), 32); return MCPT_OK;
        case MCPT_OPT_PERSISTENT: c->opt_persistent = value ? 1 : 0; return MCPT_OK;
        case MCPT_OPT_PROFILE_STAGES: c->opt_profile = value ? 1 : 0; return MCPT_OK;
        case MCPT_OPT_COUNT_WORK: c->opt_count_work = value ? 1 : 0; return MCPT_OK;
        case MCPT_OPT_FLAT_WALK: c->opt_flat_walk = value ? 1 : 0; return MCPT_OK;
        case MCPT_OPT_TOP_SAH: c->opt_top_sah = value ? 1 : 0; return MCPT_OK;
        case MCPT_OPT_TRACE_WALKER: c->opt_trace_walker = value ? 1 : 0; return MCPT_OK;
        case MCPT_OPT_UNIFIED: c->opt_unified = value ? 1 : 0; return MCPT_OK;
        case MCPT_OPT_NEE: c->opt_nee = value ? 1 : 0; return MCPT_OK;
        case MCPT_OPT_FLAT_POOL:
            if (value < 0 || value > 2) return fail(c, MCPT_E_INVALID, "flat pool must be 0, 1 or 2");
            c->opt_flat_pool = value;
            return MCPT_OK;
        case MCPT_OPT_FUSED_RAYGEN: c->opt_fused_raygen = value ? 1 : 0; return MCPT_OK;
        case MCPT_OPT_LEAF_POOL:
            if (value < 0 || value > 32) return fail(c, MCPT_E_INVALID, "leaf pool trigger must be 0 (off) or 1..32 items");
            c->opt_leaf_pool = value;
            return MCPT_OK;
        case MCPT_OPT_TRAVERSE_OCC:
            if (value != 0 && (value < 2 || value > 6))
                return fail(c, MCPT_E_INVALID, "traverse occupancy must be 0 (auto) or 2..6 CTAs per SM");
            c->opt_traverse_occ = value;
            return MCPT_OK;
    }
    return fail(c, MCPT_E_INVALID, "unknown option %d", option);
}

}  // extern "C"

#include "api_render.inl"
#include "api_scene.inl"
#include "api_static.inl"
