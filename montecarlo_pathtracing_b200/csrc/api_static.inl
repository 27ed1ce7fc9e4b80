// api_static.inl — mcpt_trace_rays_static: the ray-batch entry point for callers that cannot allocate.
// Included by api.cu (same translation unit as the array below: its address is only valid here).
//
// No counterpart in the reference. Some callers cannot call
// cudaMalloc: code under CUDA graph capture, tight serving loops, harnesses that forbid it.
// mcpt_trace_rays_device keeps its ray / hit sectors in a grow-only cudaMalloc block; this variant runs the same
// kernels (k_trace_pack -> k_traverse -> k_trace_unpack, kernels.cu) over a FIXED scratch block that is part of
// the library image: a file-scope __device__ array of MCPT_TRACE_STATIC_MAX_RAYS x 64 bytes (4 MiB), held by
// every GPU that loads the library until the process exits (with CUDA's lazy module loading the driver
// materialises it on a device when this module is first touched there: the set-up below). The call allocates
// nothing on the device: no cudaMalloc / cudaFree, no allocating library routine.
//
// The array's name is a host-side stand-in in host code, not a device address, so the address is looked up once
// per device with cudaGetSymbolAddress (that device current) and handed to the kernels as an ordinary pointer.
// Every call on a device shares the block and finds the previous call's contents in it. Nothing in it has to
// start at zero: k_trace_pack writes the ray sector of every slot < n before k_traverse reads it, and k_traverse
// (no hit queue: every record is delivered) writes the hit sector of every slot < n before k_trace_unpack reads
// it; slots >= n are never read. Calls that overlap on the GPU (two streams, two contexts on one device) race on
// the block: the caller serialises them.

namespace {

constexpr int kStaticMaxDevices = 64;
constexpr size_t kStaticPlaneBytes = size_t(MCPT_TRACE_STATIC_MAX_RAYS) * 32;  // one 32-byte sector per ray

// plane 0: ray sectors, plane 1: hit sectors (Wavefront::sec, layout.cuh)
static __device__ __align__(256) unsigned char g_trace_static_scratch[2 * kStaticPlaneBytes];

std::mutex g_static_mu;
void* g_static_addr[kStaticMaxDevices] = {};

// device address of the scratch block on `device`, which must be the current device
int static_scratch_address(mcpt_ctx* c, int device, void** out) {
    if (device < 0 || device >= kStaticMaxDevices) return fail(c, MCPT_E_UNSUPPORTED, "device ordinal %d >= %d", device, kStaticMaxDevices);
    std::lock_guard<std::mutex> lock(g_static_mu);
    if (!g_static_addr[device]) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_trace_static_scratch);
        if (e != cudaSuccess || !p) return fail(c, MCPT_E_CUDA, "cudaGetSymbolAddress(static trace scratch): %s", cudaGetErrorString(e));
        g_static_addr[device] = p;
    }
    *out = g_static_addr[device];
    return MCPT_OK;
}

}  // namespace

extern "C" int mcpt_trace_rays_static(mcpt_ctx* c, int64_t n, const float* o4, const float* d4, mcpt_hit* out_dev) {
    if (!c || n < 0 || (n > 0 && (!o4 || !d4 || !out_dev))) return MCPT_E_INVALID;
    if (!c->have_scene) return fail(c, MCPT_E_INVALID, "mcpt_trace_rays_static before mcpt_upload_scene");
    if (n > int64_t(MCPT_TRACE_STATIC_MAX_RAYS))
        return fail(c, MCPT_E_INVALID, "mcpt_trace_rays_static: %lld rays exceed the static scratch block (MCPT_TRACE_STATIC_MAX_RAYS = %d); split the batch or use mcpt_trace_rays_device",
                    (long long)n, int(MCPT_TRACE_STATIC_MAX_RAYS));
    if (n == 0) return MCPT_OK;
    CK(c, cudaSetDevice(c->device));
    void* scratch = nullptr;
    int rc = static_scratch_address(c, c->device, &scratch);
    if (rc) return rc;
    Wavefront wf{};
    wf.rec = static_cast<char*>(scratch);
    wf.plane = kStaticPlaneBytes;
    LaunchCfg lq = lcfg(c);
    lq.grid = wavefront_grid(c, uint64_t(n));
    launch_trace_wavefront(lq, c->scene, n, reinterpret_cast<const float4*>(o4), reinterpret_cast<const float4*>(d4), wf,
                           queue_cap(c, uint64_t(n)), out_dev);
    CK(c, cudaGetLastError());
    return MCPT_OK;
}
