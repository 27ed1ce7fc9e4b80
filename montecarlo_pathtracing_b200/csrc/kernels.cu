// kernels.cu — the wavefront stages as sm_100a kernels.
//
// Compiled with -fmad=false (see vecmath.cuh). All kernels are grid-stride over a
// device-resident work count, launched with a fixed grid that is a multiple of
// the SM count, so a bounce never needs a host round trip to size its launches.
#include <cfloat>
#include <cstdio>

#include "kernels.cuh"
#include "mcpt.h"
#include "shade.cuh"
#include "traverse.cuh"

namespace mcpt {

static unsigned long long g_launches = 0;
unsigned long long launches_issued() { return g_launches; }
void add_launches(unsigned long long n) { g_launches += n; }

constexpr int kBlock = 256;

static inline int grid_for(const LaunchCfg& lc, uint64_t n, int blocks_per_sm) {
    uint64_t want = (n + kBlock - 1) / kBlock;
    uint64_t cap = uint64_t(lc.sm_count) * uint64_t(blocks_per_sm);
    if (want > cap) want = cap;
    if (want < 1) want = 1;
    return int(want);
}

// ------------------------------------------------------------------ staging
// Copy the staged part of the scene (top-level nodes, primitive records, top
// levels of every mesh tree) into dynamic shared memory. Layout:
//   [top nodes: 1 + smem_top_nodes][mesh nodes: smem_mesh_nodes][prims: smem_prims]
// Index 0 of the top-node block is unused so heap indices stay 1-based.
__device__ __forceinline__ SmemScene stage_scene(const DevScene& S, unsigned char* smem) {
    DevNode* top = reinterpret_cast<DevNode*>(smem);
    DevNode* mesh = top + (S.smem_top_nodes + 1);
    DevPrim* prims = reinterpret_cast<DevPrim*>(mesh + S.smem_mesh_nodes);
    {
        // the flat leaf list (small scenes) or the top levels of the tree
        const float4* src = reinterpret_cast<const float4*>(S.n_flat > 0 ? S.flat : S.top_nodes);
        float4* dst = reinterpret_cast<float4*>(top);
        const int n4 = (S.smem_top_nodes + 1) * 2;
        for (int i = threadIdx.x; i < n4; i += blockDim.x) dst[i] = __ldg(src + i);
    }
    for (int m = 0; m < S.n_meshes; ++m) {
        const DevMesh& M = S.meshes[m];
        if (M.smem_nodes <= 0) continue;
        const float4* src = reinterpret_cast<const float4*>(M.nodes + 1);
        float4* dst = reinterpret_cast<float4*>(mesh + M.smem_off);
        const int n4 = M.smem_nodes * 2;
        for (int i = threadIdx.x; i < n4; i += blockDim.x) dst[i] = __ldg(src + i);
    }
    {
        const float4* src = reinterpret_cast<const float4*>(S.prims);
        float4* dst = reinterpret_cast<float4*>(prims);
        const int n4 = S.smem_prims * int(sizeof(DevPrim) / sizeof(float4));
        for (int i = threadIdx.x; i < n4; i += blockDim.x) dst[i] = __ldg(src + i);
    }
    __syncthreads();
    SmemScene sm;
    sm.top_nodes = top;
    sm.prims = prims;
    sm.mesh_nodes = mesh;
    return sm;
}

// ------------------------------------------------------------------- raygen
// raytracer.vert:9-22 + main.frag + raytrace()/srand() (tp/montecarlo.frag:182-188):
// Dir_in is the barycentric interpolation, over the two triangles of the
// full-screen strip, of the per-corner normalised directions; it is normalised
// again; the path starts with attenuation 0.8 and total 0 (:105-107).
__global__ void __launch_bounds__(kBlock) k_raygen(DevParams prm, CameraRays cam, Wavefront wf, int first_pass, int n_passes,
                                                    uint32_t npix) {
    const uint64_t total = uint64_t(npix) * uint64_t(n_passes);
    for (uint64_t p = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x; p < total; p += uint64_t(gridDim.x) * blockDim.x) {
        const uint32_t k = uint32_t(p / npix);
        const uint32_t pix = uint32_t(p - uint64_t(k) * npix);
        const int x = int(pix % uint32_t(prm.width));
        const int y = int(pix / uint32_t(prm.width));
        const float u = (float(x) + 0.5f) / float(prm.width);
        const float v = (float(y) + 0.5f) / float(prm.height);
        f3 d0 = mk3(cam.dir[0][0], cam.dir[0][1], cam.dir[0][2]);
        f3 d1 = mk3(cam.dir[1][0], cam.dir[1][1], cam.dir[1][2]);
        f3 d2 = mk3(cam.dir[2][0], cam.dir[2][1], cam.dir[2][2]);
        f3 d3 = mk3(cam.dir[3][0], cam.dir[3][1], cam.dir[3][2]);
        f3 dir;
        if (u + v <= 1.0f)
            dir = d0 * ((1.0f - u) - v) + d1 * u + d2 * v;
        else
            dir = d1 * (1.0f - v) + d2 * (1.0f - u) + d3 * ((u + v) - 1.0f);
        dir = normalize3(dir);
        // attenuation 0.8, total 0 and the RNG seed of a fresh path are implied (k_shade, first
        // bounce); every path writes its result exactly once, so nothing else is initialised here
        wf.ray_o[p] = make_float4(cam.ori[0], cam.ori[1], cam.ori[2], 0.0f);
        wf.ray_d[p] = make_float4(dir.x, dir.y, dir.z, 0.0f);
    }
}

void launch_raygen(const LaunchCfg& lc, const DevParams& prm, const CameraRays& cam, Wavefront wf, int first_pass,
                   int n_passes, uint32_t npix) {
    uint64_t total = uint64_t(npix) * uint64_t(n_passes);
    k_raygen<<<grid_for(lc, total, 16), kBlock, 0, lc.stream>>>(prm, cam, wf, first_pass, n_passes, npix);
    ++g_launches;
}

// ----------------------------------------------------------------- traverse
// warp-reduce the optional work counters and add them to work[0..2]
__device__ __forceinline__ void flush_work(const Work& wk, unsigned long long* work) {
    uint32_t b = wk.boxes, p = wk.prims, t = wk.tris;
    for (int off = 16; off > 0; off >>= 1) {
        b += __shfl_down_sync(0xffffffffu, b, off);
        p += __shfl_down_sync(0xffffffffu, p, off);
        t += __shfl_down_sync(0xffffffffu, t, off);
    }
    if ((threadIdx.x & 31) == 0) {
        if (b) atomicAdd(work + 0, (unsigned long long)b);
        if (p) atomicAdd(work + 1, (unsigned long long)p);
        if (t) atomicAdd(work + 2, (unsigned long long)t);
    }
}

// Append the lanes with `take` set to a device queue: warps ballot, the block's warp counts are
// combined in shared memory and ONE thread reserves the block's range with a single atomic (a
// same-address atomic per warp serialises at the L2: measured +35 % on the traversal kernel), then
// every taking lane writes at base + (takers before it). Must be reached by all threads of the
// block the same number of times; `slot` alternates 0/1 between consecutive calls so that two
// barriers per call suffice.
struct QueueAppender {
    uint32_t warp_cnt[2][kBlock / 32];
    uint32_t base[2];
};

__device__ __forceinline__ void block_append(QueueAppender& qa, int slot, bool take, uint32_t value, uint32_t* __restrict__ out,
                                             uint32_t* __restrict__ n_out) {
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const uint32_t m = __ballot_sync(0xffffffffu, take);
    if (lane == 0) qa.warp_cnt[slot][warp] = __popc(m);
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t sum = 0;
        for (int w = 0; w < kBlock / 32; ++w) {
            const uint32_t c = qa.warp_cnt[slot][w];
            qa.warp_cnt[slot][w] = sum;
            sum += c;
        }
        qa.base[slot] = sum ? atomicAdd(n_out, sum) : 0u;
    }
    __syncthreads();
    if (take) out[qa.base[slot] + qa.warp_cnt[slot][warp] + __popc(m & ((1u << lane) - 1u))] = value;
}

// Where the closest-hit search of one bounce delivers its result. With a hit queue (the main
// traversal of the Monte Carlo mode) a ray that leaves the scene ends here -- the fake sky of
// tp/montecarlo.frag:118-119 is written straight to `result` -- and the rays that hit something are
// appended to `hitq` (block_append), so the shading kernel only ever sees hits.
// Without one (interior rays, debug shaders) every ray's record goes to hit_a / hit_b.
struct HitSink {
    uint32_t* hitq;
    uint32_t* n_hit;
    int first_bounce;  // total = 0, attenuation = 0.8 are implicit (tp/montecarlo.frag:105-107)
};

__device__ __forceinline__ void deliver(const Wavefront& wf, const HitSink& hs, QueueAppender& qa, int slot, uint32_t p,
                                        bool valid, const float4& d4, const Best& best) {
    const bool hit = valid && best.h.prim >= 0;
    if (hit || (valid && !hs.hitq)) {
        wf.hit_a[p] = make_float4(best.h.pl.x, best.h.pl.y, best.h.pl.z, best.h.dist);
        wf.hit_b[p] = make_int2(best.h.prim, best.h.sub);
    } else if (valid) {
        f3 total = mk3(0.0f, 0.0f, 0.0f), attenu = mk3(0.8f, 0.8f, 0.8f);
        if (!hs.first_bounce) {
            const float4 t4 = wf.total[p], a4 = wf.attenu[p];
            total = mk3(t4.x, t4.y, t4.z);
            attenu = mk3(a4.x, a4.y, a4.z);
        }
        const f3 res = sky(total, attenu, mk3(d4.x, d4.y, d4.z));
        wf.result[p] = make_float4(res.x, res.y, res.z, 0.0f);
    }
    if (hs.hitq) block_append(qa, slot, hit, p, hs.hitq, hs.n_hit);  // block-uniform condition
}

// FLAT: warp-coherent flat walk (S.n_flat > 0); MESHES: the scene has triangle meshes.
// Every lane of a warp stays in the loop together (warp-uniform trip count).
template <bool COUNT, int MINB, bool FLAT, bool MESHES>
__global__ void __launch_bounds__(kBlock, MINB) k_traverse(DevScene S, Wavefront wf, const uint32_t* __restrict__ queue,
                                                            const uint32_t* __restrict__ n_dev, uint32_t n_host, HitSink hs,
                                                            unsigned long long* ray_counter, unsigned long long* work) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const uint32_t n = n_dev ? *n_dev : n_host;
    if (uint64_t(blockIdx.x) * blockDim.x >= n) return;
    if (blockIdx.x == 0 && threadIdx.x == 0 && ray_counter) atomicAdd(ray_counter, (unsigned long long)n);
    SmemScene sm = stage_scene(S, smem_raw);
    __shared__ QueueAppender qa;
    Work wk;
    int slot = 0;
    for (uint32_t base = blockIdx.x * blockDim.x; base < n; base += gridDim.x * blockDim.x, slot ^= 1) {
        const uint32_t i = base + threadIdx.x;
        const bool valid = i < n;
        const uint32_t p = valid ? (queue ? queue[i] : i) : 0u;
        float4 o4 = make_float4(0, 0, 0, 0), d4 = make_float4(0, 0, 0, 0);
        if (valid) {
            o4 = wf.ray_o[p];
            d4 = wf.ray_d[p];
        }
        Best best;
        if (FLAT)
            scene_walk_flat<COUNT>(wk, S, sm, mk3(o4.x, o4.y, o4.z), mk3(d4.x, d4.y, d4.z), false, valid, best);
        else if (valid)
            scene_walk<COUNT, MESHES>(wk, S, sm, mk3(o4.x, o4.y, o4.z), mk3(d4.x, d4.y, d4.z), false, best);
        deliver(wf, hs, qa, slot, p, valid, d4, best);
    }
    if (COUNT) flush_work(wk, work);
}

template <bool COUNT, int MINB>
static void launch_traverse_kind(const LaunchCfg& lc, int grid, const DevScene& S, Wavefront wf, const uint32_t* queue,
                                 const uint32_t* n_dev, uint32_t n_host, const HitSink& hs, unsigned long long* ray_counter,
                                 unsigned long long* work) {
    if (S.n_flat > 0)
        k_traverse<COUNT, MINB, true, false><<<grid, kBlock, S.smem_bytes, lc.stream>>>(S, wf, queue, n_dev, n_host, hs, ray_counter, work);
    else if (S.n_meshes == 0)
        k_traverse<COUNT, MINB, false, false><<<grid, kBlock, S.smem_bytes, lc.stream>>>(S, wf, queue, n_dev, n_host, hs, ray_counter, work);
    else
        k_traverse<COUNT, MINB, false, true><<<grid, kBlock, S.smem_bytes, lc.stream>>>(S, wf, queue, n_dev, n_host, hs, ray_counter, work);
}

void launch_traverse(const LaunchCfg& lc, const DevScene& S, Wavefront wf, const uint32_t* queue, const uint32_t* n_dev,
                     uint32_t n_host, uint32_t* hitq, uint32_t* n_hit, bool first_bounce, unsigned long long* ray_counter,
                     unsigned long long* work) {
    uint64_t n = n_dev ? 0xffffffffull : n_host;
    const int grid = grid_for(lc, n, 8);
    const HitSink hs{hitq, n_hit, first_bounce ? 1 : 0};
    if (work)
        launch_traverse_kind<true, 2>(lc, grid, S, wf, queue, n_dev, n_host, hs, ray_counter, work);
    else if (lc.traverse_occ >= 4)
        launch_traverse_kind<false, 4>(lc, grid, S, wf, queue, n_dev, n_host, hs, ray_counter, nullptr);
    else if (lc.traverse_occ == 3)
        launch_traverse_kind<false, 3>(lc, grid, S, wf, queue, n_dev, n_host, hs, ray_counter, nullptr);
    else
        launch_traverse_kind<false, 2>(lc, grid, S, wf, queue, n_dev, n_host, hs, ray_counter, nullptr);
    ++g_launches;
}

// traverse_all_bvh as a callable on caller-provided rays (mcpt_trace_rays)
__global__ void __launch_bounds__(kBlock) k_trace(DevScene S, int64_t n, const float4* __restrict__ o,
                                                   const float4* __restrict__ d, mcpt_hit* __restrict__ out, int any_hit) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    if (int64_t(blockIdx.x) * blockDim.x >= n) return;
    SmemScene sm = stage_scene(S, smem_raw);
    for (int64_t base = int64_t(blockIdx.x) * blockDim.x; base < n; base += int64_t(gridDim.x) * blockDim.x) {
        const int64_t i = base + threadIdx.x;
        const bool valid = i < n;
        float4 o4 = make_float4(0, 0, 0, 0), d4 = make_float4(0, 0, 0, 0);
        if (valid) {
            o4 = o[i];
            d4 = d[i];
        }
        Best best;
        Work wk;
        f3 O = mk3(o4.x, o4.y, o4.z);
        if (S.n_flat > 0) {
            scene_walk_flat<false>(wk, S, sm, O, mk3(d4.x, d4.y, d4.z), any_hit != 0, valid, best);
        } else if (valid) {
            scene_walk<false>(wk, S, sm, O, mk3(d4.x, d4.y, d4.z), any_hit != 0, best);
        }
        if (!valid) continue;
        mcpt_hit h;
        h.index = best.h.prim;
        h.dist = best.h.dist;
        h.pl[0] = best.h.pl.x;
        h.pl[1] = best.h.pl.y;
        h.pl[2] = best.h.pl.z;
        if (best.h.prim >= 0) {
            const DevPrim* pp = S.prims + best.h.prim;
            Aff fwd;
            fwd.r0 = pp->fwd.r0;
            fwd.r1 = pp->fwd.r1;
            fwd.r2 = pp->fwd.r2;
            f3 pg = xform_point(fwd, best.h.pl);
            h.pg[0] = pg.x;
            h.pg[1] = pg.y;
            h.pg[2] = pg.z;
            h.shape = pp->type;
            h.tri_index = (pp->type == MCPT_PRIM_MESH) ? best.h.sub : -1;
            h.dir = (pp->type == MCPT_PRIM_MESH) ? 0 : best.h.sub;
        } else {
            h.pg[0] = h.pg[1] = h.pg[2] = 0.0f;
            h.shape = -1;
            h.tri_index = -1;
            h.dir = -1;
        }
        out[i] = h;
    }
}

void launch_trace(const LaunchCfg& lc, const DevScene& S, int64_t n, const float4* o, const float4* d, void* hits_out,
                  bool any_hit) {
    k_trace<<<grid_for(lc, uint64_t(n), 8), kBlock, S.smem_bytes, lc.stream>>>(S, n, o, d, static_cast<mcpt_hit*>(hits_out),
                                                                               any_hit ? 1 : 0);
    ++g_launches;
}

// (shade / compact / accumulate / resolve follow in kernels_shade.inl)
#include "kernels_shade.inl"

}  // namespace mcpt
