// kernels.cu — the wavefront stages as sm_100a kernels.
//
// Compiled with -fmad=false (see vecmath.cuh). All kernels are grid-stride over a
// device-resident work count, launched with a fixed grid that is a multiple of
// the SM count, so a bounce never needs a host round trip to size its launches.
#include <atomic>
#include <cfloat>
#include <cstdio>
#include <cstring>

#include "kernels.cuh"
#include "mcpt.h"
#include "shade.cuh"
#include "traverse.cuh"

namespace mcpt {

constexpr int kBlock = 256;

// Kernels that take more than 48 KB of dynamic shared memory need the attribute once per DEVICE (contexts of one process may
// live on different GPUs): one bit per device ordinal, one word per kernel
// (the kernel is a template argument: kernels of the same signature must not share the word).
template <auto Kernel>
static void allow_large_smem() {
    static std::atomic<unsigned long long> done{0ull};
    int dev = 0;
    cudaGetDevice(&dev);
    const unsigned long long bit = 1ull << (dev & 63);
    if (done.load(std::memory_order_relaxed) & bit) return;
    if (cudaFuncSetAttribute(Kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024) == cudaSuccess)
        done.fetch_or(bit, std::memory_order_relaxed);
}

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
//   [top nodes: 1 + smem_top_nodes][mesh nodes: smem_mesh_nodes][prims: smem_prims][unified nodes: u_smem_nodes]
// Index 0 of the top-node block is unused so heap indices stay 1-based.
__device__ __forceinline__ SmemScene stage_scene(const DevScene& S, unsigned char* smem) {
    DevNode* top = reinterpret_cast<DevNode*>(smem);
    DevNode* mesh = top + (S.smem_top_nodes + 1);
    DevPrim* prims = reinterpret_cast<DevPrim*>(mesh + S.smem_mesh_nodes);
    DevNode4* unodes = reinterpret_cast<DevNode4*>(prims + S.smem_prims);
    if (S.n_flat > 0) {
        // the flat leaf list of a small scene, once per ray octant: copy `oct` holds each box as {planes the
        // ray enters through, planes it leaves through} for directions with sign bits oct (bit 0: x < 0, bit 1:
        // y < 0, bit 2: z < 0), so that the slab test needs no min / max per axis (scene_walk_flat)
        const int n = S.n_flat;
        for (int i = threadIdx.x; i < 8 * n; i += blockDim.x) {
            const int oct = i / n;
            const float4 lo = __ldg(&S.flat[i - oct * n].lo), hi = __ldg(&S.flat[i - oct * n].hi);
            DevNode d;
            d.lo = make_float4((oct & 1) ? hi.x : lo.x, (oct & 2) ? hi.y : lo.y, (oct & 4) ? hi.z : lo.z, lo.w);
            d.hi = make_float4((oct & 1) ? lo.x : hi.x, (oct & 2) ? lo.y : hi.y, (oct & 4) ? lo.z : hi.z, hi.w);
            top[i] = d;
        }
    } else {
        // the top levels of the tree
        const float4* src = reinterpret_cast<const float4*>(S.top_nodes);
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
    if (S.u_smem_nodes > 0) {
        // the top levels of the unified tree (its top part is stored breadth-first)
        const float4* src = reinterpret_cast<const float4*>(S.u_nodes);
        float4* dst = reinterpret_cast<float4*>(unodes);
        const int n4 = S.u_smem_nodes * int(sizeof(DevNode4) / sizeof(float4));
        for (int i = threadIdx.x; i < n4; i += blockDim.x) dst[i] = __ldg(src + i);
    }
    __syncthreads();
    SmemScene sm;
    sm.top_nodes = top;
    sm.prims = prims;
    sm.mesh_nodes = mesh;
    sm.u_nodes = unodes;
    return sm;
}

// ------------------------------------------------------------------- raygen
// raytracer.vert:9-22 + main.frag + raytrace()/srand() (tp/montecarlo.frag:182-188):
// Dir_in is the barycentric interpolation, over the two triangles of the
// full-screen strip, of the per-corner normalised directions; it is normalised
// again; the path starts with attenuation 0.8 and total 0 (:105-107).
MCPT_DEV void camera_ray(const CameraRays& cam, int width, int height, uint32_t pix, float4& o4, float4& d4) {
    const int x = int(pix % uint32_t(width));
    const int y = int(pix / uint32_t(width));
    const float u = (float(x) + 0.5f) / float(width);
    const float v = (float(y) + 0.5f) / float(height);
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
    // attenuation 0.8, total 0 and the RNG seed of a fresh path are implied (k_shade first bounce,
    // k_accumulate): nothing else is initialised here. (Leaving out the origin too -- it is the
    // camera's for every primary ray -- saved 32 B per path and no time: measured, dropped.)
    o4 = make_float4(cam.ori[0], cam.ori[1], cam.ori[2], 0.0f);
    d4 = make_float4(dir.x, dir.y, dir.z, 0.0f);
}

__global__ void __launch_bounds__(kBlock) k_raygen(DevParams prm, CameraRays cam, Wavefront wf, int first_pass, int n_passes,
                                                    uint32_t npix) {
    // the camera ray of a pixel is the same in every pass (the reference's vertex shader does not see the pass): computed
    // once, stored n_passes times (one thread per path spent more time on the index arithmetic than on its store)
    for (uint32_t pix = blockIdx.x * blockDim.x + threadIdx.x; pix < npix; pix += gridDim.x * blockDim.x) {
        float4 o4, d4;
        camera_ray(cam, prm.width, prm.height, pix, o4, d4);
        for (int k = 0; k < n_passes; ++k) wf.store_ray(uint64_t(k) * npix + pix, o4, d4);
    }
}

__global__ void __launch_bounds__(kBlock) k_camera_table(DevParams prm, CameraRays cam, float4* __restrict__ dirs, uint32_t npix) {
    for (uint32_t pix = blockIdx.x * blockDim.x + threadIdx.x; pix < npix; pix += gridDim.x * blockDim.x) {
        float4 o4, d4;
        camera_ray(cam, prm.width, prm.height, pix, o4, d4);
        dirs[pix] = d4;
    }
}

void launch_camera_table(const LaunchCfg& lc, const DevParams& prm, const CameraRays& cam, float4* dirs, uint32_t npix) {
    k_camera_table<<<grid_for(lc, npix, 16), kBlock, 0, lc.stream>>>(prm, cam, dirs, npix);
    lc.count();
}

void launch_raygen(const LaunchCfg& lc, const DevParams& prm, const CameraRays& cam, Wavefront wf, int first_pass,
                   int n_passes, uint32_t npix) {
    k_raygen<<<grid_for(lc, npix, 16), kBlock, 0, lc.stream>>>(prm, cam, wf, first_pass, n_passes, npix);
    lc.count();
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

// ---- per-CTA path queues
// A queue of path slots is split into one private region per CTA of the (fixed) wavefront grid:
// CTA b reads region b of its input queue and appends to region b of its output queues, so the
// append needs no global atomic and no barrier -- a warp ballots, lane 0 bumps the CTA's counter in
// shared memory, and the lanes write at region base + offset + rank. (One same-address global atomic
// per warp serialises at the L2, one per block needs two barriers per tile; both were measured at
// +30 % on the traversal kernel. Also measured: one ordered segment per warp inside the region,
// appended without any atomic and consumed through a prefix of the eight segment lengths -- survivors
// stay in path order, but every kernel got 5..15 % slower.) A region can never overflow: what a CTA appends is a subset of what
// it read, and the first bounce hands CTA b the tiles b, b + G, b + 2G ... of the batch, at most
// `cap` slots. Region lengths live in counts[b]; every CTA writes its own before it exits.
__device__ __forceinline__ uint32_t queue_len(const QueueIn& q) {
    if (q.items) return q.counts[blockIdx.x];
    const uint32_t n_tiles = (q.n_identity + kBlock - 1) / kBlock;
    if (blockIdx.x >= n_tiles) return 0u;
    return ((n_tiles - 1u - blockIdx.x) / gridDim.x + 1u) * kBlock;  // whole tiles; the last one is cut by queue_item
}

__device__ __forceinline__ bool queue_item(const QueueIn& q, uint32_t i, uint32_t len, uint32_t& p) {
    p = 0u;
    if (i >= len) return false;
    if (q.items) {
        p = q.items[size_t(blockIdx.x) * q.cap + i];
        return true;
    }
    p = (blockIdx.x + (i / kBlock) * gridDim.x) * kBlock + (i % kBlock);
    return p < q.n_identity;
}

// the ray of path slot p: read from its ray sector, or -- first traversal of a batch with MCPT_OPT_FUSED_RAYGEN -- made here
// and stored (k_raygen's arithmetic, so the same bits)
__device__ __forceinline__ void fetch_ray(const Wavefront& wf, const QueueIn& q, uint32_t p, float4& o4, float4& d4) {
    if (q.gen.on) {
        // pixel = p mod npix with the pass index (< 64) estimated in fp32 and corrected by one either way
        const uint32_t k = __float2uint_rz(__uint2float_rz(p) * q.gen.inv_npix);
        uint32_t pix = p - k * q.gen.npix;
        if (int32_t(pix) < 0) pix += q.gen.npix;
        else if (pix >= q.gen.npix) pix -= q.gen.npix;
        o4 = make_float4(q.gen.cam.ori[0], q.gen.cam.ori[1], q.gen.cam.ori[2], 0.0f);
        d4 = __ldg(q.gen.dirs + pix);
        if (q.gen.on == 1) wf.store_ray(p, o4, d4);
    } else {
        wf.load_ray(p, o4, d4);
    }
}

__device__ __forceinline__ void warp_append(uint32_t* s_count, const QueueOut& q, bool take, uint32_t p) {
    const uint32_t m = __ballot_sync(0xffffffffu, take);
    if (!m) return;
    const uint32_t lane = threadIdx.x & 31u;
    uint32_t off = 0;
    if (lane == 0) off = atomicAdd(s_count, uint32_t(__popc(m)));
    off = __shfl_sync(0xffffffffu, off, 0);
    if (take) q.items[size_t(blockIdx.x) * q.cap + off + __popc(m & ((1u << lane) - 1u))] = p;
}

// Where the closest-hit search of one bounce delivers its result.
//  * Main traversal of the Monte Carlo mode (hit queue): the rays that hit something get their record
//    in the hit sector and are appended to the queue, so the shading kernel only ever sees hits; a ray
//    that leaves the scene needs nothing at all: its path simply stays as it is, and k_accumulate
//    evaluates the fake sky (tp/montecarlo.frag:118-119) from the path's last total / attenuation /
//    direction. No dependent load sits between two walks of a warp.
//  * Interior rays (exit_mode): the second half of the transparent cases (tp/montecarlo.frag:150-152,
//    163-165) is done right here -- take N, P of the hit (or keep the entry hit's N, P when nothing
//    was hit: intersection_info writes nothing for shape -1) and leave through
//    refract(D, -N, 1 / refract_ind). This used to be a kernel of its own (k_shade_exit: 2.1 ms per C2
//    step at 14 % issue utilisation, waiting on its loads); here the loads overlap other warps' walks.
//  * Debug shaders (no queue, no exit): every ray's record goes to the hit sector.
struct HitSink {
    QueueOut hitq;
    int exit_mode;
    int flat_face;
    float refract_ind;
};

template <bool EXIT>
__device__ __forceinline__ void deliver(const DevScene& S, const Wavefront& wf, const HitSink& sink, uint32_t* s_count, uint32_t p,
                                        bool valid, const Best& best, f3 D) {
    if (EXIT) {
        if (valid) {
            const float4 d4 = wf.load_ray_d(p);
            float4 n4, p4;
            wf.load_keep(p, n4, p4);
            f3 N = mk3(n4.x, n4.y, n4.z), P = mk3(p4.x, p4.y, p4.z);
            intersection_info(S, best.h.prim, best.h.sub, best.h.pl, sink.flat_face, N, P);
            const f3 nO = P + N * MCPT_BIAS;
            const f3 nD = refract3(mk3(d4.x, d4.y, d4.z), -N, 1.0f / sink.refract_ind);
            wf.store_ray(p, make_float4(nO.x, nO.y, nO.z, 0.0f), make_float4(nD.x, nD.y, nD.z, d4.w));  // w: the path's RNG draw count
        }
        return;
    }
    const bool hit = valid && best.h.prim >= 0;
    if (hit || (valid && !sink.hitq.items)) {
        wf.store_hit(p, best.h.pl, best.h.prim, best.h.sub, D);
    }
    if (sink.hitq.items) warp_append(s_count, sink.hitq, hit, p);
}

// Persistent warps over the CTA's queue region: a lane whose ray is finished delivers it and, once
// S.refill_lanes lanes of the warp are idle, the idle lanes take the next entries of the region (one
// shared-memory atomic per refill). Measured: mesh scenes are fastest refilling at 8 idle lanes (1.2x
// over waiting for the whole warp); trees of analytic primitives at 32, i.e. a new tile only when the
// whole warp is done (+11 % over 8: their rays are short, and refilled lanes drift out of phase). Incoherent rays -- a few lanes deep inside a mesh while the
// others have left the scene -- no longer hold a whole warp: measured 4.4 of 32 lanes active in
// late bounces of the mesh workload with one ray per lane per tile. Every iteration each lane does
// the one step its ray needs (traverse.cuh); node steps of both levels share one code path.
// (Measured and dropped: running only the most wanted kind of step per iteration -- fewer
// instructions but fewer loads in flight, -14 %; holding the primitive step back until several lanes
// want it -- no gain; one queue region per warp instead of per CTA -- no atomics at all and ordered
// survivors, but warps of a CTA finish at different times, -4 % at C2.)

#ifndef MCPT_POPS_PER_ITER
#define MCPT_POPS_PER_ITER 2  // measured on the mesh workload: 1 -> 704, 2 -> 752, 3 -> 744, unbounded -> 731 Msamples/s
#endif
// MCPT_FLOW (default 1): one iteration of a persistent warp is a pipeline of the step kinds in the order a ray
// usually needs them -- take an entry off the stack, enter a primitive, a run of node steps, a triangle -- and a
// lane FLOWS through it: the kind of step it needs is re-read before every stage, so "pop, node, node, triangle"
// of one ray costs one iteration, not three. Every stage is executed once per iteration anyway (some lane of the
// warp needs it); chaining halves the iterations per ray and fills each stage with more lanes. Measured on the
// 1 M-triangle workload: 1106 -> 1261 Msamples/s (traverse 22.6 -> 19.7 ms, interior 5.2 -> 4.6 ms per step).
// 0: one kind of step per lane per iteration (round 1).
#ifndef MCPT_FLOW
#define MCPT_FLOW 1
#endif
// MCPT_BATCH_DELIVER (with MCPT_FLOW): a finished ray keeps its lane until the refill threshold is reached and
// all finished records are delivered together right before the refill.
#ifndef MCPT_BATCH_DELIVER
#define MCPT_BATCH_DELIVER 0
#endif
// MCPT_FLOW == 2 ("while-while"): the node steps of an iteration are a loop of their own that a warp leaves only
// when every lane stands at a leaf or is finished, so the leaf stages (primitive / mesh entry, triangle, deliver)
// run once per iteration with many lanes instead of every iteration with a few.
#ifndef MCPT_NODE_MIN_LANES
#define MCPT_NODE_MIN_LANES 0
#endif
#ifndef MCPT_WW_NODE_STEPS
#define MCPT_WW_NODE_STEPS 1
#endif

template <bool COUNT, bool MESHES, bool EXIT>
__device__ __forceinline__ void traverse_persistent(const DevScene& S, const SmemScene& sm, const Wavefront& wf, const QueueIn& qin,
                                                    uint32_t len, const HitSink& sink, uint32_t* s_next, uint32_t* s_hits,
                                                    Work& wk) {
    constexpr bool WIDE = MESHES && (MCPT_WIDE != 0);
    MCPT_WALK_STACK_DECL(st, (MESHES ? (WIDE ? kWalkStackWide : kWalkStack) : 32));
    float wsave[6];
    WalkState w;
    w.done = true;
    w.mesh_prim = -1;
    w.cur = 0;
    const uint32_t lane = threadIdx.x & 31u, below = (1u << lane) - 1u;
    bool has_ray = false, exhausted = false;
    uint32_t p = 0;
#ifndef MCPT_NODE_STEPS
#define MCPT_NODE_STEPS (WIDE ? 2 : (MESHES ? 4 : 6))  // (four-children nodes: 1 -> 930, 2 -> 1003, 3 -> 996 Msamples/s; SAH-built trees of analytic primitives, divergence-stress scene: 2 -> 3222, 4 -> 3398, 6 -> 3504 Msamples/s.) Two-children node steps per iteration of a persistent warp; measured on the mesh workload: 1 -> 762, 2 -> 901, 3 -> 934, 4 -> 946, 6 -> 930, 8 -> 900 Msamples/s
#endif
    while (true) {
#if MCPT_FLOW && MCPT_BATCH_DELIVER
        const uint32_t fin = __ballot_sync(0xffffffffu, has_ray && w.done);
        const uint32_t freeable = fin | __ballot_sync(0xffffffffu, !has_ray);
        if (freeable) {
            const bool all = freeable == 0xffffffffu;
            if (all || __popc(freeable) >= S.refill_lanes) {
                if (fin) {
                    const bool mine = has_ray && w.done;
                    deliver<EXIT>(S, wf, sink, s_hits, p, mine, w.best, w.D);
                    if (mine) has_ray = false;
                }
                if (!exhausted) {
                    uint32_t first = 0;
                    if (lane == 0) first = atomicAdd(s_next, uint32_t(__popc(freeable)));
                    first = __shfl_sync(0xffffffffu, first, 0);
                    if (!has_ray && queue_item(qin, first + __popc(freeable & below), len, p)) {
                        float4 o4, d4;
                        fetch_ray(wf, qin, p, o4, d4);
                        walk_begin<WIDE>(w, S, mk3(o4.x, o4.y, o4.z), mk3(d4.x, d4.y, d4.z));
                        has_ray = true;
                    }
                    exhausted = first + __popc(freeable) >= len;
                } else if (all) {
                    break;
                }
            }
        }
#else
        const uint32_t idle = __ballot_sync(0xffffffffu, !has_ray);
        if (idle) {
            if (!exhausted && (idle == 0xffffffffu || __popc(idle) >= S.refill_lanes)) {
                uint32_t first = 0;
                if (lane == 0) first = atomicAdd(s_next, uint32_t(__popc(idle)));
                first = __shfl_sync(0xffffffffu, first, 0);
                if (!has_ray && queue_item(qin, first + __popc(idle & below), len, p)) {
                    float4 o4, d4;
                    fetch_ray(wf, qin, p, o4, d4);
                    walk_begin<WIDE>(w, S, mk3(o4.x, o4.y, o4.z), mk3(d4.x, d4.y, d4.z));
                    has_ray = true;
                }
                exhausted = first + __popc(idle) >= len;
            } else if (idle == 0xffffffffu) {
                break;
            }
        }
#endif
#if MCPT_FLOW == 2
        // node phase: every lane pops / steps through nodes until its ray stands at a leaf (a triangle, a primitive) or is
        // finished; the leaf stages below then run ONCE, with all those lanes in them
#pragma unroll 1
        while (true) {
            const bool want = has_ray && !w.done && (w.cur >= 0 || w.cur == kTokPop);
            const uint32_t wm = __ballot_sync(0xffffffffu, want);
            if (!wm) break;
#if MCPT_NODE_MIN_LANES > 0
            // few lanes left in the tree while the others wait at their leaves: let the leaves go first
            if (__popc(wm) < MCPT_NODE_MIN_LANES && __any_sync(0xffffffffu, has_ray && !want)) break;
#endif
            if (want && w.cur == kTokPop) walk_pop_one<WIDE>(w, S, st, wsave);
            if (has_ray && !w.done && w.cur >= 0) {
                walk_node<COUNT, WIDE>(w, wk, S, st, wsave);
#pragma unroll
                for (int r = 1; r < MCPT_WW_NODE_STEPS; ++r)
                    if (w.cur >= 0) walk_node<COUNT, WIDE>(w, wk, S, st, wsave);
            }
        }
        if (has_ray && walk_kind(w) == WALK_PRIM) walk_prim<COUNT, MESHES, WIDE>(w, wk, S, sm, st, wsave);
        if (MESHES && has_ray && walk_kind(w) == WALK_TRI) walk_tri<COUNT>(w, wk, S, sm, st, wsave);
        {  // finished rays: deliver and free the lane
            const bool fin = has_ray && w.done;
            if (__any_sync(0xffffffffu, fin)) {
                deliver<EXIT>(S, wf, sink, s_hits, p, fin, w.best, w.D);
                if (fin) has_ray = false;
            }
        }
        continue;
#endif
#pragma unroll 1
        for (int r = 0; r < MCPT_POPS_PER_ITER && has_ray && w.cur == kTokPop && !w.done; ++r)
            walk_pop_one<WIDE>(w, S, st, wsave);
#if MCPT_FLOW
        if (has_ray && walk_kind(w) == WALK_PRIM) walk_prim<COUNT, MESHES, WIDE>(w, wk, S, sm, st, wsave);
        if (has_ray && walk_kind(w) == WALK_NODE) {
            walk_node<COUNT, WIDE>(w, wk, S, st, wsave);
#pragma unroll
            for (int r = 1; r < MCPT_NODE_STEPS; ++r)
                if (w.cur >= 0) walk_node<COUNT, WIDE>(w, wk, S, st, wsave);
        }
        if (MESHES && has_ray && walk_kind(w) == WALK_TRI) walk_tri<COUNT>(w, wk, S, sm, st, wsave);
#if !MCPT_BATCH_DELIVER
        {  // finished rays: deliver and free the lane
            const bool fin = has_ray && w.done;
            if (__any_sync(0xffffffffu, fin)) {
                deliver<EXIT>(S, wf, sink, s_hits, p, fin, w.best, w.D);
                if (fin) has_ray = false;
            }
        }
#endif
#else
        const int kind = has_ray ? walk_kind(w) : -1;  // WALK_POP: a dead entry was dropped, pop again next time
        // finished rays: deliver and free the lane
        if (__any_sync(0xffffffffu, kind == WALK_DONE)) {
            deliver<EXIT>(S, wf, sink, s_hits, p, kind == WALK_DONE, w.best, w.D);
            if (kind == WALK_DONE) has_ray = false;
        }
        if (kind == WALK_NODE) {
            // a run of node steps: the lanes at a triangle or a primitive wait, but the dispatch, the
            // ballots and the rarer step kinds are paid once per run (popping inside the run as well
            // was measured: -7 %)
            walk_node<COUNT, WIDE>(w, wk, S, st, wsave);
#pragma unroll
            for (int r = 1; r < MCPT_NODE_STEPS; ++r)
                if (w.cur >= 0) walk_node<COUNT, WIDE>(w, wk, S, st, wsave);
        } else if (MESHES && kind == WALK_TRI)
            walk_tri<COUNT>(w, wk, S, sm, st, wsave);
        else if (kind == WALK_PRIM)
            walk_prim<COUNT, MESHES, WIDE>(w, wk, S, sm, st, wsave);
#endif
    }
}

// Persistent warps over the unified world-space tree (traverse.cuh, UWalk): the same refill / deliver protocol as
// traverse_persistent, with two kinds of step only. One iteration: take entries off the stack, a run of node steps,
// the leaf step, deliver.
#ifndef MCPT_U_POPS
#define MCPT_U_POPS 2
#endif
#ifndef MCPT_U_NODE_STEPS
#define MCPT_U_NODE_STEPS 2
#endif
#ifndef MCPT_U_LOOP
#define MCPT_U_LOOP 2  // 1: pop -> node run -> leaf in every iteration; 2: the node phase loops until fewer than MCPT_U_MIN_LANES lanes still walk nodes while others wait at a leaf
#endif
// Measured on the 1 M-triangle workload (Msamples/s): loop 1 with 1 / 2 / 3 node steps per iteration 1438 / 1548 / 1529, with 1 / 3
// pops 1547 / 1548; loop 2 with 8 / 12 / 16 / 20 lanes 1551 / 1587 / 1575 / 1537; the top stack entry cached in two registers
// (its refill load issued one pop early) 1486 (registers); 5 / 6 CTAs per SM (48 / 40 registers) 1223 / 1033 (spills).
#ifndef MCPT_U_MIN_LANES
#define MCPT_U_MIN_LANES 12
#endif
template <bool COUNT, bool EXIT>
__device__ __forceinline__ void traverse_unified(const DevScene& S, const SmemScene& sm, const Wavefront& wf, const QueueIn& qin,
                                                 uint32_t len, const HitSink& sink, uint32_t* s_next, uint32_t* s_hits, Work& wk) {
    MCPT_WALK_STACK_DECL(st, kWalkStackWide);
    UWalk w;
    w.done = true;
    w.cur = 0;
    w.sp = 0;
    const uint32_t lane = threadIdx.x & 31u, below = (1u << lane) - 1u;
    bool has_ray = false, exhausted = false;
    uint32_t p = 0;
    const int refill = qin.items ? S.refill_lanes : S.refill_first;
    while (true) {
        const uint32_t idle = __ballot_sync(0xffffffffu, !has_ray);
        if (idle) {
            if (!exhausted && (idle == 0xffffffffu || __popc(idle) >= refill)) {
                uint32_t first = 0;
                if (lane == 0) first = atomicAdd(s_next, uint32_t(__popc(idle)));
                first = __shfl_sync(0xffffffffu, first, 0);
                if (!has_ray && queue_item(qin, first + __popc(idle & below), len, p)) {
                    float4 o4, d4;
                    fetch_ray(wf, qin, p, o4, d4);
                    uwalk_begin(w, S, mk3(o4.x, o4.y, o4.z), mk3(d4.x, d4.y, d4.z));
                    has_ray = true;
                }
                exhausted = first + __popc(idle) >= len;
            } else if (idle == 0xffffffffu) {
                break;
            }
        }
#if MCPT_U_LOOP == 2
#pragma unroll 1
        while (true) {
            const bool want = has_ray && !w.done && (w.cur >= 0 || w.cur == kTokPop);
            const uint32_t wm = __ballot_sync(0xffffffffu, want);
            if (!wm) break;
            if (__popc(wm) < MCPT_U_MIN_LANES && __any_sync(0xffffffffu, has_ray && !want)) break;
            if (want && w.cur == kTokPop) uwalk_pop_one(w, st);
            if (has_ray && !w.done && w.cur >= 0) uwalk_node<COUNT>(w, wk, S, sm, st);
        }
#else
#pragma unroll 1
        for (int r = 0; r < MCPT_U_POPS && has_ray && w.cur == kTokPop && !w.done; ++r) uwalk_pop_one(w, st);
        if (has_ray && !w.done && w.cur >= 0) {
            uwalk_node<COUNT>(w, wk, S, sm, st);
#pragma unroll
            for (int r = 1; r < MCPT_U_NODE_STEPS; ++r)
                if (w.cur >= 0) uwalk_node<COUNT>(w, wk, S, sm, st);
        }
#endif
        if (has_ray && !w.done && w.cur < 0 && w.cur != kTokPop) uwalk_leaf<COUNT>(w, wk, S, sm);
        {  // finished rays: deliver and free the lane
            const bool fin = has_ray && w.done;
            if (__any_sync(0xffffffffu, fin)) {
                deliver<EXIT>(S, wf, sink, s_hits, p, fin, w.best, w.D);
                if (fin) has_ray = false;
            }
        }
    }
}

// FLAT: warp-coherent flat walk (S.n_flat > 0); MESHES: the scene has triangle meshes; PERSIST:
// persistent warps with per-lane refill (tree scenes; mesh trees must be the GPU-built ones).
// Every lane of a warp stays in the loop together (warp-uniform trip count).
// EXIT: interior rays, the kernel builds their exit rays itself (HitSink; scenes without meshes only) -- a
// template parameter so that the main traversal does not carry that code (registers, instruction cache).
template <bool COUNT, int MINB, bool FLAT, bool MESHES, bool PERSIST, bool EXIT>
__global__ void __launch_bounds__(kBlock, MINB) k_traverse(DevScene S, Wavefront wf, QueueIn qin, HitSink sink,
                                                            unsigned long long* ray_counter, unsigned long long* work) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ uint32_t s_hits, s_next;
    const uint32_t len = queue_len(qin);
    if (len == 0) {
        if (threadIdx.x == 0 && sink.hitq.items) sink.hitq.counts[blockIdx.x] = 0u;
        return;
    }
    if (threadIdx.x == 0) s_hits = s_next = 0u;
    SmemScene sm = stage_scene(S, smem_raw);  // ends with a barrier
    Work wk;
    if (PERSIST) traverse_persistent<COUNT, MESHES, EXIT>(S, sm, wf, qin, len, sink, &s_next, &s_hits, wk);
    for (uint32_t base = 0; !PERSIST && base < len; base += kBlock) {
        uint32_t p;
        const bool valid = queue_item(qin, base + threadIdx.x, len, p);
        Best best;
        float4 o4 = make_float4(0, 0, 0, 0), d4 = make_float4(0, 0, 0, 0);
        if (valid) fetch_ray(wf, qin, p, o4, d4);
        if (FLAT)
            scene_walk_flat<COUNT>(wk, S, sm, mk3(o4.x, o4.y, o4.z), mk3(d4.x, d4.y, d4.z), false, valid, best);
        else if (valid)
            scene_walk<COUNT, MESHES>(wk, S, sm, mk3(o4.x, o4.y, o4.z), mk3(d4.x, d4.y, d4.z), false, best);
        deliver<EXIT>(S, wf, sink, &s_hits, p, valid, best, mk3(d4.x, d4.y, d4.z));
    }
    if (ray_counter && threadIdx.x == 0) {
        // rays this CTA traced: its region length, or its share of the identity queue (last tile cut at n)
        unsigned long long rays = len;
        if (!qin.items) {
            const uint32_t last_tile = blockIdx.x + (len / kBlock - 1u) * gridDim.x;
            if (last_tile * kBlock + kBlock > qin.n_identity) rays -= last_tile * kBlock + kBlock - qin.n_identity;
        }
        atomicAdd(ray_counter, rays);
    }
    if (COUNT) flush_work(wk, work);
    if (sink.hitq.items) {
        __syncthreads();
        if (threadIdx.x == 0) sink.hitq.counts[blockIdx.x] = s_hits;
    }
}

// the unified walk as a kernel of its own (queue protocol and epilogue of k_traverse)
template <bool COUNT, int MINB, bool EXIT>
__global__ void __launch_bounds__(kBlock, MINB) k_traverse_unified(DevScene S, Wavefront wf, QueueIn qin, HitSink sink,
                                                                    unsigned long long* ray_counter, unsigned long long* work) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ uint32_t s_hits, s_next;
    const uint32_t len = queue_len(qin);
    if (len == 0) {
        if (threadIdx.x == 0 && sink.hitq.items) sink.hitq.counts[blockIdx.x] = 0u;
        return;
    }
    if (threadIdx.x == 0) s_hits = s_next = 0u;
    SmemScene sm = stage_scene(S, smem_raw);  // ends with a barrier
    Work wk;
    traverse_unified<COUNT, EXIT>(S, sm, wf, qin, len, sink, &s_next, &s_hits, wk);
    if (ray_counter && threadIdx.x == 0) {
        unsigned long long rays = len;
        if (!qin.items) {
            const uint32_t last_tile = blockIdx.x + (len / kBlock - 1u) * gridDim.x;
            if (last_tile * kBlock + kBlock > qin.n_identity) rays -= last_tile * kBlock + kBlock - qin.n_identity;
        }
        atomicAdd(ray_counter, rays);
    }
    if (COUNT) flush_work(wk, work);
    if (sink.hitq.items) {
        __syncthreads();
        if (threadIdx.x == 0) sink.hitq.counts[blockIdx.x] = s_hits;
    }
}

// ---- unified walk, leaf tests pooled per warp (MCPT_OPT_LEAF_POOL)
// In traverse_unified a lane that reaches a leaf waits until the warp runs its leaf stage, and that stage runs for the
// few lanes that stand at a leaf (ncu, 1 M triangles: ~30 % of the issued instructions at 4 .. 7 of 32 lanes). Here the
// lane writes a (ray, leaf) item into a ring in the warp's slice of shared memory and WALKS ON; once S.u_pool items wait
// the whole warp runs up to 32 exact tests, each lane one item of whichever ray. The rays (origin, direction, slack)
// live in shared memory, one slot per lane; a candidate is ranked by atomicMin on a 64-bit key -- distance bits, then
// offer()'s tie order (DevScene::u_pool) -- and the winner writes the payload (local hit point, leaf) behind it; the
// owner of the ray re-reads its key after every round to shrink tmax. What is given up: between the push of an item
// and its round the ray walks with the tmax it had. A ray is finished when its stack is empty AND the ring has moved
// past its last item (sequence numbers). Same tests, same arithmetic, same order: same records.
// Measured (1 M triangles at 16 spp per step / repo scene 8, Msamples/s; one test per lane: 1584 / 688): trigger 4 / 8 / 16 / 24 / 32
// items 1356 / 1451 / 1518 / 1536 / 1531 and 561 / 612 / 634 / 651 / 651; with 2 / 3 / 4 pop passes in front of every node step
// (trigger 24) 1455 / 1422 / 1406 and 599 / 591 / 589. The rounds do run at 21 lanes instead of 4 .. 7, but the exact tests are
// only ~13 % of the issued instructions (1.4 leaf tests against 7.3 node steps per ray), the node steps stay at 16 .. 18 lanes
// (a lane without a node is mostly one whose stack entry was culled or whose warp waits for the refill, not one waiting at a
// leaf), the ring costs a ballot and a branch per iteration, and a ray walks 4 % more boxes with the tmax it had when it queued
// its item. Off by default (MCPT_OPT_LEAF_POOL = 0); kept because the records are the same and the parity tests cover it.
#ifndef MCPT_UP_POP_PASSES
#define MCPT_UP_POP_PASSES 1
#endif
constexpr int kUPoolRing = 64;
constexpr size_t kUPoolBytesPerWarp = 32 * (16 + 16 + 16 + 8) + kUPoolRing * 4;
struct UPool {
    float4* rayO;  // xyz origin, w slack
    float4* rayD;
    float4* pay;   // xyz local hit point, w = leaf | face << 26
    unsigned long long* keys;
    uint32_t* items;  // ray slot << 27 | leaf
};

template <bool COUNT>
__device__ __forceinline__ void upool_round(const DevScene& S, const SmemScene& sm, const UPool& q, uint32_t head, uint32_t n_items,
                                            uint32_t lane, Work& wk) {
    const unsigned long long key_none = (static_cast<unsigned long long>(__float_as_uint(MCPT_FLT_MAX)) << 32) | 0xffffffffull;
    const bool has = lane < n_items;
    const uint32_t it = has ? q.items[(head + lane) & uint32_t(kUPoolRing - 1)] : 0u;
    const uint32_t r = it >> 27;
    const int L = int(it & 0x7ffffffu);
    unsigned long long k1 = key_none, k2 = key_none;
    f3 Pl1 = mk3(0, 0, 0), Pl2 = mk3(0, 0, 0);
    uint32_t pw = uint32_t(L);
    if (has) {
        const float4 o4 = q.rayO[r], d4 = q.rayD[r];
        const f3 O = mk3(o4.x, o4.y, o4.z), D = mk3(d4.x, d4.y, d4.z);
        if (L < S.u_n_tris) {
            const DevTri* tp = S.u_tris + L;
            const float4 ta = ldg_tri(&tp->a), tb = ldg_tri(&tp->b), tc = ldg_tri(&tp->c);
            if (__float_as_int(ta.w) >= 0) {
                if (COUNT) ++wk.tris;
                const int pid = __float_as_int(tc.w);
                const DevPrim* pp = (pid < S.smem_prims) ? (sm.prims + pid) : (S.prims + pid);
                Aff inv;
                inv.r0 = pp->inv.r0;
                inv.r1 = pp->inv.r1;
                inv.r2 = pp->inv.r2;
                const f3 Oi = xform_point(inv, O);
                const f3 Di = normalize3(xform_vector(inv, D));
                float a;
                if (triangle_param(mk3(ta.x, ta.y, ta.z), mk3(tb.x, tb.y, tb.z), mk3(tc.x, tc.y, tc.z), Oi, Di, a)) {
                    Aff fwd;
                    fwd.r0 = pp->fwd.r0;
                    fwd.r1 = pp->fwd.r1;
                    fwd.r2 = pp->fwd.r2;
                    Pl1 = Oi + Di * a;  // offer_at (intersect.cuh)
                    const float dist = distance3(O, xform_point(fwd, Pl1));
                    const uint32_t low = (uint32_t(S.u_slot_max - pp->slot) << S.u_key_shift) | (uint32_t(S.u_rank_max - (__float_as_int(tb.w) + 1)) << 1);
                    if (dist < MCPT_FLT_MAX) k1 = (static_cast<unsigned long long>(__float_as_uint(dist)) << 32) | low;
                }
            }
        } else {
            const int prim = L - S.u_n_tris;
            const DevPrim* pp = (prim < S.smem_prims) ? (sm.prims + prim) : (S.prims + prim);
            const int type = pp->type;
            if (type > MCPT_PRIM_MESH) {
                Aff inv, fwd;
                inv.r0 = pp->inv.r0;
                inv.r1 = pp->inv.r1;
                inv.r2 = pp->inv.r2;
                fwd.r0 = pp->fwd.r0;
                fwd.r1 = pp->fwd.r1;
                fwd.r2 = pp->fwd.r2;
                if (COUNT) ++wk.prims;
                const f3 Oi = xform_point(inv, O);
                const f3 Di = normalize3(xform_vector(inv, D));
                Cand k;
                k.a1 = k.a2 = 0.0f;
                k.cl = 0;
                k.c1 = k.c2 = false;
                switch (type) {
                    case MCPT_PRIM_SPHERE: sphere_test(k, Oi, Di, o4.w * pp->inv_smin); break;
                    case MCPT_PRIM_CUBE: cube_test(k, Oi, Di); break;
                    case MCPT_PRIM_CYLINDER: cylinder_test(k, Oi, Di); break;
                    case MCPT_PRIM_CONE: cone_test(k, Oi, Di, O, S.leaf_bb + 6 * prim); break;
                    case MCPT_PRIM_ORIENTED_QUAD: quad_test(k, Oi, Di, O, S.leaf_bb + 6 * prim); break;
                    default: break;
                }
                const uint32_t low = (uint32_t(S.u_slot_max - pp->slot) << S.u_key_shift) | (uint32_t(S.u_rank_max - 1) << 1);  // rank 0
                pw = uint32_t(L) | (uint32_t(k.cl) << 26);
                if (k.c1) {
                    Pl1 = Oi + Di * k.a1;
                    const float dist = distance3(O, xform_point(fwd, Pl1));
                    if (dist < MCPT_FLT_MAX) k1 = (static_cast<unsigned long long>(__float_as_uint(dist)) << 32) | low;
                }
                if (k.c2) {
                    Pl2 = Oi + Di * k.a2;
                    const float dist = distance3(O, xform_point(fwd, Pl2));
                    if (dist < MCPT_FLT_MAX) k2 = (static_cast<unsigned long long>(__float_as_uint(dist)) << 32) | low | 1u;
                }
            }
        }
        if (k1 != key_none) atomicMin(&q.keys[r], k1);
        if (k2 != key_none) atomicMin(&q.keys[r], k2);
    }
    __syncwarp();
    if (has) {
        const unsigned long long kb = q.keys[r];
        if (k1 != key_none && kb == k1) q.pay[r] = make_float4(Pl1.x, Pl1.y, Pl1.z, __uint_as_float(pw));
        if (k2 != key_none && kb == k2) q.pay[r] = make_float4(Pl2.x, Pl2.y, Pl2.z, __uint_as_float(pw));
    }
    __syncwarp();
}

template <bool COUNT, bool EXIT>
__device__ __forceinline__ void traverse_unified_pooled(const DevScene& S, const SmemScene& sm, const Wavefront& wf, const QueueIn& qin,
                                                        uint32_t len, const HitSink& sink, uint32_t* s_next, uint32_t* s_hits,
                                                        const UPool& q, Work& wk) {
    MCPT_WALK_STACK_DECL(st, kWalkStackWide);
    const unsigned long long key_none = (static_cast<unsigned long long>(__float_as_uint(MCPT_FLT_MAX)) << 32) | 0xffffffffull;
    UWalk w;
    w.done = true;
    w.cur = 0;
    w.sp = 0;
    const uint32_t lane = threadIdx.x & 31u, below = (1u << lane) - 1u;
    bool has_ray = false, exhausted = false;
    uint32_t p = 0;
    uint32_t head = 0, tail = 0, my_end = 0;  // ring sequence numbers: items [head, tail) wait; my_end = one past this ray's last item
    const int refill = qin.items ? S.refill_lanes : S.refill_first;
    const uint32_t trigger = uint32_t(S.u_pool);
    constexpr int passes = MCPT_UP_POP_PASSES;
    while (true) {
        const uint32_t idle = __ballot_sync(0xffffffffu, !has_ray);
        if (idle) {
            if (!exhausted && (idle == 0xffffffffu || __popc(idle) >= refill)) {
                uint32_t first = 0;
                if (lane == 0) first = atomicAdd(s_next, uint32_t(__popc(idle)));
                first = __shfl_sync(0xffffffffu, first, 0);
                if (!has_ray && queue_item(qin, first + __popc(idle & below), len, p)) {
                    float4 o4, d4;
                    fetch_ray(wf, qin, p, o4, d4);
                    uwalk_begin(w, S, mk3(o4.x, o4.y, o4.z), mk3(d4.x, d4.y, d4.z));
                    q.rayO[lane] = make_float4(o4.x, o4.y, o4.z, w.slack);
                    q.rayD[lane] = make_float4(d4.x, d4.y, d4.z, 0.0f);
                    q.keys[lane] = key_none;
                    my_end = head;
                    has_ray = true;
                }
                exhausted = first + __popc(idle) >= len;
                __syncwarp();
            } else if (idle == 0xffffffffu) {
                break;
            }
        }
#pragma unroll 1
        while (true) {
            const bool want = has_ray && !w.done;
            const uint32_t wm = __ballot_sync(0xffffffffu, want);
            if (!wm) break;
            if (__popc(wm) < MCPT_U_MIN_LANES && __any_sync(0xffffffffu, has_ray && !want)) break;
            // pop passes: a lane without a node takes entries off its stack until one is a node (or the stack is empty); a
            // leaf -- from the stack or the nearest child of the last node step -- is queued at once, so that as many lanes
            // as possible enter the node step with a node
#pragma unroll 1
            for (int pass = 0; pass < passes; ++pass) {
                const bool leaf = has_ray && !w.done && w.cur < 0 && w.cur != kTokPop;
                const uint32_t lm = __ballot_sync(0xffffffffu, leaf);
                if (lm) {
                    if (leaf) {
                        const uint32_t pos = tail + uint32_t(__popc(lm & below));
                        q.items[pos & uint32_t(kUPoolRing - 1)] = (lane << 27) | uint32_t(~w.cur);
                        my_end = pos + 1u;
                        w.cur = kTokPop;
                    }
                    tail += uint32_t(__popc(lm));
                    if (tail - head >= trigger) {
                        __syncwarp();
                        do {
                            const uint32_t n_items = min(tail - head, 32u);
                            upool_round<COUNT>(S, sm, q, head, n_items, lane, wk);
                            head += n_items;
                        } while (tail - head >= trigger);
                        if (has_ray) {
                            const unsigned long long kb = q.keys[lane];
                            if (kb != key_none) w.tmax = __uint_as_float(uint32_t(kb >> 32)) * 1.0001f + w.slack;
                        }
                    }
                }
                const bool popping = has_ray && !w.done && w.cur == kTokPop;
                if (!__any_sync(0xffffffffu, popping)) break;
                if (popping) uwalk_pop_one(w, st);
            }
            if (has_ray && !w.done && w.cur >= 0) uwalk_node<COUNT>(w, wk, S, sm, st);
        }
        // rays whose walk is over but whose last items still wait: run what is queued
        if (tail != head && __any_sync(0xffffffffu, has_ray && w.done && int32_t(head - my_end) < 0)) {
            __syncwarp();
            while (tail != head) {
                const uint32_t n_items = min(tail - head, 32u);
                upool_round<COUNT>(S, sm, q, head, n_items, lane, wk);
                head += n_items;
            }
            if (has_ray) {
                const unsigned long long kb = q.keys[lane];
                if (kb != key_none) w.tmax = __uint_as_float(uint32_t(kb >> 32)) * 1.0001f + w.slack;
            }
        }
        {  // finished rays: deliver and free the lane
            const bool fin = has_ray && w.done && int32_t(head - my_end) >= 0;
            if (__any_sync(0xffffffffu, fin)) {
                Best best;
                best_reset(best);
                f3 D = mk3(0, 0, 0);
                if (fin) {
                    const unsigned long long kb = q.keys[lane];
                    const float4 d4 = q.rayD[lane];
                    D = mk3(d4.x, d4.y, d4.z);
                    if (kb != key_none) {
                        const float4 pv = q.pay[lane];
                        const uint32_t pw = __float_as_uint(pv.w);
                        const int L = int(pw & 0x3ffffffu);
                        best.h.dist = __uint_as_float(uint32_t(kb >> 32));
                        best.h.pl = mk3(pv.x, pv.y, pv.z);
                        if (L < S.u_n_tris) {
                            const DevTri* tp = S.u_tris + L;
                            best.h.prim = __float_as_int(__ldg(&tp->c).w);
                            best.h.sub = __float_as_int(__ldg(&tp->a).w);
                        } else {
                            best.h.prim = L - S.u_n_tris;
                            best.h.sub = int(pw >> 26);
                        }
                    }
                }
                deliver<EXIT>(S, wf, sink, s_hits, p, fin, best, D);
                if (fin) has_ray = false;
            }
        }
    }
}

template <bool COUNT, int MINB, bool EXIT>
__global__ void __launch_bounds__(kBlock, MINB) k_traverse_unified_pooled(DevScene S, Wavefront wf, QueueIn qin, HitSink sink,
                                                                           unsigned long long* ray_counter, unsigned long long* work) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ uint32_t s_hits, s_next;
    const uint32_t len = queue_len(qin);
    if (len == 0) {
        if (threadIdx.x == 0 && sink.hitq.items) sink.hitq.counts[blockIdx.x] = 0u;
        return;
    }
    if (threadIdx.x == 0) s_hits = s_next = 0u;
    SmemScene sm = stage_scene(S, smem_raw);  // ends with a barrier
    unsigned char* wbase = smem_raw + ((size_t(S.smem_bytes) + 15) & ~size_t(15)) + size_t(threadIdx.x >> 5) * kUPoolBytesPerWarp;
    UPool q;
    q.rayO = reinterpret_cast<float4*>(wbase);
    q.rayD = q.rayO + 32;
    q.pay = q.rayD + 32;
    q.keys = reinterpret_cast<unsigned long long*>(q.pay + 32);
    q.items = reinterpret_cast<uint32_t*>(q.keys + 32);
    Work wk;
    traverse_unified_pooled<COUNT, EXIT>(S, sm, wf, qin, len, sink, &s_next, &s_hits, q, wk);
    if (ray_counter && threadIdx.x == 0) {
        unsigned long long rays = len;
        if (!qin.items) {
            const uint32_t last_tile = blockIdx.x + (len / kBlock - 1u) * gridDim.x;
            if (last_tile * kBlock + kBlock > qin.n_identity) rays -= last_tile * kBlock + kBlock - qin.n_identity;
        }
        atomicAdd(ray_counter, rays);
    }
    if (COUNT) flush_work(wk, work);
    if (sink.hitq.items) {
        __syncthreads();
        if (threadIdx.x == 0) sink.hitq.counts[blockIdx.x] = s_hits;
    }
}

// ---- flat scenes after the first bounce: exact tests pooled per warp and per primitive TYPE
// The warp-coherent flat walk (scene_walk_flat) runs the exact test of primitive k whenever ANY of the warp's 32 rays
// passes box k. Camera rays agree on their boxes (31 of 32 lanes active); the rays of later bounces do not: every
// test body then runs for 6 .. 12 lanes (ncu, config 2: 16.9 active lanes in the second bounce, 10.8 in the fifth,
// with the exact tests -- 60 % of the issued instructions -- at ~10). Here a warp takes kPoolR rays per lane, culls all
// of them against the box list first (no test in that loop), turns every (ray, passing box) pair into a work item,
// and runs the items of one primitive TYPE 32 at a time: one instruction stream (no type switch inside a round),
// lanes filled with items of different primitives of that type and of different rays. Rays live in the warp's slice
// of shared memory; a candidate is ranked by a 64-bit key -- distance bits, then the reference's visit order -- with
// atomicMin, and the winner of a round writes the payload (local hit point, primitive, face) behind it. No CTA
// barrier is involved (the regrouping of a 256-ray tile by primitive through shared bins needed three per tile and
// lost, profiles/README.md). What is given up is the shrinking tmax of the sequential walk: every passing box is
// tested. The total order of offer() (intersect.cuh) -- smaller distance, then higher leaf slot, then the candidate
// offered first -- is the order of the keys, so the records are the same bits.
// (Measured and dropped, config 2 / scene 2 / scene 4 Msamples/s against 6314 / 2266 / 3881: the candidates of a round ranked without
// atomics -- match.any on the ray slot, two REDUX.MIN over distance and tie bits, one plain compare with the stored key -- 5061 /
// 1869 / 3116: MATCH.ANY costs far more than the compare-and-swap loop behind the 64-bit atomicMin; the exclusive prefix of the item
// counts as six independent ballots instead of five dependent shuffles 6278 / 2284 / 3894: noise.)
#ifndef MCPT_POOL_R
#define MCPT_POOL_R 2
#endif
constexpr int kPoolR = MCPT_POOL_R;
constexpr int kPoolSlots = 32 * kPoolR;
// per warp: ray origins (+ slack) and directions (two buffers of them when the next tile's rays are copied in asynchronously),
// payloads, keys, items of one primitive type (at most max_type list entries hold primitives of one type)
__host__ __device__ inline size_t pool_bytes_per_warp(int max_type, bool async) {
    const size_t b = size_t(kPoolSlots) * ((async ? 2 : 1) * (16 + 16) + 16 + 8) + size_t(kPoolSlots) * size_t(max_type) * 2;
    return (b + 15) & ~size_t(15);
}

#ifndef MCPT_CP_ASYNC_CA
#define MCPT_CP_ASYNC_CA 0  // 1: through L1 (.ca). Measured: config 2 6172 against 6298 Msamples/s with .cg (direct loads: 6151) -- the rays are read once
#endif
// 16 bytes global -> shared without a register in between (LDGSTS); completion: cp_async_wait_all of the issuing thread
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
    const uint32_t d = static_cast<uint32_t>(__cvta_generic_to_shared(smem_dst));
#if MCPT_CP_ASYNC_CA
    asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gsrc) : "memory");
#else
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gsrc) : "memory");
#endif
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

struct PoolCand {
    float a1, a2;
    int cl;
    bool c1, c2;
};

// ASYNC: the rays of tile t + 1 are copied into a second pair of ray buffers with cp.async while tile t is tested (their queue
// entries are loaded at the top of tile t, the copies are issued after its cull), so that no warp waits for a ray load
// between two tiles (ncu before: 29 % of the stall samples of this kernel sat on those loads, 68 % of the issue slots busy).
template <bool COUNT, int MINB, bool EXIT, bool ASYNC>
__global__ void __launch_bounds__(kBlock, MINB) k_traverse_flat_pooled(DevScene S, Wavefront wf, QueueIn qin, HitSink sink,
                                                                        unsigned long long* ray_counter, unsigned long long* work) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ uint32_t s_hits;
    __shared__ uint32_t s_tm[8];   // per primitive type: the list positions that hold a primitive of that type
    const uint32_t len = queue_len(qin);
    if (len == 0) {
        if (threadIdx.x == 0 && sink.hitq.items) sink.hitq.counts[blockIdx.x] = 0u;
        return;
    }
    if (threadIdx.x == 0) s_hits = 0u;
    if (threadIdx.x < 8) s_tm[threadIdx.x] = 0u;
    SmemScene sm = stage_scene(S, smem_raw);  // ends with a barrier
    const int n = S.n_flat;
    if (int(threadIdx.x) < n) {
        const int prim = __float_as_int(sm.top_nodes[threadIdx.x].lo.w);
        const int type = sm.prims[prim].type;
        if (type >= 0 && type < 8) atomicOr(&s_tm[type], 1u << threadIdx.x);
    }
    __syncthreads();
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    unsigned char* wbase = smem_raw + ((size_t(S.smem_bytes) + 15) & ~size_t(15)) + size_t(warp) * pool_bytes_per_warp(S.flat_max_type, ASYNC);
    float4* ray_buf = reinterpret_cast<float4*>(wbase);   // [buffer][O | D][slot]
    float4* pay = ray_buf + (ASYNC ? 4 : 2) * kPoolSlots;
    unsigned long long* keys = reinterpret_cast<unsigned long long*>(pay + kPoolSlots);
    unsigned short* items = reinterpret_cast<unsigned short*>(keys + kPoolSlots);
    const unsigned long long key_none = (static_cast<unsigned long long>(__float_as_uint(MCPT_FLT_MAX)) << 32) | 0xffffffffull;
    Work wk;
    const uint32_t tile_rays = kBlock * kPoolR;
    uint32_t pidn[kPoolR];
    bool validn[kPoolR];
    if (ASYNC) {  // the first tile's rays
#pragma unroll
        for (int j = 0; j < kPoolR; ++j) {
            const uint32_t slot = uint32_t(j) * 32u + lane;
            validn[j] = queue_item(qin, warp * kPoolSlots + slot, len, pidn[j]);
            if (validn[j]) {
                const PathSector* src = wf.sec(pidn[j], 0);
                cp_async16(ray_buf + slot, &src->a);
                cp_async16(ray_buf + kPoolSlots + slot, &src->b);
            }
        }
        cp_async_commit();
    }
    for (uint32_t tile = 0; tile * tile_rays < len; ++tile) {
        const uint32_t wfirst = tile * tile_rays + warp * kPoolSlots;
        float4* rayO = ray_buf + (ASYNC ? (tile & 1u) * 2u * kPoolSlots : 0u);
        float4* rayD = rayO + kPoolSlots;
        // ---- phase 1: kPoolR rays per lane, culled against the whole list
        uint32_t pid[kPoolR], mask[kPoolR];
        bool valid[kPoolR];
        if (ASYNC) {
#pragma unroll
            for (int j = 0; j < kPoolR; ++j) {
                pid[j] = pidn[j];
                valid[j] = validn[j];
                // queue entries of the next tile: in flight during the cull
                validn[j] = queue_item(qin, wfirst + tile_rays + uint32_t(j) * 32u + lane, len, pidn[j]);
            }
            cp_async_wait_all();
        }
#pragma unroll
        for (int j = 0; j < kPoolR; ++j) {
            const uint32_t slot = uint32_t(j) * 32u + lane;
            float4 o4 = make_float4(0, 0, 0, 0), d4 = make_float4(0, 0, 0, 0);
            if (ASYNC) {
                if (valid[j]) o4 = rayO[slot], d4 = rayD[slot];
            } else {
                valid[j] = queue_item(qin, wfirst + slot, len, pid[j]);
                if (valid[j]) wf.load_ray(pid[j], o4, d4);
            }
            const f3 O = mk3(o4.x, o4.y, o4.z), D = mk3(d4.x, d4.y, d4.z);
            const float dl = sqrtf(D.x * D.x + D.y * D.y + D.z * D.z);
            const bool ok = valid[j] && (dl > 0.0f) && (dl < FLT_MAX);
            const float rl = (dl < 1e-30f) ? 1.0f / dl : cull_rcp(dl);
            const f3 idir = safe_rcp3(mk3(D.x * rl, D.y * rl, D.z * rl));
            const f3 od = mk3(-O.x * idir.x, -O.y * idir.y, -O.z * idir.z);
            const float slack = 1e-4f * (fabsf(O.x) + fabsf(O.y) + fabsf(O.z)) + 1e-3f;
            const uint32_t oct = (__float_as_uint(idir.x) >> 31) | ((__float_as_uint(idir.y) >> 31) << 1) | ((__float_as_uint(idir.z) >> 31) << 2);
            const DevNode* boxes = sm.top_nodes + oct * uint32_t(n);
            uint32_t m = 0u;
            if (ok)
                for (int i = 0; i < n; ++i)
                    if (slab_sorted(boxes[i], idir, od, FLT_MAX)) m |= 1u << i;
            if (COUNT && ok) wk.boxes += n;
            mask[j] = m;
            rayO[slot] = make_float4(O.x, O.y, O.z, slack);
            rayD[slot] = make_float4(D.x, D.y, D.z, 0.0f);
            keys[slot] = key_none;
        }
        if (ASYNC) {  // the next tile's rays into the other pair of buffers (last read in the tile before this one)
            float4* nO = ray_buf + ((tile & 1u) ^ 1u) * 2u * kPoolSlots;
#pragma unroll
            for (int j = 0; j < kPoolR; ++j) {
                const uint32_t slot = uint32_t(j) * 32u + lane;
                if (validn[j]) {
                    const PathSector* src = wf.sec(pidn[j], 0);
                    cp_async16(nO + slot, &src->a);
                    cp_async16(nO + kPoolSlots + slot, &src->b);
                }
            }
            cp_async_commit();
        }
        __syncwarp();
        // ---- phase 2: the items of one primitive type at a time, 32 per round
#pragma unroll 1
        for (int type = MCPT_PRIM_SPHERE; type <= MCPT_PRIM_ORIENTED_QUAD; ++type) {
            const uint32_t tm = s_tm[type];
            if (!tm) continue;
            uint32_t mine = 0u;
#pragma unroll
            for (int j = 0; j < kPoolR; ++j) mine += uint32_t(__popc(mask[j] & tm));
            // exclusive prefix over the lanes
            uint32_t off = mine;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const uint32_t v = __shfl_up_sync(0xffffffffu, off, d);
                if (lane >= uint32_t(d)) off += v;
            }
            const uint32_t total = __shfl_sync(0xffffffffu, off, 31);
            if (!total) continue;
            off -= mine;
#pragma unroll
            for (int j = 0; j < kPoolR; ++j) {
                uint32_t m = mask[j] & tm;
                while (m) {
                    const int i = __ffs(int(m)) - 1;
                    m &= m - 1u;
                    items[off++] = (unsigned short)(((uint32_t(j) * 32u + lane) << 4) | uint32_t(i));
                }
            }
            __syncwarp();
#pragma unroll 1
            for (uint32_t base = 0; base < total; base += 32u) {
                const bool has = base + lane < total;
                const uint32_t it = has ? items[base + lane] : 0u;
                const uint32_t slot = it >> 4;
                unsigned long long k1 = key_none, k2 = key_none;
                f3 Pl1 = mk3(0, 0, 0), Pl2 = mk3(0, 0, 0);
                int packed = 0;
                if (has) {
                    const float4 o4 = rayO[slot], d4 = rayD[slot];
                    const f3 O = mk3(o4.x, o4.y, o4.z), D = mk3(d4.x, d4.y, d4.z);
                    const uint32_t oct0 = 0u;  // the primitive id sits in every octant copy of the list
                    const int prim = __float_as_int(sm.top_nodes[oct0 * uint32_t(n) + (it & 15u)].lo.w);
                    const DevPrim* pp = sm.prims + prim;
                    Aff inv, fwd;
                    inv.r0 = pp->inv.r0;
                    inv.r1 = pp->inv.r1;
                    inv.r2 = pp->inv.r2;
                    fwd.r0 = pp->fwd.r0;
                    fwd.r1 = pp->fwd.r1;
                    fwd.r2 = pp->fwd.r2;
                    const int tslot = pp->slot;
                    if (COUNT) ++wk.prims;
                    const f3 Oi = xform_point(inv, O);
                    const f3 Di = normalize3(xform_vector(inv, D));
                    Cand k;
                    k.a1 = k.a2 = 0.0f;
                    k.cl = 0;
                    k.c1 = k.c2 = false;
                    switch (type) {  // warp-uniform
                        case MCPT_PRIM_SPHERE: sphere_test(k, Oi, Di, o4.w * pp->inv_smin); break;
                        case MCPT_PRIM_CUBE: cube_test(k, Oi, Di); break;
                        case MCPT_PRIM_CYLINDER: cylinder_test(k, Oi, Di); break;
                        case MCPT_PRIM_CONE: cone_test(k, Oi, Di, O, S.leaf_bb + 6 * prim); break;
                        default: quad_test(k, Oi, Di, O, S.leaf_bb + 6 * prim); break;
                    }
                    // offer_at's arithmetic (intersect.cuh): Pl = Oi + Di a, Pg = fwd Pl, dist = |O - Pg|; offer()'s order as a key:
                    // smaller distance, then the higher leaf slot, then the candidate offered first
                    const uint32_t tie = (uint32_t(0xffff - tslot) << 8);
                    packed = prim | (k.cl << 8);
                    if (k.c1) {
                        Pl1 = Oi + Di * k.a1;
                        const float dist = distance3(O, xform_point(fwd, Pl1));
                        if (dist < MCPT_FLT_MAX) k1 = (static_cast<unsigned long long>(__float_as_uint(dist)) << 32) | tie;
                    }
                    if (k.c2) {
                        Pl2 = Oi + Di * k.a2;
                        const float dist = distance3(O, xform_point(fwd, Pl2));
                        if (dist < MCPT_FLT_MAX) k2 = (static_cast<unsigned long long>(__float_as_uint(dist)) << 32) | tie | 1u;
                    }
                    if (k1 != key_none) atomicMin(&keys[slot], k1);
                    if (k2 != key_none) atomicMin(&keys[slot], k2);
                }
                __syncwarp();
                if (has) {
                    const unsigned long long kb = keys[slot];
                    if (k1 != key_none && kb == k1) pay[slot] = make_float4(Pl1.x, Pl1.y, Pl1.z, __int_as_float(packed));
                    if (k2 != key_none && kb == k2) pay[slot] = make_float4(Pl2.x, Pl2.y, Pl2.z, __int_as_float(packed));
                }
                __syncwarp();
            }
        }
        // ---- phase 3: every ray's record
#pragma unroll
        for (int j = 0; j < kPoolR; ++j) {
            const uint32_t slot = uint32_t(j) * 32u + lane;
            Best best;
            best_reset(best);
            const unsigned long long kb = keys[slot];
            const float4 d4 = rayD[slot];
            if (kb != key_none) {
                const float4 pv = pay[slot];
                const int packed = __float_as_int(pv.w);
                best.h.dist = __uint_as_float(uint32_t(kb >> 32));
                best.h.pl = mk3(pv.x, pv.y, pv.z);
                best.h.prim = packed & 0xff;
                best.h.sub = packed >> 8;
            }
            deliver<EXIT>(S, wf, sink, &s_hits, pid[j], valid[j], best, mk3(d4.x, d4.y, d4.z));
        }
        __syncwarp();
    }
    if (ray_counter && threadIdx.x == 0) atomicAdd(ray_counter, (unsigned long long)len);
    if (COUNT) flush_work(wk, work);
    if (sink.hitq.items) {
        __syncthreads();
        if (threadIdx.x == 0) sink.hitq.counts[blockIdx.x] = s_hits;
    }
}

template <bool COUNT, int MINB>
static void launch_traverse_kind(const LaunchCfg& lc, const DevScene& S, Wavefront wf, const QueueIn& qin, const HitSink& sink,
                                 unsigned long long* ray_counter, unsigned long long* work) {
    const int grid = lc.grid;
#define MCPT_TRAVERSE(FLAT, MESHES, PERSIST)                                                                                        \
    do {                                                                                                                            \
        if (!MESHES && sink.exit_mode)                                                                                              \
            k_traverse<COUNT, MINB, FLAT, MESHES, PERSIST, !MESHES><<<grid, kBlock, S.smem_bytes, lc.stream>>>(S, wf, qin, sink,    \
                                                                                                              ray_counter, work);  \
        else                                                                                                                        \
            k_traverse<COUNT, MINB, FLAT, MESHES, PERSIST, false><<<grid, kBlock, S.smem_bytes, lc.stream>>>(S, wf, qin, sink,      \
                                                                                                            ray_counter, work);    \
    } while (0)
    if (S.n_flat > 0 && S.flat_pool && qin.items && !sink.exit_mode) {
        // after the first bounce: exact tests pooled per warp and primitive type (k_traverse_flat_pooled). Interior rays keep
        // the sequential walk: they start inside a primitive whose far side bounds the search at once (measured: config 2
        // 1.81 vs 1.95 ms per step, scene 2 1.23 vs 1.28, scene 4 0.73 vs 0.76)
        const size_t base = (size_t(S.smem_bytes) + 15) & ~size_t(15);
        const size_t smem_async = base + (kBlock / 32) * pool_bytes_per_warp(S.flat_max_type, true);
        const size_t smem_plain = base + (kBlock / 32) * pool_bytes_per_warp(S.flat_max_type, false);
        allow_large_smem<k_traverse_flat_pooled<COUNT, MINB, false, false>>();
        allow_large_smem<k_traverse_flat_pooled<COUNT, MINB, false, true>>();
        // the second pair of ray buffers only where MINB CTAs of it still fit one SM (228 KB, 1 KB reserved per CTA)
        if (S.flat_pool >= 2 && MINB * (smem_async + 1024 + 64) <= size_t(228) * 1024)
            k_traverse_flat_pooled<COUNT, MINB, false, true><<<grid, kBlock, smem_async, lc.stream>>>(S, wf, qin, sink, ray_counter, work);
        else
            k_traverse_flat_pooled<COUNT, MINB, false, false><<<grid, kBlock, smem_plain, lc.stream>>>(S, wf, qin, sink, ray_counter, work);
    } else if (S.n_flat > 0)
        MCPT_TRAVERSE(true, false, false);
    else if (S.n_meshes == 0) {
        if (S.persistent) MCPT_TRAVERSE(false, false, true);
        else MCPT_TRAVERSE(false, false, false);
    } else {
        if (S.persistent) MCPT_TRAVERSE(false, true, true);
        else MCPT_TRAVERSE(false, true, false);
    }
#undef MCPT_TRAVERSE
}

void launch_traverse(const LaunchCfg& lc, const DevScene& S, Wavefront wf, const QueueIn& qin, const QueueOut& hitq,
                     const DevParams* exit_prm, unsigned long long* ray_counter, unsigned long long* work) {
    HitSink sink{hitq, 0, 0, 1.0f};
    if (exit_prm) {
        sink.exit_mode = 1;
        sink.flat_face = exit_prm->flat_face;
        sink.refract_ind = exit_prm->refract_ind;
    }
    if (S.u_nodes && S.persistent && S.u_pool > 0) {
        const size_t smem = ((size_t(S.smem_bytes) + 15) & ~size_t(15)) + (kBlock / 32) * kUPoolBytesPerWarp;
#define MCPT_TRAVERSE_UP(COUNT, MINB)                                                                                                   \
    do {                                                                                                                                \
        allow_large_smem<k_traverse_unified_pooled<COUNT, MINB, true>>();                                                               \
        allow_large_smem<k_traverse_unified_pooled<COUNT, MINB, false>>();                                                              \
        if (sink.exit_mode)                                                                                                             \
            k_traverse_unified_pooled<COUNT, MINB, true><<<lc.grid, kBlock, smem, lc.stream>>>(S, wf, qin, sink, ray_counter, work);    \
        else                                                                                                                            \
            k_traverse_unified_pooled<COUNT, MINB, false><<<lc.grid, kBlock, smem, lc.stream>>>(S, wf, qin, sink, ray_counter, work);   \
    } while (0)
        if (work) MCPT_TRAVERSE_UP(true, 2);
        else if (lc.traverse_occ >= 4) MCPT_TRAVERSE_UP(false, 4);
        else if (lc.traverse_occ == 3) MCPT_TRAVERSE_UP(false, 3);
        else MCPT_TRAVERSE_UP(false, 2);
#undef MCPT_TRAVERSE_UP
        lc.count();
        return;
    }
    if (S.u_nodes && S.persistent) {
#define MCPT_TRAVERSE_U(COUNT, MINB)                                                                                              \
    do {                                                                                                                          \
        if (sink.exit_mode)                                                                                                       \
            k_traverse_unified<COUNT, MINB, true><<<lc.grid, kBlock, S.smem_bytes, lc.stream>>>(S, wf, qin, sink, ray_counter, work);  \
        else                                                                                                                      \
            k_traverse_unified<COUNT, MINB, false><<<lc.grid, kBlock, S.smem_bytes, lc.stream>>>(S, wf, qin, sink, ray_counter, work); \
    } while (0)
        if (work) MCPT_TRAVERSE_U(true, 2);
        else if (lc.traverse_occ >= 6) MCPT_TRAVERSE_U(false, 6);
        else if (lc.traverse_occ == 5) MCPT_TRAVERSE_U(false, 5);
        else if (lc.traverse_occ == 4) MCPT_TRAVERSE_U(false, 4);
        else if (lc.traverse_occ == 3) MCPT_TRAVERSE_U(false, 3);
        else MCPT_TRAVERSE_U(false, 2);
#undef MCPT_TRAVERSE_U
        lc.count();
        return;
    }
    if (work)
        launch_traverse_kind<true, 2>(lc, S, wf, qin, sink, ray_counter, work);
    else if (lc.traverse_occ >= 4)
        launch_traverse_kind<false, 4>(lc, S, wf, qin, sink, ray_counter, nullptr);
    else if (lc.traverse_occ == 3)
        launch_traverse_kind<false, 3>(lc, S, wf, qin, sink, ray_counter, nullptr);
    else
        launch_traverse_kind<false, 2>(lc, S, wf, qin, sink, ray_counter, nullptr);
    lc.count();
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
    lc.count();
}

// ---- the same callable through the PRODUCTION traversal kernel
// The ray batches of the parity tests go through k_traverse itself -- flat list, heap walk, persistent warps over
// the SAH top2 / top4 trees and the GPU-built mesh trees, whichever mcpt_upload_scene selected -- exactly as a
// render launches it for the debug shaders: identity queue, no hit queue, every record to the hit sector
// (deliver()). Two thin kernels move the caller's rays into ray sectors and the hit sectors into mcpt_hit
// records. The hit sector carries {pl, prim, sub}; dist and pg are functions of those and of the ray origin
// (offer_at: Pg = fwd * Pl, dist = |O - Pg|) and are re-evaluated here with the same unfused arithmetic.
__global__ void __launch_bounds__(kBlock) k_trace_pack(int64_t n, const float4* __restrict__ o, const float4* __restrict__ d,
                                                        Wavefront wf) {
    for (int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += int64_t(gridDim.x) * blockDim.x) {
        const float4 o4 = o[i], d4 = d[i];
        wf.store_ray(size_t(i), make_float4(o4.x, o4.y, o4.z, 0.0f), make_float4(d4.x, d4.y, d4.z, 0.0f));
    }
}

__global__ void __launch_bounds__(kBlock) k_trace_unpack(DevScene S, int64_t n, const float4* __restrict__ o, Wavefront wf,
                                                          mcpt_hit* __restrict__ out) {
    for (int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += int64_t(gridDim.x) * blockDim.x) {
        f3 pl, D;
        int prim, sub;
        wf.load_hit(size_t(i), pl, prim, sub, D);
        const float4 o4 = o[i];
        mcpt_hit h;
        h.index = prim;
        h.pl[0] = pl.x;
        h.pl[1] = pl.y;
        h.pl[2] = pl.z;
        if (prim >= 0) {
            const DevPrim* pp = S.prims + prim;
            Aff fwd;
            fwd.r0 = pp->fwd.r0;
            fwd.r1 = pp->fwd.r1;
            fwd.r2 = pp->fwd.r2;
            const f3 pg = xform_point(fwd, pl);
            h.dist = distance3(mk3(o4.x, o4.y, o4.z), pg);
            h.pg[0] = pg.x;
            h.pg[1] = pg.y;
            h.pg[2] = pg.z;
            h.shape = pp->type;
            h.tri_index = (pp->type == MCPT_PRIM_MESH) ? sub : -1;
            h.dir = (pp->type == MCPT_PRIM_MESH) ? 0 : sub;
        } else {
            h.dist = MCPT_FLT_MAX;
            h.pg[0] = h.pg[1] = h.pg[2] = 0.0f;
            h.shape = -1;
            h.tri_index = -1;
            h.dir = -1;
        }
        out[i] = h;
    }
}

void launch_trace_wavefront(const LaunchCfg& lq, const DevScene& S, int64_t n, const float4* o, const float4* d, Wavefront wf,
                            uint32_t cap, void* hits_out) {
    k_trace_pack<<<grid_for(lq, uint64_t(n), 16), kBlock, 0, lq.stream>>>(n, o, d, wf);
    lq.count();
    const QueueIn identity{nullptr, nullptr, uint32_t(n), cap};
    const QueueOut none{nullptr, nullptr, 0};
    launch_traverse(lq, S, wf, identity, none, nullptr, nullptr, nullptr);
    k_trace_unpack<<<grid_for(lq, uint64_t(n), 16), kBlock, 0, lq.stream>>>(S, n, o, wf, static_cast<mcpt_hit*>(hits_out));
    lq.count();
}

// (shade / compact / accumulate / resolve follow in kernels_shade.inl)
#include "kernels_shade.inl"

}  // namespace mcpt
