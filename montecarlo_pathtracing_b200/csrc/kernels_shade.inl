// kernels_shade.inl — included by kernels.cu inside namespace mcpt.
// shade (+ queue append) / shade_exit (mesh scenes) / accumulate / resolve stages.

// -------------------------------------------------------------------- shade
// One iteration of random_path's loop after traverse_all_bvh
// (tp/montecarlo.frag:117-175), or one of the two debug shaders
// (tp/montecarlo_mat.frag:5-20, tp/montecarlo_mat_tr.frag:5-20) when mode != 0.
// In the Monte Carlo mode the queue holds only rays that hit something (a ray that leaves the scene keeps its
// path state and k_accumulate evaluates the fake sky). Paths that go on are appended to `next` -- and to `interior` when the new ray starts
// inside a transparent primitive -- by warp_append into this CTA's regions of the two queues
// (kernels.cu): that is the whole compaction, there is no separate pass and no flag array.
// (Measured and dropped: prefetch.global.L2 of the next path's state at the end of an iteration, +3 % time;
// a software pipeline that requests the next path's queue entry and sectors before the current path is
// shaded: +2..9 % time at 64 or 80 registers -- the gathers are limited by DRAM efficiency on scattered
// 32-byte sectors (45-54 % of peak), not by exposed latency; the same pipeline through shared memory with cp.async
// (hit and radiance sectors of every thread's next path as four 16-byte LDGSTS, no registers): TWICE the time (C2 7.8 ->
// 15.4 ms per step, scene 8 12.4 -> 30.9) -- a 32-byte sector then costs two requests to the L2. Filing the rays of the next bounce in two
// classes -- leaving upwards / downwards -- at the two ends of the queue region: the flat traversal got 9 %
// faster on warps of one class (C2 11.6 -> 10.6 ms per step), but the hit queue then walks the path slots
// in two sparser runs and this kernel lost more through DRAM locality (7.8 -> 9.2 ms).)
#ifndef MCPT_SHADE_MINB
#define MCPT_SHADE_MINB 4  // 64 registers, 4 CTAs per SM: the kernel waits on its first loads (measured +2 %)
// (Measured and dropped: compiling the FAST instantiation with FMA contraction. 48 % of its instructions
// are separate FMUL / FADD, but the kernel got only 1.4 % faster, while the fixed-seed agreement with the
// reference's 16k-spp image fell from RMSE 1.3e-5 to 2.9e-3: fused rounding sends the chaotic paths
// elsewhere. Not worth the parity evidence.)
#endif
template <bool FAST>
__global__ void __launch_bounds__(kBlock, MCPT_SHADE_MINB) k_shade(DevScene S, DevParams prm, Wavefront wf, QueueIn qin, QueueOut next,
                                                   QueueOut interior, int first_bounce, int last_bounce, int first_pass,
                                                   uint32_t npix, unsigned long long* seg_counter) {
    __shared__ uint32_t s_count[2];
    const uint32_t len = queue_len(qin);
    if (len == 0) {
        if (threadIdx.x == 0 && next.items) next.counts[blockIdx.x] = interior.counts[blockIdx.x] = 0u;
        return;
    }
    if (threadIdx.x < 2) s_count[threadIdx.x] = 0u;
    __syncthreads();
    const uint32_t lane = threadIdx.x & 31u;
    const float tag_alive = __uint_as_float(2u * prm.batch), tag_ended = __uint_as_float(2u * prm.batch + 1u);
    unsigned int my_segments = 0;
    for (uint32_t base = 0; base < len; base += kBlock) {
        uint32_t kind = 0, p;  // kind: 0 ended, 1 goes on, 2 goes on after an interior traversal
        if (queue_item(qin, base + threadIdx.x, len, p)) {
            f3 pl, D;
            int2 hb;
            wf.load_hit(p, pl, hb.x, hb.y, D);  // the direction of the ray that hit comes with the record
            f3 O = mk3(0.0f, 0.0f, 0.0f);
            if (!FAST) {
                float4 o4, d4;
                wf.load_ray(p, o4, d4);  // only E = normalize(O - P) needs the origin (see bounce())
                O = mk3(o4.x, o4.y, o4.z);
            }
            // a fresh path starts with total = 0 and attenuation 0.8 (tp/montecarlo.frag:105-107), no draws
            f3 total = mk3(0.0f, 0.0f, 0.0f), attenu = mk3(0.8f, 0.8f, 0.8f);
            uint32_t n_draws = 0u;
            if (!first_bounce) {
                float4 a4, t4;
                wf.load_rad(p, a4, t4);
                total = mk3(t4.x, t4.y, t4.z);
                attenu = mk3(a4.x, a4.y, a4.z);
                n_draws = __float_as_uint(t4.w);
            }
            if (prm.mode != 0) {
                // debug shaders: a single traversal, no bounce
                f3 res = mk3(0.0f, 0.0f, 0.2f);
                if (hb.x >= 0) {
                    uint3 seed = path_seed(p, npix, prm.width, prm.height, first_pass, prm.date, n_draws);
                    f3 N = mk3(0, 0, 0), P = mk3(0, 0, 0);
                    intersection_info(S, hb.x, hb.y, pl, prm.flat_face, N, P);
                    if (prm.mode == 1) {
                        float r0 = random_float(seed), r1 = random_float(seed), r2 = random_float(seed);
                        res = mk3(fabsf(N.x) * r0, fabsf(N.y) * r1, fabsf(N.z) * r2);
                    } else {
                        float4 col = __ldg(S.color + hb.x);
                        float r = random_float(seed);
                        res = mk3(col.x * r, col.y * r, col.z * r);
                    }
                }
                wf.store_rad(p, make_float4(0.0f, 0.0f, 0.0f, tag_ended), make_float4(res.x, res.y, res.z, 0.0f));
            } else {
                ++my_segments;
                const uint3 seed0 = path_seed(p, npix, prm.width, prm.height, first_pass, prm.date, n_draws);
                uint3 seed = seed0;
                f3 N = mk3(0, 0, 0), P = mk3(0, 0, 0);
                intersection_info(S, hb.x, hb.y, pl, prm.flat_face, N, P);
                const float4 mat = __ldg(S.mat + hb.x);
                const float4 col = __ldg(S.color + hb.x);
                BounceOut bo = bounce<FAST>(O, D, N, P, mat, col, attenu, total, seed, prm.refract_ind);
                if (bo.kind == 0) {
                    wf.store_rad(p, make_float4(0.0f, 0.0f, 0.0f, tag_ended), make_float4(total.x, total.y, total.z, 0.0f));
                } else if (last_bounce) {
                    // a path still alive after the last iteration returns vec3(0) (tp/montecarlo.frag:178)
                    wf.store_rad(p, make_float4(0.0f, 0.0f, 0.0f, tag_ended), make_float4(0.0f, 0.0f, 0.0f, 0.0f));
                } else {
                    const uint32_t nd = n_draws + draws_between(seed0, seed);
                    wf.store_ray(p, make_float4(bo.nO.x, bo.nO.y, bo.nO.z, 0.0f),
                                 make_float4(bo.nD.x, bo.nD.y, bo.nD.z, __uint_as_float(nd)));
                    wf.store_rad(p, make_float4(bo.new_att.x, bo.new_att.y, bo.new_att.z, tag_alive),
                                 make_float4(total.x, total.y, total.z, __uint_as_float(nd)));
                    if (bo.kind == 2) wf.store_keep(p, make_float4(N.x, N.y, N.z, 0.0f), make_float4(P.x, P.y, P.z, 0.0f));
                    kind = uint32_t(bo.kind);
                }
            }
        }
        if (next.items) {
            warp_append(&s_count[0], next, kind != 0u, p);
            warp_append(&s_count[1], interior, kind == 2u, p);
        }
    }
    if (seg_counter) {
        // one atomic per warp
        for (int off = 16; off > 0; off >>= 1) my_segments += __shfl_down_sync(0xffffffffu, my_segments, off);
        if (lane == 0 && my_segments) atomicAdd(seg_counter, (unsigned long long)my_segments);
    }
    if (next.items) {
        __syncthreads();
        if (threadIdx.x == 0) {
            next.counts[blockIdx.x] = s_count[0];
            interior.counts[blockIdx.x] = s_count[1];
        }
    }
}

void launch_shade(const LaunchCfg& lc, const DevScene& S, const DevParams& prm, Wavefront wf, const QueueIn& qin,
                  const QueueOut& next, const QueueOut& interior, bool first_bounce, bool last_bounce, int first_pass,
                  uint32_t npix, bool fast_math, unsigned long long* seg_counter) {
    const QueueOut none{nullptr, nullptr, 0};
    const QueueOut& nx = last_bounce ? none : next;
    const QueueOut& in = last_bounce ? none : interior;
    if (fast_math)
        k_shade<true><<<lc.grid, kBlock, 0, lc.stream>>>(S, prm, wf, qin, nx, in, first_bounce ? 1 : 0, last_bounce ? 1 : 0,
                                                         first_pass, npix, seg_counter);
    else
        k_shade<false><<<lc.grid, kBlock, 0, lc.stream>>>(S, prm, wf, qin, nx, in, first_bounce ? 1 : 0, last_bounce ? 1 : 0,
                                                          first_pass, npix, seg_counter);
    lc.count();
}

// ---------------------------------------------------- next-event estimation (MCPT_OPT_NEE)
// A NEW estimator on top of the reference's random_path (which has none: its light sampler is commented out,
// shaders/raytracer_func.frag:908-956; the data it would use -- emissive records first, mat_info.a -- is what
// DevLight is built from). In random_path a path only DELIVERS its `total` when it ends: on an emissive surface
// (emissivity > 0.5), or in the sky. The event "the next ray hits light L" is here also estimated directly: at every
// vertex whose continuation ray is drawn from a cone (diffuse and reflective cases), one point y on one light quad is
// sampled, and (total + what the hit on L would add along that direction) is delivered at once if y is visible, weighted
// against the continuation ray by the balance heuristic of the two densities (the cone's, cone_pdf, and the light
// sampler's converted to solid angle); a continuation ray that does hit a sampled light delivers its value with the
// complementary weight. The expectation of every pixel is unchanged; the bounce rays themselves are the SAME rays as
// without NEE (the light sampler draws from a stream of its own), only the bookkeeping differs.
// Extra state per path slot, three 32-byte sectors (`nee`): 0 shadow ray, 1 its hit record (written by the traversal
// kernel), 2 {delivered sum.xyz, density of the continuation ray | pending contribution.xyz, expected primitive}.
MCPT_DEV int find_light(const DevScene& S, int prim) {
    for (int i = 0; i < S.n_lights; ++i)
        if (__float_as_int(__ldg(&S.lights[i].n).w) == prim) return i;
    return -1;
}

template <bool FAST>
__global__ void __launch_bounds__(kBlock, 3) k_shade_nee(DevScene S, DevParams prm, Wavefront wf, Wavefront nee, QueueIn qin, QueueOut next,
                                                         QueueOut interior, QueueOut shadow, int first_bounce, int last_bounce,
                                                         int first_pass, uint32_t npix, unsigned long long* seg_counter) {
    __shared__ uint32_t s_count[3];
    const uint32_t len = queue_len(qin);
    if (len == 0) {
        if (threadIdx.x == 0) {
            if (next.items) next.counts[blockIdx.x] = interior.counts[blockIdx.x] = 0u;
            shadow.counts[blockIdx.x] = 0u;
        }
        return;
    }
    if (threadIdx.x < 3) s_count[threadIdx.x] = 0u;
    __syncthreads();
    const uint32_t lane = threadIdx.x & 31u;
    const float tag_alive = __uint_as_float(2u * prm.batch), tag_ended = __uint_as_float(2u * prm.batch + 1u);
    unsigned int my_segments = 0;
    for (uint32_t base = 0; base < len; base += kBlock) {
        uint32_t kind = 0, p;
        bool shadow_job = false;
        if (queue_item(qin, base + threadIdx.x, len, p)) {
            f3 pl, D;
            int2 hb;
            wf.load_hit(p, pl, hb.x, hb.y, D);
            float4 o4, d4;
            wf.load_ray(p, o4, d4);
            const f3 O = mk3(o4.x, o4.y, o4.z);
            f3 total = mk3(0.0f, 0.0f, 0.0f), attenu = mk3(0.8f, 0.8f, 0.8f);
            uint32_t n_draws = 0u;
            if (!first_bounce) {
                float4 a4, t4;
                wf.load_rad(p, a4, t4);
                total = mk3(t4.x, t4.y, t4.z);
                attenu = mk3(a4.x, a4.y, a4.z);
                n_draws = __float_as_uint(t4.w);
            }
            PathSector ns = *nee.sec(p, 2);   // zeroed at the start of the batch
            const float p_prev = first_bounce ? 0.0f : ns.a.w;
            ++my_segments;
            const uint3 seed0 = path_seed(p, npix, prm.width, prm.height, first_pass, prm.date, n_draws);
            uint3 seed = seed0;
            f3 N = mk3(0, 0, 0), P = mk3(0, 0, 0);
            intersection_info(S, hb.x, hb.y, pl, prm.flat_face, N, P);
            const float4 mat = __ldg(S.mat + hb.x);
            const float4 col = __ldg(S.color + hb.x);
            NeeInfo ni;
            BounceOut bo = bounce<FAST>(O, D, N, P, mat, col, attenu, total, seed, prm.refract_ind, &ni);
            if (bo.kind == 0) {
                // the path ends on an emissive surface. If that surface is a sampled light and the ray came from a cone
                // that next-event estimation also covered, the value carries the continuation ray's share of the weight
                float w_b = 1.0f;
                if (p_prev > 0.0f) {
                    const int li = find_light(S, hb.x);
                    if (li >= 0) {
                        const DevLight L = S.lights[li];
                        const f3 w = P - O;
                        const float d2 = dot3(w, w);
                        const float cy = fabsf(dot3(mk3(L.n.x, L.n.y, L.n.z), w)) * rsqrtf(fmaxf(d2, 1e-30f));
                        const float p_l = L.ex.w * L.c.w * d2 / fmaxf(cy, 1e-8f);
                        w_b = p_prev / (p_prev + p_l);
                    }
                }
                wf.store_rad(p, make_float4(0.0f, 0.0f, 0.0f, tag_ended), make_float4(total.x * w_b, total.y * w_b, total.z * w_b, 0.0f));
            } else if (last_bounce) {
                wf.store_rad(p, make_float4(0.0f, 0.0f, 0.0f, tag_ended), make_float4(0.0f, 0.0f, 0.0f, 0.0f));
            } else {
                const uint32_t nd = n_draws + draws_between(seed0, seed);
                wf.store_ray(p, make_float4(bo.nO.x, bo.nO.y, bo.nO.z, 0.0f), make_float4(bo.nD.x, bo.nD.y, bo.nD.z, __uint_as_float(nd)));
                wf.store_rad(p, make_float4(bo.new_att.x, bo.new_att.y, bo.new_att.z, tag_alive),
                             make_float4(total.x, total.y, total.z, __uint_as_float(nd)));
                if (bo.kind == 2) wf.store_keep(p, make_float4(N.x, N.y, N.z, 0.0f), make_float4(P.x, P.y, P.z, 0.0f));
                kind = uint32_t(bo.kind);
                float p_next = 0.0f;
                if (ni.cone != 0 && S.n_lights > 0 && ni.rho * ni.rho > 1e-3f) {
                    p_next = cone_pdf(dot3(ni.axis, Tm<FAST>::normalize_(bo.nD)), ni.rho);
                    // the light sampler's own random numbers: the bounce rays must not depend on MCPT_OPT_NEE
                    uint3 ls = make_uint3(seed0.x ^ 0x9E3779B9u, seed0.y ^ 0x85EBCA6Bu, seed0.z ^ 0xC2B2AE35u);
                    const float r0 = random_float(ls), r1 = random_float(ls), r2 = random_float(ls);
                    int li = 0;
                    while (li + 1 < S.n_lights && r0 >= __ldg(&S.lights[li].ey).w) ++li;
                    const DevLight L = S.lights[li];
                    const f3 y = mk3(L.c.x, L.c.y, L.c.z) + mk3(L.ex.x, L.ex.y, L.ex.z) * (2.0f * r1 - 1.0f) + mk3(L.ey.x, L.ey.y, L.ey.z) * (2.0f * r2 - 1.0f);
                    const f3 w = y - bo.nO;
                    const float d2 = dot3(w, w);
                    if (d2 > 1e-12f) {
                        const f3 wn = w * rsqrtf(d2);
                        const float cy = fabsf(dot3(mk3(L.n.x, L.n.y, L.n.z), wn));
                        const float pc = cone_pdf(dot3(ni.axis, wn), ni.rho);
                        if (pc > 0.0f && cy > 1e-6f) {
                            const float p_l = L.ex.w * L.c.w * d2 / cy;
                            float spec = 1.0f;
                            if (ni.cone == 1) spec = Tm<FAST>::pow_(fmaxf(0.0f, dot3(ni.E, reflect3(-wn, ni.N))), ni.se);
                            const f3 att = ni.base + ni.gain * spec;
                            const f3 value = total + mk3(L.emit.x, L.emit.y, L.emit.z) + att * L.emit.w;
                            const float wgt = pc / (pc + p_l);
                            ns.b = make_float4(value.x * wgt, value.y * wgt, value.z * wgt, L.n.w);
                            nee.store_ray(p, make_float4(bo.nO.x, bo.nO.y, bo.nO.z, 0.0f), make_float4(wn.x, wn.y, wn.z, 0.0f));
                            shadow_job = true;
                        }
                    }
                }
                ns.a.w = p_next;
                *nee.sec(p, 2) = ns;
            }
        }
        if (next.items) {
            warp_append(&s_count[0], next, kind != 0u, p);
            warp_append(&s_count[1], interior, kind == 2u, p);
        }
        warp_append(&s_count[2], shadow, shadow_job, p);
    }
    if (seg_counter) {
        for (int off = 16; off > 0; off >>= 1) my_segments += __shfl_down_sync(0xffffffffu, my_segments, off);
        if (lane == 0 && my_segments) atomicAdd(seg_counter, (unsigned long long)my_segments);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        if (next.items) {
            next.counts[blockIdx.x] = s_count[0];
            interior.counts[blockIdx.x] = s_count[1];
        }
        shadow.counts[blockIdx.x] = s_count[2];
    }
}

// after the shadow rays were traced: a sampled light point is visible when the closest hit along the shadow ray is
// that light; its pending contribution is then delivered
__global__ void __launch_bounds__(kBlock) k_nee_resolve(Wavefront nee, QueueIn qin) {
    const uint32_t len = queue_len(qin);
    for (uint32_t i = threadIdx.x; i < len; i += kBlock) {
        uint32_t p;
        if (!queue_item(qin, i, len, p)) continue;
        f3 pl, D;
        int prim, sub;
        nee.load_hit(p, pl, prim, sub, D);
        PathSector ns = *nee.sec(p, 2);
        if (prim >= 0 && prim == __float_as_int(ns.b.w)) {
            ns.a.x += ns.b.x;
            ns.a.y += ns.b.y;
            ns.a.z += ns.b.z;
            *nee.sec(p, 2) = ns;
        }
    }
}

// k_accumulate with the delivered next-event sums added per path
__global__ void __launch_bounds__(kBlock) k_accumulate_nee(Wavefront wf, Wavefront nee, float4* __restrict__ accum, uint32_t npix, int n_passes,
                                                            uint32_t batch) {
    for (uint32_t pix = blockIdx.x * blockDim.x + threadIdx.x; pix < npix; pix += gridDim.x * blockDim.x) {
        float4 a = accum[pix];
        for (int k = 0; k < n_passes; ++k) {
            const uint64_t p = uint64_t(k) * npix + pix;
            float4 a4, t4;
            wf.load_rad(p, a4, t4);
            const uint32_t tag = __float_as_uint(a4.w);
            f3 r;
            if (tag == 2u * batch + 1u) {
                r = mk3(t4.x, t4.y, t4.z);
            } else {
                f3 total = mk3(0.0f, 0.0f, 0.0f), attenu = mk3(0.8f, 0.8f, 0.8f);
                if (tag == 2u * batch) {
                    total = mk3(t4.x, t4.y, t4.z);
                    attenu = mk3(a4.x, a4.y, a4.z);
                }
                const float4 d4 = wf.load_ray_d(p);
                r = sky(total, attenu, mk3(d4.x, d4.y, d4.z));
            }
            const float4 s4 = nee.sec(p, 2)->a;
            a.x += r.x + s4.x;
            a.y += r.y + s4.y;
            a.z += r.z + s4.z;
        }
        accum[pix] = a;
    }
}

void launch_shade_nee(const LaunchCfg& lc, const DevScene& S, const DevParams& prm, Wavefront wf, Wavefront nee, const QueueIn& qin,
                      const QueueOut& next, const QueueOut& interior, const QueueOut& shadow, bool first_bounce, bool last_bounce,
                      int first_pass, uint32_t npix, bool fast_math, unsigned long long* seg_counter) {
    const QueueOut none{nullptr, nullptr, 0};
    const QueueOut& nx = last_bounce ? none : next;
    const QueueOut& in = last_bounce ? none : interior;
    if (fast_math)
        k_shade_nee<true><<<lc.grid, kBlock, 0, lc.stream>>>(S, prm, wf, nee, qin, nx, in, shadow, first_bounce ? 1 : 0, last_bounce ? 1 : 0,
                                                             first_pass, npix, seg_counter);
    else
        k_shade_nee<false><<<lc.grid, kBlock, 0, lc.stream>>>(S, prm, wf, nee, qin, nx, in, shadow, first_bounce ? 1 : 0, last_bounce ? 1 : 0,
                                                              first_pass, npix, seg_counter);
    lc.count();
}

void launch_nee_resolve(const LaunchCfg& lc, Wavefront nee, const QueueIn& qin) {
    k_nee_resolve<<<lc.grid, kBlock, 0, lc.stream>>>(nee, qin);
    lc.count();
}

void launch_accumulate_nee(const LaunchCfg& lc, Wavefront wf, Wavefront nee, float4* accum, uint32_t npix, int n_passes, uint32_t batch) {
    k_accumulate_nee<<<grid_for(lc, npix, 16), kBlock, 0, lc.stream>>>(wf, nee, accum, npix, n_passes, batch);
    lc.count();
}

// --------------------------------------------------------------- shade_exit
// Scenes with meshes only (elsewhere the interior traversal does this itself, deliver() in kernels.cu:
// inside the persistent mesh walk the smooth-normal fetches of a few finished lanes would hold the warp).
// Second half of the transparent cases (tp/montecarlo.frag:150-152,163-165): the
// interior ray has been traced; take N,P of that hit (or keep the entry hit's
// N,P when nothing was hit: intersection_info writes nothing for shape -1) and
// leave through refract(D, -N, 1/refract_ind).
__global__ void __launch_bounds__(kBlock) k_shade_exit(DevScene S, DevParams prm, Wavefront wf, QueueIn qin) {
    const uint32_t len = queue_len(qin);
    for (uint32_t i = threadIdx.x; i < len; i += kBlock) {
        uint32_t p;
        if (!queue_item(qin, i, len, p)) continue;
        float4 n4, p4;
        f3 pl, Dhit;
        int2 hb;
        wf.load_hit(p, pl, hb.x, hb.y, Dhit);
        const float4 d4 = wf.load_ray_d(p);
        wf.load_keep(p, n4, p4);
        f3 N = mk3(n4.x, n4.y, n4.z), P = mk3(p4.x, p4.y, p4.z);
        intersection_info(S, hb.x, hb.y, pl, prm.flat_face, N, P);
        f3 nO = P + N * MCPT_BIAS;
        f3 nD = refract3(mk3(d4.x, d4.y, d4.z), -N, 1.0f / prm.refract_ind);
        wf.store_ray(p, make_float4(nO.x, nO.y, nO.z, 0.0f), make_float4(nD.x, nD.y, nD.z, d4.w));  // w: the path's RNG draw count
    }
}

void launch_shade_exit(const LaunchCfg& lc, const DevScene& S, const DevParams& prm, Wavefront wf, const QueueIn& qin) {
    k_shade_exit<<<lc.grid, kBlock, 0, lc.stream>>>(S, prm, wf, qin);
    lc.count();
}

// --------------------------------------------------------------- accumulate
#ifndef MCPT_ACC_UNROLL
#define MCPT_ACC_UNROLL 4
#endif
// glBlendFunc(GL_ONE, GL_ONE) into the RGB32F target, one draw per pass in pass
// order (montecarlo.cpp:450-466): accum = ((accum + r0) + r1) + ...
// The value random_path returns is read off the path's final state: `total` when the path ended
// in k_shade (emissive hit, absorbed, bounce limit), else the ray it was left with escaped and the
// fake sky applies (tp/montecarlo.frag:118-119) with the path's total and attenuation -- the
// initial 0 and 0.8 (:105-107) if it was never shaded in this batch.
__global__ void __launch_bounds__(kBlock) k_accumulate(Wavefront wf, float4* __restrict__ accum, uint32_t npix,
                                                        int n_passes, uint32_t batch, const float4* __restrict__ cam_dirs) {
    for (uint32_t pix = blockIdx.x * blockDim.x + threadIdx.x; pix < npix; pix += gridDim.x * blockDim.x) {
        float4 a = accum[pix];
        // a path that was never shaded still has its camera ray: with RayGen::on == 2 that ray was never stored, its
        // direction is the pixel's entry of the camera table
        float4 cam_d = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        if (cam_dirs) cam_d = __ldg(cam_dirs + pix);
        // kAccUnroll passes at a time: their radiance sectors are requested together, then the directions of the ones that
        // escaped, and only then are they added -- in pass order (one pass per trip left 32 dependent round trips to DRAM
        // per pixel: 0.65 ms per 33 M paths)
        constexpr int kAccUnroll = MCPT_ACC_UNROLL;
        for (int k0 = 0; k0 < n_passes; k0 += kAccUnroll) {
            float4 a4[kAccUnroll], t4[kAccUnroll], d4[kAccUnroll];
#pragma unroll
            for (int j = 0; j < kAccUnroll; ++j)
                if (k0 + j < n_passes) wf.load_rad(uint64_t(k0 + j) * npix + pix, a4[j], t4[j]);
#pragma unroll
            for (int j = 0; j < kAccUnroll; ++j)
                if (k0 + j < n_passes && __float_as_uint(a4[j].w) != 2u * batch + 1u) {
                    if (cam_dirs && __float_as_uint(a4[j].w) != 2u * batch) d4[j] = cam_d;
                    else d4[j] = wf.load_ray_d(uint64_t(k0 + j) * npix + pix);
                }
#pragma unroll
            for (int j = 0; j < kAccUnroll; ++j) {
                if (k0 + j >= n_passes) break;
                const uint32_t tag = __float_as_uint(a4[j].w);
                f3 r;
                if (tag == 2u * batch + 1u) {
                    r = mk3(t4[j].x, t4[j].y, t4[j].z);
                } else {
                    f3 total = mk3(0.0f, 0.0f, 0.0f), attenu = mk3(0.8f, 0.8f, 0.8f);
                    if (tag == 2u * batch) {
                        total = mk3(t4[j].x, t4[j].y, t4[j].z);
                        attenu = mk3(a4[j].x, a4[j].y, a4[j].z);
                    }
                    r = sky(total, attenu, mk3(d4[j].x, d4[j].y, d4[j].z));
                }
                a.x += r.x;
                a.y += r.y;
                a.z += r.z;
            }
        }
        accum[pix] = a;
    }
}

void launch_accumulate(const LaunchCfg& lc, Wavefront wf, float4* accum, uint32_t npix, int n_passes, uint32_t batch, const float4* cam_dirs) {
    k_accumulate<<<grid_for(lc, npix, 16), kBlock, 0, lc.stream>>>(wf, accum, npix, n_passes, batch, cam_dirs);
    lc.count();
}

// ------------------------------------------------------------------ resolve
// fs_frag (montecarlo.cpp:59-70): frag = texture(TU, tc).rgb / nb with NEAREST
// filtering, and shaders/average.frag:7-10: texelFetch(gl_FragCoord.xy /
// subsampling) * k. Destination is (w*sub) x (h*sub); RGBA8 clamps like the
// 8-bit back buffer (round to nearest, no gamma).
__global__ void __launch_bounds__(kBlock) k_resolve(const float4* __restrict__ accum, int w, int h, float divisor, int sub,
                                                     float* __restrict__ dst_f, uchar4* __restrict__ dst_u8) {
    const int W = w * sub, H = h * sub;
    const uint32_t total = uint32_t(W) * uint32_t(H);
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        const int X = int(i % uint32_t(W)), Y = int(i / uint32_t(W));
        const float4 a = accum[uint32_t(Y / sub) * uint32_t(w) + uint32_t(X / sub)];
        const float r = a.x / divisor, g = a.y / divisor, b = a.z / divisor;
        if (dst_f) {
            dst_f[3 * size_t(i) + 0] = r;
            dst_f[3 * size_t(i) + 1] = g;
            dst_f[3 * size_t(i) + 2] = b;
        } else {
            auto q = [](float v) -> unsigned char {
                v = fminf(fmaxf(v, 0.0f), 1.0f);  // NaN -> 0
                return (unsigned char)(__float2int_rn(v * 255.0f));
            };
            dst_u8[i] = make_uchar4(q(r), q(g), q(b), 255);
        }
    }
}

void launch_resolve(const LaunchCfg& lc, const float4* accum, int w, int h, float divisor, int subsampling, void* dst,
                    bool rgba8) {
    uint64_t total = uint64_t(w) * subsampling * uint64_t(h) * subsampling;
    k_resolve<<<grid_for(lc, total, 16), kBlock, 0, lc.stream>>>(accum, w, h, divisor, subsampling,
                                                                 rgba8 ? nullptr : static_cast<float*>(dst),
                                                                 rgba8 ? static_cast<uchar4*>(dst) : nullptr);
    lc.count();
}

// ---- the bounce sampler on its own (mcpt_sample_directions)
// n consecutive random_ray(D, roughness) draws of ONE RNG stream -- what tp/hsphere.vert:1-41 (DrawSampling, the
// reference's picture of the sampling cone) and the sampler fixtures draw sequentially: draw i starts from the state
// after 2 i random_float() calls, which is seed0 + 2 i (11, 43, 67), so the draws are independent threads here.
template <bool FAST>
__global__ void __launch_bounds__(kBlock) k_sample_dirs(float3 D, float roughness, uint3 seed0, int64_t n, float* __restrict__ out) {
    for (int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += int64_t(gridDim.x) * blockDim.x) {
        const uint32_t k = 2u * uint32_t(i);
        uint3 s = make_uint3(seed0.x + 11u * k, seed0.y + 43u * k, seed0.z + 67u * k);
        const f3 r = random_ray<FAST>(mk3(D.x, D.y, D.z), roughness, s);
        out[3 * i] = r.x;
        out[3 * i + 1] = r.y;
        out[3 * i + 2] = r.z;
    }
}

void launch_sample_dirs(const LaunchCfg& lc, const float D[3], float roughness, float tcx, float tcy, int numero_pass, float date,
                        int64_t n, float* out_dev, bool fast_math) {
    // srand() on the host: the same float products as rng_seed (raytracer_func.frag:105-110)
    const float k = float(1 + numero_pass);
    const float wx = tcx * k * 1.125f, wy = (float(numero_pass) * 3.14f + date) * k * 1.125f, wz = tcy * k * 1.125f;
    uint3 seed0;
    std::memcpy(&seed0.x, &wx, 4);
    std::memcpy(&seed0.y, &wy, 4);
    std::memcpy(&seed0.z, &wz, 4);
    const float3 d3 = make_float3(D[0], D[1], D[2]);
    if (fast_math)
        k_sample_dirs<true><<<grid_for(lc, uint64_t(n), 8), kBlock, 0, lc.stream>>>(d3, roughness, seed0, n, out_dev);
    else
        k_sample_dirs<false><<<grid_for(lc, uint64_t(n), 8), kBlock, 0, lc.stream>>>(d3, roughness, seed0, n, out_dev);
    lc.count();
}
