// kernels_shade.inl — included by kernels.cu inside namespace mcpt.
// shade / compact / shade_exit / accumulate / resolve stages.

// -------------------------------------------------------------------- shade
// One iteration of random_path's loop after traverse_all_bvh
// (tp/montecarlo.frag:117-175), or one of the two debug shaders
// (tp/montecarlo_mat.frag:5-20, tp/montecarlo_mat_tr.frag:5-20) when mode != 0.
// Writes flag[i] for queue position i: 0 ended, 1 continue, 2 continue after an
// interior traversal.
template <bool FAST>
__global__ void __launch_bounds__(kBlock) k_shade(DevScene S, DevParams prm, Wavefront wf, const uint32_t* __restrict__ queue,
                                                   const uint32_t* __restrict__ n_dev, uint32_t n_host, int first_bounce,
                                                   int last_bounce, int first_pass, uint32_t npix,
                                                   unsigned long long* seg_counter) {
    const uint32_t n = n_dev ? *n_dev : n_host;
    unsigned int my_segments = 0;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint32_t p = queue ? queue[i] : i;
        const float4 ha = wf.hit_a[p];
        const int2 hb = wf.hit_b[p];
        const float4 d4 = wf.ray_d[p];
        float4 o4 = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        if (!FAST) o4 = wf.ray_o[p];  // only E = normalize(O - P) needs it (see bounce())
        const f3 O = mk3(o4.x, o4.y, o4.z), D = mk3(d4.x, d4.y, d4.z);
        const uint32_t n_draws = first_bounce ? 0u : __float_as_uint(d4.w);
        // a fresh path starts with total = 0 and attenuation 0.8 (tp/montecarlo.frag:105-107)
        f3 total = mk3(0.0f, 0.0f, 0.0f), attenu = mk3(0.8f, 0.8f, 0.8f);
        if (!first_bounce && prm.mode == 0) {
            const float4 t4 = wf.total[p];
            total = mk3(t4.x, t4.y, t4.z);
            const float4 a4 = wf.attenu[p];
            attenu = mk3(a4.x, a4.y, a4.z);
        }
        uint8_t flag = 0;
        if (prm.mode != 0) {
            // debug shaders: a single traversal, no bounce
            f3 res = mk3(0.0f, 0.0f, 0.2f);
            if (hb.x >= 0) {
                uint3 seed = path_seed(p, npix, prm.width, prm.height, first_pass, prm.date, n_draws);
                f3 N = mk3(0, 0, 0), P = mk3(0, 0, 0);
                intersection_info(S, hb.x, hb.y, mk3(ha.x, ha.y, ha.z), prm.flat_face, N, P);
                if (prm.mode == 1) {
                    float r0 = random_float(seed), r1 = random_float(seed), r2 = random_float(seed);
                    res = mk3(fabsf(N.x) * r0, fabsf(N.y) * r1, fabsf(N.z) * r2);
                } else {
                    float4 col = __ldg(S.color + hb.x);
                    float r = random_float(seed);
                    res = mk3(col.x * r, col.y * r, col.z * r);
                }
            }
            wf.result[p] = make_float4(res.x, res.y, res.z, 0.0f);
        } else if (hb.x < 0) {
            f3 res = sky(total, attenu, D);
            wf.result[p] = make_float4(res.x, res.y, res.z, 0.0f);
        } else {
            ++my_segments;
            const uint3 seed0 = path_seed(p, npix, prm.width, prm.height, first_pass, prm.date, n_draws);
            uint3 seed = seed0;
            f3 N = mk3(0, 0, 0), P = mk3(0, 0, 0);
            intersection_info(S, hb.x, hb.y, mk3(ha.x, ha.y, ha.z), prm.flat_face, N, P);
            const float4 mat = __ldg(S.mat + hb.x);
            const float4 col = __ldg(S.color + hb.x);
            BounceOut bo = bounce<FAST>(O, D, N, P, mat, col, attenu, total, seed, prm.refract_ind);
            if (bo.kind == 0) {
                wf.result[p] = make_float4(total.x, total.y, total.z, 0.0f);
            } else if (last_bounce) {
                // a path still alive after the last iteration returns vec3(0) (tp/montecarlo.frag:178)
                wf.result[p] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
            } else {
                const uint32_t nd = n_draws + draws_between(seed0, seed);
                wf.ray_o[p] = make_float4(bo.nO.x, bo.nO.y, bo.nO.z, 0.0f);
                wf.ray_d[p] = make_float4(bo.nD.x, bo.nD.y, bo.nD.z, __uint_as_float(nd));
                wf.attenu[p] = make_float4(bo.new_att.x, bo.new_att.y, bo.new_att.z, 0.0f);
                wf.total[p] = make_float4(total.x, total.y, total.z, 0.0f);
                if (bo.kind == 2) {
                    wf.keep_n[p] = make_float4(N.x, N.y, N.z, 0.0f);
                    wf.keep_p[p] = make_float4(P.x, P.y, P.z, 0.0f);
                }
                flag = uint8_t(bo.kind | ((hb.x & 15) << 2));  // kind + grouping key (see k_compact_grouped)
            }
        }
        wf.flag[i] = flag;
    }
    if (seg_counter) {
        // one atomic per warp
        for (int off = 16; off > 0; off >>= 1) my_segments += __shfl_down_sync(0xffffffffu, my_segments, off);
        if ((threadIdx.x & 31) == 0 && my_segments) atomicAdd(seg_counter, (unsigned long long)my_segments);
    }
}

void launch_shade(const LaunchCfg& lc, const DevScene& S, const DevParams& prm, Wavefront wf, const QueueSet& q,
                  bool first_bounce, bool last_bounce, int first_pass, uint32_t npix, bool fast_math) {
    uint64_t n = q.n_cur ? 0xffffffffull : q.n_cur_host;
    int grid = grid_for(lc, n, 8);
    if (fast_math)
        k_shade<true><<<grid, kBlock, 0, lc.stream>>>(S, prm, wf, q.cur, q.n_cur, q.n_cur_host, first_bounce ? 1 : 0,
                                                      last_bounce ? 1 : 0, first_pass, npix, q.seg_counter);
    else
        k_shade<false><<<grid, kBlock, 0, lc.stream>>>(S, prm, wf, q.cur, q.n_cur, q.n_cur_host, first_bounce ? 1 : 0,
                                                       last_bounce ? 1 : 0, first_pass, npix, q.seg_counter);
    ++g_launches;
}

// ------------------------------------------------------------------ compact
// Warp-ballot stream compaction of the current queue into (a) the queue of
// paths entering the next bounce and (b) the queue of interior rays. Each warp
// ballots its flags, warps of a block combine their counts in shared memory, one
// thread reserves the block's output range with a single atomic per queue, and
// every surviving lane writes at base + (survivors before it). Order is kept
// inside a block's tile, which keeps neighbouring pixels neighbours.
__global__ void __launch_bounds__(kBlock) k_compact(Wavefront wf, const uint32_t* __restrict__ queue,
                                                     const uint32_t* __restrict__ n_dev, uint32_t n_host,
                                                     uint32_t* __restrict__ next, uint32_t* __restrict__ n_next,
                                                     uint32_t* __restrict__ interior, uint32_t* __restrict__ n_interior) {
    __shared__ uint32_t warp_alive[kBlock / 32], warp_int[kBlock / 32];
    __shared__ uint32_t base_alive, base_int;
    const uint32_t n = n_dev ? *n_dev : n_host;
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const uint32_t n_tiles = (n + kBlock - 1) / kBlock;
    for (uint32_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const uint32_t i = tile * kBlock + threadIdx.x;
        uint8_t f = 0;
        uint32_t p = 0;
        if (i < n) {
            f = wf.flag[i] & 3u;
            p = queue ? queue[i] : i;
        }
        const uint32_t m_alive = __ballot_sync(0xffffffffu, f != 0);
        const uint32_t m_int = __ballot_sync(0xffffffffu, f == 2);
        if (lane == 0) {
            warp_alive[warp] = __popc(m_alive);
            warp_int[warp] = __popc(m_int);
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            uint32_t sa = 0, si = 0;
            for (int w = 0; w < kBlock / 32; ++w) {
                uint32_t a = warp_alive[w], b = warp_int[w];
                warp_alive[w] = sa;
                warp_int[w] = si;
                sa += a;
                si += b;
            }
            base_alive = sa ? atomicAdd(n_next, sa) : 0u;
            base_int = si ? atomicAdd(n_interior, si) : 0u;
        }
        __syncthreads();
        const uint32_t below = (1u << lane) - 1u;
        if (f != 0) next[base_alive + warp_alive[warp] + __popc(m_alive & below)] = p;
        if (f == 2) interior[base_int + warp_int[warp] + __popc(m_int & below)] = p;
        __syncthreads();
    }
}

// Grouping variant for scenes of <= 16 primitives (the flat-walk scenes): survivors of a 1024-entry
// tile leave grouped by the primitive they just hit (= the primitive the next ray starts on), in
// stable order inside a group. Rays that start on the same primitive pass the same leaf boxes first
// (their own primitive's box always), so the warp-coherent walk and the material branches of the
// next shade run with fuller warps. Per warp __match_any_sync ranks lanes inside their key group;
// per-tile offsets come from a [warp][key] count table in shared memory.
constexpr int kGroupBlock = 1024;
constexpr int kGroupKeys = 16;

__global__ void __launch_bounds__(kGroupBlock) k_compact_grouped(Wavefront wf, const uint32_t* __restrict__ queue,
                                                                  const uint32_t* __restrict__ n_dev, uint32_t n_host,
                                                                  uint32_t* __restrict__ next, uint32_t* __restrict__ n_next,
                                                                  uint32_t* __restrict__ interior,
                                                                  uint32_t* __restrict__ n_interior) {
    constexpr int W = kGroupBlock / 32;
    __shared__ uint32_t cnt[2][W][kGroupKeys];   // [alive | interior][warp][key] -> exclusive offsets
    __shared__ uint32_t key_base[2][kGroupKeys];
    __shared__ uint32_t tile_base[2];
    const uint32_t n = n_dev ? *n_dev : n_host;
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const uint32_t n_tiles = (n + kGroupBlock - 1) / kGroupBlock;
    const uint32_t below = (1u << lane) - 1u;
    for (uint32_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        for (int t = threadIdx.x; t < 2 * W * kGroupKeys; t += kGroupBlock) (&cnt[0][0][0])[t] = 0u;
        __syncthreads();
        const uint32_t i = tile * kGroupBlock + threadIdx.x;
        uint32_t f = 0, p = 0;
        if (i < n) {
            f = wf.flag[i];
            p = queue ? queue[i] : i;
        }
        const bool alive = (f & 3u) != 0u, inner = (f & 3u) == 2u;
        const uint32_t key = f >> 2;
        // lanes with the same key (dead lanes share the sentinel 255)
        const uint32_t g_alive = __match_any_sync(0xffffffffu, alive ? key : 255u);
        const uint32_t g_inner = __match_any_sync(0xffffffffu, inner ? key : 255u);
        const uint32_t r_alive = __popc(g_alive & below), r_inner = __popc(g_inner & below);
        if (alive && r_alive == 0) cnt[0][warp][key] = __popc(g_alive);
        if (inner && r_inner == 0) cnt[1][warp][key] = __popc(g_inner);
        __syncthreads();
        if (threadIdx.x < 2 * kGroupKeys) {  // one thread per (queue, key): exclusive scan over the warps
            const int q = threadIdx.x / kGroupKeys, k = threadIdx.x % kGroupKeys;
            uint32_t off = 0;
            for (int w = 0; w < W; ++w) {
                const uint32_t c = cnt[q][w][k];
                cnt[q][w][k] = off;
                off += c;
            }
            key_base[q][k] = off;  // total of this key, turned into an offset below
        }
        __syncthreads();
        if (threadIdx.x < 2) {
            const int q = threadIdx.x;
            uint32_t off = 0;
            for (int k = 0; k < kGroupKeys; ++k) {
                const uint32_t c = key_base[q][k];
                key_base[q][k] = off;
                off += c;
            }
            tile_base[q] = off ? atomicAdd(q == 0 ? n_next : n_interior, off) : 0u;
        }
        __syncthreads();
        if (alive) next[tile_base[0] + key_base[0][key] + cnt[0][warp][key] + r_alive] = p;
        if (inner) interior[tile_base[1] + key_base[1][key] + cnt[1][warp][key] + r_inner] = p;
        __syncthreads();
    }
}

void launch_compact(const LaunchCfg& lc, Wavefront wf, const QueueSet& q, bool grouped) {
    uint64_t n = q.n_cur ? 0xffffffffull : q.n_cur_host;
    if (grouped) {
        uint64_t want = (n + kGroupBlock - 1) / kGroupBlock, cap = uint64_t(lc.sm_count) * 2;
        int grid = int(want < cap ? (want ? want : 1) : cap);
        k_compact_grouped<<<grid, kGroupBlock, 0, lc.stream>>>(wf, q.cur, q.n_cur, q.n_cur_host, q.next, q.n_next, q.interior,
                                                               q.n_interior);
    } else {
        k_compact<<<grid_for(lc, n, 8), kBlock, 0, lc.stream>>>(wf, q.cur, q.n_cur, q.n_cur_host, q.next, q.n_next, q.interior,
                                                                q.n_interior);
    }
    ++g_launches;
}

// --------------------------------------------------------------- shade_exit
// Second half of the transparent cases (tp/montecarlo.frag:150-152,163-165): the
// interior ray has been traced; take N,P of that hit (or keep the entry hit's
// N,P when nothing was hit: intersection_info writes nothing for shape -1) and
// leave through refract(D, -N, 1/refract_ind).
__global__ void __launch_bounds__(kBlock) k_shade_exit(DevScene S, DevParams prm, Wavefront wf,
                                                        const uint32_t* __restrict__ queue,
                                                        const uint32_t* __restrict__ n_dev) {
    const uint32_t n = *n_dev;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint32_t p = queue[i];
        const float4 ha = wf.hit_a[p];
        const int2 hb = wf.hit_b[p];
        const float4 d4 = wf.ray_d[p];
        const float4 n4 = wf.keep_n[p], p4 = wf.keep_p[p];
        f3 N = mk3(n4.x, n4.y, n4.z), P = mk3(p4.x, p4.y, p4.z);
        intersection_info(S, hb.x, hb.y, mk3(ha.x, ha.y, ha.z), prm.flat_face, N, P);
        f3 nO = P + N * MCPT_BIAS;
        f3 nD = refract3(mk3(d4.x, d4.y, d4.z), -N, 1.0f / prm.refract_ind);
        wf.ray_o[p] = make_float4(nO.x, nO.y, nO.z, 0.0f);
        wf.ray_d[p] = make_float4(nD.x, nD.y, nD.z, d4.w);  // w: the path's RNG draw count
    }
}

void launch_shade_exit(const LaunchCfg& lc, const DevScene& S, const DevParams& prm, Wavefront wf, const uint32_t* queue,
                       const uint32_t* n_dev) {
    k_shade_exit<<<grid_for(lc, 0xffffffffull, 8), kBlock, 0, lc.stream>>>(S, prm, wf, queue, n_dev);
    ++g_launches;
}

// --------------------------------------------------------------- accumulate
// glBlendFunc(GL_ONE, GL_ONE) into the RGB32F target, one draw per pass in pass
// order (montecarlo.cpp:450-466): accum = ((accum + r0) + r1) + ...
__global__ void __launch_bounds__(kBlock) k_accumulate(Wavefront wf, float4* __restrict__ accum, uint32_t npix,
                                                        int n_passes) {
    for (uint32_t pix = blockIdx.x * blockDim.x + threadIdx.x; pix < npix; pix += gridDim.x * blockDim.x) {
        float4 a = accum[pix];
        for (int k = 0; k < n_passes; ++k) {
            float4 r = wf.result[uint64_t(k) * npix + pix];
            a.x += r.x;
            a.y += r.y;
            a.z += r.z;
        }
        accum[pix] = a;
    }
}

void launch_accumulate(const LaunchCfg& lc, Wavefront wf, float4* accum, uint32_t npix, int n_passes) {
    k_accumulate<<<grid_for(lc, npix, 16), kBlock, 0, lc.stream>>>(wf, accum, npix, n_passes);
    ++g_launches;
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
    ++g_launches;
}
