// api_render.inl — viewer-facing half of the C ABI: resize / camera / params /
// clear / render / resolve / accumulation access / NCCL reduce / ray batches.
// Included by api.cu.

namespace {

// The queue-driven kernels of a batch run on a fixed grid of G CTAs, each owning one region of
// every queue (kernels.cu): G = 8 CTAs per SM, fewer when the batch has fewer 256-path tiles.
constexpr int kBlockPaths = 256;
int wavefront_grid_max(const mcpt_ctx* c) { return c->sm_count * 8; }
int wavefront_grid(const mcpt_ctx* c, uint64_t slots) {
    const uint64_t tiles = (slots + kBlockPaths - 1) / kBlockPaths;
    return int(std::max<uint64_t>(1, std::min<uint64_t>(tiles, uint64_t(wavefront_grid_max(c)))));
}
// region capacity: the most tiles any CTA is handed in the first bounce
uint32_t queue_cap(const mcpt_ctx* c, uint64_t slots) {
    const uint64_t tiles = (slots + kBlockPaths - 1) / kBlockPaths, g = uint64_t(wavefront_grid(c, slots));
    return uint32_t((tiles + g - 1) / g * kBlockPaths);
}
size_t queue_entries(const mcpt_ctx* c, uint64_t slots) { return size_t(slots) + size_t(kBlockPaths) * size_t(wavefront_grid_max(c)); }

// wavefront buffers for `slots` path slots
int ensure_wavefront(mcpt_ctx* c, uint64_t slots, int bounces) {
    if (slots > 0x7fffff00ull) return fail(c, MCPT_E_INVALID, "batch of %llu paths is too large", (unsigned long long)slots);
    if (slots > c->wf_capacity) {
        CK(c, cudaStreamSynchronize(c->stream));
        free_all(c, c->wf_allocs);
        c->wf_capacity = 0;
        c->nee_capacity = 0;
        c->counters = nullptr;
        c->counters_len = 0;
        Wavefront& w = c->wf;
        int rc;
        char* records = nullptr;
        if ((rc = dev_alloc(c, c->wf_allocs, &records, size_t(slots) * kPathRecordBytes)) ||
            (rc = dev_alloc(c, c->wf_allocs, &c->queue[0], queue_entries(c, slots))) ||
            (rc = dev_alloc(c, c->wf_allocs, &c->queue[1], queue_entries(c, slots))) ||
            (rc = dev_alloc(c, c->wf_allocs, &c->queue[2], queue_entries(c, slots)))) {
            free_all(c, c->wf_allocs);
            trim_cache(c);
            return rc;
        }
        // path tags live in attenu.w: start from a state no batch generation can match
        w.rec = records;
        w.plane = size_t(slots) * 32;
        CK(c, cudaMemsetAsync(w.rec, 0, size_t(slots) * kPathRecordBytes, c->stream));
        c->batch_gen = 0;
        c->wf_capacity = slots;
        trim_cache(c);
    }
    (void)bounces;
    if (c->opt_nee && c->nee_capacity < c->wf_capacity) {
        // next-event estimation: three more sectors per slot and the shadow queue (allocated on first use)
        CK(c, cudaStreamSynchronize(c->stream));
        char* rec = nullptr;
        int rc;
        if ((rc = dev_alloc(c, c->wf_allocs, &rec, size_t(c->wf_capacity) * 96)) ||
            (rc = dev_alloc(c, c->wf_allocs, &c->queue[3], queue_entries(c, c->wf_capacity))))
            return rc;
        c->nee.rec = rec;
        c->nee.plane = size_t(c->wf_capacity) * 32;
        c->nee_capacity = c->wf_capacity;
    }
    const int need = 4 * wavefront_grid_max(c);  // region lengths of the four queues
    if (need > c->counters_len) {
        int rc = dev_alloc(c, c->wf_allocs, &c->counters, size_t(need));
        if (rc) return rc;
        c->counters_len = need;
    }
    return MCPT_OK;
}

int passes_per_batch(const mcpt_ctx* c, uint32_t npix) {
    if (c->opt_passes_per_batch > 0) return c->opt_passes_per_batch;
    // ~64M paths in flight (9 GB of wavefront state out of 180 GB): late bounces keep enough rays per launch
    // to fill 148 SMs, and a step needs fewer launches (measured on config 2: 16M paths 5.33, 32M 5.58, 64M 5.73,
    // 128M 5.76 Gsamples/s)
    uint64_t target = 64ull << 20;
    uint64_t ppb = target / std::max<uint32_t>(npix, 1);
    return int(std::min<uint64_t>(std::max<uint64_t>(ppb, 1), 64));
}

// RAII bracket of one stage launch with two events when profiling is on
struct StageSpan {
    mcpt_ctx* c;
    int idx = -1;
    StageSpan(mcpt_ctx* ctx, int stage) : c(ctx) {
        if (!c->opt_profile) return;
        auto get = [&]() {
            cudaEvent_t e = nullptr;
            if (!c->free_events.empty()) {
                e = c->free_events.back();
                c->free_events.pop_back();
            } else {
                cudaEventCreate(&e);
            }
            return e;
        };
        mcpt_ctx::Span sp{stage, get(), get()};
        cudaEventRecord(sp.a, c->stream);
        c->spans.push_back(sp);
        idx = int(c->spans.size()) - 1;
    }
    void end() {
        if (idx >= 0) cudaEventRecord(c->spans[idx].b, c->stream);
        idx = -1;
    }
    ~StageSpan() { end(); }
};

// raytracer.vert:9-22 evaluated for the four strip vertices, in fp32
void camera_rays(const float iv[16], const float ipv[16], CameraRays& out) {
    // P4 = invV * vec4(0,0,0,1)
    float P[3];
    for (int r = 0; r < 3; ++r) P[r] = iv[0 + r] * 0.0f + iv[4 + r] * 0.0f + iv[8 + r] * 0.0f + iv[12 + r] * 1.0f;
    for (int r = 0; r < 3; ++r) out.ori[r] = P[r];
    for (int vid = 0; vid < 4; ++vid) {
        float tcx = float(vid % 2), tcy = float(vid / 2);
        float cx = 2.0f * tcx - 1.0f, cy = 2.0f * tcy - 1.0f;
        float Q[4];
        for (int r = 0; r < 4; ++r) Q[r] = ipv[0 + r] * cx + ipv[4 + r] * cy + ipv[8 + r] * 1.0f + ipv[12 + r] * 1.0f;
        float d[3];
        for (int r = 0; r < 3; ++r) d[r] = Q[r] / Q[3] - P[r];
        float inv = 1.0f / sqrtf(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
        for (int r = 0; r < 3; ++r) out.dir[vid][r] = d[r] * inv;
    }
}

}  // namespace

extern "C" {

int mcpt_resize(mcpt_ctx* c, int w, int h) {
    if (!c) return MCPT_E_INVALID;
    if (w <= 0 || h <= 0 || int64_t(w) * h > (int64_t(1) << 28)) return fail(c, MCPT_E_INVALID, "bad size %dx%d", w, h);
    CK(c, cudaSetDevice(c->device));
    CK(c, cudaStreamSynchronize(c->stream));
    if (c->accum_owned && c->accum) cudaFree(c->accum);
    c->accum = nullptr;
    c->accum_owned = false;
    c->accum_bytes = size_t(w) * h * sizeof(float4);
    void* p = nullptr;
    cudaError_t e = cudaMalloc(&p, c->accum_bytes);
    if (e != cudaSuccess) return fail(c, MCPT_E_NOMEM, "accumulation buffer: %s", cudaGetErrorString(e));
    c->accum = static_cast<float4*>(p);
    c->accum_owned = true;
    c->prm.width = w;
    c->prm.height = h;
    CK(c, cudaMemsetAsync(c->accum, 0, c->accum_bytes, c->stream));
    return MCPT_OK;
}

int mcpt_set_camera(mcpt_ctx* c, const float inv_v[16], const float inv_pv[16]) {
    if (!c || !inv_v || !inv_pv) return MCPT_E_INVALID;
    std::memcpy(c->prm.inv_v, inv_v, sizeof c->prm.inv_v);
    std::memcpy(c->prm.inv_pv, inv_pv, sizeof c->prm.inv_pv);
    camera_rays(inv_v, inv_pv, c->cam);
    c->have_camera = true;
    return MCPT_OK;
}

int mcpt_set_params(mcpt_ctx* c, const mcpt_params* p) {
    if (!c || !p) return MCPT_E_INVALID;
    if (p->nb_bounces < 0 || p->nb_bounces > 4096) return fail(c, MCPT_E_INVALID, "nb_bounces %d out of range", p->nb_bounces);
    if (p->mode < 0 || p->mode > 2) return fail(c, MCPT_E_INVALID, "mode %d out of range", p->mode);
    c->user_params = *p;
    c->prm.nb_bounces = p->nb_bounces;
    c->prm.refract_ind = p->refract_ind;
    c->prm.date = p->date;
    c->prm.flat_face = p->flat_face;
    c->prm.mode = p->mode;
    return MCPT_OK;
}

int mcpt_clear(mcpt_ctx* c) {
    if (!c) return MCPT_E_INVALID;
    if (!c->accum) return fail(c, MCPT_E_INVALID, "mcpt_clear before mcpt_resize");
    CK(c, cudaSetDevice(c->device));
    CK(c, cudaMemsetAsync(c->accum, 0, c->accum_bytes, c->stream));
    return MCPT_OK;
}

int mcpt_render(mcpt_ctx* c, int first_pass, int n_passes) {
    if (!c) return MCPT_E_INVALID;
    if (!c->have_scene) return fail(c, MCPT_E_INVALID, "mcpt_render before mcpt_upload_scene");
    if (!c->accum) return fail(c, MCPT_E_INVALID, "mcpt_render before mcpt_resize");
    if (!c->have_camera) return fail(c, MCPT_E_INVALID, "mcpt_render before mcpt_set_camera");
    if (n_passes < 0) return fail(c, MCPT_E_INVALID, "n_passes < 0");
    if (n_passes == 0) return MCPT_OK;
    // NB_BOUNCES = 0: random_path's loop body never runs and every path returns vec3(0)
    if (c->prm.mode == 0 && c->prm.nb_bounces == 0) return MCPT_OK;
    CK(c, cudaSetDevice(c->device));
    const uint32_t npix = uint32_t(c->prm.width) * uint32_t(c->prm.height);
    const int ppb = std::min(passes_per_batch(c, npix), n_passes);
    // mode 1/2 are single-traversal debug shaders; mode 0 loops NB_BOUNCES times
    const int iters = (c->prm.mode != 0) ? 1 : c->prm.nb_bounces;
    int rc = ensure_wavefront(c, uint64_t(npix) * ppb, iters);
    if (rc) return rc;
    const LaunchCfg lc = lcfg(c);
    const DevScene& S = c->scene;
    unsigned long long* work = c->opt_count_work ? c->dev_stats + 2 : nullptr;

    if (c->opt_fused_raygen) {
        // the camera direction of every pixel, once per call: the first traversal of every batch reads it (16 B per ray out
        // of the L2) instead of evaluating the vertex shader's interpolation per path
        if (c->cam_dirs_px < npix) {
            if (c->cam_dirs) cudaFree(c->cam_dirs);
            c->cam_dirs = nullptr;
            c->cam_dirs_px = 0;
            void* p = nullptr;
            if (cudaMalloc(&p, size_t(npix) * sizeof(float4)) != cudaSuccess) return fail(c, MCPT_E_NOMEM, "camera direction table");
            c->cam_dirs = static_cast<float4*>(p);
            c->cam_dirs_px = npix;
        }
        launch_camera_table(lc, c->prm, c->cam, c->cam_dirs, npix);
    }
    for (int b0 = 0; b0 < n_passes; b0 += ppb) {
        const int nb = std::min(ppb, n_passes - b0);
        const uint32_t slots = npix * uint32_t(nb);
        // a fresh generation per batch: stale tags of earlier batches never match (layout.cuh)
        if (++c->batch_gen >= 0x7ffffff0u) {
            CK(c, cudaMemsetAsync(c->wf.rec, 0, size_t(c->wf_capacity) * kPathRecordBytes, c->stream));
            c->batch_gen = 1;
        }
        c->prm.batch = c->batch_gen;
        // camera rays: by k_raygen, or (MCPT_OPT_FUSED_RAYGEN, default) by the first traversal itself -- it computes the ray of
        // every slot it takes and stores it, so the ray sectors are the same afterwards and no kernel only writes them
        const bool fused_raygen = c->opt_fused_raygen != 0;
        if (!fused_raygen) {
            StageSpan sp(c, MCPT_STAGE_RAYGEN);
            launch_raygen(lc, c->prm, c->cam, c->wf, first_pass + b0, nb, npix);
        }
        // queue[0]: paths entering the bounce, rewritten in place with the survivors by shade (it is
        // dead once traverse has run); queue[1]: rays that hit something; queue[2]: interior rays.
        // counters[q * G_max + b] = length of CTA b's region of queue q.
        LaunchCfg lq = lc;
        lq.grid = wavefront_grid(c, slots);
        const uint32_t cap = queue_cap(c, slots);
        const int gmax = wavefront_grid_max(c);
        const QueueOut q_next{c->queue[0], c->counters + 0 * gmax, cap}, q_hit{c->queue[1], c->counters + 1 * gmax, cap},
            q_int{c->queue[2], c->counters + 2 * gmax, cap}, q_none{nullptr, nullptr, 0};
        QueueIn in_identity{nullptr, nullptr, slots, cap};
        const bool mc = (c->prm.mode == 0);
        const bool nee = mc && c->opt_nee && S.n_lights > 0;
        // nothing reads a primary ray back except the fake sky (paths that were never shaded: k_accumulate then takes the
        // direction from the camera table) -- unless the shading uses libm (it rebuilds E from the origin), a debug shader or
        // next-event estimation runs: then every ray is stored
        if (fused_raygen) in_identity.gen = RayGen{c->cam, c->cam_dirs, c->prm.width, c->prm.height, npix, 1.0f / float(npix), (mc && !nee && c->opt_fast_math) ? 2 : 1};
        const QueueIn in_next{q_next.items, q_next.counts, 0, cap}, in_hit{q_hit.items, q_hit.counts, 0, cap},
            in_int{q_int.items, q_int.counts, 0, cap};
        const QueueOut q_shadow{c->queue[3], c->counters + 3 * gmax, cap};
        const QueueIn in_shadow{q_shadow.items, q_shadow.counts, 0, cap};
        if (nee) CK(c, cudaMemsetAsync(c->nee.rec + 2 * c->nee.plane, 0, size_t(slots) * 32, c->stream));  // sums and densities of this batch
        for (int it = 0; it < iters; ++it) {
            const bool last = (it == iters - 1);
            const QueueIn& cur = (it == 0) ? in_identity : in_next;
            {
                StageSpan sp(c, MCPT_STAGE_TRAVERSE);
                launch_traverse(lq, S, c->wf, cur, mc ? q_hit : q_none, nullptr, c->dev_stats + 0, work);
            }
            {
                StageSpan sp(c, MCPT_STAGE_SHADE);
                if (nee)
                    launch_shade_nee(lq, S, c->prm, c->wf, c->nee, in_hit, q_next, q_int, q_shadow, it == 0, last, first_pass + b0, npix,
                                     c->opt_fast_math != 0, c->dev_stats + 1);
                else
                    launch_shade(lq, S, c->prm, c->wf, mc ? in_hit : cur, q_next, q_int, it == 0, last, first_pass + b0, npix,
                                 c->opt_fast_math != 0, c->dev_stats + 1);
            }
            if (last) break;
            if (nee) {
                // the shadow rays of this bounce: the production traversal kernel over their own ray / hit sectors
                // (every record is delivered, as for the debug shaders), then the visible light samples are delivered
                StageSpan sp(c, MCPT_STAGE_TRAVERSE_INTERIOR);
                launch_traverse(lq, S, c->nee, in_shadow, q_none, nullptr, c->dev_stats + 0, work);
                launch_nee_resolve(lq, c->nee, in_shadow);
            }
            // interior rays: traced and turned into their exit rays by the same kernel, except in scenes
            // with meshes (measured: +4 % at C2, +7 % on the divergence-stress scene, but -5 % on the mesh
            // workload, where the exit's normal fetches stall the persistent warps)
            const bool fused_exit = S.n_meshes == 0;
            {
                StageSpan sp(c, MCPT_STAGE_TRAVERSE_INTERIOR);
                launch_traverse(lq, S, c->wf, in_int, q_none, fused_exit ? &c->prm : nullptr, c->dev_stats + 0, work);
            }
            if (!fused_exit) {
                StageSpan sp(c, MCPT_STAGE_SHADE_EXIT);
                launch_shade_exit(lq, S, c->prm, c->wf, in_int);
            }
        }
        {
            StageSpan sp(c, MCPT_STAGE_ACCUMULATE);
            if (nee)
                launch_accumulate_nee(lc, c->wf, c->nee, c->accum, npix, nb, c->prm.batch);
            else
                launch_accumulate(lc, c->wf, c->accum, npix, nb, c->prm.batch, in_identity.gen.on == 2 ? c->cam_dirs : nullptr);
        }
        c->paths_started += slots;
    }
    CK(c, cudaGetLastError());
    return MCPT_OK;
}

int mcpt_resolve(mcpt_ctx* c, float divisor, int subsampling, void* dst, int dst_kind) {
    if (!c || !dst) return MCPT_E_INVALID;
    if (!c->accum) return fail(c, MCPT_E_INVALID, "mcpt_resolve before mcpt_resize");
    if (subsampling < 1 || subsampling > 64) return fail(c, MCPT_E_INVALID, "subsampling %d out of range", subsampling);
    if (dst_kind < 0 || dst_kind > 3) return fail(c, MCPT_E_INVALID, "bad dst_kind %d", dst_kind);
    CK(c, cudaSetDevice(c->device));
    const bool rgba8 = (dst_kind == MCPT_DST_HOST_RGBA8 || dst_kind == MCPT_DST_DEVICE_RGBA8);
    const bool host = (dst_kind == MCPT_DST_HOST_RGB32F || dst_kind == MCPT_DST_HOST_RGBA8);
    const size_t n = size_t(c->prm.width) * subsampling * size_t(c->prm.height) * subsampling;
    const size_t bytes = n * (rgba8 ? 4 : 3 * sizeof(float));
    const LaunchCfg lc = lcfg(c);
    if (!host) {
        launch_resolve(lc, c->accum, c->prm.width, c->prm.height, divisor, subsampling, dst, rgba8);
        CK(c, cudaGetLastError());
        return MCPT_OK;
    }
    // staging buffer for host destinations: kept across calls (grow-only), no per-frame cudaMalloc
    if (bytes > c->resolve_tmp_bytes) {
        if (c->resolve_tmp) cudaFree(c->resolve_tmp);
        c->resolve_tmp = nullptr;
        c->resolve_tmp_bytes = 0;
        CK(c, cudaMalloc(&c->resolve_tmp, bytes));
        c->resolve_tmp_bytes = bytes;
    }
    void* tmp = c->resolve_tmp;
    launch_resolve(lc, c->accum, c->prm.width, c->prm.height, divisor, subsampling, tmp, rgba8);
    cudaError_t e = cudaMemcpyAsync(dst, tmp, bytes, cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    if (e != cudaSuccess) return fail(c, MCPT_E_CUDA, "resolve readback: %s", cudaGetErrorString(e));
    return MCPT_OK;
}

int mcpt_get_accum(mcpt_ctx* c, float** dev_ptr, size_t* bytes) {
    if (!c || !dev_ptr || !bytes) return MCPT_E_INVALID;
    if (!c->accum) return fail(c, MCPT_E_INVALID, "no accumulation buffer (call mcpt_resize)");
    *dev_ptr = reinterpret_cast<float*>(c->accum);
    *bytes = c->accum_bytes;
    return MCPT_OK;
}

int mcpt_bind_accum(mcpt_ctx* c, float* dev_ptr, size_t bytes) {
    if (!c || !dev_ptr) return MCPT_E_INVALID;
    if (c->prm.width <= 0) return fail(c, MCPT_E_INVALID, "mcpt_bind_accum before mcpt_resize");
    if (bytes < c->accum_bytes) return fail(c, MCPT_E_INVALID, "buffer of %zu bytes < %zu needed", bytes, c->accum_bytes);
    if ((reinterpret_cast<uintptr_t>(dev_ptr) & 15u) != 0) return fail(c, MCPT_E_INVALID, "buffer must be 16-byte aligned");
    CK(c, cudaSetDevice(c->device));
    CK(c, cudaStreamSynchronize(c->stream));
    if (c->accum_owned && c->accum) cudaFree(c->accum);
    c->accum = reinterpret_cast<float4*>(dev_ptr);
    c->accum_owned = false;
    return MCPT_OK;
}

int mcpt_read_accum(mcpt_ctx* c, float* host_rgba, size_t bytes) {
    if (!c || !host_rgba) return MCPT_E_INVALID;
    if (!c->accum || bytes < c->accum_bytes) return fail(c, MCPT_E_INVALID, "bad accumulation read");
    CK(c, cudaSetDevice(c->device));
    CK(c, cudaMemcpyAsync(host_rgba, c->accum, c->accum_bytes, cudaMemcpyDeviceToHost, c->stream));
    CK(c, cudaStreamSynchronize(c->stream));
    return MCPT_OK;
}

int mcpt_write_accum(mcpt_ctx* c, const float* host_rgba, size_t bytes) {
    if (!c || !host_rgba) return MCPT_E_INVALID;
    if (!c->accum || bytes < c->accum_bytes) return fail(c, MCPT_E_INVALID, "bad accumulation write");
    CK(c, cudaSetDevice(c->device));
    CK(c, cudaMemcpyAsync(c->accum, host_rgba, c->accum_bytes, cudaMemcpyHostToDevice, c->stream));
    CK(c, cudaStreamSynchronize(c->stream));
    return MCPT_OK;
}

int mcpt_reduce_nccl(mcpt_ctx* c, void* nccl_comm, int root) {
    if (!c || !nccl_comm) return MCPT_E_INVALID;
    if (!c->accum) return fail(c, MCPT_E_INVALID, "no accumulation buffer");
    // ncclResult_t ncclReduce(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, int, ncclComm_t, cudaStream_t)
    typedef int (*reduce_fn)(const void*, void*, size_t, int, int, int, void*, cudaStream_t);
    static reduce_fn fn = nullptr;
    if (!fn) {
        fn = reinterpret_cast<reduce_fn>(dlsym(RTLD_DEFAULT, "ncclReduce"));
        if (!fn) {
            void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
            if (h) fn = reinterpret_cast<reduce_fn>(dlsym(h, "ncclReduce"));
        }
        if (!fn) return fail(c, MCPT_E_NCCL, "ncclReduce not found (no libnccl loaded)");
    }
    CK(c, cudaSetDevice(c->device));
    const int nccl_float32 = 7, nccl_sum = 0;
    int r = fn(c->accum, c->accum, c->accum_bytes / sizeof(float), nccl_float32, nccl_sum, root, nccl_comm, c->stream);
    if (r != 0) return fail(c, MCPT_E_NCCL, "ncclReduce failed with %d", r);
    return MCPT_OK;
}

int mcpt_trace_rays_device(mcpt_ctx* c, int64_t n, const float* o4, const float* d4, mcpt_hit* out_dev, int any_hit) {
    if (!c || n < 0 || (n > 0 && (!o4 || !d4 || !out_dev))) return MCPT_E_INVALID;
    if (!c->have_scene) return fail(c, MCPT_E_INVALID, "mcpt_trace_rays before mcpt_upload_scene");
    if (n == 0) return MCPT_OK;
    CK(c, cudaSetDevice(c->device));
    const float4 *o = reinterpret_cast<const float4*>(o4), *d = reinterpret_cast<const float4*>(d4);
    // The one-ray-per-lane walk in the reference's visit order needs the per-mesh trees; a scene uploaded as ONE
    // world-space tree (MCPT_OPT_UNIFIED) has none, and every batch goes through the production kernel: any_hit then
    // reports the closest hit, which exists exactly when just_hit_bvh would find one.
    const bool unified = c->scene.u_nodes != nullptr;
    if ((any_hit || c->opt_trace_walker != 0) && !unified) {
        launch_trace(lcfg(c), c->scene, n, o, d, out_dev, any_hit != 0);
        CK(c, cudaGetLastError());
        return MCPT_OK;
    }
    // the production kernel over a private pair of sector planes (ray, hit)
    if (n > 0x7fffff00ll) return fail(c, MCPT_E_INVALID, "batch of %lld rays is too large", (long long)n);
    if (uint64_t(n) > c->trace_wf_slots) {
        CK(c, cudaStreamSynchronize(c->stream));
        if (c->trace_wf) cudaFree(c->trace_wf);
        c->trace_wf = nullptr;
        c->trace_wf_slots = 0;
        void* p = nullptr;
        cudaError_t e = cudaMalloc(&p, size_t(n) * 64);
        if (e != cudaSuccess) return fail(c, MCPT_E_NOMEM, "trace sectors: %s", cudaGetErrorString(e));
        c->trace_wf = static_cast<char*>(p);
        c->trace_wf_slots = uint64_t(n);
    }
    Wavefront wf{};
    wf.rec = c->trace_wf;
    wf.plane = size_t(c->trace_wf_slots) * 32;
    LaunchCfg lq = lcfg(c);
    lq.grid = wavefront_grid(c, uint64_t(n));
    launch_trace_wavefront(lq, c->scene, n, o, d, wf, queue_cap(c, uint64_t(n)), out_dev);
    CK(c, cudaGetLastError());
    return MCPT_OK;
}

int mcpt_trace_rays(mcpt_ctx* c, int64_t n, const float* o, const float* d, mcpt_hit* out, int any_hit) {
    if (!c || n < 0 || (n > 0 && (!o || !d || !out))) return MCPT_E_INVALID;
    if (!c->have_scene) return fail(c, MCPT_E_INVALID, "mcpt_trace_rays before mcpt_upload_scene");
    if (n == 0) return MCPT_OK;
    CK(c, cudaSetDevice(c->device));
    std::vector<float4> ho, hd;
    try {
        ho.resize(size_t(n));
        hd.resize(size_t(n));
    } catch (const std::exception&) {
        return fail(c, MCPT_E_NOMEM, "host staging of %lld rays", (long long)n);
    }
    for (int64_t i = 0; i < n; ++i) {
        ho[i] = make_float4(o[3 * i], o[3 * i + 1], o[3 * i + 2], 0.0f);
        hd[i] = make_float4(d[3 * i], d[3 * i + 1], d[3 * i + 2], 0.0f);
    }
    float4 *dev_o = nullptr, *dev_d = nullptr;
    mcpt_hit* dev_h = nullptr;
    int rc = MCPT_OK;
    cudaError_t e = cudaMalloc(&dev_o, n * sizeof(float4));
    if (e == cudaSuccess) e = cudaMalloc(&dev_d, n * sizeof(float4));
    if (e == cudaSuccess) e = cudaMalloc(&dev_h, n * sizeof(mcpt_hit));
    if (e == cudaSuccess) e = cudaMemcpyAsync(dev_o, ho.data(), n * sizeof(float4), cudaMemcpyHostToDevice, c->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(dev_d, hd.data(), n * sizeof(float4), cudaMemcpyHostToDevice, c->stream);
    if (e == cudaSuccess) rc = mcpt_trace_rays_device(c, n, &dev_o->x, &dev_d->x, dev_h, any_hit);
    if (e == cudaSuccess && rc == MCPT_OK) e = cudaMemcpyAsync(out, dev_h, n * sizeof(mcpt_hit), cudaMemcpyDeviceToHost, c->stream);
    cudaError_t e2 = cudaStreamSynchronize(c->stream);
    if (e == cudaSuccess) e = e2;
    cudaFree(dev_o);
    cudaFree(dev_d);
    cudaFree(dev_h);
    if (rc != MCPT_OK) return rc;
    if (e != cudaSuccess) return fail(c, MCPT_E_CUDA, "trace_rays: %s", cudaGetErrorString(e));
    return MCPT_OK;
}

int mcpt_sample_directions(mcpt_ctx* c, const float D[3], float roughness, float tcx, float tcy, int numero_pass, float date,
                           int64_t n, float* out_host) {
    if (!c || !D || n < 0 || (n > 0 && !out_host)) return MCPT_E_INVALID;
    if (n == 0) return MCPT_OK;
    if (n > (int64_t(1) << 30)) return fail(c, MCPT_E_INVALID, "too many draws: the RNG state advances modulo 2^32");
    CK(c, cudaSetDevice(c->device));
    float* dev = nullptr;
    cudaError_t e = cudaMalloc(&dev, size_t(n) * 3 * sizeof(float));
    if (e != cudaSuccess) return fail(c, MCPT_E_NOMEM, "sample buffer: %s", cudaGetErrorString(e));
    launch_sample_dirs(lcfg(c), D, roughness, tcx, tcy, numero_pass, date, n, dev, c->opt_fast_math != 0);
    e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemcpyAsync(out_host, dev, size_t(n) * 3 * sizeof(float), cudaMemcpyDeviceToHost, c->stream);
    const cudaError_t e2 = cudaStreamSynchronize(c->stream);
    cudaFree(dev);
    if (e == cudaSuccess) e = e2;
    if (e != cudaSuccess) return fail(c, MCPT_E_CUDA, "sample_directions: %s", cudaGetErrorString(e));
    return MCPT_OK;
}

int mcpt_get_stats(mcpt_ctx* c, mcpt_stats* out) {
    if (!c || !out) return MCPT_E_INVALID;
    CK(c, cudaSetDevice(c->device));
    unsigned long long h[kStatWords] = {0, 0, 0, 0, 0};
    CK(c, cudaMemcpyAsync(h, c->dev_stats, sizeof h, cudaMemcpyDeviceToHost, c->stream));
    CK(c, cudaStreamSynchronize(c->stream));
    out->paths = c->paths_started;
    out->rays = h[0];
    out->segments = h[1];
    out->launches = c->launches;
    out->boxes = h[2];
    out->prim_tests = h[3];
    out->tri_tests = h[4];
    return MCPT_OK;
}

int mcpt_get_stage_times(mcpt_ctx* c, mcpt_stage_times* out) {
    if (!c || !out) return MCPT_E_INVALID;
    CK(c, cudaSetDevice(c->device));
    CK(c, cudaStreamSynchronize(c->stream));
    std::memset(out, 0, sizeof *out);
    const bool dump = std::getenv("MCPT_DUMP_SPANS") != nullptr;  // diagnostic: every launch's stage and time on stderr
    for (auto& sp : c->spans) {
        float ms = 0.0f;
        if (cudaEventElapsedTime(&ms, sp.a, sp.b) == cudaSuccess) out->ms[sp.stage] += ms;
        if (dump) std::fprintf(stderr, "span %d %.4f\n", sp.stage, ms);
        out->launches[sp.stage] += 1;
        c->free_events.push_back(sp.a);
        c->free_events.push_back(sp.b);
    }
    c->spans.clear();
    return MCPT_OK;
}

int mcpt_reset_stats(mcpt_ctx* c) {
    if (!c) return MCPT_E_INVALID;
    CK(c, cudaSetDevice(c->device));
    CK(c, cudaMemsetAsync(c->dev_stats, 0, kStatWords * sizeof(unsigned long long), c->stream));
    c->paths_started = 0;
    c->launches = 0;
    return MCPT_OK;
}

}  // extern "C"
