// api_scene.inl — mcpt_upload_scene: validate the reference-layout tables and
// re-pack them into the device layout of layout.cuh. Included by api.cu.

namespace {

template <class T>
int upload_vec(mcpt_ctx* c, const std::vector<T>& h, T** out) {
    int rc = dev_alloc(c, c->scene_allocs, out, h.size());
    if (rc) return rc;
    if (!h.empty()) CK(c, cudaMemcpyAsync(*out, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice, c->stream));
    return MCPT_OK;
}

template <class T>
int upload_raw(mcpt_ctx* c, const T* h, size_t count, T** out) {
    int rc = dev_alloc(c, c->scene_allocs, out, count);
    if (rc) return rc;
    if (count) CK(c, cudaMemcpyAsync(*out, h, count * sizeof(T), cudaMemcpyHostToDevice, c->stream));
    return MCPT_OK;
}

int validate_bvh(mcpt_ctx* c, const mcpt_bvh_desc& b, int n_items, const char* what) {
    if (!b.bb || !b.ind) return fail(c, MCPT_E_INVALID, "%s: null BVH arrays", what);
    if (b.depth < 1 || b.depth > 29) return fail(c, MCPT_E_INVALID, "%s: BVH depth %d out of range [1,29]", what, b.depth);
    const int64_t n_leaves = int64_t(1) << b.depth;
    for (int64_t s = 0; s < n_leaves; ++s)
        if (b.ind[s] < -1 || b.ind[s] >= n_items)
            return fail(c, MCPT_E_INVALID, "%s: leaf slot %lld holds index %d (have %d items)", what, (long long)s, b.ind[s],
                        n_items);
    return MCPT_OK;
}

}  // namespace

extern "C" namespace {

// ---- host-side builders of the persistent walk's top-level tree
struct TopItem {
    float lo[3], hi[3];
    int prim;
};

inline void set_child2(DevNode2& N, int j, const DevNode& b, int tok) {
    float4& lo = j ? N.lo1 : N.lo0;
    float4& hi = j ? N.hi1 : N.hi0;
    lo.x = b.lo.x; lo.y = b.lo.y; lo.z = b.lo.z;
    hi.x = b.hi.x; hi.y = b.hi.y; hi.z = b.hi.z;
    std::memcpy(j ? &N.hi0.w : &N.lo0.w, &tok, sizeof(int));  // tokens live in lo0.w / hi0.w
    if (j) lo.w = hi.w = 0.0f;
}

inline DevNode box_of(const std::vector<TopItem>& it, int begin, int end) {
    DevNode b = empty_node();
    for (int i = begin; i < end; ++i) {
        b.lo.x = std::min(b.lo.x, it[i].lo[0]); b.lo.y = std::min(b.lo.y, it[i].lo[1]); b.lo.z = std::min(b.lo.z, it[i].lo[2]);
        b.hi.x = std::max(b.hi.x, it[i].hi[0]); b.hi.y = std::max(b.hi.y, it[i].hi[1]); b.hi.z = std::max(b.hi.z, it[i].hi[2]);
    }
    return b;
}

inline double half_area(const float lo[3], const float hi[3]) {
    const double x = double(hi[0]) - lo[0], y = double(hi[1]) - lo[1], z = double(hi[2]) - lo[2];
    return (x < 0 || y < 0 || z < 0) ? 0.0 : x * y + y * z + z * x;
}

// world box of the local box [lo, hi] under the primitive's forward transform, inflated like every other box
inline DevNode world_box(const Aff& fwd, const float lo[3], const float hi[3]) {
    mcpt_bb b{{FLT_MAX, FLT_MAX, FLT_MAX}, {-FLT_MAX, -FLT_MAX, -FLT_MAX}};
    for (int k = 0; k < 8; ++k) {
        const float x = (k & 1) ? hi[0] : lo[0], y = (k & 2) ? hi[1] : lo[1], z = (k & 4) ? hi[2] : lo[2];
        const float w[3] = {fwd.r0.x * x + fwd.r0.y * y + fwd.r0.z * z + fwd.r0.w, fwd.r1.x * x + fwd.r1.y * y + fwd.r1.z * z + fwd.r1.w,
                            fwd.r2.x * x + fwd.r2.y * y + fwd.r2.z * z + fwd.r2.w};
        for (int a = 0; a < 3; ++a) {
            b.min[a] = std::min(b.min[a], w[a]);
            b.max[a] = std::max(b.max[a], w[a]);
        }
    }
    return pack_node(b, -1);
}

// A rotated, elongated primitive has a loose axis-aligned box (a 80 x 10 beam at 45 degrees: 6x the area of
// the beam). Such a primitive enters the tree as kTopSlices boxes instead, the world boxes of slabs of its
// bounding cube (the reference's own bound, unit cube +-1.005, bvh_gpu/scene.cpp:18-42) cut along its longest
// axis: their union still contains the primitive, every slab is a leaf of its own with the same primitive
// token (a ray through two slabs offers the same candidate twice, which changes nothing). Only for
// primitives whose intersector demands a > EPSILON (sphere, cube, cylinder): the quad and the cone accept
// hits behind the origin when the origin is inside the reference's leaf box, which a sliced bound could miss.
constexpr int kTopSlices = 8;
bool slice_loose_primitive(const DevPrim& P, const DevNode& full, int prim, std::vector<TopItem>& items) {
    if (P.type != MCPT_PRIM_SPHERE && P.type != MCPT_PRIM_CUBE && P.type != MCPT_PRIM_CYLINDER) return false;
    const float e = 1.005f;
    const float col[3] = {P.fwd.r0.x * P.fwd.r0.x + P.fwd.r1.x * P.fwd.r1.x + P.fwd.r2.x * P.fwd.r2.x,
                          P.fwd.r0.y * P.fwd.r0.y + P.fwd.r1.y * P.fwd.r1.y + P.fwd.r2.y * P.fwd.r2.y,
                          P.fwd.r0.z * P.fwd.r0.z + P.fwd.r1.z * P.fwd.r1.z + P.fwd.r2.z * P.fwd.r2.z};
    const int ax = (col[0] >= col[1] && col[0] >= col[2]) ? 0 : (col[1] >= col[2] ? 1 : 2);
    DevNode part[kTopSlices];
    double sum = 0.0;
    for (int k = 0; k < kTopSlices; ++k) {
        float lo[3] = {-e, -e, -e}, hi[3] = {e, e, e};
        lo[ax] = -e + 2.0f * e * float(k) / float(kTopSlices);
        hi[ax] = -e + 2.0f * e * float(k + 1) / float(kTopSlices);
        part[k] = world_box(P.fwd, lo, hi);
        // never larger than the reference's own leaf box
        part[k].lo.x = std::max(part[k].lo.x, full.lo.x); part[k].lo.y = std::max(part[k].lo.y, full.lo.y); part[k].lo.z = std::max(part[k].lo.z, full.lo.z);
        part[k].hi.x = std::min(part[k].hi.x, full.hi.x); part[k].hi.y = std::min(part[k].hi.y, full.hi.y); part[k].hi.z = std::min(part[k].hi.z, full.hi.z);
        const float l3[3] = {part[k].lo.x, part[k].lo.y, part[k].lo.z}, h3[3] = {part[k].hi.x, part[k].hi.y, part[k].hi.z};
        sum += half_area(l3, h3);
    }
    const float fl[3] = {full.lo.x, full.lo.y, full.lo.z}, fh[3] = {full.hi.x, full.hi.y, full.hi.z};
    if (!(sum < 0.7 * half_area(fl, fh))) return false;  // not loose: one box
    for (int k = 0; k < kTopSlices; ++k)
        items.push_back(TopItem{{part[k].lo.x, part[k].lo.y, part[k].lo.z}, {part[k].hi.x, part[k].hi.y, part[k].hi.z}, prim});
    return true;
}

// Full-sweep SAH over the items' boxes, one item per leaf. The items are sorted ONCE per axis by box centre (ties in
// item order); a node owns the same range [begin, end) of the three orders, sweeps each for the cheapest split
// (cost = area(left) * n_left + area(right) * n_right, first axis and first position win ties), and hands its
// children their sub-ranges by a stable partition of the other two orders: O(n log n) in all (the first version
// re-sorted every node's range four times: 11 ms for the 3163 leaves of repo scene 8, 0.9 ms now). Below `depth_left`
// levels of head room the split is the median of the longest axis, which bounds the depth (the walk's stack holds one
// entry per level).
struct TopSahBuilder {
    const std::vector<TopItem>& it;
    std::vector<DevNode2>& nodes;
    std::vector<int> ord[3], tmp;
    std::vector<char> side;
    std::vector<double> right;

    TopSahBuilder(const std::vector<TopItem>& items, std::vector<DevNode2>& out) : it(items), nodes(out) {
        const int n = int(it.size());
        for (int ax = 0; ax < 3; ++ax) {
            ord[ax].resize(n);
            for (int i = 0; i < n; ++i) ord[ax][i] = i;
            std::stable_sort(ord[ax].begin(), ord[ax].end(), [&](int a, int b) { return it[a].lo[ax] + it[a].hi[ax] < it[b].lo[ax] + it[b].hi[ax]; });
        }
        tmp.resize(n);
        side.assign(n, 0);
        right.assign(n + 1, 0.0);
    }

    DevNode box_of_range(int ax, int begin, int end) const {
        DevNode b = empty_node();
        for (int i = begin; i < end; ++i) {
            const TopItem& t = it[ord[ax][i]];
            b.lo.x = std::min(b.lo.x, t.lo[0]); b.lo.y = std::min(b.lo.y, t.lo[1]); b.lo.z = std::min(b.lo.z, t.lo[2]);
            b.hi.x = std::max(b.hi.x, t.hi[0]); b.hi.y = std::max(b.hi.y, t.hi[1]); b.hi.z = std::max(b.hi.z, t.hi[2]);
        }
        return b;
    }

    int build(int begin, int end, int depth_left) {
        const int n = end - begin;
        if (n == 1) return ~it[ord[0][begin]].prim;
        const int idx = int(nodes.size());
        nodes.push_back(DevNode2{});
        int best_axis = -1, best_split = begin + n / 2;
        int balanced_levels = 0;
        while ((1 << balanced_levels) < n) ++balanced_levels;
        if (depth_left > balanced_levels) {
            double best = 1e300;
            for (int ax = 0; ax < 3; ++ax) {
                const int* o = ord[ax].data() + begin;
                float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
                for (int i = n - 1; i >= 1; --i) {
                    for (int k = 0; k < 3; ++k) {
                        lo[k] = std::min(lo[k], it[o[i]].lo[k]);
                        hi[k] = std::max(hi[k], it[o[i]].hi[k]);
                    }
                    right[i] = half_area(lo, hi);
                }
                float llo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, lhi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
                for (int i = 1; i < n; ++i) {
                    for (int k = 0; k < 3; ++k) {
                        llo[k] = std::min(llo[k], it[o[i - 1]].lo[k]);
                        lhi[k] = std::max(lhi[k], it[o[i - 1]].hi[k]);
                    }
                    const double cost = half_area(llo, lhi) * i + right[i] * (n - i);
                    if (cost < best) {
                        best = cost;
                        best_axis = ax;
                        best_split = begin + i;
                    }
                }
            }
        }
        if (best_axis < 0) {  // no head room left (or degenerate boxes): median of the longest axis
            const DevNode b = box_of_range(0, begin, end);
            const float ex = b.hi.x - b.lo.x, ey = b.hi.y - b.lo.y, ez = b.hi.z - b.lo.z;
            best_axis = (ex >= ey && ex >= ez) ? 0 : (ey >= ez ? 1 : 2);
            best_split = begin + n / 2;
        }
        for (int i = begin; i < end; ++i) side[ord[best_axis][i]] = (i < best_split) ? 0 : 1;
        for (int ax = 0; ax < 3; ++ax) {
            if (ax == best_axis) continue;
            int l = begin, r = best_split;
            for (int i = begin; i < end; ++i) tmp[(side[ord[ax][i]] ? r : l)++] = ord[ax][i];
            std::copy(tmp.begin() + begin, tmp.begin() + end, ord[ax].begin() + begin);
        }
        const DevNode b0 = box_of_range(0, begin, best_split), b1 = box_of_range(0, best_split, end);
        const int t0 = build(begin, best_split, depth_left - 1);
        const int t1 = build(best_split, end, depth_left - 1);
        DevNode2 N{};
        set_child2(N, 0, b0, t0);
        set_child2(N, 1, b1, t1);
        nodes[idx] = N;
        return idx;
    }
};

void build_top_sah(std::vector<TopItem>& items, std::vector<DevNode2>& nodes) {
    nodes.clear();
    if (items.size() == 1) {  // a single primitive: the root holds that leaf and an empty child
        DevNode2 N{};
        set_child2(N, 0, box_of(items, 0, 1), ~items[0].prim);
        set_child2(N, 1, empty_node(), INT_MIN);
        nodes.push_back(N);
        return;
    }
    nodes.reserve(items.size());
    TopSahBuilder b(items, nodes);
    b.build(0, int(items.size()), 30);  // <= 30 levels: one stack entry per level (kWalkStack)
}

// fold every second level of a two-children tree (root = node 0) into four-children nodes
void fold_top4(const std::vector<DevNode2>& n2, std::vector<DevNode4>& n4) {
    for (DevNode4& N : n4) {
        std::memset(&N, 0, sizeof N);
        N.tok = make_int4(INT_MIN, INT_MIN, INT_MIN, INT_MIN);
    }
    if (n2.empty()) return;
    auto child = [&](const DevNode2& N, int j, DevNode& b, int& tok) {
        const float4& lo = j ? N.lo1 : N.lo0;
        const float4& hi = j ? N.hi1 : N.hi0;
        b.lo = lo;
        b.hi = hi;
        std::memcpy(&tok, j ? &N.hi0.w : &N.lo0.w, sizeof(int));
    };
    std::vector<int> todo{0};
    while (!todo.empty()) {
        const int k = todo.back();
        todo.pop_back();
        DevNode4& W = n4[k];
        int m = 0;
        auto put = [&](const DevNode& b, int tok) {
            (&W.lox.x)[m] = b.lo.x; (&W.loy.x)[m] = b.lo.y; (&W.loz.x)[m] = b.lo.z;
            (&W.hix.x)[m] = b.hi.x; (&W.hiy.x)[m] = b.hi.y; (&W.hiz.x)[m] = b.hi.z;
            (&W.tok.x)[m] = tok;
            ++m;
            if (tok >= 0) todo.push_back(tok);
        };
        for (int j = 0; j < 2; ++j) {
            DevNode b;
            int tok;
            child(n2[k], j, b, tok);
            if (tok == INT_MIN) continue;
            if (tok >= 0) {
                for (int jj = 0; jj < 2; ++jj) {
                    DevNode bb;
                    int t2;
                    child(n2[tok], jj, bb, t2);
                    if (t2 != INT_MIN) put(bb, t2);
                }
            } else {
                put(b, tok);
            }
        }
    }
}

// The two trees of the persistent walk from the uploaded tables: `prims` with type, fwd and slot filled in, `top`
// the reference's heap-order tree with inflated boxes (index 1-based, leaf level = primitive ids).
void build_walk_trees(const mcpt_scene_desc* d, const std::vector<DevPrim>& prims, const std::vector<DevNode>& top, bool top_sah,
                      std::vector<DevNode2>& top2, std::vector<DevNode4>& top4) {
    const int td = d->root.depth;
    const uint32_t t_first_leaf = 1u << td;
    // ---- the top-level tree of the persistent walk. Hit records do not depend on the tree (candidates are
    // ranked by the reference's exact distance and visit order, intersect.cuh), so the persistent warps walk a
    // tree of their own instead of the reference's median split with its cycling axis (bvh_gpu/bvh.cpp:34-93):
    // a full-sweep SAH build over the primitives' (inflated) leaf boxes, one primitive per leaf, in the
    // two-children node format of the mesh trees (top2, root = node 0) and folded into four-children nodes
    // (top4). Measured on repo scene 8 (895 primitives): 69 -> see profiles/README.md box tests per ray.
    // MCPT_OPT_TOP_SAH = 0 keeps the reference's topology.
    top2.clear();
    {
        std::vector<TopItem> items;
        for (uint32_t sl = 0; sl < t_first_leaf; ++sl) {
            const int prim = d->root.ind[sl];
            if (prim < 0 || prims[prim].slot != int(sl)) continue;  // each primitive once, at its first slot
            const DevNode& nb = top[t_first_leaf + sl];
            if (top_sah && slice_loose_primitive(prims[prim], nb, prim, items)) continue;
            items.push_back(TopItem{{nb.lo.x, nb.lo.y, nb.lo.z}, {nb.hi.x, nb.hi.y, nb.hi.z}, prim});
        }
        if (top_sah && !items.empty()) {
            build_top_sah(items, top2);
        } else {
            // the reference's own topology: node k-1 holds the boxes of heap children 2k and 2k+1
            top2.resize(std::max<uint32_t>(t_first_leaf - 1, 1u));
            if (td == 0) {
                const int prim = d->root.ind[0];
                set_child2(top2[0], 0, top[1], prim >= 0 ? ~prim : INT_MIN);
                set_child2(top2[0], 1, empty_node(), INT_MIN);
            }
            for (uint32_t k = 1; k < t_first_leaf; ++k)
                for (int j = 0; j < 2; ++j) {
                    const uint32_t ch = 2 * k + uint32_t(j);
                    int tok;
                    if (ch < t_first_leaf)
                        tok = int(ch - 1);
                    else {
                        const int prim = d->root.ind[ch - t_first_leaf];
                        tok = prim >= 0 ? ~prim : INT_MIN;  // kTokEmpty
                    }
                    set_child2(top2[k - 1], j, top[ch], tok);
                }
        }
    }
    // ---- and with every second level folded away (DevNode4, scenes with meshes): a node at even depth
    // holds its grandchildren (or a child, where that child is a leaf); stored at the index of the
    // two-children node it replaces, the entries of odd depths stay unused
    top4.assign(top2.size(), DevNode4{});
    fold_top4(top2, top4);
}

}  // namespace

static int upload_scene_impl(mcpt_ctx* c, const mcpt_scene_desc* d);

// the C ABI never throws: host allocation failures of the re-pack become MCPT_E_NOMEM
int mcpt_upload_scene(mcpt_ctx* c, const mcpt_scene_desc* d) {
    if (!c || !d) return MCPT_E_INVALID;
    try {
        return upload_scene_impl(c, d);
    } catch (const std::bad_alloc&) {
        c->have_scene = false;
        return fail(c, MCPT_E_NOMEM, "host allocation failed while re-packing the scene");
    } catch (const std::exception& e) {
        c->have_scene = false;
        return fail(c, MCPT_E_INVALID, "scene upload: %s", e.what());
    }
}

static int upload_scene_impl(mcpt_ctx* c, const mcpt_scene_desc* d) {
    CK(c, cudaSetDevice(c->device));
    // diagnostic: MCPT_DUMP_UPLOAD=1 prints the host-side timeline of the call on stderr
    const bool dump = std::getenv("MCPT_DUMP_UPLOAD") != nullptr;
    const auto t_start = std::chrono::steady_clock::now();
    auto mark = [&](const char* what) {
        if (dump) std::fprintf(stderr, "upload %-28s %8.3f ms\n", what, std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_start).count());
    };
    if (d->n_prims < 0 || (d->n_prims > 0 && !d->prim_data)) return fail(c, MCPT_E_INVALID, "bad primitive table");
    if (d->n_meshes < 0 || (d->n_meshes > 0 && !d->meshes)) return fail(c, MCPT_E_INVALID, "bad mesh table");
    int rc = validate_bvh(c, d->root, d->n_prims, "root");
    if (rc) return rc;
    for (int m = 0; m < d->n_meshes; ++m) {
        const mcpt_mesh_desc& md = d->meshes[m];
        if (!md.tri_idx || !md.pos || !md.nrm) return fail(c, MCPT_E_INVALID, "mesh %d: null arrays", m);
        const bool no_tree = !md.bvh.bb && !md.bvh.ind;
        if (no_tree && !c->opt_mesh_bvh) return fail(c, MCPT_E_INVALID, "mesh %d: no BVH uploaded; that needs MCPT_OPT_MESH_BVH = 1", m);
        if (!no_tree) {
            // depth and pointers here; the leaf slots and the vertex indices are range-checked on the device below
            if (!md.bvh.bb || !md.bvh.ind) return fail(c, MCPT_E_INVALID, "mesh %d: null BVH arrays", m);
            if (md.bvh.depth < 1 || md.bvh.depth > 29) return fail(c, MCPT_E_INVALID, "mesh %d: BVH depth %d out of range [1,29]", m, md.bvh.depth);
            if (!c->opt_mesh_bvh && (rc = validate_bvh(c, md.bvh, int(md.n_tris), "mesh"))) return rc;
        }
        if (md.n_tris > 0x7fffff00u / 3u) return fail(c, MCPT_E_INVALID, "mesh %d: too many triangles", m);
    }

    mark("validated");
    CK(c, cudaStreamSynchronize(c->stream));
    free_all(c, c->scene_allocs);
    c->have_scene = false;
    DevScene S{};
    mark("previous scene released");

    // ---- primitive records + shading tables
    std::vector<DevPrim> prims(d->n_prims);
    std::vector<float4> color(d->n_prims), mat(d->n_prims);
    for (int i = 0; i < d->n_prims; ++i) {
        const mcpt_prim_data& pd = d->prim_data[i];
        DevPrim& p = prims[i];
        std::memset(&p, 0, sizeof p);
        p.type = int(pd.type[0]);  // int(texel.x), raytracer_func.frag:171-178
        p.mesh = -1;
        p.slot = -1;
        p.inv = rows_of(pd.inv_transfo);
        if (p.type == MCPT_PRIM_MESH) {
            p.fwd = rows_of(pd.mesh_transfo);
            const int line = int(pd.type[1]);
            for (int m = 0; m < d->n_meshes; ++m)
                if (d->meshes[m].mesh_line == line) p.mesh = m;
            if (p.mesh < 0) return fail(c, MCPT_E_INVALID, "primitive %d: no mesh with mesh_line %d", i, line);
        } else {
            p.fwd = rows_of(pd.transfo);
            if (p.type > MCPT_PRIM_ORIENTED_QUAD) p.type = -1;  // the shader's switch has no such case
        }
        {
            const float cx = std::sqrt(p.fwd.r0.x * p.fwd.r0.x + p.fwd.r1.x * p.fwd.r1.x + p.fwd.r2.x * p.fwd.r2.x);
            const float cy = std::sqrt(p.fwd.r0.y * p.fwd.r0.y + p.fwd.r1.y * p.fwd.r1.y + p.fwd.r2.y * p.fwd.r2.y);
            const float cz = std::sqrt(p.fwd.r0.z * p.fwd.r0.z + p.fwd.r1.z * p.fwd.r1.z + p.fwd.r2.z * p.fwd.r2.z);
            const float smin = std::min(cx, std::min(cy, cz));
            p.inv_smin = (smin > 1e-20f && std::isfinite(smin)) ? 1.0f / smin : 1e30f;  // degenerate: always both roots
        }
        color[i] = make_float4(pd.color[0], pd.color[1], pd.color[2], pd.color[3]);
        mat[i] = make_float4(pd.mat_info[0], pd.mat_info[1], pd.mat_info[2], pd.mat_info[3]);
    }

    // ---- top-level tree: 1-based heap array, leaf level carries the primitive id
    const int td = d->root.depth;
    const uint32_t t_first_leaf = 1u << td, t_nodes = 2u << td;  // indices 1 .. t_nodes-1
    std::vector<DevNode> top(t_nodes, empty_node());
    std::vector<float> leaf_bb(6 * size_t(std::max(d->n_prims, 1)), 0.0f);
    for (uint32_t k = 1; k < t_nodes; ++k) {
        int aux = -1;
        bool empty = false;
        if (k >= t_first_leaf) {
            aux = d->root.ind[k - t_first_leaf];
            empty = aux < 0;
            if (!empty && prims[aux].slot < 0) {
                prims[aux].slot = int(k - t_first_leaf);
                std::memcpy(&leaf_bb[6 * size_t(aux)], &d->root.bb[k - 1], 6 * sizeof(float));
            }
        }
        if (!empty) top[k] = pack_node(d->root.bb[k - 1], aux);
    }

    // ---- the trees of the persistent walk (build_walk_trees above)
    std::vector<DevNode2> top2;
    std::vector<DevNode4> top4;
    build_walk_trees(d, prims, top, c->opt_top_sah != 0, top2, top4);

    // ---- small scenes: flat list of leaf boxes for the warp-coherent walk
    std::vector<DevNode> flat;
    if (c->opt_flat_walk && d->n_prims > 0 && d->n_prims <= 16 && d->n_meshes == 0)
        for (uint32_t k = t_nodes - 1; k >= t_first_leaf; --k)  // the reference's visit order: high slots first
            if (d->root.ind[k - t_first_leaf] >= 0) flat.push_back(top[k]);

    mark("top-level trees built (host)");
    // ---- meshes. Nothing here loops over a mesh on the host: the arrays go to the device as they are, the vertex
    // indices and the reference tree's leaf slots are range-checked there (flag read back once at the end), the
    // tie-break ranks are scattered from the leaf slots there, and the tree builds reduce their own vertex boxes.
    std::vector<DevMesh> meshes(d->n_meshes);
    std::vector<void*> build_allocs;  // scratch of the GPU tree builds, parked again at the end
    struct ParkOnExit {               // ... also on the error returns below
        mcpt_ctx* c;
        std::vector<void*>& v;
        ~ParkOnExit() {
            if (v.empty()) return;
            for (cudaStream_t s : c->side)
                if (s) cudaStreamSynchronize(s);
            cudaStreamSynchronize(c->stream);
            free_all(c, v);
        }
    } park_build_allocs{c, build_allocs};
    constexpr int kSide = mcpt_ctx::kSideStreams;
    char* build_ws = nullptr;          // one workspace per side stream, back to back
    size_t build_ws_bytes = 0;
    int* d_flag = nullptr;
    if (d->n_meshes > 0) {
        if ((rc = dev_alloc(c, build_allocs, &d_flag, 1))) return rc;
        CK(c, cudaMemsetAsync(d_flag, 0, sizeof(int), c->stream));
    }
    if (c->opt_mesh_bvh) {
        for (int m = 0; m < d->n_meshes; ++m) {
            const int nt = int(d->meshes[m].n_tris);
            if (nt > 0)
                build_ws_bytes = std::max(build_ws_bytes, lbvh_workspace_bytes(nt, (nt + kMeshLeafTris - 1) / kMeshLeafTris, int(d->meshes[m].n_verts)));
        }
        build_ws_bytes = (build_ws_bytes + 255) & ~size_t(255);
        if (build_ws_bytes && (rc = dev_alloc(c, build_allocs, &build_ws, build_ws_bytes * kSide))) return rc;
    }
    // one world-space tree over the top level and every mesh instance (traverse.cuh, UWalk)?
    // ... also for large trees of analytic primitives (more than kUnifiedMinLeaves leaves): the four-children walk with
    // its lean lane state and a refill at 12 idle lanes beats the two-children walk that waits for the whole warp there
    // (repo scenes 3 / 7 / 8, 410 / 491 / 895 primitives: 4175 -> 4320, 1777 -> 2047, 606 -> 691 Msamples/s), while small
    // trees lose (122 primitives: scene 5 6662 -> 5730, divergence stress 3626 -> 3363) and keep the two-children walk
    constexpr size_t kUnifiedMinLeaves = 256;
    size_t top_leaves = 0;
    for (const DevNode4& N : top4)
        for (int j = 0; j < 4; ++j) top_leaves += ((&N.tok.x)[j] < 0 && (&N.tok.x)[j] != INT_MIN) ? 1 : 0;
    const bool unified = c->opt_unified && c->opt_mesh_bvh && c->opt_persistent && !c->opt_trace_walker &&
                         (d->n_meshes > 0 || top_leaves > kUnifiedMinLeaves) && kMeshLeafTris == 1 && flat.empty() && d->n_prims > 0;
    // ---- plan of the unified world-space tree: [top-level four-children nodes | subtree of instance 0 | instance 1 | ...]
    struct Inst {
        int prim, mesh, node_base, leaf_base;
    };
    std::vector<Inst> inst;
    DevNode4* d_unodes = nullptr;
    DevTri* d_utris = nullptr;
    int u_n_tris = 0, n_top_compact = 0;
    float u_pad_rel = 0.0f, u_pad_abs = 0.0f, u_radius = 0.0f;
    if (unified) {
        std::vector<int> inst_of_prim(d->n_prims, -1);
        size_t n_nodes = top4.size(), n_tris_total = 0;
        for (int i = 0; i < d->n_prims; ++i) {
            if (prims[i].type != MCPT_PRIM_MESH || prims[i].slot < 0) continue;
            const int nt = int(d->meshes[prims[i].mesh].n_tris);
            if (nt <= 0) continue;
            inst_of_prim[i] = int(inst.size());
            inst.push_back(Inst{i, prims[i].mesh, int(n_nodes), int(n_tris_total)});
            n_nodes += size_t(std::max(nt - 1, 1));
            n_tris_total += size_t(nt);
        }
        if (n_nodes > 0x3fffffffu || n_tris_total + size_t(d->n_prims) > 0x3fffffffu)
            return fail(c, MCPT_E_UNSUPPORTED, "unified tree too large (%zu nodes, %zu triangle slots): set MCPT_OPT_UNIFIED = 0", n_nodes, n_tris_total);
        u_n_tris = int(n_tris_total);
        // leaf tokens of the top-level tree: a mesh instance becomes a link to the root of its subtree, an analytic
        // primitive moves behind the triangle slots
        // ... stored in breadth-first order without the unused entries of top4, so that "the first N nodes" are the
        // top levels (what the traversal kernel stages in shared memory)
        std::vector<DevNode4> utop;
        {
            std::vector<int> order{0}, new_id(top4.size(), -1);
            new_id[0] = 0;
            for (size_t hd = 0; hd < order.size(); ++hd) {
                const DevNode4& N = top4[order[hd]];
                for (int j = 0; j < 4; ++j) {
                    const int tok = (&N.tok.x)[j];
                    if (tok >= 0 && new_id[tok] < 0) {
                        new_id[tok] = int(order.size());
                        order.push_back(tok);
                    }
                }
            }
            utop.resize(order.size());
            for (size_t k = 0; k < order.size(); ++k) {
                utop[k] = top4[order[k]];
                for (int j = 0; j < 4; ++j) {
                    int& tok = (&utop[k].tok.x)[j];
                    if (tok >= 0) tok = new_id[tok];
                }
            }
        }
        n_top_compact = int(utop.size());
        // the instances' subtrees follow the compacted top part
        {
            const int shift = int(top4.size()) - n_top_compact;
            for (Inst& in : inst) in.node_base -= shift;
            n_nodes -= size_t(shift);
        }
        for (DevNode4& N : utop)
            for (int j = 0; j < 4; ++j) {
                int& tok = (&N.tok.x)[j];
                if (tok >= 0 || tok == INT_MIN) continue;
                const int prim = ~tok;
                if (prims[prim].type == MCPT_PRIM_MESH)
                    tok = inst_of_prim[prim] >= 0 ? inst[inst_of_prim[prim]].node_base : INT_MIN;
                else
                    tok = ~(u_n_tris + prim);
            }
        if ((rc = dev_alloc(c, c->scene_allocs, &d_unodes, n_nodes)) || (rc = dev_alloc(c, c->scene_allocs, &d_utris, std::max<size_t>(n_tris_total, 1))))
            return rc;
        if (!utop.empty()) CK(c, cudaMemcpyAsync(d_unodes, utop.data(), utop.size() * sizeof(DevNode4), cudaMemcpyHostToDevice, c->stream));
        CK(c, cudaStreamSynchronize(c->stream));  // utop is a temporary
        // how far fwd * inv is from the identity (the uploaded inverse is a rounded one): linear part as the largest
        // absolute row sum, translation part as the largest component
        double pad_rel = 0.0, pad_abs = 0.0, radius = 0.0;
        for (const Inst& in : inst) {
            const DevPrim& P = prims[in.prim];
            const float4 f[3] = {P.fwd.r0, P.fwd.r1, P.fwd.r2}, v[3] = {P.inv.r0, P.inv.r1, P.inv.r2};
            for (int r = 0; r < 3; ++r) {
                double row = 0.0;
                const double fr[3] = {f[r].x, f[r].y, f[r].z};
                for (int cc = 0; cc < 4; ++cc) {
                    const double vc[3] = {cc == 0 ? v[0].x : cc == 1 ? v[0].y : cc == 2 ? v[0].z : v[0].w,
                                          cc == 0 ? v[1].x : cc == 1 ? v[1].y : cc == 2 ? v[1].z : v[1].w,
                                          cc == 0 ? v[2].x : cc == 1 ? v[2].y : cc == 2 ? v[2].z : v[2].w};
                    double e_rc = fr[0] * vc[0] + fr[1] * vc[1] + fr[2] * vc[2] + (cc == 3 ? double(f[r].w) : 0.0) - (cc == r ? 1.0 : 0.0);
                    if (cc < 3) row += std::fabs(e_rc);
                    else pad_abs = std::max(pad_abs, std::fabs(e_rc));
                }
                pad_rel = std::max(pad_rel, row);
            }
            const DevNode& nb = top[t_first_leaf + uint32_t(P.slot)];  // the instance's world box (reference leaf box)
            radius = std::max({radius, std::fabs(double(nb.lo.x)), std::fabs(double(nb.lo.y)), std::fabs(double(nb.lo.z)),
                               std::fabs(double(nb.hi.x)), std::fabs(double(nb.hi.y)), std::fabs(double(nb.hi.z))});
        }
        if (!std::isfinite(pad_rel) || !std::isfinite(pad_abs)) pad_rel = pad_abs = 1.0;
        // twice the bound, plus a few ulps of the quantities involved
        u_pad_rel = float(4.0 * pad_rel + 1e-6);
        u_pad_abs = float(2.0 * pad_abs + 1e-6 * radius);
        u_radius = float(3.0 * radius);
    }
    // The tree builds run on side streams: a build is a chain of small dependent kernels (150 us of device time, most
    // of it launch latency), so the builds of different meshes / instances overlap each other and the host -> device
    // copies of the meshes that follow. Fork: the side stream waits for the mesh's data on the context's stream;
    // join: the context's stream waits for every side stream at the end.
    if ((rc = ensure_side_streams(c))) return rc;
    int n_builds = 0;
    StageSpan build_span(c, MCPT_STAGE_BVH_BUILD);
    auto build_on_side = [&](const LbvhBuild& b) -> int {
        const int k = n_builds++ % kSide;
        CK(c, cudaEventRecord(c->ev_fork, c->stream));
        CK(c, cudaStreamWaitEvent(c->side[k], c->ev_fork, 0));
        unsigned long long launched = 0;
        const cudaError_t e = build_mesh_lbvh(c->side[k], b, build_ws + size_t(k) * build_ws_bytes, build_ws_bytes, &launched);
        c->launches += launched;
        CK(c, e);
        return MCPT_OK;
    };
    std::vector<int32_t*> d_slots(d->n_meshes, nullptr);  // per mesh: the reference's leaf slot of every triangle (or null)
    for (int m = 0; m < d->n_meshes; ++m) {
        const mcpt_mesh_desc& md = d->meshes[m];
        const bool no_tree = !md.bvh.bb && !md.bvh.ind;
        const int depth = no_tree ? 0 : md.bvh.depth;
        const uint32_t first_leaf = 1u << depth;
        DevMesh& M = meshes[m];
        std::memset(&M, 0, sizeof M);
        uint32_t* di = nullptr;
        float *dp = nullptr, *dnr = nullptr;
        // straight from the caller's arrays (no host-side copies)
        if ((rc = upload_raw(c, md.tri_idx, 3 * size_t(md.n_tris), &di)) || (rc = upload_raw(c, md.pos, 3 * size_t(md.n_verts), &dp)) ||
            (rc = upload_raw(c, md.nrm, 3 * size_t(md.n_verts), &dnr)))
            return rc;
        check_indices(c->stream, di, 3 * size_t(md.n_tris), md.n_verts, d_flag);
        c->launches += md.n_tris ? 1 : 0;
        M.idx = di;
        M.pos = dp;
        M.nrm = dnr;
        M.depth = depth;
        M.n_tris = int(md.n_tris);
        if (c->opt_mesh_bvh && md.n_tris > 0) {
            // the tree is rebuilt on the GPU (bvh_build.cu); an uploaded reference tree only supplies each triangle's
            // leaf slot, which ranks equal-distance candidates like the reference's own visit order
            if (!no_tree) {
                int32_t* d_ind = nullptr;
                if ((rc = upload_raw(c, md.bvh.ind, size_t(first_leaf), &d_ind)) || (rc = dev_alloc(c, build_allocs, &d_slots[m], size_t(md.n_tris))))
                    return rc;
                c->scene_allocs.pop_back();  // d_ind is scratch of this upload, not part of the scene
                build_allocs.push_back(d_ind);
                CK(c, slots_from_leaves(c->stream, d_ind, int(first_leaf), int(md.n_tris), d_slots[m], d_flag));
                c->launches += 1;
            }
            if (unified) {
                for (const Inst& in : inst) {
                    if (in.mesh != m) continue;
                    LbvhBuild b;
                    b.d_pos = dp;
                    b.d_idx = di;
                    b.n_tris = int(md.n_tris);
                    b.n_verts = int(md.n_verts);
                    b.d_slot_of_tri = d_slots[m];
                    b.fwd = &prims[in.prim].fwd;
                    b.d_nodes4 = d_unodes + in.node_base;
                    b.d_tris = d_utris + in.leaf_base;
                    b.n_leaves = int(md.n_tris);
                    b.node_base = in.node_base;
                    b.leaf_base = in.leaf_base;
                    b.prim = in.prim;
                    if ((rc = build_on_side(b))) return rc;
                }
                continue;
            }
            const int n_leaves = int((md.n_tris + kMeshLeafTris - 1) / kMeshLeafTris);
            DevNode2* dn2 = nullptr;
            DevNode4* dn4 = nullptr;
            DevTri* dt2 = nullptr;
            if ((rc = dev_alloc(c, c->scene_allocs, &dn2, size_t(std::max(n_leaves - 1, 1)))) ||
                (rc = dev_alloc(c, c->scene_allocs, &dn4, size_t(std::max(n_leaves - 1, 1)))) ||
                (rc = dev_alloc(c, c->scene_allocs, &dt2, size_t(n_leaves) * kMeshLeafTris)))
                return rc;
            {
                LbvhBuild b;
                b.d_pos = dp;
                b.d_idx = di;
                b.n_tris = int(md.n_tris);
                b.n_verts = int(md.n_verts);
                b.d_slot_of_tri = d_slots[m];
                b.d_nodes = dn2;
                b.d_nodes4 = dn4;
                b.d_tris = dt2;
                b.n_leaves = n_leaves;
                if ((rc = build_on_side(b))) return rc;
            }
            M.nodes2 = dn2;
            M.nodes4 = dn4;
            M.tris2 = dt2;
            M.n_leaves2 = n_leaves;
            continue;
        }
        if (md.n_tris == 0 && no_tree) continue;
        std::vector<DevNode> nodes(first_leaf, empty_node());  // interior nodes 1 .. first_leaf-1
        for (uint32_t k = 1; k < first_leaf; ++k) nodes[k] = pack_node(md.bvh.bb[k - 1], -1);
        std::vector<DevTri> tris(first_leaf);
        for (uint32_t s = 0; s < first_leaf; ++s) {
            int t = md.bvh.ind[s];
            DevTri& T = tris[s];
            T.a = T.b = T.c = make_float4(0, 0, 0, 0);
            if (t >= 0) {
                const uint32_t* ix = md.tri_idx + 3 * size_t(t);
                for (int q = 0; q < 3; ++q)
                    if (ix[q] >= md.n_verts) return fail(c, MCPT_E_INVALID, "mesh %d: vertex index out of range", m);
                const float *A = md.pos + 3 * size_t(ix[0]), *B = md.pos + 3 * size_t(ix[1]), *C = md.pos + 3 * size_t(ix[2]);
                T.a = make_float4(A[0], A[1], A[2], 0);
                T.b = make_float4(B[0], B[1], B[2], 0);
                T.c = make_float4(C[0], C[1], C[2], 0);
            }
            std::memcpy(&T.a.w, &t, sizeof t);
        }
        DevNode* dn = nullptr;
        DevTri* dt = nullptr;
        if ((rc = upload_vec(c, nodes, &dn)) || (rc = upload_vec(c, tris, &dt))) return rc;
        M.nodes = dn;
        M.tris = dt;
    }

    mark("meshes copied, builds enqueued");
    // join the side streams
    for (int k = 0; k < std::min(n_builds, kSide); ++k) {
        CK(c, cudaEventRecord(c->ev_join[k], c->side[k]));
        CK(c, cudaStreamWaitEvent(c->stream, c->ev_join[k], 0));
    }
    build_span.end();

    // ---- shared-memory staging plan (<= 40 KB so no opt-in is needed and
    // several CTAs stay resident per SM)
    S.smem_top_nodes = 0;
    S.smem_prims = 0;
    S.smem_mesh_nodes = 0;
    // persistent warps need every mesh tree in the GPU-built form
    // (measured: 1.8x on the 1 M-triangle workload, +20 % on a tree of 122 analytic primitives)
    const bool persistent = c->opt_persistent && flat.empty() && (d->n_meshes == 0 || c->opt_mesh_bvh);
    if (c->opt_smem_staging) {
        uint32_t keep = t_nodes - 1;            // all top-level nodes ...
        while (keep > 511) keep >>= 1;          // ... or the top levels that fit 16 KB
        // the heap-order tree is what the one-ray-per-lane walk reads; persistent warps walk top2 / top4 / u_nodes
        // (read through L1) and stage the primitive records only
        S.smem_top_nodes = persistent ? 0 : int(keep);
        S.smem_prims = std::min(d->n_prims, 64);
        if (d->n_meshes > 0 && !persistent) {
            int share = 512 / d->n_meshes;  // nodes per mesh out of a 16 KB budget
            int off = 0;
            for (DevMesh& M : meshes) {
                if (M.nodes2 || !M.nodes) continue;  // GPU-built trees are read through L1/L2 only
                int cnt = (1 << M.depth) - 1;
                while (cnt > share) cnt >>= 1;  // 2^L - 1 stays of that form
                M.smem_off = off;
                M.smem_nodes = cnt;
                off += cnt;
            }
            S.smem_mesh_nodes = off;
        }
    }
    S.u_smem_nodes = 0;
    if (unified && c->opt_smem_staging >= 2) {
        // MCPT_OPT_SMEM_STAGING = 2: also the top levels of the unified tree, what fits 32 KB next to the primitive records
        // (4 CTAs per SM stay resident). Off by default -- measured with 0 / 64 / 256 staged nodes (Msamples/s): repo scene 8
        // 686 / 661 / 644, scene 7 2046 / 2005 / 2031, scene 3 4344 / 4275 / 4341, 1 M triangles 1587 / 1563 / 1558: the top
        // levels already live in L1, and the walk pays a branch per node for the second address space.
        S.u_smem_nodes = std::min(n_top_compact, 256);
    }
    if (!flat.empty()) {
        S.smem_top_nodes = 8 * int(flat.size()) - 1;  // the flat list, once per ray octant (stage_scene), takes the tree's place ...
        S.smem_prims = d->n_prims;                // ... and so do all (<= 16) primitive records
    }
    S.smem_bytes = int(sizeof(DevNode)) * (S.smem_top_nodes + 1 + S.smem_mesh_nodes) + int(sizeof(DevPrim)) * S.smem_prims +
                   int(sizeof(DevNode4)) * S.u_smem_nodes;

    // ---- lights of next-event estimation: emissive quads (emissivity > 0.5: the ones that END a path) among the
    // emissive-first records, picked with probability proportional to the reference's area figure mat_info.a
    std::vector<DevLight> lights;
    {
        double wsum = 0.0;
        const int n_emi = std::min(std::max(d->nb_emissives, 0), d->n_prims);
        for (int i = 0; i < n_emi && lights.size() < 256; ++i) {
            const DevPrim& P = prims[i];
            if (P.type != MCPT_PRIM_ORIENTED_QUAD || !(mat[i].z > 0.5f) || P.slot < 0) continue;
            const float ex[3] = {P.fwd.r0.x, P.fwd.r1.x, P.fwd.r2.x}, ey[3] = {P.fwd.r0.y, P.fwd.r1.y, P.fwd.r2.y};
            const float nx = ex[1] * ey[2] - ex[2] * ey[1], ny = ex[2] * ey[0] - ex[0] * ey[2], nz = ex[0] * ey[1] - ex[1] * ey[0];
            const float nl = std::sqrt(nx * nx + ny * ny + nz * nz);
            if (!(nl > 1e-20f) || !std::isfinite(nl)) continue;
            const float area = 4.0f * nl;                       // the unit quad spans [-1, 1]^2
            const float w = mat[i].w > 0.0f ? mat[i].w : area;  // mat_info.a
            DevLight L;
            L.c = make_float4(P.fwd.r0.w, P.fwd.r1.w, P.fwd.r2.w, 1.0f / area);
            L.ex = make_float4(ex[0], ex[1], ex[2], w);
            L.ey = make_float4(ey[0], ey[1], ey[2], 0.0f);
            L.n = make_float4(nx / nl, ny / nl, nz / nl, 0.0f);
            std::memcpy(&L.n.w, &i, sizeof(int));
            L.emit = make_float4(0.1f * color[i].x, 0.1f * color[i].y, 0.1f * color[i].z, mat[i].z * (1.0f - mat[i].x) * color[i].w);
            lights.push_back(L);
            wsum += w;
        }
        double cum = 0.0;
        for (DevLight& L : lights) {
            L.ex.w = float(L.ex.w / wsum);
            cum += L.ex.w;
            L.ey.w = float(cum);
        }
        if (!lights.empty()) lights.back().ey.w = 1.0f;
    }
    DevLight* d_lights = nullptr;
    if ((rc = upload_vec(c, lights, &d_lights))) return rc;
    S.lights = lights.empty() ? nullptr : d_lights;
    S.n_lights = int(lights.size());

    DevNode* d_top = nullptr;
    DevPrim* d_prims = nullptr;
    float4 *d_color = nullptr, *d_mat = nullptr;
    DevMesh* d_meshes = nullptr;
    float* d_leaf_bb = nullptr;
    DevNode* d_flat = nullptr;
    DevNode2* d_top2 = nullptr;
    DevNode4* d_top4 = nullptr;
    if ((rc = upload_vec(c, top2, &d_top2)) || (rc = upload_vec(c, top4, &d_top4))) return rc;
    S.top2 = d_top2;
    S.top4 = d_top4;
    S.u_nodes = d_unodes;
    S.u_tris = d_utris;
    S.u_n_tris = u_n_tris;
    S.u_pad_rel = u_pad_rel;
    S.u_pad_abs = u_pad_abs;
    S.u_radius = u_radius;
    {   // key layout of the pooled leaf tests (layout.cuh: u_pool)
        S.u_pool = 0, S.u_key_shift = 0, S.u_slot_max = 0, S.u_rank_max = 0;
        if (unified && c->opt_leaf_pool > 0) {
            // a triangle's rank is its leaf slot in the uploaded reference tree (< 2^depth) or its index (< n_tris), -1 if absent
            long long rank_max = 1, slot_max = 0;
            for (const Inst& in : inst) {
                const mcpt_mesh_desc& md = d->meshes[in.mesh];
                long long r = (long long)md.n_tris;
                if (md.bvh.ind && md.bvh.depth >= 0 && md.bvh.depth < 31) r = std::max(r, 1ll << md.bvh.depth);
                rank_max = std::max(rank_max, r + 1);
            }
            for (int i = 0; i < d->n_prims; ++i) slot_max = std::max(slot_max, (long long)prims[i].slot);
            int rb = 0, sb = 0;
            while ((1ll << rb) <= rank_max) ++rb;
            while ((1ll << sb) <= slot_max) ++sb;
            if (rb + sb + 1 <= 32 && size_t(u_n_tris) + size_t(d->n_prims) < (1u << 26)) {
                S.u_pool = std::min(c->opt_leaf_pool, 32);
                S.u_key_shift = rb + 1;
                S.u_slot_max = int(slot_max);
                S.u_rank_max = int(rank_max);
            }
        }
    }
    if ((rc = upload_vec(c, top, &d_top)) || (rc = upload_vec(c, prims, &d_prims)) || (rc = upload_vec(c, color, &d_color)) ||
        (rc = upload_vec(c, mat, &d_mat)) || (rc = upload_vec(c, meshes, &d_meshes)) || (rc = upload_vec(c, leaf_bb, &d_leaf_bb)) ||
        (rc = upload_vec(c, flat, &d_flat)))
        return rc;
    S.top_nodes = d_top;
    S.prims = d_prims;
    S.color = d_color;
    S.mat = d_mat;
    S.meshes = d_meshes;
    S.leaf_bb = d_leaf_bb;
    S.flat = flat.empty() ? nullptr : d_flat;
    S.n_flat = int(flat.size());
    S.flat_pool = c->opt_flat_pool;
    {
        int per_type[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        for (const DevNode& nd : flat) {
            int prim;
            std::memcpy(&prim, &nd.lo.w, sizeof(int));
            if (prim >= 0 && prim < d->n_prims && prims[prim].type >= 0 && prims[prim].type < 8) ++per_type[prims[prim].type];
        }
        S.flat_max_type = std::max(1, *std::max_element(per_type, per_type + 8));
    }
    S.top_depth = td;
    S.n_prims = d->n_prims;
    S.n_meshes = d->n_meshes;
    S.persistent = persistent ? 1 : 0;
    S.refill_lanes = c->opt_refill_lanes > 0 ? c->opt_refill_lanes : (d->n_meshes > 0 ? 8 : (unified ? 12 : 32));
    // camera rays: the 32 rays of a warp are neighbours in the image and finish together; they are replaced together
    // (the unified walk on the 1 M-triangle workload: 8 / 16 / 24 / 32 idle lanes -> 1548 / 1564 / 1571 / 1578 Msamples/s)
    S.refill_first = (c->opt_refill_lanes > 0) ? c->opt_refill_lanes : 32;
    // every copy above was enqueued on the context's stream from host temporaries of this call (pageable memory:
    // the runtime stages them, but the order against later work on c->stream is only defined in-stream)
    mark("tables enqueued");
    int h_flag = 0;
    if (d_flag) CK(c, cudaMemcpyAsync(&h_flag, d_flag, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    CK(c, cudaStreamSynchronize(c->stream));
    mark("device done");
    if (h_flag & 1) return fail(c, MCPT_E_INVALID, "mesh: vertex index out of range");
    if (h_flag & 2) return fail(c, MCPT_E_INVALID, "mesh: leaf slot of the uploaded BVH holds an index out of range");
    // what the previous scene held and this one did not take goes back to the driver; the scratch of the tree builds
    // stays parked for the next upload (an animated viewer re-uploads every frame: no cudaMalloc / cudaFree then)
    trim_cache(c);
    if (!build_allocs.empty()) free_all(c, build_allocs);
    c->scene = S;
    c->have_scene = true;
    mark("done");
    return MCPT_OK;
}

// ---- host-only inspection of the walk trees (CPU tests)
static int inspect_walk_tree_impl(const mcpt_scene_desc* d, int top_sah, mcpt_tree_stats* out);
int mcpt_inspect_walk_tree(const mcpt_scene_desc* d, int top_sah, mcpt_tree_stats* out) {
    try {
        return inspect_walk_tree_impl(d, top_sah, out);
    } catch (const std::exception&) {
        return MCPT_E_NOMEM;
    }
}
static int inspect_walk_tree_impl(const mcpt_scene_desc* d, int top_sah, mcpt_tree_stats* out) {
    if (!d || !out || d->n_prims < 0 || (d->n_prims > 0 && !d->prim_data) || !d->root.bb || !d->root.ind || d->root.depth < 0 ||
        d->root.depth > 30)
        return MCPT_E_INVALID;
    std::memset(out, 0, sizeof *out);
    const int td = d->root.depth;
    const uint32_t t_first_leaf = 1u << td, t_nodes = 2u << td;
    std::vector<DevPrim> prims(d->n_prims);
    for (int i = 0; i < d->n_prims; ++i) {
        const mcpt_prim_data& pd = d->prim_data[i];
        DevPrim& p = prims[i];
        std::memset(&p, 0, sizeof p);
        p.type = int(pd.type[0]);
        p.slot = -1;
        p.fwd = rows_of(p.type == MCPT_PRIM_MESH ? pd.mesh_transfo : pd.transfo);
        if (p.type > MCPT_PRIM_ORIENTED_QUAD) p.type = -1;
    }
    std::vector<DevNode> top(t_nodes, empty_node());
    for (uint32_t k = 1; k < t_nodes; ++k) {
        int aux = -1;
        if (k >= t_first_leaf) {
            aux = d->root.ind[k - t_first_leaf];
            if (aux >= d->n_prims) return MCPT_E_INVALID;
            if (aux < 0) continue;
            if (prims[aux].slot < 0) prims[aux].slot = int(k - t_first_leaf);
        }
        top[k] = pack_node(d->root.bb[k - 1], aux);
    }
    std::vector<DevNode2> top2;
    std::vector<DevNode4> top4;
    const auto t_build = std::chrono::steady_clock::now();
    build_walk_trees(d, prims, top, top_sah != 0, top2, top4);
    if (std::getenv("MCPT_DUMP_UPLOAD"))
        std::fprintf(stderr, "walk trees of %d primitives built on the host in %.3f ms\n", d->n_prims,
                     std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_build).count());
    out->nodes = int(top2.size());
    for (int i = 0; i < d->n_prims; ++i) out->prims_expected += prims[i].slot >= 0 ? 1 : 0;

    auto child = [&](int node, int j, float lo[3], float hi[3]) {
        const DevNode2& N = top2[node];
        const float4& l = j ? N.lo1 : N.lo0;
        const float4& h = j ? N.hi1 : N.hi0;
        lo[0] = l.x; lo[1] = l.y; lo[2] = l.z;
        hi[0] = h.x; hi[1] = h.y; hi[2] = h.z;
        int tok;
        std::memcpy(&tok, j ? &N.hi0.w : &N.lo0.w, sizeof tok);
        return tok;
    };
    std::vector<std::vector<std::array<float, 6>>> boxes_of(d->n_prims);
    double root_area = 0.0;
    // iterative post-order: (node, depth); returns the box of each subtree and the leaves below it
    struct Ret { float lo[3], hi[3]; int leaves; };
    std::function<Ret(int, int)> visit = [&](int node, int depth) -> Ret {
        Ret r{{FLT_MAX, FLT_MAX, FLT_MAX}, {-FLT_MAX, -FLT_MAX, -FLT_MAX}, 0};
        out->depth = std::max(out->depth, depth + 1);
        for (int j = 0; j < 2; ++j) {
            float lo[3], hi[3];
            const int tok = child(node, j, lo, hi);
            if (tok == INT_MIN) continue;
            int below = 1;
            if (tok >= 0) {
                if (tok >= int(top2.size()) || depth > 64) { ++out->violations; continue; }
                const Ret sub = visit(tok, depth + 1);
                below = sub.leaves;
                for (int k = 0; k < 3; ++k)
                    if (sub.lo[k] < lo[k] || sub.hi[k] > hi[k]) { ++out->violations; break; }
            } else {
                const int prim = ~tok;
                if (prim < 0 || prim >= d->n_prims) { ++out->violations; continue; }
                ++out->leaves;
                boxes_of[prim].push_back({lo[0], lo[1], lo[2], hi[0], hi[1], hi[2]});
            }
            out->sah_cost += half_area(lo, hi) * below;
            r.leaves += below;
            for (int k = 0; k < 3; ++k) {
                r.lo[k] = std::min(r.lo[k], lo[k]);
                r.hi[k] = std::max(r.hi[k], hi[k]);
            }
        }
        return r;
    };
    if (!top2.empty()) {
        const Ret r = visit(0, 0);
        root_area = half_area(r.lo, r.hi);
    }
    if (root_area > 0.0) out->sah_cost /= root_area;
    // every primitive: points of its bounding cube (the reference's bound of the unit shape) lie in one of its leaf boxes
    for (int i = 0; i < d->n_prims; ++i) {
        if (boxes_of[i].empty()) continue;
        ++out->prims_reached;
        if (prims[i].type <= MCPT_PRIM_MESH) continue;
        const Aff& f = prims[i].fwd;
        for (int a = -2; a <= 2; ++a)
            for (int b = -2; b <= 2; ++b)
                for (int c2 = -2; c2 <= 2; ++c2) {
                    const float x = 0.5f * a, y = 0.5f * b, z = (prims[i].type == MCPT_PRIM_ORIENTED_QUAD) ? 0.0f : 0.5f * c2;
                    const float w[3] = {f.r0.x * x + f.r0.y * y + f.r0.z * z + f.r0.w, f.r1.x * x + f.r1.y * y + f.r1.z * z + f.r1.w,
                                        f.r2.x * x + f.r2.y * y + f.r2.z * z + f.r2.w};
                    bool inside = false;
                    for (const auto& bx : boxes_of[i])
                        inside = inside || (w[0] >= bx[0] && w[1] >= bx[1] && w[2] >= bx[2] && w[0] <= bx[3] && w[1] <= bx[4] && w[2] <= bx[5]);
                    if (!inside) ++out->violations;
                }
    }
    // the four-children fold reaches the same leaves
    std::function<void(int, int)> visit4 = [&](int node, int depth) {
        out->depth4 = std::max(out->depth4, depth + 1);
        const DevNode4& W = top4[node];
        for (int j = 0; j < 4; ++j) {
            const int tok = (&W.tok.x)[j];
            if (tok == INT_MIN) continue;
            if (tok >= 0) {
                if (tok >= int(top4.size()) || depth > 64) { ++out->violations; continue; }
                visit4(tok, depth + 1);
            } else {
                ++out->leaves4;
            }
        }
    };
    if (!top4.empty()) visit4(0, 0);
    return MCPT_OK;
}
