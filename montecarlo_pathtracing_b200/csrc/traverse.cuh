// traverse.cuh — closest-hit traversal of the two-level scene
// (replaces intersect_bvh / Mesh_intersect, shaders/raytracer_func.frag:642-678,
// 734-781, and the AABB tests intersect_bv / intersect_bvm, :273-352).
//
// Design (B200): the reference walks the tree with an int[29] stack in local
// memory, visits children in fixed left/right order and pays two dependent texel
// fetches plus three vector divides per box. Here there are three walkers, chosen
// per scene at upload (all rank candidates by the exact world-space distance of
// intersect.cuh with the reference's visit order as tie-break, so they return the
// same record bit for bit):
//   * scene_walk_flat -- <= 16 analytic primitives: the leaf boxes as one list in
//     shared memory, stepped through by all lanes of a warp together;
//   * the resumable walk (walk_begin / walk_node / walk_tri / walk_prim) -- every
//     other scene, under persistent warps (k_traverse, PERSIST): a top-level tree
//     of the library's own (full-sweep SAH at upload, api_scene.inl) and the GPU-built
//     mesh trees, both in the 64-byte two-children node format -- or, on scenes with
//     meshes, folded into 128-byte four-children nodes -- nearest child first, the
//     others pushed with their entry parameters on one explicit stack;
//   * scene_walk + mesh_walk / mesh_walk_lbvh -- one ray per lane over the
//     reference's implicit heap order, which needs no stack at all: the node's heap
//     index plus a 32-bit trail ("a sibling is still pending at this level") live in
//     two registers per thread; children are visited near-first using the split axis
//     that the builder cycles per level (bvh_gpu/bvh.cpp:45-51). Used by
//     mcpt_trace_rays and when MCPT_OPT_PERSISTENT / MCPT_OPT_MESH_BVH are off.
// The slab test runs on precomputed (approximate) reciprocals with FMAs. It only
// culls: boxes are inflated at upload so it never rejects a box the exact arithmetic
// of the reference would accept.
#pragma once
#include <climits>

#include "intersect.cuh"

namespace mcpt {

// optional per-thread work counters (MCPT_OPT_COUNT_WORK): boxes tested, primitive
// records tested, triangles tested
struct Work {
    uint32_t boxes = 0, prims = 0, tris = 0;
};

struct SmemScene {
    const DevNode* top_nodes;   // staged top-level nodes, index 1-based like global
    const DevPrim* prims;       // staged primitive records
    const DevNode* mesh_nodes;  // staged per-mesh top levels, concatenated
    const DevNode4* u_nodes;    // staged top levels of the unified tree (DevScene::u_smem_nodes of them)
};

MCPT_DEV DevNode load_node(const DevNode* p) {
    DevNode n;
    n.lo = __ldg(&p->lo);
    n.hi = __ldg(&p->hi);
    return n;
}
MCPT_DEV DevNode load_node_smem(const DevNode* p) { return *p; }

// conservative slab test on [0, tmax]; idir = 1/dir, od = -origin*idir
MCPT_DEV bool slab(const DevNode& n, f3 idir, f3 od, float tmax) {
    float tx0 = __fmaf_rn(n.lo.x, idir.x, od.x), tx1 = __fmaf_rn(n.hi.x, idir.x, od.x);
    float ty0 = __fmaf_rn(n.lo.y, idir.y, od.y), ty1 = __fmaf_rn(n.hi.y, idir.y, od.y);
    float tz0 = __fmaf_rn(n.lo.z, idir.z, od.z), tz1 = __fmaf_rn(n.hi.z, idir.z, od.z);
    float tn = fmaxf(fmaxf(fminf(tx0, tx1), fminf(ty0, ty1)), fmaxf(fminf(tz0, tz1), 0.0f));
    float tf = fminf(fminf(fmaxf(tx0, tx1), fmaxf(ty0, ty1)), fminf(fmaxf(tz0, tz1), tmax));
    return tn <= tf;
}

// the same test on a box whose planes are already ordered for the ray's octant (lo: entry, hi: exit):
// a product with a reciprocal of the matching sign is monotonic, so the per-axis min / max drop out
MCPT_DEV bool slab_sorted(const DevNode& n, f3 idir, f3 od, float tmax) {
    const float tn = fmaxf(fmaxf(__fmaf_rn(n.lo.x, idir.x, od.x), __fmaf_rn(n.lo.y, idir.y, od.y)), fmaxf(__fmaf_rn(n.lo.z, idir.z, od.z), 0.0f));
    const float tf = fminf(fminf(__fmaf_rn(n.hi.x, idir.x, od.x), __fmaf_rn(n.hi.y, idir.y, od.y)), fminf(__fmaf_rn(n.hi.z, idir.z, od.z), tmax));
    return tn <= tf;
}

// Reciprocals that feed only the culling frame (idir, od, tmax) are MUFU approximations: their
// relative error (2^-23) is two orders of magnitude below the inflation of the boxes (1e-5) and the
// slack on tmax (1e-4), and no hit record depends on them. The IEEE divide they replace was 8 % of
// the flat traversal's instructions.
MCPT_DEV float cull_rcp(float d) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(d));
    return r;
}
MCPT_DEV float safe_rcp(float d) {
    // A zero component would give 1/0 = inf and inf - inf = NaN inside the FMA slab
    // test; 1e-30 keeps the products finite and the ray "parallel" for any box.
    if (fabsf(d) < 1e-30f) d = copysignf(1e-30f, d);
    return cull_rcp(d);
}
MCPT_DEV f3 safe_rcp3(f3 d) { return mk3(safe_rcp(d.x), safe_rcp(d.y), safe_rcp(d.z)); }

// Walk one mesh (second level). Oi/Di: ray in mesh-local space (Di normalised by
// intersect_prim, raytracer_func.frag:688); Ow: world origin; fwd: mesh -> world.
template <bool COUNT>
MCPT_DEV void mesh_walk(Work& wk, const DevMesh& M, const DevNode* smem_mesh_nodes, f3 Oi, f3 Di, f3 Ow, const Aff& fwd, int prim,
                        int tslot, float slack, Best& best) {
    // world length of one unit of local parameter: dist ~= a * scale
    f3 wd = mk3(fwd.r0.x * Di.x + fwd.r0.y * Di.y + fwd.r0.z * Di.z, fwd.r1.x * Di.x + fwd.r1.y * Di.y + fwd.r1.z * Di.z,
                fwd.r2.x * Di.x + fwd.r2.y * Di.y + fwd.r2.z * Di.z);
    float scale = sqrtf(wd.x * wd.x + wd.y * wd.y + wd.z * wd.z);
    if (!(scale > 0.0f)) return;  // zero or NaN direction: the shader finds nothing either
    const float inv_scale = cull_rcp(scale);
    const f3 idir = safe_rcp3(Di);
    const f3 od = mk3(-Oi.x * idir.x, -Oi.y * idir.y, -Oi.z * idir.z);
    float tmax = (best.h.dist * 1.0001f + slack) * inv_scale;

    const int depth = M.depth;
    const uint32_t first_leaf = 1u << depth;
    const uint32_t staged = uint32_t(M.smem_nodes);  // nodes with heap index <= staged live in smem
    const DevNode* snodes = smem_mesh_nodes + M.smem_off - 1;  // 1-based
    const DevNode* gnodes = M.nodes;                           // 1-based (entry 0 unused)
    // near-first: level l splits along axis l % 3 with the smaller centres on the left
    const uint32_t neg = (Di.x < 0.0f ? 1u : 0u) | (Di.y < 0.0f ? 2u : 0u) | (Di.z < 0.0f ? 4u : 0u);

    uint32_t k = 1, trail = 0;
    int level = 0, axis = 0;
    while (true) {
        bool descend = false;
        if (k < first_leaf) {
            DevNode n = (k <= staged) ? load_node_smem(snodes + k) : load_node(gnodes + k);
            descend = slab(n, idir, od, tmax);
            if (COUNT) ++wk.boxes;
        } else {
            const DevTri* tp = M.tris + (k - first_leaf);
            float4 ta = __ldg(&tp->a);
            int tri = __float_as_int(ta.w);
            if (tri >= 0) {
                if (COUNT) ++wk.tris;
                float4 tb = __ldg(&tp->b), tc = __ldg(&tp->c);
                float a;
                if (triangle_test(best, fwd, mk3(ta.x, ta.y, ta.z), mk3(tb.x, tb.y, tb.z), mk3(tc.x, tc.y, tc.z), Oi, Di, Ow,
                                  prim, tri, tslot, int(k - first_leaf), a))
                    tmax = (best.h.dist * 1.0001f + slack) * inv_scale;
            }
        }
        if (descend) {
            k = 2u * k + ((neg >> axis) & 1u);
            trail = (trail << 1) | 1u;
            ++level;
            axis = (axis == 2) ? 0 : axis + 1;
            continue;
        }
        // pop: climb while no sibling is pending
        while (level > 0 && !(trail & 1u)) {
            trail >>= 1;
            k >>= 1;
            --level;
            axis = (axis == 0) ? 2 : axis - 1;
        }
        if (level == 0) break;
        k ^= 1u;
        trail &= ~1u;
    }
}

// conservative slab test that also returns the entry parameter (>= 0)
MCPT_DEV bool slab_entry(const float4& lo, const float4& hi, f3 idir, f3 od, float tmax, float& tn_out) {
    float tx0 = __fmaf_rn(lo.x, idir.x, od.x), tx1 = __fmaf_rn(hi.x, idir.x, od.x);
    float ty0 = __fmaf_rn(lo.y, idir.y, od.y), ty1 = __fmaf_rn(hi.y, idir.y, od.y);
    float tz0 = __fmaf_rn(lo.z, idir.z, od.z), tz1 = __fmaf_rn(hi.z, idir.z, od.z);
    float tn = fmaxf(fmaxf(fminf(tx0, tx1), fminf(ty0, ty1)), fmaxf(fminf(tz0, tz1), 0.0f));
    float tf = fminf(fminf(fmaxf(tx0, tx1), fmaxf(ty0, ty1)), fminf(fmaxf(tz0, tz1), tmax));
    tn_out = tn;
    return tn <= tf;
}

// Walk a GPU-built mesh tree (bvh_build.cu): 64-byte nodes holding both child boxes, nearer child
// first, the farther one pushed with its entry parameter so that it is dropped without a fetch
// once a closer hit is known; kMeshLeafTris triangles per leaf. The radix tree is at most 64 levels deep (30
// Morton bits + index bits), which bounds the stack. Candidates go through the same triangle_test
// / offer_at as the reference-tree walk with the reference's leaf slot as rank, so the winner is
// the one the reference's own traversal keeps.
constexpr int kLbvhStack = 64;

template <bool COUNT>
MCPT_DEV void mesh_walk_lbvh(Work& wk, const DevMesh& M, f3 Oi, f3 Di, f3 Ow, const Aff& fwd, int prim, int tslot, float slack,
                             Best& best) {
    f3 wd = mk3(fwd.r0.x * Di.x + fwd.r0.y * Di.y + fwd.r0.z * Di.z, fwd.r1.x * Di.x + fwd.r1.y * Di.y + fwd.r1.z * Di.z,
                fwd.r2.x * Di.x + fwd.r2.y * Di.y + fwd.r2.z * Di.z);
    float scale = sqrtf(wd.x * wd.x + wd.y * wd.y + wd.z * wd.z);
    if (!(scale > 0.0f)) return;  // zero or NaN direction: the shader finds nothing either
    const float inv_scale = cull_rcp(scale);
    const f3 idir = safe_rcp3(Di);
    const f3 od = mk3(-Oi.x * idir.x, -Oi.y * idir.y, -Oi.z * idir.z);
    float tmax = (best.h.dist * 1.0001f + slack) * inv_scale;

    int stack_node[kLbvhStack];
    float stack_t[kLbvhStack];
    int sp = 0;
    int cur = 0;  // the root is node 0
    while (true) {
        if (cur >= 0) {
            const DevNode2* np = M.nodes2 + cur;
            const float4 lo0 = __ldg(&np->lo0), hi0 = __ldg(&np->hi0), lo1 = __ldg(&np->lo1), hi1 = __ldg(&np->hi1);
            float t0, t1;
            const bool h0 = slab_entry(lo0, hi0, idir, od, tmax, t0);
            const bool h1 = slab_entry(lo1, hi1, idir, od, tmax, t1);
            if (COUNT) wk.boxes += 2;
            const int c0 = __float_as_int(lo0.w), c1 = __float_as_int(hi0.w);
            if (h0 && h1) {
                const bool swap = t1 < t0;
                stack_node[sp] = swap ? c0 : c1;
                stack_t[sp] = swap ? t0 : t1;
                ++sp;
                cur = swap ? c1 : c0;
                continue;
            }
            if (h0 || h1) {
                cur = h0 ? c0 : c1;
                continue;
            }
        } else {
            const DevTri* tp = M.tris2 + kMeshLeafTris * (~cur);
#pragma unroll 1
            for (int j = 0; j < kMeshLeafTris; ++j, ++tp) {
                const float4 ta = __ldg(&tp->a);
                const int tri = __float_as_int(ta.w);
                if (tri < 0) continue;
                if (COUNT) ++wk.tris;
                const float4 tb = __ldg(&tp->b), tc = __ldg(&tp->c);
                float a;
                if (triangle_test(best, fwd, mk3(ta.x, ta.y, ta.z), mk3(tb.x, tb.y, tb.z), mk3(tc.x, tc.y, tc.z), Oi, Di, Ow,
                                  prim, tri, tslot, __float_as_int(tb.w), a))
                    tmax = (best.h.dist * 1.0001f + slack) * inv_scale;
            }
        }
        // pop the nearest pending subtree that can still hold a closer hit
        do {
            if (sp == 0) return;
            --sp;
            cur = stack_node[sp];
        } while (stack_t[sp] > tmax);
    }
}

// intersect_prim, raytracer_func.frag:681-705. MESHES = false compiles the mesh walk out
// (scenes made of analytic primitives only: fewer live registers).
// ALL_STAGED: every primitive record is in shared memory (flat scenes): plain LDS instead of a
// generic load behind a pointer select.
template <bool COUNT, bool MESHES = true, bool ALL_STAGED = false>
MCPT_DEV void prim_test(Work& wk, const DevScene& S, const SmemScene& sm, int p, f3 O, f3 D, float slack, Best& best) {
    const DevPrim* pp = (ALL_STAGED || p < S.smem_prims) ? (sm.prims + p) : (S.prims + p);
    const int type = pp->type;
    if (type < 0) return;
    Aff inv, fwd;
    inv.r0 = pp->inv.r0;
    inv.r1 = pp->inv.r1;
    inv.r2 = pp->inv.r2;
    fwd.r0 = pp->fwd.r0;
    fwd.r1 = pp->fwd.r1;
    fwd.r2 = pp->fwd.r2;
    const int tslot = pp->slot;
    if (COUNT) ++wk.prims;
    f3 Oi = xform_point(inv, O);
    f3 Di = normalize3(xform_vector(inv, D));
    Cand k;
    k.a1 = k.a2 = 0.0f;
    k.cl = 0;
    k.c1 = k.c2 = false;
    switch (type) {
        case MCPT_PRIM_MESH:
            if (MESHES) {
                const DevMesh& M = S.meshes[pp->mesh];
                if (M.nodes2)
                    mesh_walk_lbvh<COUNT>(wk, M, Oi, Di, O, fwd, p, tslot, slack, best);
                else
                    mesh_walk<COUNT>(wk, M, sm.mesh_nodes, Oi, Di, O, fwd, p, tslot, slack, best);
            }
            break;
        case MCPT_PRIM_SPHERE: sphere_test(k, Oi, Di, slack * pp->inv_smin); break;
        case MCPT_PRIM_CUBE: cube_test(k, Oi, Di); break;
        case MCPT_PRIM_CYLINDER: cylinder_test(k, Oi, Di); break;
        case MCPT_PRIM_CONE: cone_test(k, Oi, Di, O, S.leaf_bb + 6 * p); break;
        case MCPT_PRIM_ORIENTED_QUAD: quad_test(k, Oi, Di, O, S.leaf_bb + 6 * p); break;
        default: break;
    }
    if (k.c1) offer_at(best, fwd, Oi, Di, k.a1, O, p, k.cl, tslot, 0);
    if (k.c2) offer_at(best, fwd, Oi, Di, k.a2, O, p, k.cl, tslot, 0);
}

// traverse_all_bvh, raytracer_func.frag:777-781. D need not be unit length.
// any_hit: just_hit_bvh (:771-775) — stop at the first accepted candidate.
template <bool COUNT, bool MESHES = true>
MCPT_DEV void scene_walk(Work& wk, const DevScene& S, const SmemScene& sm, f3 O, f3 D, bool any_hit, Best& best) {
    best_reset(best);
    const float len = sqrtf(D.x * D.x + D.y * D.y + D.z * D.z);
    if (!(len > 0.0f) || !(len < FLT_MAX)) return;  // zero / NaN / inf direction: nothing is hit
    const float rl = (len < 1e-30f) ? 1.0f / len : cull_rcp(len);  // flush-to-zero would turn a denormal length into inf
    const f3 dn = mk3(D.x * rl, D.y * rl, D.z * rl);  // culling only
    const f3 idir = safe_rcp3(dn);
    const f3 od = mk3(-O.x * idir.x, -O.y * idir.y, -O.z * idir.z);
    const float slack = 1e-4f * (fabsf(O.x) + fabsf(O.y) + fabsf(O.z)) + 1e-3f;
    float tmax = FLT_MAX;

    const int depth = S.top_depth;
    const uint32_t first_leaf = 1u << depth;
    const uint32_t staged = uint32_t(S.smem_top_nodes);
    const DevNode* snodes = sm.top_nodes;  // 1-based
    const DevNode* gnodes = S.top_nodes;   // 1-based
    const uint32_t neg = (dn.x < 0.0f ? 1u : 0u) | (dn.y < 0.0f ? 2u : 0u) | (dn.z < 0.0f ? 4u : 0u);

    uint32_t k = 1, trail = 0;
    int level = 0, axis = 0;
    while (true) {
        DevNode n = (k <= staged) ? load_node_smem(snodes + k) : load_node(gnodes + k);
        bool inside = slab(n, idir, od, tmax);
        if (COUNT) ++wk.boxes;
        if (inside && k >= first_leaf) {
            int p = __float_as_int(n.lo.w);
            if (p >= 0) {
                prim_test<COUNT, MESHES>(wk, S, sm, p, O, D, slack, best);
                if (best.h.prim >= 0) {
                    if (any_hit) return;
                    tmax = best.h.dist * 1.0001f + slack;
                }
            }
            inside = false;
        }
        if (inside) {
            k = 2u * k + ((neg >> axis) & 1u);
            trail = (trail << 1) | 1u;
            ++level;
            axis = (axis == 2) ? 0 : axis + 1;
            continue;
        }
        while (level > 0 && !(trail & 1u)) {
            trail >>= 1;
            k >>= 1;
            --level;
            axis = (axis == 0) ? 2 : axis - 1;
        }
        if (level == 0) break;
        k ^= 1u;
        trail &= ~1u;
    }
}


// ------------------------------------------------------------- resumable walk
// The closest-hit search cut into single steps, for persistent warps (k_traverse, PERSIST): a lane
// whose ray is finished takes the next ray of its warp's queue region while its neighbours go on
// walking. To keep the lanes of a warp in the same instruction stream as often as possible, BOTH
// levels of the scene are walked by the same node step: the top-level tree is built at upload
// (SAH) in the 64-byte two-children node format of the mesh trees (DevScene::top2), and one explicit
// stack serves both levels. Entering a mesh pushes a marker, switches the culling frame (idir, od,
// tmax) to mesh-local space and parks the world frame in `wsave`; popping the marker restores it.
// Three kinds of step remain:
//   WALK_NODE  one node of either level: both (or all four, DevNode4) child boxes, nearest first,
//              the others pushed with their entry parameters (dropped when popped if a closer hit
//              is known by then)
//   WALK_TRI   one triangle of a mesh leaf
//   WALK_PRIM  a top-level leaf: the analytic intersector, or the entry into a mesh
// Candidates go through the same arithmetic and the same total order (distance, then the
// reference's visit order) as in scene_walk, so the result is the same record bit for bit.
enum { WALK_DONE = 0, WALK_NODE = 1, WALK_TRI = 2, WALK_PRIM = 3, WALK_POP = 4 };

constexpr int kTokEmpty = INT_MIN;         // child slot without content
constexpr int kTokExitMesh = INT_MIN + 1;  // stack marker: back to the world frame
constexpr int kTokPop = INT_MIN + 2;       // w.cur: the next thing to do is to take an entry off the stack
// MCPT_WIDE (default 1): on scenes with meshes the persistent walk steps through four-children nodes
// (DevNode4: every second level of both trees folded away at upload) -- half as many dependent node
// fetches, loop iterations and stack operations per ray: +5 % on the 1 M-triangle workload (1003 vs 955
// Msamples/s). Trees of analytic primitives stay on two-children nodes (measured: -6 % with wide nodes on
// the divergence-stress scene; their rays are short and the primitive step dominates).
#ifndef MCPT_WIDE
#define MCPT_WIDE 1
#endif
// wide: up to three siblings are pushed per level: 15 levels of the top-level tree (depth <= 30), the
// mesh-exit marker, 32 levels of a radix tree (<= 64 two-children levels)
constexpr int kWalkStackWide = 3 * 15 + 1 + 3 * (kLbvhStack / 2);
constexpr int kWalkStack = 30 + 1 + kLbvhStack;

// The explicit stack of the resumable walk lives in local memory. MCPT_STACK64 (default): one 8-byte
// entry {token, entry parameter} -- one local load per pop, one store per push; 0: two 4-byte arrays.
#ifndef MCPT_STACK64
#define MCPT_STACK64 1
#endif
#if MCPT_STACK64
struct WalkStack {
    int2* e;
    MCPT_DEV void put(int i, int tok, float t) const { e[i] = make_int2(tok, __float_as_int(t)); }
    MCPT_DEV void get(int i, int& tok, float& t) const {
        const int2 v = e[i];
        tok = v.x;
        t = __int_as_float(v.y);
    }
};
#define MCPT_WALK_STACK_DECL(name, n) int2 name##_e[n]; WalkStack name{name##_e}
#else
struct WalkStack {
    int* tok_;
    float* t_;
    MCPT_DEV void put(int i, int tok, float t) const {
        tok_[i] = tok;
        t_[i] = t;
    }
    MCPT_DEV void get(int i, int& tok, float& t) const {
        tok = tok_[i];
        t = t_[i];
    }
};
#define MCPT_WALK_STACK_DECL(name, n) int name##_tok[n]; float name##_t[n]; WalkStack name{name##_tok, name##_t}
#endif

struct WalkState {
    f3 O, D;         // world ray as given (D not normalised)
    float slack;
    f3 idir, od;     // culling frame
    float tmax;
    const void* nodes;  // node array of the current frame (DevNode2 or DevNode4)
    int cur;         // token: >= 0 node, < 0 leaf (~primitive in the world frame, ~leaf inside a mesh)
    int sp;
    bool done;
    int mesh_prim;   // >= 0 while inside that mesh primitive
    int tri_j;
    f3 Oi, Di;
    float inv_scale;
    int tslot;
    const DevTri* tris2;
    Best best;
};

template <bool WIDE>
MCPT_DEV void walk_begin(WalkState& w, const DevScene& S, f3 O, f3 D) {
    best_reset(w.best);
    w.O = O;
    w.D = D;
    w.mesh_prim = -1;
    w.done = true;
    const float len = sqrtf(D.x * D.x + D.y * D.y + D.z * D.z);
    if (!(len > 0.0f) || !(len < FLT_MAX)) return;  // zero / NaN / inf direction: nothing is hit
    const float rl = (len < 1e-30f) ? 1.0f / len : cull_rcp(len);  // flush-to-zero would turn a denormal length into inf
    const f3 dn = mk3(D.x * rl, D.y * rl, D.z * rl);  // culling only
    w.idir = safe_rcp3(dn);
    w.od = mk3(-O.x * w.idir.x, -O.y * w.idir.y, -O.z * w.idir.z);
    w.slack = 1e-4f * (fabsf(O.x) + fabsf(O.y) + fabsf(O.z)) + 1e-3f;
    w.tmax = FLT_MAX;
    w.nodes = WIDE ? (const void*)S.top4 : (const void*)S.top2;
    w.cur = 0;
    w.sp = 0;
    w.tri_j = 0;
    w.done = false;
}

// the step this ray needs next; WALK_DONE: finished (w.best holds the result)
MCPT_DEV int walk_kind(const WalkState& w) {
    if (w.done) return WALK_DONE;
    if (w.cur >= 0) return WALK_NODE;
    if (w.cur == kTokPop) return WALK_POP;
    return w.mesh_prim >= 0 ? WALK_TRI : WALK_PRIM;
}

// Take ONE entry off the stack (w.cur == kTokPop): the mesh-exit marker restores the world frame, an
// entry that can no longer hold a closer hit is dropped, anything else becomes the current token.
// One entry per iteration of the warp's loop, not a loop of its own: a lane that has several dead
// entries to drop would otherwise hold the other 31 (measured: 20 % of the kernel's instructions at
// 2.5 active lanes).
template <bool WIDE>
MCPT_DEV void walk_pop_one(WalkState& w, const DevScene& S, const WalkStack& st, const float* wsave) {
    if (w.sp == 0) {
        w.done = true;
        return;
    }
    --w.sp;
    int tok;
    float tok_t;
    st.get(w.sp, tok, tok_t);
    if (tok == kTokExitMesh) {
        w.mesh_prim = -1;
        w.nodes = WIDE ? (const void*)S.top4 : (const void*)S.top2;
        w.idir = mk3(wsave[0], wsave[1], wsave[2]);
        w.od = mk3(wsave[3], wsave[4], wsave[5]);
        w.tmax = (w.best.h.prim >= 0) ? (w.best.h.dist * 1.0001f + w.slack) : FLT_MAX;
    } else if (tok_t <= w.tmax) {
        w.cur = tok;
        w.tri_j = 0;
    }
}

// one child of a four-children node: entry parameter of the conservative slab test, +inf if culled
MCPT_DEV float slab_entry4(float lx, float ly, float lz, float hx, float hy, float hz, int tok, f3 idir, f3 od, float tmax) {
    const float tx0 = __fmaf_rn(lx, idir.x, od.x), tx1 = __fmaf_rn(hx, idir.x, od.x);
    const float ty0 = __fmaf_rn(ly, idir.y, od.y), ty1 = __fmaf_rn(hy, idir.y, od.y);
    const float tz0 = __fmaf_rn(lz, idir.z, od.z), tz1 = __fmaf_rn(hz, idir.z, od.z);
    const float tn = fmaxf(fmaxf(fminf(tx0, tx1), fminf(ty0, ty1)), fmaxf(fminf(tz0, tz1), 0.0f));
    const float tf = fminf(fminf(fmaxf(tx0, tx1), fmaxf(ty0, ty1)), fminf(fmaxf(tz0, tz1), tmax));
    return (tn <= tf && tok != kTokEmpty) ? tn : __int_as_float(0x7f800000);
}
MCPT_DEV void order2(float& ta, int& ca, float& tb, int& cb) {
    const bool sw = tb < ta;
    const float t = sw ? ta : tb;
    const int c = sw ? ca : cb;
    ta = sw ? tb : ta;
    ca = sw ? cb : ca;
    tb = t;
    cb = c;
}

// One four-children node: the nearest child that can hold a closer hit becomes the current token, the
// others are pushed farthest first with their entry parameters.
template <bool COUNT>
MCPT_DEV void walk_node4(WalkState& w, Work& wk, const DevScene& S, const WalkStack& st, const float* wsave) {
    const DevNode4* np = static_cast<const DevNode4*>(w.nodes) + w.cur;
    const float4 lox = __ldg(&np->lox), loy = __ldg(&np->loy), loz = __ldg(&np->loz);
    const float4 hix = __ldg(&np->hix), hiy = __ldg(&np->hiy), hiz = __ldg(&np->hiz);
    const int4 tok = __ldg(&np->tok);
    int c0 = tok.x, c1 = tok.y, c2 = tok.z, c3 = tok.w;
    float t0 = slab_entry4(lox.x, loy.x, loz.x, hix.x, hiy.x, hiz.x, c0, w.idir, w.od, w.tmax);
    float t1 = slab_entry4(lox.y, loy.y, loz.y, hix.y, hiy.y, hiz.y, c1, w.idir, w.od, w.tmax);
    float t2 = slab_entry4(lox.z, loy.z, loz.z, hix.z, hiy.z, hiz.z, c2, w.idir, w.od, w.tmax);
    float t3 = slab_entry4(lox.w, loy.w, loz.w, hix.w, hiy.w, hiz.w, c3, w.idir, w.od, w.tmax);
    if (COUNT) wk.boxes += 4;
    w.tri_j = 0;
    order2(t0, c0, t1, c1);
    order2(t2, c2, t3, c3);
    order2(t0, c0, t2, c2);
    order2(t1, c1, t3, c3);
    order2(t1, c1, t2, c2);
    const float inf = __int_as_float(0x7f800000);
    if (t3 < inf) st.put(w.sp++, c3, t3);
    if (t2 < inf) st.put(w.sp++, c2, t2);
    if (t1 < inf) st.put(w.sp++, c1, t1);
    w.cur = (t0 < inf) ? c0 : kTokPop;
}
template <bool COUNT>
MCPT_DEV void walk_node2(WalkState& w, Work& wk, const DevScene& S, const WalkStack& st, const float* wsave) {
    const DevNode2* np = static_cast<const DevNode2*>(w.nodes) + w.cur;
    const float4 lo0 = __ldg(&np->lo0), hi0 = __ldg(&np->hi0), lo1 = __ldg(&np->lo1), hi1 = __ldg(&np->hi1);
    const int c0 = __float_as_int(lo0.w), c1 = __float_as_int(hi0.w);
    float t0, t1;
    const bool h0 = slab_entry(lo0, hi0, w.idir, w.od, w.tmax, t0) && c0 != kTokEmpty;
    const bool h1 = slab_entry(lo1, hi1, w.idir, w.od, w.tmax, t1) && c1 != kTokEmpty;
    if (COUNT) wk.boxes += 2;
    w.tri_j = 0;
    if (h0 && h1) {
        const bool swap = t1 < t0;
        st.put(w.sp, swap ? c0 : c1, swap ? t0 : t1);
        ++w.sp;
        w.cur = swap ? c1 : c0;
    } else if (h0 || h1) {
        w.cur = h0 ? c0 : c1;
    } else {
        w.cur = kTokPop;
    }
}

template <bool COUNT, bool WIDE>
MCPT_DEV void walk_node(WalkState& w, Work& wk, const DevScene& S, const WalkStack& st, const float* wsave) {
    if (WIDE)
        walk_node4<COUNT>(w, wk, S, st, wsave);
    else
        walk_node2<COUNT>(w, wk, S, st, wsave);
}

template <bool COUNT>
MCPT_DEV void walk_tri(WalkState& w, Work& wk, const DevScene& S, const SmemScene& sm, const WalkStack& st, const float* wsave) {
    const DevTri* tp = w.tris2 + kMeshLeafTris * (~w.cur) + w.tri_j;
    const float4 ta = __ldg(&tp->a);
    const int tri = __float_as_int(ta.w);
    if (tri >= 0) {
        if (COUNT) ++wk.tris;
        const float4 tb = __ldg(&tp->b), tc = __ldg(&tp->c);
        float a;
        if (triangle_param(mk3(ta.x, ta.y, ta.z), mk3(tb.x, tb.y, tb.z), mk3(tc.x, tc.y, tc.z), w.Oi, w.Di, a)) {
            const DevPrim* pp = (w.mesh_prim < S.smem_prims) ? (sm.prims + w.mesh_prim) : (S.prims + w.mesh_prim);
            Aff fwd;
            fwd.r0 = pp->fwd.r0;
            fwd.r1 = pp->fwd.r1;
            fwd.r2 = pp->fwd.r2;
            if (offer_at(w.best, fwd, w.Oi, w.Di, a, w.O, w.mesh_prim, tri, w.tslot, __float_as_int(tb.w)))
                w.tmax = (w.best.h.dist * 1.0001f + w.slack) * w.inv_scale;
        }
    }
    if (++w.tri_j == kMeshLeafTris) w.cur = kTokPop;
}

template <bool COUNT, bool MESHES, bool WIDE>
MCPT_DEV void walk_prim(WalkState& w, Work& wk, const DevScene& S, const SmemScene& sm, const WalkStack& st, float* wsave) {
    const int pid = ~w.cur;
    const DevPrim* pp = (pid < S.smem_prims) ? (sm.prims + pid) : (S.prims + pid);
    if (MESHES && pp->type == MCPT_PRIM_MESH) {
        // intersect_prim + Mesh_intersect prologue (raytracer_func.frag:681-705)
        Aff inv, fwd;
        inv.r0 = pp->inv.r0;
        inv.r1 = pp->inv.r1;
        inv.r2 = pp->inv.r2;
        fwd.r0 = pp->fwd.r0;
        fwd.r1 = pp->fwd.r1;
        fwd.r2 = pp->fwd.r2;
        if (COUNT) ++wk.prims;
        const f3 Oi = xform_point(inv, w.O);
        const f3 Di = normalize3(xform_vector(inv, w.D));
        const f3 wd = mk3(fwd.r0.x * Di.x + fwd.r0.y * Di.y + fwd.r0.z * Di.z, fwd.r1.x * Di.x + fwd.r1.y * Di.y + fwd.r1.z * Di.z,
                          fwd.r2.x * Di.x + fwd.r2.y * Di.y + fwd.r2.z * Di.z);
        const float scale = sqrtf(wd.x * wd.x + wd.y * wd.y + wd.z * wd.z);
        if (scale > 0.0f) {  // else zero or NaN direction: the shader finds nothing either
            const DevMesh& M = S.meshes[pp->mesh];
            st.put(w.sp, kTokExitMesh, 0.0f);
            ++w.sp;
            w.nodes = WIDE ? (const void*)M.nodes4 : (const void*)M.nodes2;
            w.tris2 = M.tris2;
            w.mesh_prim = pid;
            w.tslot = pp->slot;
            w.Oi = Oi;
            w.Di = Di;
            w.inv_scale = cull_rcp(scale);
            wsave[0] = w.idir.x;
            wsave[1] = w.idir.y;
            wsave[2] = w.idir.z;
            wsave[3] = w.od.x;
            wsave[4] = w.od.y;
            wsave[5] = w.od.z;
            w.idir = safe_rcp3(Di);
            w.od = mk3(-Oi.x * w.idir.x, -Oi.y * w.idir.y, -Oi.z * w.idir.z);
            w.tmax = (w.best.h.dist * 1.0001f + w.slack) * w.inv_scale;
            w.cur = 0;
            w.tri_j = 0;
            return;
        }
    } else {
        prim_test<COUNT, false>(wk, S, sm, pid, w.O, w.D, w.slack, w.best);
        if (w.best.h.prim >= 0) w.tmax = w.best.h.dist * 1.0001f + w.slack;
    }
    w.cur = kTokPop;
}

// ---------------------------------------------------------- unified world-space walk
// Scenes with meshes, MCPT_OPT_UNIFIED (default): ONE tree of four-children nodes in world space (DevScene::u_nodes) --
// the top-level SAH tree with, below each placement of a mesh, a subtree fitted to the world positions of that
// instance's triangles. A ray never changes frame while it walks: no mesh-entry step, no exit marker on the stack, no
// parked world frame, and the state a lane carries shrinks from ~48 to ~30 registers. The exact tests still run in the
// primitive's own frame (Oi = inv * O, Di = normalize(inv * D), rebuilt at the leaf from the primitive record), with the
// reference's arithmetic, so hit records are the same bits as with the two-level walk. Two kinds of step remain:
//   node   four child boxes; the ray keeps, per axis, the index of the plane it enters through, so entry and exit
//          planes are loaded directly (no per-axis min / max) and every box is padded per ray (DevScene::u_pad_*)
//          by folding the padding into two offset vectors (odn, odf);
//   leaf   one triangle of one instance (transform + Moeller-Trumbore + offer), or one analytic primitive.
#ifndef MCPT_U_TOPREG
#define MCPT_U_TOPREG 0
#endif
#ifndef MCPT_U_PREFETCH
#define MCPT_U_PREFETCH 0  // bit 0: prefetch the triangle slot of a leaf when it is reached; bit 1: the second nearest child node
// (measured on the 1 M-triangle workload / repo scene 8, Msamples/s: off 1586 / 688, bit 0 1544 / 667, bit 1 1549 / 666, both 1529 / 655;
// triangle slots loaded around L1, MCPT_TRI_NO_L1: 1560 / 688 -- the loads these would hide are not what the warps wait for)
#endif
struct UWalk {
    f3 O, D;           // world ray as given
    f3 idir, odn, odf; // culling frame: t_entry = plane * idir + odn, t_exit = plane * idir + odf
    float tmax, slack;
    uint32_t nx, ny, nz;  // float4 index of the entry plane per axis inside a DevNode4 (exit plane: ^ 1)
    int cur, sp;
    bool done;
#if MCPT_U_TOPREG
    int top_tok;     // entry sp - 1 of the stack lives in these two registers (memory holds the ones below): the
    float top_t;     // load that refills them is issued when an entry is taken and used one pop later
#endif
    Best best;
};

#if MCPT_U_TOPREG
MCPT_DEV void upush(UWalk& w, const WalkStack& st, int tok, float t) {
    if (w.sp > 0) st.put(w.sp - 1, w.top_tok, w.top_t);
    w.top_tok = tok;
    w.top_t = t;
    ++w.sp;
}
#else
MCPT_DEV void upush(UWalk& w, const WalkStack& st, int tok, float t) { st.put(w.sp++, tok, t); }
#endif

MCPT_DEV void uwalk_begin(UWalk& w, const DevScene& S, f3 O, f3 D) {
    best_reset(w.best);
    w.O = O;
    w.D = D;
    w.done = true;
    w.cur = 0;
    w.sp = 0;
    const float len = sqrtf(D.x * D.x + D.y * D.y + D.z * D.z);
    if (!(len > 0.0f) || !(len < FLT_MAX)) return;  // zero / NaN / inf direction: nothing is hit
    const float rl = (len < 1e-30f) ? 1.0f / len : cull_rcp(len);
    const f3 dn = mk3(D.x * rl, D.y * rl, D.z * rl);  // culling only
    w.idir = safe_rcp3(dn);
    const f3 od = mk3(-O.x * w.idir.x, -O.y * w.idir.y, -O.z * w.idir.z);
    const float o1 = fabsf(O.x) + fabsf(O.y) + fabsf(O.z);
    w.slack = 1e-4f * o1 + 1e-3f;
    const float eps = S.u_pad_rel * (o1 + S.u_radius) + S.u_pad_abs;
    const f3 pe = mk3(eps * fabsf(w.idir.x), eps * fabsf(w.idir.y), eps * fabsf(w.idir.z));
    w.odn = mk3(od.x - pe.x, od.y - pe.y, od.z - pe.z);
    w.odf = mk3(od.x + pe.x, od.y + pe.y, od.z + pe.z);
    w.nx = (__float_as_uint(w.idir.x) >> 31);        // direction < 0: the ray enters through the max plane
    w.ny = 2u + (__float_as_uint(w.idir.y) >> 31);
    w.nz = 4u + (__float_as_uint(w.idir.z) >> 31);
    w.tmax = FLT_MAX;
    w.done = false;
}

MCPT_DEV float slab_entry_u(float nx, float ny, float nz, float fx, float fy, float fz, int tok, const UWalk& w) {
    const float tn = fmaxf(fmaxf(__fmaf_rn(nx, w.idir.x, w.odn.x), __fmaf_rn(ny, w.idir.y, w.odn.y)), fmaxf(__fmaf_rn(nz, w.idir.z, w.odn.z), 0.0f));
    const float tf = fminf(fminf(__fmaf_rn(fx, w.idir.x, w.odf.x), __fmaf_rn(fy, w.idir.y, w.odf.y)), fminf(__fmaf_rn(fz, w.idir.z, w.odf.z), w.tmax));
    return (tn <= tf && tok != kTokEmpty) ? tn : __int_as_float(0x7f800000);
}

template <bool COUNT>
MCPT_DEV void uwalk_node(UWalk& w, Work& wk, const DevScene& S, const SmemScene& sm, const WalkStack& st) {
    float4 nxp, fxp, nyp, fyp, nzp, fzp;
    int4 tok;
    if (w.cur < S.u_smem_nodes) {  // the top levels: shared memory
        const float4* q = reinterpret_cast<const float4*>(sm.u_nodes + w.cur);
        nxp = q[w.nx], fxp = q[w.nx ^ 1u];
        nyp = q[w.ny], fyp = q[w.ny ^ 1u];
        nzp = q[w.nz], fzp = q[w.nz ^ 1u];
        tok = *reinterpret_cast<const int4*>(q + 6);
    } else {
        const float4* q = reinterpret_cast<const float4*>(S.u_nodes + w.cur);
        nxp = __ldg(q + w.nx), fxp = __ldg(q + (w.nx ^ 1u));
        nyp = __ldg(q + w.ny), fyp = __ldg(q + (w.ny ^ 1u));
        nzp = __ldg(q + w.nz), fzp = __ldg(q + (w.nz ^ 1u));
        tok = __ldg(reinterpret_cast<const int4*>(q + 6));
    }
    int c0 = tok.x, c1 = tok.y, c2 = tok.z, c3 = tok.w;
    float t0 = slab_entry_u(nxp.x, nyp.x, nzp.x, fxp.x, fyp.x, fzp.x, c0, w);
    float t1 = slab_entry_u(nxp.y, nyp.y, nzp.y, fxp.y, fyp.y, fzp.y, c1, w);
    float t2 = slab_entry_u(nxp.z, nyp.z, nzp.z, fxp.z, fyp.z, fzp.z, c2, w);
    float t3 = slab_entry_u(nxp.w, nyp.w, nzp.w, fxp.w, fyp.w, fzp.w, c3, w);
    if (COUNT) wk.boxes += 4;
    order2(t0, c0, t1, c1);
    order2(t2, c2, t3, c3);
    order2(t0, c0, t2, c2);
    order2(t1, c1, t3, c3);
    order2(t1, c1, t2, c2);
    const float inf = __int_as_float(0x7f800000);
    if (t3 < inf) upush(w, st, c3, t3);
    if (t2 < inf) upush(w, st, c2, t2);
    if (t1 < inf) upush(w, st, c1, t1);
    w.cur = (t0 < inf) ? c0 : kTokPop;
#if MCPT_U_PREFETCH & 1
    // the triangle of a leaf that was just reached: the lane waits for the other lanes' node steps before the leaf stage
    // runs, the L2 -> L1 transfer of the slot overlaps them
    if (t0 < inf && c0 < 0 && ~c0 < S.u_n_tris) {
        const char* tp = reinterpret_cast<const char*>(S.u_tris + ~c0);
        asm volatile("prefetch.global.L1 [%0];" ::"l"(tp));
        asm volatile("prefetch.global.L1 [%0];" ::"l"(tp + 32));
    }
#endif
#if MCPT_U_PREFETCH & 2
    // the second nearest child, which is the entry most likely to be taken off the stack next
    if (t1 < inf && c1 >= 0) asm volatile("prefetch.global.L1 [%0];" ::"l"(reinterpret_cast<const char*>(S.u_nodes + c1)));
#endif
}

// one entry off the stack: dropped if it can no longer hold a closer hit
MCPT_DEV void uwalk_pop_one(UWalk& w, const WalkStack& st) {
    if (w.sp == 0) {
        w.done = true;
        return;
    }
    --w.sp;
    int tok;
    float tok_t;
#if MCPT_U_TOPREG
    tok = w.top_tok;
    tok_t = w.top_t;
    if (w.sp > 0) st.get(w.sp - 1, w.top_tok, w.top_t);
#else
    st.get(w.sp, tok, tok_t);
#endif
    if (tok_t <= w.tmax) w.cur = tok;
}

// A triangle slot is read once per ray that reaches it and 48 MB of them stream through a 256 KB L1 that also has to
// hold the nodes: MCPT_TRI_NO_L1 = 1 loads them around L1 (they stay in L2).
#ifndef MCPT_TRI_NO_L1
#define MCPT_TRI_NO_L1 0
#endif
MCPT_DEV float4 ldg_tri(const float4* p) {
#if MCPT_TRI_NO_L1
    float4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    return v;
#else
    return __ldg(p);
#endif
}

// leaf token: a triangle slot of a mesh instance, or an analytic primitive
template <bool COUNT>
MCPT_DEV void uwalk_leaf(UWalk& w, Work& wk, const DevScene& S, const SmemScene& sm) {
    const int L = ~w.cur;
    if (L < S.u_n_tris) {
        const DevTri* tp = S.u_tris + L;
        const float4 ta = ldg_tri(&tp->a), tb = ldg_tri(&tp->b), tc = ldg_tri(&tp->c);
        const int tri = __float_as_int(ta.w);
        if (tri >= 0) {
            if (COUNT) ++wk.tris;
            const int pid = __float_as_int(tc.w);
            const DevPrim* pp = (pid < S.smem_prims) ? (sm.prims + pid) : (S.prims + pid);
            Aff inv;
            inv.r0 = pp->inv.r0;
            inv.r1 = pp->inv.r1;
            inv.r2 = pp->inv.r2;
            // intersect_prim + Mesh_intersect prologue (raytracer_func.frag:681-705)
            const f3 Oi = xform_point(inv, w.O);
            const f3 Di = normalize3(xform_vector(inv, w.D));
            float a;
            if (triangle_param(mk3(ta.x, ta.y, ta.z), mk3(tb.x, tb.y, tb.z), mk3(tc.x, tc.y, tc.z), Oi, Di, a)) {
                Aff fwd;
                fwd.r0 = pp->fwd.r0;
                fwd.r1 = pp->fwd.r1;
                fwd.r2 = pp->fwd.r2;
                offer_at(w.best, fwd, Oi, Di, a, w.O, pid, tri, pp->slot, __float_as_int(tb.w));
            }
        }
    } else {
        prim_test<COUNT, false>(wk, S, sm, L - S.u_n_tris, w.O, w.D, w.slack, w.best);
    }
    if (w.best.h.prim >= 0) w.tmax = w.best.h.dist * 1.0001f + w.slack;
    w.cur = kTokPop;
}

// Warp-coherent walk for scenes with a handful of primitives (the reference's box
// scenes and the four-sphere scene): the leaf boxes form one flat list that every
// lane of the warp steps through TOGETHER. Each lane culls the box against its own
// ray; the exact primitive test runs under that predicate, and only if some lane of
// the warp needs it. All 32 lanes therefore execute the same instruction stream for
// the same primitive (same type: no switch divergence, no per-lane loop trip counts,
// one shared-memory broadcast per box). The winner is chosen by the same total order
// as in the tree walk, so the result does not depend on the visiting order.
// (Measured and dropped: regrouping the exact tests of a 256-ray tile by primitive through shared-memory
// bins after the first bounce -- cull, barrier, test 32 rays per primitive at full width, barrier, fold. On
// scattered rays the warp-coherent walk runs 2.9 test bodies per warp at 14 of 32 lanes (ncu, C2 second
// bounce) where a ray needs 1.1, and the bins did cut the instructions by 40 %, but three barriers per tile
// and a second, 10 %-full round of items put the warps of a CTA in lock step: C2 13.6 vs 12.1 ms of
// traversal per step (scene 1: 2.08 vs 2.20 ms).)
// Must be called by all 32 lanes of a warp; `valid` masks lanes without a ray.
template <bool COUNT>
MCPT_DEV void scene_walk_flat(Work& wk, const DevScene& S, const SmemScene& sm, f3 O, f3 D, bool any_hit, bool valid,
                              Best& best) {
    best_reset(best);
    const float len = sqrtf(D.x * D.x + D.y * D.y + D.z * D.z);
    valid = valid && (len > 0.0f) && (len < FLT_MAX);
    const float rl = (len < 1e-30f) ? 1.0f / len : cull_rcp(len);  // flush-to-zero would turn a denormal length into inf
    const f3 dn = mk3(D.x * rl, D.y * rl, D.z * rl);  // culling only
    const f3 idir = safe_rcp3(dn);
    const f3 od = mk3(-O.x * idir.x, -O.y * idir.y, -O.z * idir.z);
    const float slack = 1e-4f * (fabsf(O.x) + fabsf(O.y) + fabsf(O.z)) + 1e-3f;
    float tmax = FLT_MAX;
    const int n = S.n_flat;
    // the staged copy of the list for this ray's octant: lo = entry planes, hi = exit planes (stage_scene)
    const uint32_t oct = (__float_as_uint(idir.x) >> 31) | ((__float_as_uint(idir.y) >> 31) << 1) | ((__float_as_uint(idir.z) >> 31) << 2);
    const DevNode* boxes = sm.top_nodes + oct * uint32_t(n);
    for (int i = 0; i < n; ++i) {
        const DevNode nd = boxes[i];
        const bool pass = valid && slab_sorted(nd, idir, od, tmax);
        if (COUNT && valid) ++wk.boxes;
        if (__any_sync(0xffffffffu, pass)) {
            if (pass) {
                prim_test<COUNT, false, true>(wk, S, sm, __float_as_int(nd.lo.w), O, D, slack, best);
                if (best.h.prim >= 0) {
                    tmax = best.h.dist * 1.0001f + slack;
                    if (any_hit) valid = false;
                }
            }
        }
    }
}

}  // namespace mcpt
