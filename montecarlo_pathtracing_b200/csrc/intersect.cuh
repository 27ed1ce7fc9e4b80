// intersect.cuh — ray / primitive tests in primitive-local space.
//
// Numerically these are the tests of shaders/raytracer_func.frag:354-640
// (Triangle_, Sphere_, OrientedQuad_, Cube_, Cylinder_, Cone_intersect): the same
// quantities, the same comparisons, the same EPSILON, evaluated operation by
// operation in IEEE binary32 without contraction, so that the world-space
// distance that orders candidates comes out identical to a CPU evaluation of the
// shader. The structure around them is different: candidates arrive from a
// stack-free, front-to-back traversal over compact nodes (traverse.cuh), and a
// candidate replaces the current closest hit by an explicit total order
// (distance, then the reference's visit order) so the winner does not depend on
// the order in which this library happens to visit the tree.
#pragma once
#include <cfloat>

#include "layout.cuh"

namespace mcpt {

// the shader's own constants (raytracer_func.frag:8,13-14); its FLT_MAX is 3.402823e38, not the IEEE one
#define MCPT_FLT_MAX 3.402823e38f
#define MCPT_EPSILON 1e-10f
#define MCPT_BIAS 1e-2f

// The reference keeps the first candidate it meets among equal distances
// (strict '<' at raytracer_func.frag:384,413,...). It meets top-level leaves in
// decreasing slot order (right child is popped first, :759-767), and triangles of
// a mesh in decreasing leaf-slot order (:668-676). `tslot`/`rank` carry those
// two slots; candidates of one analytic primitive are offered in shader order.
struct Best {
    Hit h;
    int32_t tslot;
};

MCPT_DEV void best_reset(Best& b) {
    b.h.dist = MCPT_FLT_MAX;
    b.h.pl = mk3(0.0f, 0.0f, 0.0f);
    b.h.prim = -1;
    b.h.sub = -1;
    b.h.rank = -1;
    b.tslot = -1;
}

MCPT_DEV bool offer(Best& b, float dist, f3 pl, int prim, int sub, int tslot, int rank) {
    bool take = dist < b.h.dist;
    if (!take && dist == b.h.dist && b.h.prim >= 0)
        take = (tslot > b.tslot) || (tslot == b.tslot && rank > b.h.rank);
    if (take) {
        b.h.dist = dist;
        b.h.pl = pl;
        b.h.prim = prim;
        b.h.sub = sub;
        b.h.rank = rank;
        b.tslot = tslot;
    }
    return take;
}

// candidate at parameter a along the local ray: Pl = O + a*D, Pg = trf * Pl,
// dist = |Ol - Pg| (e.g. raytracer_func.frag:409-413)
MCPT_DEV bool offer_at(Best& b, const Aff& fwd, f3 O, f3 D, float a, f3 Ow, int prim, int sub, int tslot, int rank) {
    f3 Pl = O + D * a;
    f3 Pg = xform_point(fwd, Pl);
    float dist = distance3(Ow, Pg);
    return offer(b, dist, Pl, prim, sub, tslot, rank);
}

// The reference only reaches a primitive whose leaf box passes intersect_bv
// (raytracer_func.frag:314-352): origin strictly inside the box, or the box entered
// from the front. For every intersector that demands a > EPSILON a valid hit implies
// that, so the conservative slab test upstream is enough. The quad and the cone
// quadric accept hits BEHIND the origin (no a > 0 test); those are reachable in the
// reference only when the origin is inside the leaf box, which is evaluated here
// with the shader's own arithmetic on the un-inflated box.
MCPT_DEV bool origin_inside_leaf_box(const float* __restrict__ bb, f3 O) {
    f3 mn = mk3(__ldg(bb + 0), __ldg(bb + 1), __ldg(bb + 2));
    f3 mx = mk3(__ldg(bb + 3), __ldg(bb + 4), __ldg(bb + 5));
    f3 center = (mn + mx) / 2.0f;
    f3 width = (mx - mn) * 0.5f;
    f3 Oi = mk3((O.x - center.x) / width.x, (O.y - center.y) / width.y, (O.z - center.z) / width.z);
    return fabsf(Oi.x) < 1.0f && fabsf(Oi.y) < 1.0f && fabsf(Oi.z) < 1.0f;
}

// raytracer_func.frag:398-441
// The intersectors only find the candidate parameters (a1 first, then a2, in shader order); prim_test
// offers them in ONE place after its type switch, so the closest-hit registers are updated by a single
// copy of offer_at instead of one inlined copy per primitive type.
struct Cand {
    float a1, a2;
    int cl;        // face id `dir` of the candidate(s)
    bool c1, c2;
};

// tol: the rounding noise of a world-space distance from this ray's origin, in local units (prim_test: the
// culling slack, ~1e-4 |Ow|, over the smallest scale of the primitive's transform)
MCPT_DEV void sphere_test(Cand& k, f3 O, f3 D, float tol) {
    float OO = dot3(O, O);
    float OD = dot3(O, D);
    float D2 = dot3(D, D);
    float delta4 = OD * OD - D2 * (OO - 1.0f);
    if (delta4 > 0.0f) {
        float sq = sqrtf(delta4);
        float a = -(OD + sq) / D2;
        if (a > MCPT_EPSILON) {
            k.a1 = a;
            k.c1 = true;
            // The far root loses against the near one under the shader's strict '<' whenever the two are
            // separated by more than the rounding noise of dist = |Ow - Pg| (which grows with |Ow|, not with
            // the hit distance: hence tol); it is evaluated only when they are not.
            if (a > 1e-3f && 2.0f * sq > (1e-3f * a + tol) * D2) return;
        }
        a = -(OD - sq) / D2;
        if (a > MCPT_EPSILON) {
            k.a2 = a;
            k.c2 = true;
        }
    }
}

// raytracer_func.frag:443-470. One-sided (D.z < -EPSILON); note there is no
// a > 0 test in the shader, so there is none here.
MCPT_DEV void quad_test(Cand& k, f3 O, f3 D, f3 Ow, const float* leaf_bb) {
    if (D.z > -MCPT_EPSILON) return;
    float a = -O.z / D.z;
    if (!(a > 0.0f) && !origin_inside_leaf_box(leaf_bb, Ow)) return;  // hit behind the origin
    f3 Pl = O + D * a;
    if (fabsf(Pl.x) > 1.0f || fabsf(Pl.y) > 1.0f) return;
    k.a1 = a;  // offer_at rebuilds Pl = O + D * a with the same two roundings
    k.c1 = true;
}

// raytracer_func.frag:472-512: nearest of the six faces, face id -> dir
// (Measured and dropped: skipping the divide of faces whose numerator and denominator differ in sign -- no gain.)
MCPT_DEV void cube_test(Cand& k, f3 O, f3 D) {
    float al = MCPT_FLT_MAX;
    int cl = 0;
#pragma unroll
    for (int c = 0; c < 6; ++c) {
        const int c0 = c / 2, c1 = (c0 + 1) % 3, c2 = (c0 + 2) % 3;
        float d0 = comp(D, c0);
        if (fabsf(d0) > MCPT_EPSILON) {
            const float cd = (c % 2) ? 1.0f : -1.0f;
            float a = (cd - comp(O, c0)) / d0;
            if ((a > MCPT_EPSILON) && (fabsf(comp(O, c1) + a * comp(D, c1)) <= 1.0f) &&
                (fabsf(comp(O, c2) + a * comp(D, c2)) <= 1.0f)) {
                if (a < al) {
                    al = a;
                    cl = c;
                }
            }
        }
    }
    if (al < MCPT_FLT_MAX) {
        k.a1 = al;
        k.cl = cl;
        k.c1 = true;
    }
}

// raytracer_func.frag:515-577: caps z = -1 (dir 0), z = +1 (dir 1), side (dir 2)
MCPT_DEV void cylinder_test(Cand& k, f3 O, f3 D) {
    int cl = -1;
    float al = MCPT_FLT_MAX;
    if (fabsf(D.z) > MCPT_EPSILON) {
        float a = (-1.0f - O.z) / D.z;
        if (a > MCPT_EPSILON) {
            float rx = O.x + a * D.x, ry = O.y + a * D.y;
            if ((dot2(rx, ry, rx, ry) < 1.0f) && (a < al)) {
                cl = 0;
                al = a;
            }
        }
        a = (1.0f - O.z) / D.z;
        if (a > MCPT_EPSILON) {
            float rx = O.x + a * D.x, ry = O.y + a * D.y;
            if ((dot2(rx, ry, rx, ry) < 1.0f) && (a < al)) {
                cl = 1;
                al = a;
            }
        }
    }
    float O2 = dot2(O.x, O.y, O.x, O.y);
    float OD = dot2(O.x, O.y, D.x, D.y);
    float D2 = dot2(D.x, D.y, D.x, D.y);
    float delta4 = OD * OD - D2 * (O2 - 1.0f);
    if (delta4 > 0.0f) {
        float a = -(OD + sqrtf(delta4)) / D2;
        if ((a > MCPT_EPSILON) && (a < al)) {
            float z = O.z + a * D.z;
            if (fabsf(z) < 1.0f) {
                cl = 2;
                al = a;
            }
        }
    }
    if (al < MCPT_FLT_MAX) {
        k.a1 = al;
        k.cl = cl;
        k.c1 = true;
    }
}

// raytracer_func.frag:579-640: cap z = -1 (dir 0) and the quadric with k = 0.8,
// apex at z = +1 (dir 2); as in the shader the quadric roots are not tested for
// positivity.
MCPT_DEV void cone_test(Cand& k, f3 O, f3 D, f3 Ow, const float* leaf_bb) {
    int cl = -1;
    float tl = MCPT_FLT_MAX;
    if (fabsf(D.z) > MCPT_EPSILON) {
        float t0 = (-1.0f - O.z) / D.z;
        if (t0 > MCPT_EPSILON) {
            float rx = O.x + t0 * D.x, ry = O.y + t0 * D.y;
            if ((dot2(rx, ry, rx, ry) < 1.0f) && (t0 < tl)) {
                cl = 0;
                tl = t0;
            }
        }
    }
    f3 co = O;
    co.z -= 1.0f;
    float qa = D.z * D.z - 0.8f;
    float qb = 2.0f * (D.z * co.z - dot3(D, co) * 0.8f);
    float qc = co.z * co.z - dot3(co, co) * 0.8f;
    float det = qb * qb - 4.0f * qa * qc;
    if (det > 0.0f) {
        det = sqrtf(det);
        float t1 = (-qb - det) / (2.0f * qa);
        if (fabsf(O.z + t1 * D.z) > 1.0f) t1 = MCPT_FLT_MAX;
        float t2 = (-qb + det) / (2.0f * qa);
        if (fabsf(O.z + t2 * D.z) > 1.0f) t2 = MCPT_FLT_MAX;
        float t = (t2 < t1) ? t2 : t1;  // GLSL min(t1,t2)
        if (t < tl) {
            cl = 2;
            tl = t;
        }
    }
    if (tl < MCPT_FLT_MAX) {
        if (!(tl > 0.0f) && !origin_inside_leaf_box(leaf_bb, Ow)) return;  // hit behind the origin
        k.a1 = tl;
        k.cl = cl;
        k.c1 = true;
    }
}

// raytracer_func.frag:354-396 (Moeller-Trumbore, u,v in [0,1], a > EPSILON)
// the parameter test alone (Triangle_intersect up to `a > EPSILON`): true when the candidate at
// parameter a_out has to be offered
MCPT_DEV bool triangle_param(f3 A, f3 B, f3 C, f3 O, f3 D, float& a_out) {
    f3 edge1 = B - A;
    f3 edge2 = C - A;
    f3 h = cross3(D, edge2);
    float det = dot3(edge1, h);
    if (fabsf(det) < MCPT_EPSILON) return false;
    float invdet = 1.0f / det;
    f3 s = O - A;
    float u = dot3(s, h) * invdet;
    if (u < 0.0f || u > 1.0f) return false;
    f3 q = cross3(s, edge1);
    float v = dot3(D, q) * invdet;
    if (v < 0.0f || (u + v) > 1.0f) return false;
    a_out = dot3(edge2, q) * invdet;
    return a_out > MCPT_EPSILON;
}

MCPT_DEV bool triangle_test(Best& b, const Aff& fwd, f3 A, f3 B, f3 C, f3 O, f3 D, f3 Ow, int prim, int tri, int tslot,
                            int rank, float& a_out) {
    f3 edge1 = B - A;
    f3 edge2 = C - A;
    f3 h = cross3(D, edge2);
    float det = dot3(edge1, h);
    if (fabsf(det) < MCPT_EPSILON) return false;
    float invdet = 1.0f / det;
    f3 s = O - A;
    float u = dot3(s, h) * invdet;
    if (u < 0.0f || u > 1.0f) return false;
    f3 q = cross3(s, edge1);
    float v = dot3(D, q) * invdet;
    if (v < 0.0f || (u + v) > 1.0f) return false;
    float a = dot3(edge2, q) * invdet;
    if (a > MCPT_EPSILON) {
        a_out = a;
        return offer_at(b, fwd, O, D, a, Ow, prim, tri, tslot, rank);
    }
    return false;
}

}  // namespace mcpt
