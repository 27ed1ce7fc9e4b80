// shade.cuh — per-hit work of the bounce loop: RNG, geometric normal, material
// fetch, bounce sampling and the four material cases.
//
// Follows tp/montecarlo.frag:49-179 (sample_hemisphere, random_ray, rSchlick,
// random_path) and shaders/raytracer_func.frag:90-124 (xxhash32 RNG), :783-907
// (normal reconstruction, colour / material fetch). The RNG is integer and
// bit-exact; sin/cos/log/pow go through CUDA's libm (or MUFU with
// MCPT_OPT_FAST_MATH), so radiance agrees with a CPU evaluation to rounding, not
// bit for bit.
#pragma once
#include "intersect.cuh"

namespace mcpt {

// ---- RNG, raytracer_func.frag:90-124 ---------------------------------------
MCPT_DEV uint32_t xxhash32(uint3 p) {
    const uint32_t PRIME32_2 = 2246822519U, PRIME32_3 = 3266489917U;
    const uint32_t PRIME32_4 = 668265263U, PRIME32_5 = 374761393U;
    uint32_t h32 = p.z + PRIME32_5 + p.x * PRIME32_3;
    h32 = PRIME32_4 * ((h32 << 17) | (h32 >> (32 - 17)));
    h32 += p.y * PRIME32_3;
    h32 = PRIME32_4 * ((h32 << 17) | (h32 >> (32 - 17)));
    h32 = PRIME32_2 * (h32 ^ (h32 >> 15));
    h32 = PRIME32_3 * (h32 ^ (h32 >> 13));
    return h32 ^ (h32 >> 16);
}

// srand(): seed from screen_tc, numero_pass and date (:105-110)
MCPT_DEV uint3 rng_seed(float tcx, float tcy, int numero_pass, float date) {
    float k = float(1 + numero_pass);
    float wx = tcx * k * 1.125f;
    float wy = (float(numero_pass) * 3.14f + date) * k * 1.125f;
    float wz = tcy * k * 1.125f;
    return make_uint3(__float_as_uint(wx), __float_as_uint(wy), __float_as_uint(wz));
}

// The xxhash32 state after n draws is seed0 + n * (11, 43, 67) (mod 2^32), and seed0 is a function
// of the pixel, the pass and `date` only. So the wavefront stores the draw COUNT (4 bytes, in the
// spare lane of the direction) instead of the 12-byte state, and shade rebuilds the state.
// (Measured and dropped: screen_tc of every pixel from a table built next to the camera table instead of the two integer and two
// IEEE divisions below -- the extra gather costs k_shade more than the arithmetic: 7.8 -> 8.25 ms per config-2 step.)
MCPT_DEV uint3 path_seed(uint32_t p, uint32_t npix, int width, int height, int first_pass, float date, uint32_t n_draws) {
    const uint32_t k = p / npix;
    const uint32_t pix = p - k * npix;
    const int x = int(pix % uint32_t(width)), y = int(pix / uint32_t(width));
    const float u = (float(x) + 0.5f) / float(width);
    const float v = (float(y) + 0.5f) / float(height);
    uint3 s = rng_seed(u, v, first_pass + int(k), date);
    s.x += 11u * n_draws;
    s.y += 43u * n_draws;
    s.z += 67u * n_draws;
    return s;
}
// draws consumed between two states: (x1 - x0) / 11 in Z/2^32 (11^-1 = 0xBA2E8BA3)
MCPT_DEV uint32_t draws_between(uint3 s0, uint3 s1) { return (s1.x - s0.x) * 0xBA2E8BA3u; }

MCPT_DEV float random_float(uint3& s) {
    uint32_t m = xxhash32(s);
    m &= 0x007FFFFFu;
    m |= 0x3F800000u;
    float f = __uint_as_float(m);
    s.x += 11u;
    s.y += 43u;
    s.z += 67u;
    return f - 1.0f;
}

// ---- transcendental back-ends ---------------------------------------------
template <bool FAST>
struct Tm {
    static MCPT_DEV float sin_(float x) { return FAST ? __sinf(x) : sinf(x); }
    static MCPT_DEV float cos_(float x) { return FAST ? __cosf(x) : cosf(x); }
    static MCPT_DEV float log_(float x) { return FAST ? __logf(x) : logf(x); }
    static MCPT_DEV float pow_(float x, float y) { return FAST ? __powf(x, y) : powf(x, y); }
    // 1/sqrt(x): MUFU.RSQ (<= 2 ulp) or the IEEE sqrt + divide pair of normalize3
    static MCPT_DEV float rsqrt_(float x) { return FAST ? rsqrtf(x) : 1.0f / sqrtf(x); }
    static MCPT_DEV f3 normalize_(f3 a) { return a * rsqrt_(dot3(a, a)); }
};

#define MCPT_PI 3.14159274101257324f /* float(2*acos(0)) */

// sample_hemisphere, tp/montecarlo.frag:49-70
template <bool FAST>
MCPT_DEV f3 sample_hemisphere(float roughness, uint3& seed) {
    float alpha = roughness * roughness;
    float beta = 2.0f * MCPT_PI * random_float(seed);
    float tanTheta2 = -alpha * alpha * Tm<FAST>::log_(1.0f - random_float(seed));
    float cosTheta = Tm<FAST>::rsqrt_(1.0f + tanTheta2);
    float sinTheta = sqrtf(fmaxf(0.0f, 1.0f - cosTheta * cosTheta));
    f3 localDir = mk3(Tm<FAST>::cos_(beta) * sinTheta, Tm<FAST>::sin_(beta) * sinTheta, cosTheta);
    return Tm<FAST>::normalize_(localDir);
}

// random_ray, tp/montecarlo.frag:72-89: frame (U,V,D) around D, M * sample
template <bool FAST>
MCPT_DEV f3 random_ray(f3 D, float roughness, uint3& seed) {
    f3 W = Tm<FAST>::normalize_(mk3(D.x, D.y + 5.0f, D.z + 3.0f));
    f3 U = Tm<FAST>::normalize_(cross3(D, W));
    f3 V = Tm<FAST>::normalize_(cross3(D, U));
    f3 s = sample_hemisphere<FAST>(roughness, seed);
    // mat3(U,V,D) * s = U*s.x + V*s.y + D*s.z
    f3 r = mk3(U.x * s.x + V.x * s.y + D.x * s.z, U.y * s.x + V.y * s.y + D.y * s.z, U.z * s.x + V.z * s.y + D.z * s.z);
    return Tm<FAST>::normalize_(r);
}

// rSchlick, tp/montecarlo.frag:91-98 (I is the incoming direction, as written)
MCPT_DEV float r_schlick(f3 I, f3 N, float refract_ind) {
    float r0 = (refract_ind - 1.0f) / (refract_ind + 1.0f);
    r0 *= r0;
    float x = 1.0f - dot3(N, I);
    float v = r0 + (1.0f - r0) * x * x * x * x * x;
    return fminf(fmaxf(v, 0.0f), 1.0f);
}

// ---- geometric info of a hit, raytracer_func.frag:783-897 -------------------
// Returns false when the shape writes nothing (shape < 0): N, P keep their values.
MCPT_DEV bool intersection_info(const DevScene& S, int prim, int sub, f3 pl, int flat_face, f3& N, f3& Pg) {
    if (prim < 0) return false;
    const DevPrim* pp = S.prims + prim;
    Aff fwd;
    fwd.r0 = __ldg(&pp->fwd.r0);
    fwd.r1 = __ldg(&pp->fwd.r1);
    fwd.r2 = __ldg(&pp->fwd.r2);
    const int type = __ldg(&pp->type);
    f3 P = xform_point(fwd, pl);  // closest_intersection.pg
    f3 tip;                        // local point whose image gives the normal
    switch (type) {
        case MCPT_PRIM_MESH: {
            const DevMesh& M = S.meshes[__ldg(&pp->mesh)];
            uint32_t i0 = __ldg(M.idx + 3 * sub), i1 = __ldg(M.idx + 3 * sub + 1), i2 = __ldg(M.idx + 3 * sub + 2);
            f3 A = mk3(__ldg(M.pos + 3 * i0), __ldg(M.pos + 3 * i0 + 1), __ldg(M.pos + 3 * i0 + 2));
            f3 B = mk3(__ldg(M.pos + 3 * i1), __ldg(M.pos + 3 * i1 + 1), __ldg(M.pos + 3 * i1 + 2));
            f3 C = mk3(__ldg(M.pos + 3 * i2), __ldg(M.pos + 3 * i2 + 1), __ldg(M.pos + 3 * i2 + 2));
            if (flat_face) {
                tip = pl + cross3(B - A, C - A);
            } else {
                f3 nA = mk3(__ldg(M.nrm + 3 * i0), __ldg(M.nrm + 3 * i0 + 1), __ldg(M.nrm + 3 * i0 + 2));
                f3 nB = mk3(__ldg(M.nrm + 3 * i1), __ldg(M.nrm + 3 * i1 + 1), __ldg(M.nrm + 3 * i1 + 2));
                f3 nC = mk3(__ldg(M.nrm + 3 * i2), __ldg(M.nrm + 3 * i2 + 1), __ldg(M.nrm + 3 * i2 + 2));
                f3 PA = A - pl, PB = B - pl, PC = C - pl;
                float tA = length3(cross3(PB, PC));
                float tB = length3(cross3(PA, PC));
                float tC = length3(cross3(PA, PB));
                f3 No = nA * tA + nB * tB + nC * tC;
                tip = pl + No;
            }
            break;
        }
        case MCPT_PRIM_SPHERE: tip = pl * 2.0f; break;
        case MCPT_PRIM_CUBE: {
            f3 No = mk3(0.0f, 0.0f, 0.0f);
            float s = (sub % 2 != 0) ? 1.0f : -1.0f;
            int ax = sub / 2;
            if (ax == 0) No.x = s; else if (ax == 1) No.y = s; else No.z = s;
            tip = pl + No;
            break;
        }
        case MCPT_PRIM_CYLINDER: {
            f3 No = mk3(0.0f, 0.0f, 0.0f);
            if (sub < 2) No.z = (sub % 2 != 0) ? 1.0f : -1.0f;
            else No = mk3(pl.x, pl.y, 0.0f);
            tip = pl + No;
            break;
        }
        case MCPT_PRIM_CONE: {
            if (sub == 0) {
                tip = mk3(pl.x, pl.y, pl.z - 1.0f);
            } else if (sub == 2) {
                f3 No = mk3(pl.x, pl.y, sqrtf(pl.x * pl.x + pl.y * pl.y) / 2.0f);
                tip = pl + No;
            } else {
                Pg = P;
                if (sub == 1) N = mk3(0.0f, 0.0f, 0.0f);
                return true;
            }
            break;
        }
        case MCPT_PRIM_ORIENTED_QUAD: tip = pl + mk3(0.0f, 0.0f, 1.0f); break;
        default: return false;
    }
    Pg = P;
    N = normalize3(xform_point(fwd, tip) - P);
    return true;
}

// ---- one iteration of random_path's loop body after a confirmed hit ---------
// tp/montecarlo.frag:121-175. Outputs:
//   kind 0: path ends, `total` is the value returned            (:174-175)
//   kind 1: continue with ray (nO, nD) and attenuation new_att   (:141-144,156-159,169-172)
//   kind 2: interior ray (nO, nD) must be traced first; the exit ray is derived
//           from that hit by finish_interior()                   (:146-153,160-166)
struct BounceOut {
    int kind;
    f3 nO, nD, new_att;
};

// What next-event estimation needs to know about the ray a bounce continues with (MCPT_OPT_NEE): the cone it was drawn
// from -- random_ray(axis, rho) -- and the attenuation the path would carry along ANY direction w of that cone:
// new_att(w) = base + gain * spec(w), spec(w) = pow(max(0, dot(E, reflect(-w, N))), se) in the diffuse case (the
// shader's `spec` is a function of the sampled ray there), spec = 1 in the reflective case (gain already holds it).
struct NeeInfo {
    int cone;  // 0: no cone (interior ray, path end), 1: diffuse case, 2: reflective case
    f3 axis, E, N, base, gain;
    float rho, se;
};

// solid-angle density of random_ray(axis, rho) at a unit direction whose cosine to the axis is c (0 behind the
// axis plane): sample_hemisphere draws tan^2(theta) = -a^2 ln(1 - u), a = rho^2, and a uniform azimuth
MCPT_DEV float cone_pdf(float c, float rho) {
    const float a2 = rho * rho * rho * rho;
    if (!(c > 1e-6f) || !(a2 > 1e-12f)) return 0.0f;
    const float c2 = c * c, t2 = (1.0f - c2) / c2;
    return __expf(-t2 / a2) / (MCPT_PI * a2 * c2 * c);
}

template <bool FAST>
MCPT_DEV BounceOut bounce(f3 O, f3 D, f3 N, f3 P, float4 mat, float4 col, f3 attenu, f3& total, uint3& seed, float refract_ind,
                          NeeInfo* ni = nullptr) {
    BounceOut out;
    if (ni) ni->cone = 0;
    f3 c3 = mk3(col.x, col.y, col.z);
    f3 ray = random_ray<FAST>(N, 1.0f - mat.y, seed);
    float rs = r_schlick(D, N, refract_ind);
    f3 R = reflect3(-ray, N);
    // E = normalize(O - P) is the reversed ray direction (P lies on the ray from O along D). The
    // fast variant uses -D / |D| and never reads the ray origin back from the wavefront.
    f3 E = FAST ? -Tm<FAST>::normalize_(D) : Tm<FAST>::normalize_(O - P);
    float se = mixf(100.0f, 2.0f, mat.y);
    float spec = Tm<FAST>::pow_(fmaxf(0.0f, dot3(E, R)), se);

    // total += col.xyz * 0.1 + attenu * mat.b * (1.0-mat.r) * col.a
    total = total + (c3 * 0.1f + attenu * mat.z * (1.0f - mat.x) * col.w);

    if (!(mat.z <= 0.5f)) {
        out.kind = 0;
        return out;
    }
    f3 mixed = mix3(attenu, c3, mat.x);
    // Which of the four material cases (tp/montecarlo.frag:139-172) applies is decided first, so that the
    // reflective and the transparent formulas -- and the second random_ray -- exist once in the code,
    // not once per case that uses them. The draw order is the shader's: the reflect / refract choice of
    // the transparent+reflective case comes before the draws of the reflected ray.
    bool reflective = false, interior = false, refract_entry = false;
    if (mat.x > 0.0f && col.w == 1.0f) {
        reflective = true;
    } else if (col.w < 1.0f && mat.x == 0.0f) {
        interior = true;
        refract_entry = true;
    } else if (col.w < 1.0f && mat.x > 0.0f) {
        float r = random_float(seed);
        if (r > 0.5f)
            reflective = true;
        else
            interior = true;  // D is not refracted on entry (:160-165)
    }
    if (reflective) {
        out.new_att = c3 * attenu + attenu * col.w * rs * spec * mixed;
        out.nO = P + N * MCPT_BIAS;
        out.nD = random_ray<FAST>(reflect3(D, N), 1.0f - mat.x * mat.y, seed);
        out.kind = 1;
        if (ni) {
            ni->cone = 2;
            ni->axis = Tm<FAST>::normalize_(reflect3(D, N));
            ni->rho = 1.0f - mat.x * mat.y;
            ni->base = out.new_att;
            ni->gain = mk3(0.0f, 0.0f, 0.0f);
            ni->E = E;
            ni->N = N;
            ni->se = se;
        }
    } else if (interior) {
        out.new_att = c3 * attenu + attenu * (1.0f - col.w) * (1.0f - rs) * spec * mixed;
        out.nO = P - N * MCPT_BIAS;
        out.nD = refract_entry ? refract3(D, N, refract_ind) : D;
        out.kind = 2;
    } else {
        out.new_att = c3 * attenu + attenu * spec * mixed;
        out.nO = P + N * MCPT_BIAS;
        out.nD = ray;
        out.kind = 1;
        if (ni) {
            ni->cone = 1;
            ni->axis = N;
            ni->rho = 1.0f - mat.y;
            ni->base = c3 * attenu;
            ni->gain = attenu * mixed;
            ni->E = E;
            ni->N = N;
            ni->se = se;
        }
    }
    return out;
}

// fake sky, tp/montecarlo.frag:118-119
MCPT_DEV f3 sky(f3 total, f3 attenu, f3 D) {
    f3 c = mix3(mk3(0.5f, 0.5f, 0.9f), mk3(1.0f, 1.0f, 0.8f), fmaxf(0.0f, D.z));
    return total + attenu * c;
}

}  // namespace mcpt
