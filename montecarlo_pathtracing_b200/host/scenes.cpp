// scenes.cpp — see scenes.h. Scene content follows
// MontecarloGPU/montecarlo.cpp:33-46 (palette), :143-179 (menger), :629-795.
#include "scenes.h"

#include <cmath>

namespace mcpt {
namespace {

using namespace Transfo;

// palette, montecarlo.cpp:33-44
const Vec4 ROUGE(0.9f, 0, 0, 1), VERT(0, 0.9f, 0, 1), BLEU(0, 0, 0.9f, 1), JAUNE(0.9f, 0.9f, 0, 1);
const Vec4 CYAN(0, 0.9f, 0.9f, 1), MAGENTA(0.9f, 0, 0.9f, 1), BLANC(0.9f, 0.9f, 0.9f, 1), NOIR(0, 0, 0, 1);

Vec4 OPA(Vec4 c, float o) {
    c.w = o;
    return c;
}

Mat4 quad_at(float tx, float ty, float tz, char axis, float deg, float sx, float sy) {
    Mat4 R = (axis == 'x') ? rotateX(deg) : (axis == 'y') ? rotateY(deg) : Mat4();
    return translate(tx, ty, tz) * R * scale(sx, sy, 1);
}

// Menger-sponge generator, montecarlo.cpp:143-179: 20 sub-cubes per level at
// offsets +-2/3 (edge and corner cells of a 3x3x3 lattice), each scaled sc/3.
void menger(BVH_GPU_Scene& s, const Mat4& m, int d, float sc, const Material& mater) {
    const float x = 2.0f / 3.0f, y = sc / 3.0f;
    static const signed char cell[20][3] = {
        {1, 1, 0},  {-1, 1, 0},  {-1, -1, 0},  {1, -1, 0},  {1, 0, 1},  {-1, 0, 1},  {-1, 0, -1},
        {1, 0, -1}, {0, 1, 1},   {0, -1, 1},   {0, -1, -1}, {0, 1, -1}, {1, 1, 1},   {-1, 1, 1},
        {-1, -1, 1}, {1, -1, 1}, {1, 1, -1},   {-1, 1, -1}, {-1, -1, -1}, {1, -1, -1}};
    for (const auto& c : cell) {
        Mat4 mm = m * (translate(x * c[0], x * c[1], x * c[2]) * scale(y));
        if (d > 0)
            menger(s, mm, d - 1, sc, mater);
        else
            s.add_cube(mm, mater);
    }
}

// the closed coloured box shared by scenes 1, 2 and 4
void scene_box_diffuse(BVH_GPU_Scene& s, float li) {
    s.add_orientedQuad(quad_at(0, 0, -100, ' ', 0, 100, 100), Material(BLANC));
    s.add_orientedQuad(quad_at(0, 0, 100, 'x', 180, 100, 100), Material(BLANC));
    s.add_orientedQuad(quad_at(0, 100, 0, 'x', 90, 100, 100), Material(CYAN));
    s.add_orientedQuad(quad_at(0, -100, 0, 'x', -90, 100, 100), Material(JAUNE));
    s.add_orientedQuad(quad_at(-100, 0, 0, 'y', 90, 100, 100), Material(ROUGE));
    s.add_orientedQuad(quad_at(100, 0, 0, 'y', -90, 100, 100), Material(VERT));
    s.add_cube(translate(70, 20, -40) * rotateZ(20) * scale(20, 20, 60), Material(BLANC));
    s.add_cube(translate(-70, 40, -40) * rotateZ(-20) * scale(20, 20, 60), Material(BLANC));
    s.add_orientedQuad(quad_at(0, 0, 99, 'x', 180, 40, 40), Material::light(BLANC, 10 * li));
}

void box_objects(BVH_GPU_Scene& s, float a_big, float a_right, bool right_opa) {
    s.add_cube(translate(70, 20, -60) * rotateZ(20) * scale(20, 20, 40), Material(ROUGE));
    s.add_cube(translate(-70, 40, -60) * rotateZ(-20) * scale(20, 20, 40), Material(VERT));
    s.add_sphere(translate(0, 50, -80) * scale(20), Material(MAGENTA, 0.8f, 0.995f));
    s.add_sphere(translate(0, -30, 0) * scale(40), Material(OPA(JAUNE, a_big), 0.65f, 1));
    s.add_sphere(translate(70, 20, 5) * scale(20), Material(right_opa ? OPA(ROUGE, a_right) : ROUGE, 0.8f, 0.95f));
    s.add_sphere(translate(-70, 40, 5) * scale(20), Material(VERT, 0.7f, 0.9f));
}

void scene_box_balls(BVH_GPU_Scene& s, float li) {
    s.add_orientedQuad(quad_at(0, 0, -100, ' ', 0, 100, 100), Material(BLANC));
    s.add_orientedQuad(quad_at(0, 0, 100, 'x', 180, 100, 100), Material(BLANC));
    s.add_orientedQuad(quad_at(0, 100, 0, 'x', 90, 100, 100), Material(CYAN));
    s.add_orientedQuad(quad_at(0, 99, 0, 'x', 90, 40, 60), Material(BLANC, 1, 1));
    s.add_orientedQuad(quad_at(0, -100, 0, 'x', -90, 100, 100), Material(BLANC));
    s.add_orientedQuad(quad_at(-100, 0, 0, 'y', 90, 100, 100), Material(BLANC));
    s.add_orientedQuad(quad_at(100, 0, 0, 'y', -90, 100, 100), Material(BLANC));
    box_objects(s, 0.5f, 0.2f, true);
    s.add_orientedQuad(quad_at(0, 0, 99, 'x', 180, 40, 40), Material::light(BLANC, 12.0f * li));
}

void scene_box_no_top(BVH_GPU_Scene& s, float li) {
    s.add_orientedQuad(quad_at(0, 0, -100, ' ', 0, 100, 100), Material(BLANC));
    s.add_orientedQuad(quad_at(0, 100, 0, 'x', 90, 100, 100), Material(CYAN));
    s.add_orientedQuad(quad_at(0, 99, 0, 'x', 90, 40, 60), Material(BLANC, 1, 1));
    s.add_orientedQuad(quad_at(0, -100, 0, 'x', -90, 100, 100), Material(BLANC));
    s.add_orientedQuad(quad_at(-100, 0, 0, 'y', 90, 100, 100), Material(BLANC));
    s.add_orientedQuad(quad_at(100, 0, 0, 'y', -90, 100, 100), Material(BLANC));
    box_objects(s, 0.1f, 1.0f, false);
    s.add_orientedQuad(quad_at(99, -10, -40, 'y', -90, 60, 5), Material::light(BLANC, 10 * li));
}

void scene_menger(BVH_GPU_Scene& s, float) {
    s.add_orientedQuad(translate(0, 0, -100) * scale(9000, 9000, 1), Material(BLANC, 0.8f, 0.999f));
    menger(s, translate(0, 0, -50) * rotateZ(15) * scale(50), 1, 0.9f, Material(MAGENTA));
    const Vec4 cyl_col[4] = {BLEU, VERT, ROUGE, JAUNE};
    const float sx[4] = {80, -80, -80, 80}, sy[4] = {80, 80, -80, -80};
    for (int k = 0; k < 4; ++k)
        s.add_cylinder(translate(sx[k], sy[k], -75) * scale(15, 15, 25), Material(cyl_col[k]));
    s.add_sphere(translate(80, 80, -30) * scale(20), Material(CYAN, 0.6f, 0.998f));
    s.add_sphere(translate(-80, 80, -30) * scale(20), Material(OPA(VERT, 0.1f), 0.7f, 0.5f));
    s.add_sphere(translate(-80, -80, -30) * scale(20), Material(ROUGE, 0.95f, 0.97f));
    s.add_sphere(translate(80, -80, -30) * scale(20), Material(OPA(JAUNE, 0.25f), 0.5f, 0.999f));
    s.add_sphere(translate(0, 0, -50) * scale(20), Material(BLANC, 1, 1));
}

void scene_materials(BVH_GPU_Scene& s, float) {
    s.add_cube(translate(0, 0, -50) * scale(9000, 9000, 1), Material(BLANC));
    for (int j = -5; j <= 5; j++)
        for (int i = -5; i <= 5; i++)
            s.add_sphere(translate(float(30 * i), float(30 * j), -41) * scale(8),
                         Material(ROUGE, float(1.0f - 0.075 * (i + 5)), 1.0f - 0.01f * (j + 5)));
}

void scene_4boules(BVH_GPU_Scene& s, float li) {
    s.add_cube(translate(0, 0, -51) * scale(9000, 9000, 1), Material(BLANC, 0.2f, 0.99999f));
    s.add_sphere(translate(110, 0, 0) * scale(50), Material(OPA(MAGENTA, 0.01f), 0.7f, 0.99f));
    s.add_sphere(translate(-110, 0, 0) * scale(50), Material(OPA(ROUGE, 0.15f), 0.5f, 0.5f));
    s.add_sphere(translate(0, 110, 0) * scale(50), Material(OPA(CYAN, 0.05f), 0.8f, 0.7f));
    s.add_sphere(translate(0, -110, 0) * scale(50), Material(OPA(VERT, 0.25f), 0.7f, 0.9f));
    s.add_orientedQuad(translate(200, 0, 100) * rotateY(-110) * scale(20, 20, 1), Material::light(BLANC, 20 * li));
}

void scene_menger_lights(BVH_GPU_Scene& s, float li) {
    s.add_cube(translate(0, 0, -10) * scale(9975, 9975, 1), Material(BLANC, 0.5f, 0.9f));
    menger(s, translate(0, 0, 42) * rotateZ(15) * scale(50.0f), 1, 0.9f, Material(ROUGE));
    menger(s, translate(-105, 0, 11) * scale(20.0f), 0, 0.7f, Material(BLEU));
    menger(s, translate(0, -105, 11) * scale(20.0f), 0, 0.7f, Material(CYAN));
    menger(s, translate(0, 105, 11) * scale(20.0f), 0, 0.7f, Material(MAGENTA));
    menger(s, translate(105, 0, 11) * scale(20.0f), 0, 0.7f, Material(JAUNE));
    s.add_sphere(translate(-100, -100, 5) * scale(15), Material(Vec4(1, 1, 1, 0.3f), 0.99f, 0.6f));
    s.add_sphere(translate(-100, 100, 5) * scale(15), Material(Vec4(1, 0, 1, 0.2f), 0.8f, 0.4f));
    s.add_sphere(translate(100, 100, 5) * scale(15), Material(Vec4(1, 1, 0, 0.4f), 0.6f, 0.2f));
    s.add_sphere(translate(100, -100, 5) * scale(15), Material(Vec4(0, 1, 0, 0.1f), 0.4f, 0.1f));
    s.add_cube(translate(0, 0, 500) * scale(1000, 1000, 1), Material(NOIR));  // dark roof
    s.add_sphere(translate(0, 0, 42) * scale(10), Material::light(BLANC, 10 * li));
    s.add_sphere(translate(-105, 0, 11) * scale(5), Material::light(BLANC, 10 * li));
    s.add_sphere(translate(105, 0, 11) * scale(5), Material::light(BLANC, 10 * li));
    s.add_sphere(translate(0, 105, 11) * scale(5), Material::light(BLANC, 10 * li));
    s.add_sphere(translate(0, -105, 11) * scale(5), Material::light(BLANC, 10 * li));
}

void scene_colonnes(BVH_GPU_Scene& s, float) {
    Vec4 ground = 0.6f * BLANC + 0.4f * VERT;
    s.add_orientedQuad(translate(0, 0, -100) * scale(90000, 90000, 1), Material(ground, 0.7f, 0.9999f));
    for (int i = -1000; i <= 1000; i += 250)
        for (int j = -1000; j <= 1000; j += 250) {
            const float fi = float(i), fj = float(j);
            s.add_cylinder(translate(fi, fj, -98) * scale(60, 60, 2), Material(BLANC));
            s.add_cylinder(translate(fi, fj, -93) * scale(50, 50, 3), Material(BLANC));
            s.add_cylinder(translate(fi, fj, -85) * scale(30, 30, 5), Material(BLANC));
            s.add_cylinder(translate(fi, fj, 0) * scale(20, 20, 80), Material(BLANC));
            s.add_cube(translate(fi, fj, 90) * scale(30, 30, 10), Material(BLANC));
            for (int a = 45; a < 360; a += 90)
                s.add_cube(translate(fi, fj, 105) * rotateZ(float(a)) * translate(90, 0, 0) * scale(80, 10, 5),
                           Material(BLANC));
            s.add_cylinder(translate(fi + 125, fj + 125, 115) * scale(75, 75, 5), Material(BLANC));
            s.add_cylinder(translate(fi, fj, 115) * scale(65, 65, 5), Material(BLANC));
        }
    s.add_sphere(translate(150, 375, -70) * scale(30), Material(JAUNE, 0.5f, 0.999f));
    s.add_sphere(translate(100, 125, -70) * scale(30), Material(OPA(CYAN, 0.2f), 0.5f, 0.9f));
    s.add_cube(translate(125, -125, -80) * rotateZ(45) * scale(20), Material(ROUGE, 0.1f, 0.2f));
}

inline uint32_t hash32(uint32_t x) {
    x ^= x >> 16;
    x *= 0x7feb352dU;
    x ^= x >> 15;
    x *= 0x846ca68bU;
    x ^= x >> 16;
    return x;
}
inline float hash01(uint32_t a, uint32_t b) { return float(hash32(a * 0x9e3779b9U + hash32(b)) >> 8) * (1.0f / 16777216.0f); }

}  // namespace

bool build_repo_scene(BVH_GPU_Scene& s, int id, float li) {
    switch (id) {
        case 1: scene_box_diffuse(s, li); return true;
        case 2: scene_box_balls(s, li); return true;
        case 3: scene_menger(s, li); return true;
        case 4: scene_box_no_top(s, li); return true;
        case 5: scene_materials(s, li); return true;
        case 6: scene_4boules(s, li); return true;
        case 7: scene_menger_lights(s, li); return true;
        case 8: scene_colonnes(s, li); return true;
    }
    return false;
}

void build_scene6_roughness(BVH_GPU_Scene& s, float li, float roughness) {
    scene_4boules(s, li);
    ScenePrimitives& sp = s.primitives();
    for (int i = 0; i < sp.nb(); ++i) {
        PrimData& pd = sp.prim_data_mut()[i];
        if (!(pd.mat_info[2] > 0.0f)) pd.mat_info[1] = roughness;
    }
}

void build_divergence_stress(BVH_GPU_Scene& s, float li) {
    scene_materials(s, li);
    ScenePrimitives& sp = s.primitives();
    for (int i = 0; i < sp.nb(); ++i) {
        PrimData& pd = sp.prim_data_mut()[i];
        pd.color[3] = 0.3f;
        pd.mat_info[0] = 0.8f;
        pd.mat_info[1] = 0.1f;
    }
    s.add_orientedQuad(translate(200, 0, 100) * rotateY(-110) * scale(20, 20, 1), Material::light(BLANC, 20 * li));
}

SP_Mesh make_displaced_sphere(int tris_log2, uint32_t seed, float radius, float amplitude) {
    if (tris_log2 < 3) tris_log2 = 3;
    const int quads = 1 << (tris_log2 - 1);
    const int n_lat = 1 << ((tris_log2 - 1) / 2);
    const int n_lon = quads / n_lat;
    auto m = std::make_shared<Mesh>();
    m->vertices_.reserve(size_t(n_lat + 1) * (n_lon + 1));
    // low-frequency radial bumps built from a few hashed harmonics
    float amp[6], ph[6];
    int fu[6], fv[6];
    for (int k = 0; k < 6; ++k) {
        amp[k] = amplitude * (0.3f + 0.7f * hash01(seed, 4 * k)) / float(k + 1);
        ph[k] = 6.2831853f * hash01(seed, 4 * k + 1);
        fu[k] = 1 + int(hash01(seed, 4 * k + 2) * 5.0f);
        fv[k] = 1 + int(hash01(seed, 4 * k + 3) * 5.0f);
    }
    for (int a = 0; a <= n_lat; ++a) {
        float theta = 3.14159265f * float(a) / float(n_lat);
        for (int b = 0; b <= n_lon; ++b) {
            float phi = 6.2831853f * float(b % n_lon) / float(n_lon);
            float r = radius;
            for (int k = 0; k < 6; ++k) r += amp[k] * std::sin(float(fu[k]) * theta * 2.0f + ph[k]) * std::cos(float(fv[k]) * phi + ph[k]);
            r += 0.02f * amplitude * (hash01(seed ^ 0x5bd1e995U, uint32_t(a * (n_lon + 1) + (b % n_lon))) - 0.5f);
            float st = std::sin(theta), ct = std::cos(theta);
            m->vertices_.push_back(Vec3(r * st * std::cos(phi), r * st * std::sin(phi), r * ct));
        }
    }
    m->tri_indices.reserve(size_t(quads) * 6);
    const uint32_t row = uint32_t(n_lon + 1);
    for (int a = 0; a < n_lat; ++a)
        for (int b = 0; b < n_lon; ++b) {
            uint32_t v00 = a * row + b, v01 = v00 + 1, v10 = v00 + row, v11 = v10 + 1;
            m->tri_indices.insert(m->tri_indices.end(), {v00, v10, v11, v00, v11, v01});
        }
    m->compute_bb();
    m->compute_normals();
    return m;
}

void build_synthetic_meshes(BVH_GPU_Scene& s, int n_meshes, int tris_log2, uint32_t seed) {
    if (n_meshes < 1) n_meshes = 1;
    // a 9000^2 floor and one quad light (emissivity 12) so the top level has >= 2 records
    s.add_orientedQuad(translate(0, 0, -100) * scale(9000, 9000, 1), Material(BLANC, 0.0f, 0.9f));
    s.add_orientedQuad(translate(0, 0, 140) * rotateX(180) * scale(60, 60, 1), Material::light(BLANC, 12.0f));
    const Vec4 base[4] = {ROUGE, CYAN, JAUNE, VERT};
    int side = int(std::ceil(std::sqrt(double(n_meshes))));
    float pitch = 200.0f / float(side);
    for (int k = 0; k < n_meshes; ++k) {
        SP_Mesh mesh = make_displaced_sphere(tris_log2, seed + 7919u * uint32_t(k), 1.0f, 0.25f);
        int id = s.add_mesh(mesh);
        int gx = k % side, gy = k / side;
        float r = 0.38f * pitch * (0.8f + 0.4f * hash01(seed, 100 + k));
        float cx = -100.0f + pitch * (gx + 0.5f), cy = -100.0f + pitch * (gy + 0.5f);
        float cz = -100.0f + 1.3f * r + 60.0f * hash01(seed, 200 + k);
        Mat4 trf = translate(cx, cy, cz) * rotateZ(360.0f * hash01(seed, 300 + k)) * rotateX(360.0f * hash01(seed, 400 + k)) * scale(r);
        Vec4 col = base[k & 3];
        switch ((k >> 0) & 3) {  // cycle the four material cases of tp/montecarlo.frag:139-172
            case 0: s.place_mesh(id, trf, Material(col, 0.0f, 0.8f)); break;            // diffuse
            case 1: s.place_mesh(id, trf, Material(col, 0.8f, 0.3f)); break;            // reflective
            case 2: s.place_mesh(id, trf, Material(OPA(col, 0.3f), 0.0f, 0.5f)); break; // transparent
            default: s.place_mesh(id, trf, Material(OPA(col, 0.3f), 0.7f, 0.2f)); break; // both
        }
    }
}

}  // namespace mcpt
