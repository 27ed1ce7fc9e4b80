// oracle/ref_tables.cpp — TEST INFRASTRUCTURE (CPU oracle). See ref_tables.h.
#include "ref_tables.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <limits>

namespace oracle {

// ---- easycppogl/gl_eigen.cpp:30-121 equivalents live in oracle_capi.cpp (the
// tests pass finished matrices); here only product / inverse / apply.
M4 m4_identity() {
    M4 r;
    std::memset(r.m, 0, sizeof r.m);
    r.m[0] = r.m[5] = r.m[10] = r.m[15] = 1.0f;
    return r;
}

M4 m4_mul(const M4& a, const M4& b) {
    M4 r;
    for (int col = 0; col < 4; ++col)
        for (int row = 0; row < 4; ++row) {
            float s = 0.0f;
            for (int k = 0; k < 4; ++k) s += a.m[4 * k + row] * b.m[4 * col + k];
            r.m[4 * col + row] = s;
        }
    return r;
}

void m4_apply(const M4& a, const float v[4], float out[4]) {
    for (int row = 0; row < 4; ++row)
        out[row] = a.m[row] * v[0] + a.m[4 + row] * v[1] + a.m[8 + row] * v[2] + a.m[12 + row] * v[3];
}

// trf.inverse() (bvh_gpu/scene.cpp:48,62). Eigen is not available; the inverse is
// the adjugate over the determinant evaluated in double and rounded to float.
M4 m4_inverse(const M4& A) {
    double a[4][4], c[4][4];
    for (int col = 0; col < 4; ++col)
        for (int row = 0; row < 4; ++row) a[row][col] = A.m[4 * col + row];
    auto minor3 = [&](int r, int cc) {
        int rs[3], cs[3], n = 0;
        for (int i = 0; i < 4; ++i)
            if (i != r) rs[n++] = i;
        n = 0;
        for (int i = 0; i < 4; ++i)
            if (i != cc) cs[n++] = i;
        return a[rs[0]][cs[0]] * (a[rs[1]][cs[1]] * a[rs[2]][cs[2]] - a[rs[1]][cs[2]] * a[rs[2]][cs[1]]) -
               a[rs[0]][cs[1]] * (a[rs[1]][cs[0]] * a[rs[2]][cs[2]] - a[rs[1]][cs[2]] * a[rs[2]][cs[0]]) +
               a[rs[0]][cs[2]] * (a[rs[1]][cs[0]] * a[rs[2]][cs[1]] - a[rs[1]][cs[1]] * a[rs[2]][cs[0]]);
    };
    for (int r = 0; r < 4; ++r)
        for (int cc = 0; cc < 4; ++cc) c[r][cc] = (((r + cc) & 1) ? -1.0 : 1.0) * minor3(r, cc);
    double det = 0.0;
    for (int cc = 0; cc < 4; ++cc) det += a[0][cc] * c[0][cc];
    M4 R;
    for (int row = 0; row < 4; ++row)
        for (int col = 0; col < 4; ++col) R.m[4 * col + row] = float(c[col][row] / det);  // adj = cofactor^T
    return R;
}

// ---- bvh_gpu/scene.cpp:91-100
static Box merge(const Box& a, const Box& b) {
    Box r;
    for (int i = 0; i < 3; ++i) {
        r.mn[i] = std::min(a.mn[i], b.mn[i]);
        r.mx[i] = std::max(a.mx[i], b.mx[i]);
    }
    return r;
}

// ---- bvh_gpu/bvh.cpp:18-93, loop structure kept as in the reference
void KdTree::compute() {
    std::vector<int> splt, splt2;
    splt.push_back(0);
    splt.push_back(int(ids.size()));
    int d = 0;
    // the reference reads out of bounds for fewer than 2 items (bvh.cpp:44,59-68);
    // the oracle gives those a depth-1 tree.
    depth = ids.size() < 2 ? 1 : int(ceilf(log2f(float(ids.size()))));
    for (int i = 1; i < depth; ++i) {
        splt2.push_back(splt[0]);
        int nb = int(splt.size());
        for (int s = 1; s < nb; ++s) {
            int64_t j0 = splt[s - 1], j2 = splt[s], j1 = (j0 + j2) / 2;
            std::nth_element(ids.begin() + j0, ids.begin() + j1, ids.begin() + j2,
                             [&](int a, int b) { return centers[3 * a + d] < centers[3 * b + d]; });
            splt2.push_back(int32_t(j1));
            splt2.push_back(int32_t(j2));
        }
        splt2.swap(splt);
        splt2.clear();
        d = (d + 1) % 3;
    }
    int sz = 1 << depth;
    ind_bvh.assign(sz, -1);
    sz = 2 * sz - 1;
    Box none;
    for (int i = 0; i < 3; ++i) {
        none.mn[i] = std::numeric_limits<float>::max();
        none.mx[i] = -std::numeric_limits<float>::max();
    }
    BB_bvh.assign(sz, none);
    int j = sz - 1;
    int k = int(ind_bvh.size()) - 1;
    for (int i = int(splt.size()) - 1; i > 0; --i) {
        int a = splt[i - 1];
        int cnt = splt[i] - a;
        if (cnt == 1) {
            int id = ids[a];
            ind_bvh[k] = -1;
            ind_bvh[k - 1] = id;
            BB_bvh[j] = bbs[id];
            BB_bvh[j - 1] = bbs[id];
        } else if (cnt >= 2) {
            int id = ids[a + 1];
            ind_bvh[k] = id;
            BB_bvh[j] = bbs[id];
            id = ids[a];
            ind_bvh[k - 1] = id;
            BB_bvh[j - 1] = bbs[id];
        }
        k -= 2;
        j -= 2;
    }
    k = sz - 1;
    while (k >= 2) {
        BB_bvh[(k - 2) / 2] = merge(BB_bvh[k], BB_bvh[k - 1]);
        k -= 2;
    }
}

void Tables::clear() { *this = Tables(); }

// ---- ScenePrimitives::add_* (scene.h:128-172) + add_prim (scene.cpp:44-53)
static void apply3(const M4& t, float x, float y, float z, float out[3]) {
    float v[4] = {x, y, z, 1.0f}, r[4];
    m4_apply(t, v, r);
    out[0] = r[0];
    out[1] = r[1];
    out[2] = r[2];
}
static void sub3(const float a[3], const float b[3], float r[3]) {
    for (int i = 0; i < 3; ++i) r[i] = a[i] - b[i];
}
static void cross3(const float a[3], const float b[3], float r[3]) {
    r[0] = a[1] * b[2] - a[2] * b[1];
    r[1] = a[2] * b[0] - a[0] * b[2];
    r[2] = a[0] * b[1] - a[1] * b[0];
}
static float dot3(const float a[3], const float b[3]) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
static float norm3(const float a[3]) { return std::sqrt(dot3(a, a)); }

int Tables::add_prim(int type, const M4& trf, const float color[4], float shin, float rough, float emi) {
    float area = 0.0f;
    float o[3], px[3], py[3], pz[3], U[3], V[3], W[3], c1[3], c2[3], c3[3];
    const float zlo = (type == 5) ? 0.0f : -1.0f;
    apply3(trf, -1, -1, zlo, o);
    apply3(trf, 1, -1, zlo, px);
    apply3(trf, -1, 1, zlo, py);
    apply3(trf, -1, -1, 1, pz);
    sub3(px, o, U);
    sub3(py, o, V);
    sub3(pz, o, W);
    switch (type) {
        case 1: {
            float col0[3] = {trf.m[0], trf.m[1], trf.m[2]};
            float r = norm3(col0);
            area = float(2.0 * M_PI) * r * r;
            break;
        }
        case 2:
            cross3(U, V, c1);
            cross3(U, W, c2);
            cross3(W, V, c3);
            area = 2.0f * (norm3(c1) + norm3(c2) + norm3(c3));
            break;
        case 3: area = (dot3(U, U) + dot3(V, V)) / 4.0f * sqrtf(2.0f) * float(M_PI) * norm3(W); break;
        case 4: area = 0.0f; break;
        case 5:
            cross3(U, V, c1);
            area = norm3(c1);
            break;
    }
    PrimRec p;
    std::memset(&p, 0, sizeof p);
    std::memcpy(p.transfo, trf.m, sizeof p.transfo);
    M4 inv = m4_inverse(trf);
    std::memcpy(p.inv_transfo, inv.m, sizeof p.inv_transfo);
    std::memcpy(p.inv_mesh_bb_transfo, trf.m, sizeof p.transfo);
    p.type[0] = float(type);
    std::memcpy(p.color, color, sizeof p.color);
    p.mat_info[0] = shin;
    p.mat_info[1] = rough;
    p.mat_info[2] = emi;
    p.mat_info[3] = area;
    prims.push_back(p);
    return int(prims.size()) - 1;
}

// ---- BVH_GPU_Scene::add_mesh (gpu_bvh_scene.cpp:51-74) + SceneMesh::prim_bb (scene.cpp:106-122)
int Tables::add_mesh(const MeshIn& m) {
    meshes.push_back(m);
    KdTree kd;
    int nt = int(m.tri.size() / 3);
    kd.bbs.resize(nt);
    kd.centers.resize(3 * size_t(nt));
    for (int t = 0; t < nt; ++t) {
        kd.ids.push_back(t);
        const float* P0 = &m.pos[3 * m.tri[3 * t]];
        const float* P1 = &m.pos[3 * m.tri[3 * t + 1]];
        const float* P2 = &m.pos[3 * m.tri[3 * t + 2]];
        Box& bb = kd.bbs[t];
        for (int i = 0; i < 3; ++i) {
            bb.mn[i] = std::min({P0[i], P1[i], P2[i]});
            bb.mx[i] = std::max({P0[i], P1[i], P2[i]});
            bb.mn[i] -= 0.001f;
            bb.mx[i] += 0.001f;
            kd.centers[3 * t + i] = (bb.mn[i] + bb.mx[i]) / 2.0f;
        }
    }
    kd.compute();
    mesh_bvh.push_back(kd);
    mesh_vert_begins.push_back(next_vert_row);
    mesh_tri_begins.push_back(next_tri_row);
    int nv = int(m.pos.size() / 3);
    next_vert_row += nv / TEX_WIDTH + ((nv % TEX_WIDTH != 0) ? 1 : 0);
    // The reference reserves ceil(nb_triangles/1024) rows but writes nb_triangles+1
    // texels (header + triangles, gpu_bvh_scene.cpp:60-62,102-107), overflowing when
    // nb_triangles is a multiple of 1024; the oracle reserves room for the header.
    int nt1 = nt + 1;
    next_tri_row += nt1 / TEX_WIDTH + ((nt1 % TEX_WIDTH != 0) ? 1 : 0);
    return int(meshes.size()) - 1;
}

// ---- ScenePrimitives::add_mesh (scene.cpp:56-67), BoundingBox::matrix (mesh.h:67-71)
int Tables::place_mesh(int id, const M4& trf, const float color[4], float shin, float rough, float emi) {
    const MeshIn& m = meshes[id];
    M4 bbm = m4_identity();
    for (int i = 0; i < 3; ++i) {
        bbm.m[5 * i] = (m.bb.mx[i] - m.bb.mn[i]) / 2.0f;
        bbm.m[12 + i] = (m.bb.mn[i] + m.bb.mx[i]) / 2.0f;
    }
    PrimRec p;
    std::memset(&p, 0, sizeof p);
    M4 trbb = m4_mul(trf, bbm);
    std::memcpy(p.transfo, trbb.m, sizeof p.transfo);
    M4 inv = m4_inverse(trf);
    std::memcpy(p.inv_transfo, inv.m, sizeof p.inv_transfo);
    std::memcpy(p.inv_mesh_bb_transfo, trf.m, sizeof p.transfo);
    p.type[0] = 0.0f;
    p.type[1] = float(mesh_tri_begins[id]);
    std::memcpy(p.color, color, sizeof p.color);
    p.mat_info[0] = shin;
    p.mat_info[1] = rough;
    p.mat_info[2] = emi;
    p.mat_info[3] = 0.0f;
    prims.push_back(p);
    return int(prims.size()) - 1;
}

static int rows_for(int texels) { return texels / TEX_WIDTH + ((texels % TEX_WIDTH != 0) ? 1 : 0); }

// ---- BVH_GPU_Scene::finalize (gpu_bvh_scene.cpp:121-187)
void Tables::finalize() {
    // sortEmissiveFirst, scene.cpp:70-88
    {
        auto next_emi = prims.begin();
        while (next_emi != prims.end() && next_emi->mat_info[2] > 0.0f) ++next_emi;
        auto it = next_emi;
        while (it != prims.end()) {
            if (it->mat_info[2] > 0.0f) {
                PrimRec tmp = *next_emi;
                *next_emi++ = *it;
                *it = tmp;
            }
            ++it;
        }
        nb_emissives = int(next_emi - prims.begin());
    }
    // root tree: BVH_KDtree::init over ScenePrimitives::prim_bb (scene.cpp:18-42)
    root = KdTree();
    int n = int(prims.size());
    root.bbs.resize(n);
    root.centers.resize(3 * size_t(n));
    for (int p = 0; p < n; ++p) {
        root.ids.push_back(p);
        Box& bb = root.bbs[p];
        for (int i = 0; i < 3; ++i) {
            bb.mn[i] = std::numeric_limits<float>::max();
            bb.mx[i] = std::numeric_limits<float>::lowest();
        }
        M4 T;
        std::memcpy(T.m, prims[p].transfo, sizeof T.m);
        for (uint32_t v = 0; v < 8; ++v) {
            float x = float(v & 1) * 2.01f - 1.005f;
            float y = float(v >> 1 & 1) * 2.01f - 1.005f;
            float z = float(v >> 2 & 1) * 2.01f - 1.005f;
            if (prims[p].type[0] == 5.0f) z /= std::abs(z) * 1000.0f;
            float in[4] = {x, y, z, 1.0f}, B[4];
            m4_apply(T, in, B);
            for (int i = 0; i < 3; ++i) {
                if (B[i] < bb.mn[i]) bb.mn[i] = B[i];
                if (B[i] > bb.mx[i]) bb.mx[i] = B[i];
            }
        }
        for (int i = 0; i < 3; ++i) root.centers[3 * p + i] = (bb.mn[i] + bb.mx[i]) / 2.0f;
    }
    root.compute();

    // row offsets: root first, then each mesh tree (gpu_bvh_scene.cpp:127-141)
    std::vector<const KdTree*> trees{&root};
    for (const KdTree& k : mesh_bvh) trees.push_back(&k);
    bb_begins.assign(trees.size() + 1, 0);
    ind_begins.assign(trees.size() + 1, 0);
    for (size_t t = 0; t < trees.size(); ++t) {
        bb_begins[t + 1] = bb_begins[t] + rows_for(2 * int(trees[t]->BB_bvh.size()));
        ind_begins[t + 1] = ind_begins[t] + rows_for(int(trees[t]->ind_bvh.size()));
    }
    tex_bb = Tex();
    tex_bb.comps = 3;
    tex_bb.height = bb_begins.back();
    tex_bb.f.assign(size_t(TEX_WIDTH) * tex_bb.height * 3, 0.0f);
    tex_ind = Tex();
    tex_ind.comps = 1;
    tex_ind.height = ind_begins.back();
    tex_ind.i.assign(size_t(TEX_WIDTH) * tex_ind.height, 0);
    for (size_t t = 0; t < trees.size(); ++t) {  // send_bvh, :31-48
        std::memcpy(&tex_bb.f[size_t(bb_begins[t]) * TEX_WIDTH * 3], trees[t]->BB_bvh.data(), trees[t]->BB_bvh.size() * sizeof(Box));
        std::memcpy(&tex_ind.i[size_t(ind_begins[t]) * TEX_WIDTH], trees[t]->ind_bvh.data(), trees[t]->ind_bvh.size() * sizeof(int));
    }
    // primitive table: 64 records per 1024-texel row (:153-160)
    tex_prim = Tex();
    tex_prim.comps = 4;
    tex_prim.height = n / 64 + ((n % 64 != 0) ? 1 : 0);
    tex_prim.f.assign(size_t(TEX_WIDTH) * tex_prim.height * 4, 0.0f);
    if (n) std::memcpy(tex_prim.f.data(), prims.data(), size_t(n) * sizeof(PrimRec));
    // mesh tables (:162-175, send_mesh :87-118)
    tex_tri = Tex();
    tex_tri.comps = 3;
    tex_tri.height = next_tri_row;
    tex_tri.i.assign(size_t(TEX_WIDTH) * tex_tri.height * 3, 0);
    tex_p = Tex();
    tex_p.comps = 3;
    tex_p.height = next_vert_row;
    tex_p.f.assign(size_t(TEX_WIDTH) * tex_p.height * 3, 0.0f);
    tex_n = tex_p;
    for (size_t m = 0; m < meshes.size(); ++m) {
        const MeshIn& M = meshes[m];
        int32_t* t = &tex_tri.i[size_t(mesh_tri_begins[m]) * TEX_WIDTH * 3];
        t[0] = bb_begins[m + 1];                               // line in BVH_BB
        t[1] = ind_begins[m + 1];                              // line in BVH_IND
        t[2] = int32_t(pow(2.0, mesh_bvh[m].depth) - 1.0);     // 2^depth - 1
        uint32_t d = uint32_t(TEX_WIDTH) * uint32_t(mesh_vert_begins[m]);
        for (size_t i = 0; i < M.tri.size(); ++i) t[3 + i] = int32_t(M.tri[i] + d);
        std::memcpy(&tex_p.f[size_t(mesh_vert_begins[m]) * TEX_WIDTH * 3], M.pos.data(), M.pos.size() * sizeof(float));
        std::memcpy(&tex_n.f[size_t(mesh_vert_begins[m]) * TEX_WIDTH * 3], M.nrm.data(), M.nrm.size() * sizeof(float));
    }
}

}  // namespace oracle
