// bvh_build.cu — GPU construction of the per-mesh acceleration structure.
//
// The reference builds every mesh tree on the CPU by median split with a cycling axis
// (BVH_KDtree::compute, bvh_gpu/bvh.cpp:34-93): a complete binary tree with one triangle per leaf
// slot, whose boxes overlap heavily on curved surfaces. Hit results do not depend on the tree
// (candidates are ranked by the reference's exact distance and visit order, intersect.cuh), so the
// library is free to traverse a better one. This file builds an LBVH on the device at upload time:
//   1. Morton code (30 bit) of every triangle centroid inside the mesh box          k_morton
//   2. radix sort of (code, triangle)                                               cub::DeviceRadixSort
//   3. leaves = runs of kMeshLeafTris (1) consecutive triangles in Morton order, stored as DevTri
//      slots with the triangle id and the reference's leaf slot (tie-break rank)     k_leaves
//   4. Karras 2012 radix-tree hierarchy over the leaves, duplicates split by index   k_hierarchy
//   5. bottom-up box fit with one atomic arrival counter per node; the second arrival writes the
//      64-byte node {child boxes, child tokens}                                      k_fit
//   6. every second level folded into four-children nodes for the persistent walk   k_collapse4
// Boxes are widened like the top-level ones (pack_node, api.cu) so that the FMA slab test of the
// traversal only culls.
#include <cub/device/device_radix_sort.cuh>
#include <cuda_runtime.h>

#include <algorithm>
#include <cfloat>
#include <climits>
#include <cstdint>

#include "bvh_build.cuh"

namespace mcpt {

namespace {

constexpr int kB = 256;
constexpr int kLeafTris = kMeshLeafTris;

__device__ __forceinline__ uint32_t expand10(uint32_t v) {
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}

// order-preserving map float -> uint32 (atomicMin / atomicMax on floats of either sign)
__device__ __forceinline__ uint32_t f2ord(float f) {
    const uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ord2f(uint32_t u) { return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u); }

__global__ void k_bounds_init(uint32_t* __restrict__ bounds) {
    if (threadIdx.x < 3) bounds[threadIdx.x] = 0xffffffffu;
    else if (threadIdx.x < 6) bounds[threadIdx.x] = 0u;
}

// Box of the vertices the tree is built over (bounds[0..2] = min, [3..5] = max, order-encoded), and -- for a tree in
// world space (the unified tree of a mesh instance) -- those vertices themselves: wpos = fwd * pos. No host pass
// over the vertex array is needed at upload.
__global__ void k_xform_bounds(const float* __restrict__ pos, int n_verts, Aff fwd, int transform, float* __restrict__ wpos,
                               uint32_t* __restrict__ bounds) {
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (int v = blockIdx.x * blockDim.x + threadIdx.x; v < n_verts; v += gridDim.x * blockDim.x) {
        float x = pos[3 * v], y = pos[3 * v + 1], z = pos[3 * v + 2];
        if (transform) {
            const float wx = fwd.r0.x * x + fwd.r0.y * y + fwd.r0.z * z + fwd.r0.w;
            const float wy = fwd.r1.x * x + fwd.r1.y * y + fwd.r1.z * z + fwd.r1.w;
            const float wz = fwd.r2.x * x + fwd.r2.y * y + fwd.r2.z * z + fwd.r2.w;
            x = wx; y = wy; z = wz;
            wpos[3 * v] = x; wpos[3 * v + 1] = y; wpos[3 * v + 2] = z;
        }
        lo[0] = fminf(lo[0], x); lo[1] = fminf(lo[1], y); lo[2] = fminf(lo[2], z);
        hi[0] = fmaxf(hi[0], x); hi[1] = fmaxf(hi[1], y); hi[2] = fmaxf(hi[2], z);
    }
    for (int k = 0; k < 3; ++k)
        for (int off = 16; off > 0; off >>= 1) {
            lo[k] = fminf(lo[k], __shfl_xor_sync(0xffffffffu, lo[k], off));
            hi[k] = fmaxf(hi[k], __shfl_xor_sync(0xffffffffu, hi[k], off));
        }
    if ((threadIdx.x & 31) == 0)
        for (int k = 0; k < 3; ++k) {
            if (lo[k] <= hi[k]) {
                atomicMin(bounds + k, f2ord(lo[k]));
                atomicMax(bounds + 3 + k, f2ord(hi[k]));
            }
        }
}

// (vertex indices are clamped wherever a build kernel follows them: an out-of-range index is reported by
// check_indices, it must not fault first)
__global__ void k_morton(const float* __restrict__ pos, const uint32_t* __restrict__ idx, int n_tris, uint32_t vmax, const uint32_t* __restrict__ bounds,
                         uint32_t* __restrict__ keys, uint32_t* __restrict__ vals) {
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_tris) return;
    const float3 lo = make_float3(ord2f(bounds[0]), ord2f(bounds[1]), ord2f(bounds[2]));
    const float3 inv_ext = make_float3(1.0f / fmaxf(ord2f(bounds[3]) - lo.x, 1e-30f), 1.0f / fmaxf(ord2f(bounds[4]) - lo.y, 1e-30f),
                                       1.0f / fmaxf(ord2f(bounds[5]) - lo.z, 1e-30f));
    float c[3];
    for (int k = 0; k < 3; ++k) {
        float a = pos[3 * min(idx[3 * t], vmax) + k], b = pos[3 * min(idx[3 * t + 1], vmax) + k], d = pos[3 * min(idx[3 * t + 2], vmax) + k];
        c[k] = (a + b + d) * (1.0f / 3.0f);
    }
    float fx = fminf(fmaxf((c[0] - lo.x) * inv_ext.x * 1024.0f, 0.0f), 1023.0f);
    float fy = fminf(fmaxf((c[1] - lo.y) * inv_ext.y * 1024.0f, 0.0f), 1023.0f);
    float fz = fminf(fmaxf((c[2] - lo.z) * inv_ext.z * 1024.0f, 0.0f), 1023.0f);
    keys[t] = (expand10(uint32_t(fx)) << 2) | (expand10(uint32_t(fy)) << 1) | expand10(uint32_t(fz));
    vals[t] = uint32_t(t);
}

// leaf c holds sorted triangles [L*c, L*c + L), L = kMeshLeafTris; writes the DevTri slots, the leaf box and the leaf key
// pos: the vertices stored in the DevTri slots (mesh-local: the exact test runs there); bpos: the vertices the boxes are
// fitted to (= pos, or the world positions of a mesh instance). slot_of_tri == null: no reference tree was uploaded,
// the tie-break rank of a triangle is its index. prim: stored in c.w (the unified tree finds the instance there).
__global__ void k_leaves(const float* __restrict__ pos, const float* __restrict__ bpos, const uint32_t* __restrict__ idx,
                         const int32_t* __restrict__ slot_of_tri, const uint32_t* __restrict__ sorted_keys,
                         const uint32_t* __restrict__ sorted_tris, int n_tris, uint32_t vmax, int n_leaves, int prim, DevTri* __restrict__ tris,
                         float* __restrict__ box_lo, float* __restrict__ box_hi, uint32_t* __restrict__ leaf_keys) {
    int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= n_leaves) return;
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (int j = 0; j < kLeafTris; ++j) {
        int s = kLeafTris * c + j;
        DevTri T;
        T.a = T.b = T.c = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        int tri = -1, rank = -1;
        if (s < n_tris) {
            tri = int(sorted_tris[s]);
            rank = slot_of_tri ? slot_of_tri[tri] : tri;
        }
        if (rank < 0) tri = -1;  // not in the reference tree: never tested there either
        if (tri >= 0) {
            float v[3][3];
            for (int q = 0; q < 3; ++q)
                for (int k = 0; k < 3; ++k) {
                    const uint32_t vi = 3 * min(idx[3 * tri + q], vmax) + k;
                    v[q][k] = pos[vi];
                    const float b = bpos[vi];
                    lo[k] = fminf(lo[k], b);
                    hi[k] = fmaxf(hi[k], b);
                }
            T.a = make_float4(v[0][0], v[0][1], v[0][2], 0.0f);
            T.b = make_float4(v[1][0], v[1][1], v[1][2], 0.0f);
            T.c = make_float4(v[2][0], v[2][1], v[2][2], 0.0f);
        }
        T.a.w = __int_as_float(tri);
        T.b.w = __int_as_float(rank);
        T.c.w = __int_as_float(prim);
        tris[s] = T;
    }
    for (int k = 0; k < 3; ++k) {
        box_lo[3 * c + k] = lo[k];
        box_hi[3 * c + k] = hi[k];
    }
    leaf_keys[c] = sorted_keys[kLeafTris * c];
}

// common-prefix length of leaves i and j (Karras 2012), index as tie-break for equal codes
__device__ __forceinline__ int delta(const uint32_t* __restrict__ keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    uint32_t a = keys[i], b = keys[j];
    if (a == b) return 32 + __clz(uint32_t(i) ^ uint32_t(j));
    return __clz(a ^ b);
}

// internal node i of the radix tree over n leaves: children and parent links.
// Node numbering in `parent`: internal nodes 0..n-2, leaves n-1 .. 2n-2.
__global__ void k_hierarchy(const uint32_t* __restrict__ keys, int n, int2* __restrict__ children, int* __restrict__ parent) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    int d = (delta(keys, n, i, i + 1) - delta(keys, n, i, i - 1)) >= 0 ? 1 : -1;
    int dmin = delta(keys, n, i, i - d);
    int lmax = 2;
    while (delta(keys, n, i, i + lmax * d) > dmin) lmax *= 2;
    int l = 0;
    for (int t = lmax / 2; t >= 1; t /= 2)
        if (delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
    int j = i + l * d;
    int dnode = delta(keys, n, i, j);
    int s = 0, t = l;
    do {
        t = (t + 1) >> 1;
        if (delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
    } while (t > 1);
    int gamma = i + s * d + min(d, 0);
    int lo = min(i, j), hi = max(i, j);
    int left = (lo == gamma) ? (n - 1 + gamma) : gamma;            // leaf or internal
    int right = (hi == gamma + 1) ? (n - 1 + gamma + 1) : (gamma + 1);
    children[i] = make_int2(left, right);
    parent[left] = i;
    parent[right] = i;
    if (i == 0) parent[0] = -1;
}

__device__ __forceinline__ void widen(float& lo, float& hi) {
    if (lo <= hi) {
        float mag = fmaxf(fabsf(lo), fabsf(hi));
        float pad = 1e-5f * fmaxf(mag, hi - lo) + 1e-7f;
        lo -= pad;
        hi += pad;
    }
}

// one thread per leaf climbs; the second thread to reach a node fits it and goes on
__global__ void k_fit(int n, const int2* __restrict__ children, const int* __restrict__ parent, float* __restrict__ box_lo,
                      float* __restrict__ box_hi, int* __restrict__ arrivals, DevNode2* __restrict__ nodes) {
    int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= n) return;
    // boxes: leaves at [0, n), internal node i at n + i
    int node = parent[n - 1 + c];
    while (node >= 0) {
        __threadfence();
        if (atomicAdd(&arrivals[node], 1) == 0) return;  // first arrival: the sibling subtree is not ready
        __threadfence();
        int2 ch = children[node];
        int b0 = (ch.x >= n - 1) ? (ch.x - (n - 1)) : (n + ch.x);
        int b1 = (ch.y >= n - 1) ? (ch.y - (n - 1)) : (n + ch.y);
        float l0[3], h0[3], l1[3], h1[3];
        for (int k = 0; k < 3; ++k) {
            // written by other SMs during this launch: read at L2, not through a stale L1 line
            l0[k] = __ldcg(box_lo + 3 * b0 + k);
            h0[k] = __ldcg(box_hi + 3 * b0 + k);
            l1[k] = __ldcg(box_lo + 3 * b1 + k);
            h1[k] = __ldcg(box_hi + 3 * b1 + k);
            box_lo[3 * (n + node) + k] = fminf(l0[k], l1[k]);
            box_hi[3 * (n + node) + k] = fmaxf(h0[k], h1[k]);
        }
        for (int k = 0; k < 3; ++k) {
            widen(l0[k], h0[k]);
            widen(l1[k], h1[k]);
        }
        // child token: >= 0 internal node index, < 0 leaf: ~leaf
        int t0 = (ch.x >= n - 1) ? ~(ch.x - (n - 1)) : ch.x;
        int t1 = (ch.y >= n - 1) ? ~(ch.y - (n - 1)) : ch.y;
        DevNode2 N;
        N.lo0 = make_float4(l0[0], l0[1], l0[2], __int_as_float(t0));
        N.hi0 = make_float4(h0[0], h0[1], h0[2], __int_as_float(t1));
        N.lo1 = make_float4(l1[0], l1[1], l1[2], 0.0f);
        N.hi1 = make_float4(h1[0], h1[1], h1[2], 0.0f);
        nodes[node] = N;
        node = parent[node];
    }
}

// a mesh with a single leaf: one node whose two children are that leaf (an empty child cannot be
// expressed as a box the symmetric slab test rejects; offering a candidate twice changes nothing)
__global__ void k_single(const float* __restrict__ box_lo, const float* __restrict__ box_hi, DevNode2* __restrict__ nodes) {
    float l[3], h[3];
    for (int k = 0; k < 3; ++k) {
        l[k] = box_lo[k];
        h[k] = box_hi[k];
        widen(l[k], h[k]);
    }
    DevNode2 N;
    N.lo0 = make_float4(l[0], l[1], l[2], __int_as_float(~0));
    N.hi0 = make_float4(h[0], h[1], h[2], __int_as_float(~0));
    N.lo1 = make_float4(l[0], l[1], l[2], 0.0f);
    N.hi1 = make_float4(h[0], h[1], h[2], 0.0f);
    nodes[0] = N;
}

// Fold every second level of the fitted two-children tree into its parent (DevNode4, layout.cuh): a node
// at even depth becomes a four-children node holding its grandchildren (or a child, where that child
// is a leaf). Nodes at odd depth are not written: the wide walk never visits them. The boxes are the
// already widened ones of nodes2.
constexpr int kTokEmptyBuild = INT_MIN;  // = kTokEmpty (traverse.cuh)

struct Slot4 {
    float lo[3], hi[3];
    int tok;
};

// node_base / leaf_base: where this tree's nodes and triangle slots sit in the arrays the walk indexes (0 for a mesh
// tree of its own; the unified tree holds every instance's subtree in one node array and one slot array)
__device__ __forceinline__ void write_node4(DevNode4* __restrict__ out, const Slot4* s, int n_slots, int node_base, int leaf_base) {
    float lo[3][4], hi[3][4];
    int tok[4];
    for (int j = 0; j < 4; ++j) {
        const bool used = j < n_slots;
        for (int k = 0; k < 3; ++k) {
            lo[k][j] = used ? s[j].lo[k] : FLT_MAX;  // an unused slot is rejected by its token (the walk checks it)
            hi[k][j] = used ? s[j].hi[k] : -FLT_MAX;
        }
        tok[j] = used ? (s[j].tok >= 0 ? s[j].tok + node_base : s[j].tok - leaf_base) : kTokEmptyBuild;
    }
    DevNode4 N;
    N.lox = make_float4(lo[0][0], lo[0][1], lo[0][2], lo[0][3]);
    N.loy = make_float4(lo[1][0], lo[1][1], lo[1][2], lo[1][3]);
    N.loz = make_float4(lo[2][0], lo[2][1], lo[2][2], lo[2][3]);
    N.hix = make_float4(hi[0][0], hi[0][1], hi[0][2], hi[0][3]);
    N.hiy = make_float4(hi[1][0], hi[1][1], hi[1][2], hi[1][3]);
    N.hiz = make_float4(hi[2][0], hi[2][1], hi[2][2], hi[2][3]);
    N.tok = make_int4(tok[0], tok[1], tok[2], tok[3]);
    N.pad = make_int4(0, 0, 0, 0);
    *out = N;
}

__device__ __forceinline__ int unpack2(const DevNode2& N, Slot4* s) {
    s[0].lo[0] = N.lo0.x; s[0].lo[1] = N.lo0.y; s[0].lo[2] = N.lo0.z;
    s[0].hi[0] = N.hi0.x; s[0].hi[1] = N.hi0.y; s[0].hi[2] = N.hi0.z;
    s[1].lo[0] = N.lo1.x; s[1].lo[1] = N.lo1.y; s[1].lo[2] = N.lo1.z;
    s[1].hi[0] = N.hi1.x; s[1].hi[1] = N.hi1.y; s[1].hi[2] = N.hi1.z;
    s[0].tok = __float_as_int(N.lo0.w);
    s[1].tok = __float_as_int(N.hi0.w);
    return 2;
}

__global__ void k_collapse4(int n, const int* __restrict__ parent, const DevNode2* __restrict__ nodes2, DevNode4* __restrict__ nodes4,
                            int node_base, int leaf_base) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    int depth = 0;
    for (int a = parent[i]; a >= 0; a = parent[a]) ++depth;
    if (depth & 1) return;
    Slot4 ch[2], out[4];
    unpack2(nodes2[i], ch);
    int m = 0;
    for (int j = 0; j < 2; ++j) {
        if (ch[j].tok >= 0) {
            unpack2(nodes2[ch[j].tok], out + m);
            m += 2;
        } else {
            out[m++] = ch[j];
        }
    }
    write_node4(nodes4 + i, out, m, node_base, leaf_base);
}

// single-leaf mesh: one wide node with that leaf as its only child
__global__ void k_single4(const DevNode2* __restrict__ nodes2, DevNode4* __restrict__ nodes4, int node_base, int leaf_base) {
    Slot4 ch[2];
    unpack2(nodes2[0], ch);
    write_node4(nodes4, ch, 1, node_base, leaf_base);
}

}  // namespace

namespace {

inline size_t align256(size_t b) { return (b + 255) & ~size_t(255); }

// carve the build's scratch arrays out of one workspace (same order in the size query and the build)
struct Scratch {
    uint32_t *keys, *vals, *keys2, *vals2, *leaf_keys, *bounds;
    float *box_lo, *box_hi, *wpos;
    int2* children;
    int *parent, *arrivals;
    DevNode2* nodes2;
    void* sort_tmp;
    size_t sort_bytes, total;
    Scratch(void* base, int n_tris, int n, int n_verts, size_t sort_b) : sort_bytes(sort_b) {
        size_t off = 0;
        auto take = [&](size_t bytes) {
            void* p = base ? static_cast<char*>(base) + off : nullptr;
            off += align256(bytes ? bytes : 1);
            return p;
        };
        keys = static_cast<uint32_t*>(take(sizeof(uint32_t) * n_tris));
        vals = static_cast<uint32_t*>(take(sizeof(uint32_t) * n_tris));
        keys2 = static_cast<uint32_t*>(take(sizeof(uint32_t) * n_tris));
        vals2 = static_cast<uint32_t*>(take(sizeof(uint32_t) * n_tris));
        leaf_keys = static_cast<uint32_t*>(take(sizeof(uint32_t) * n));
        bounds = static_cast<uint32_t*>(take(sizeof(uint32_t) * 8));
        box_lo = static_cast<float*>(take(sizeof(float) * 3 * (2 * size_t(n))));
        box_hi = static_cast<float*>(take(sizeof(float) * 3 * (2 * size_t(n))));
        wpos = static_cast<float*>(take(sizeof(float) * 3 * size_t(n_verts)));
        children = static_cast<int2*>(take(sizeof(int2) * n));
        parent = static_cast<int*>(take(sizeof(int) * 2 * size_t(n)));
        arrivals = static_cast<int*>(take(sizeof(int) * n));
        nodes2 = static_cast<DevNode2*>(take(sizeof(DevNode2) * size_t(n > 1 ? n - 1 : 1)));
        sort_tmp = take(sort_bytes);
        total = off;
    }
};

size_t sort_temp_bytes(int n_tris) {
    size_t b = 0;
    uint32_t* z = nullptr;
    cub::DeviceRadixSort::SortPairs(nullptr, b, z, z, z, z, n_tris, 0, 30, nullptr);
    return b;
}

}  // namespace

namespace {

// flag[0] |= 1: a vertex index out of range; |= 2: a leaf slot of the uploaded reference tree out of range
__global__ void k_check_idx(const uint32_t* __restrict__ idx, size_t n, uint32_t n_verts, int* __restrict__ flag) {
    bool bad = false;
    for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += size_t(gridDim.x) * blockDim.x) bad |= idx[i] >= n_verts;
    if (__any_sync(0xffffffffu, bad) && (threadIdx.x & 31) == 0) atomicOr(flag, 1);
}

// slot_of_tri[ind[s]] = s: the reference tree's leaf slot of every triangle (a triangle sits in one slot; the highest wins otherwise, as on the host)
__global__ void k_slots(const int32_t* __restrict__ ind, int n_slots, int n_tris, int32_t* __restrict__ slot_of_tri, int* __restrict__ flag) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_slots) return;
    const int t = ind[s];
    if (t < -1 || t >= n_tris) {
        atomicOr(flag, 2);
        return;
    }
    if (t >= 0) atomicMax(slot_of_tri + t, s);
}

}  // namespace

void check_indices(cudaStream_t stream, const uint32_t* d_idx, size_t n, uint32_t n_verts, int* d_flag) {
    if (!n) return;
    const int grid = int(std::min<size_t>((n + kB - 1) / kB, 1184));
    k_check_idx<<<grid, kB, 0, stream>>>(d_idx, n, n_verts, d_flag);
}

cudaError_t slots_from_leaves(cudaStream_t stream, const int32_t* d_ind, int n_slots, int n_tris, int32_t* d_slot_of_tri, int* d_flag) {
    cudaError_t e = cudaMemsetAsync(d_slot_of_tri, 0xff, sizeof(int32_t) * size_t(n_tris), stream);  // -1
    if (e != cudaSuccess) return e;
    k_slots<<<(n_slots + kB - 1) / kB, kB, 0, stream>>>(d_ind, n_slots, n_tris, d_slot_of_tri, d_flag);
    return cudaGetLastError();
}

size_t lbvh_workspace_bytes(int n_tris, int n_leaves, int n_verts) {
    return Scratch(nullptr, n_tris, n_leaves, n_verts, sort_temp_bytes(n_tris)).total;
}

cudaError_t build_mesh_lbvh(cudaStream_t stream, const LbvhBuild& b, void* workspace, size_t workspace_bytes,
                            unsigned long long* launches) {
    const int n = b.n_leaves, n_tris = b.n_tris;
    Scratch w(workspace, n_tris, n, b.n_verts, sort_temp_bytes(n_tris));
    if (!workspace || w.total > workspace_bytes) return cudaErrorInvalidValue;
    DevNode2* nodes2 = b.d_nodes ? b.d_nodes : w.nodes2;
    Aff fwd{};
    if (b.fwd) fwd = *b.fwd;
    const float* bpos = b.fwd ? w.wpos : b.d_pos;
    k_bounds_init<<<1, 32, 0, stream>>>(w.bounds);
    const int vgrid = std::min((b.n_verts + kB - 1) / kB, 1184);
    k_xform_bounds<<<std::max(vgrid, 1), kB, 0, stream>>>(b.d_pos, b.n_verts, fwd, b.fwd ? 1 : 0, w.wpos, w.bounds);
    const uint32_t vmax = uint32_t(std::max(b.n_verts, 1) - 1);
    k_morton<<<(n_tris + kB - 1) / kB, kB, 0, stream>>>(bpos, b.d_idx, n_tris, vmax, w.bounds, w.keys, w.vals);
    size_t sort_bytes = w.sort_bytes;
    cudaError_t e = cub::DeviceRadixSort::SortPairs(w.sort_tmp, sort_bytes, w.keys, w.keys2, w.vals, w.vals2, n_tris, 0, 30, stream);
    if (e != cudaSuccess) return e;
    k_leaves<<<(n + kB - 1) / kB, kB, 0, stream>>>(b.d_pos, bpos, b.d_idx, b.d_slot_of_tri, w.keys2, w.vals2, n_tris, vmax, n, b.prim,
                                                   b.d_tris, w.box_lo, w.box_hi, w.leaf_keys);
    if (n == 1) {
        k_single<<<1, 1, 0, stream>>>(w.box_lo, w.box_hi, nodes2);
        k_single4<<<1, 1, 0, stream>>>(nodes2, b.d_nodes4, b.node_base, b.leaf_base);
        if (launches) *launches += 6;
    } else {
        e = cudaMemsetAsync(w.arrivals, 0, sizeof(int) * n, stream);
        if (e != cudaSuccess) return e;
        k_hierarchy<<<(n - 1 + kB - 1) / kB, kB, 0, stream>>>(w.leaf_keys, n, w.children, w.parent);
        k_fit<<<(n + kB - 1) / kB, kB, 0, stream>>>(n, w.children, w.parent, w.box_lo, w.box_hi, w.arrivals, nodes2);
        k_collapse4<<<(n - 1 + kB - 1) / kB, kB, 0, stream>>>(n, w.parent, nodes2, b.d_nodes4, b.node_base, b.leaf_base);
        if (launches) *launches += 7;
    }
    return cudaGetLastError();
}

}  // namespace mcpt
