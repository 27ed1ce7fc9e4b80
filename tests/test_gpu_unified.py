"""GPU parity of the unified world-space tree (MCPT_OPT_UNIFIED, the default on scenes with meshes) and of mesh
uploads that carry no host-built tree.

* The unified walk and the two-level walk return the same hit records bit for bit, and both equal the CPU oracle.
* A mesh placed several times (instances) gets one world-space subtree per placement.
* Without the reference's median-split mesh trees (mcpt_mesh_desc.bvh empty) the library still returns the same
  records, except where two triangles of one mesh are hit at EXACTLY the same world distance: the reference keeps the
  one it visits first (the higher leaf slot of its tree, which depends on std::nth_element's arrangement), the
  library then keeps the higher triangle index. The test counts those ties and checks nothing else differs.
"""
import numpy as np
import pytest

from common import DATE, both, camera_rays, compare_hits, random_rays
from montecarlo_pathtracing_b200 import Context, McptError, Scene, capi
from oracle.binding import Oracle

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    c = Context(0)
    yield c
    c.set_option(capi.OPT_UNIFIED, 1)
    c.close()


def _trace(ctx, desc, ro, rd, unified):
    ctx.set_option(capi.OPT_UNIFIED, unified)
    try:
        ctx.upload_scene(desc)
        return ctx.trace_rays(ro, rd)
    finally:
        ctx.set_option(capi.OPT_UNIFIED, 1)


@pytest.mark.parametrize("n_meshes,tris_log2", [(1, 3), (3, 8), (4, 12), (16, 10)])
def test_unified_equals_two_level_equals_oracle(ctx, n_meshes, tris_log2):
    s, o = both("mesh", n_meshes=n_meshes, tris_log2=tris_log2)
    ro, rd = random_rays(o.prim_data(), 200_000, seed=23, clip=150.0)
    iv, ipv = capi.default_camera(320, 180)
    o.set_uniforms(8, 1.5, DATE, 0, iv, ipv)
    co, cd = camera_rays(o, 320, 180)
    ro, rd = np.concatenate([ro, co]), np.concatenate([rd, cd])
    ref = o.trace(ro, rd)
    uni = _trace(ctx, o.scene_desc(), ro, rd, 1)
    two = _trace(ctx, o.scene_desc(), ro, rd, 0)
    assert uni.tobytes() == two.tobytes()
    assert len(compare_hits(uni, ref)) == 0
    assert np.array_equal(uni["dist"].view(np.uint32), ref["dist"].view(np.uint32))
    assert np.array_equal(uni["pl"].view(np.uint32), ref["pl"].view(np.uint32))
    assert (ref["tri_index"] >= 0).sum() > 1000


def _instanced_scene():
    """One mesh placed three times (rotated, scaled, far from the origin), a second mesh once, a floor."""
    rng = np.random.default_rng(99)
    n = 24
    u, v = np.meshgrid(np.linspace(0, np.pi, n), np.linspace(0, 2 * np.pi, 2 * n), indexing="ij")
    r = 1.0 + 0.15 * rng.standard_normal(u.shape)
    pos = np.stack([r * np.sin(u) * np.cos(v), r * np.sin(u) * np.sin(v), r * np.cos(u)], -1).reshape(-1, 3).astype(np.float32)
    tri = []
    for i in range(n - 1):
        for j in range(2 * n - 1):
            a = i * 2 * n + j
            tri += [[a, a + 2 * n, a + 1], [a + 1, a + 2 * n, a + 2 * n + 1]]
    tri = np.asarray(tri, np.uint32)
    s = Scene()
    m0 = s.add_mesh(pos, tri)
    m1 = s.add_mesh((pos * np.float32([1.0, 0.5, 2.0])).astype(np.float32), tri[::2].copy())
    s.place_mesh(m0, capi.transfo("tZs", -60, 10, 0, 30, 25, 25, 25), (0.8, 0.7, 0.6))
    s.place_mesh(m0, capi.transfo("tXs", 70, -20, 15, 75, 12, 30, 18), (0.2, 0.7, 0.6))
    s.place_mesh(m1, capi.transfo("tYs", 0, 60, 30, 40, 20, 20, 20), (0.8, 0.2, 0.6))
    s.place_mesh(m0, capi.transfo("tZs", 900, 850, -700, 10, 40, 40, 40), (0.5, 0.5, 0.9))      # far from the origin
    s.add_orientedQuad(capi.transfo("tZs", 0, 0, -60, 0, 2000, 2000, 1), (0.5, 0.5, 0.5))
    s.add_sphere(capi.transfo("tZs", 10, -70, 0, 0, 20, 20, 20), (0.9, 0.9, 0.9))
    raw = s.prim_data()
    s.finalize()
    o = Oracle().load_unsorted(raw, s.tables()["meshes"]).finalize()
    return s, o


def test_instances_of_one_mesh(ctx):
    s, o = _instanced_scene()
    ro, rd = random_rays(o.prim_data(), 150_000, seed=4, clip=1200.0)
    # shares of rays aimed at each placement; the far one from far away (large |O|: the per-ray padding of the unified boxes)
    rng = np.random.default_rng(8)
    k = 25_000
    for j, (c, r, span) in enumerate([((-60, 10, 0), 30, 300), ((70, -20, 15), 30, 300), ((0, 60, 30), 40, 300), ((900, 850, -700), 45, 1500)]):
        tgt = np.float32(c) + rng.uniform(-r, r, (k, 3)).astype(np.float32)
        org = (np.float32(c) if span < 1000 else np.float32([0, 0, 0])) + rng.uniform(-span, span, (k, 3)).astype(np.float32)
        ro[j * k:(j + 1) * k], rd[j * k:(j + 1) * k] = org, (tgt - org)
    ref = o.trace(ro, rd)
    uni = _trace(ctx, o.scene_desc(), ro, rd, 1)
    two = _trace(ctx, o.scene_desc(), ro, rd, 0)
    assert uni.tobytes() == two.tobytes()
    assert len(compare_hits(uni, ref)) == 0
    assert np.array_equal(uni["dist"].view(np.uint32), ref["dist"].view(np.uint32))
    hit_prims = set(np.unique(ref["index"][ref["tri_index"] >= 0]).tolist())
    assert len(hit_prims) == 4, hit_prims            # all four placements are hit
    assert min(int((ref["index"] == p).sum()) for p in hit_prims) > 2000


@pytest.mark.parametrize("n_meshes,tris_log2", [(2, 7), (4, 12)])
def test_upload_without_host_trees(ctx, n_meshes, tris_log2):
    s, o = both("mesh", n_meshes=n_meshes, tris_log2=tris_log2)
    ro, rd = random_rays(o.prim_data(), 1 << 20, seed=31, clip=150.0)
    with_tree = _trace(ctx, s.desc(), ro, rd, 1)
    no_tree = _trace(ctx, s.desc(mesh_trees=False), ro, rd, 1)
    no_tree_two_level = _trace(ctx, s.desc(mesh_trees=False), ro, rd, 0)
    assert no_tree.tobytes() == no_tree_two_level.tobytes()
    diff = np.nonzero(with_tree.view(np.uint8).reshape(len(ro), -1) != no_tree.view(np.uint8).reshape(len(ro), -1))[0]
    diff = np.unique(diff)
    # every difference is an exact-distance tie between two triangles of the same mesh primitive
    a, b = with_tree[diff], no_tree[diff]
    assert np.array_equal(a["index"], b["index"]) and np.array_equal(a["dist"].view(np.uint32), b["dist"].view(np.uint32))
    assert np.all(a["tri_index"] != b["tri_index"]) and np.all(b["tri_index"] > a["tri_index"]) or len(diff) == 0
    n_hits = int((with_tree["tri_index"] >= 0).sum())
    print(f"{len(diff)} exact-distance ties among {n_hits} triangle hits of {len(ro)} rays")
    assert len(diff) <= 1e-4 * n_hits


def test_bad_vertex_index_is_reported_not_followed(ctx):
    s, o = both("mesh", n_meshes=1, tris_log2=6)
    t = s.tables()
    m = t["meshes"][0]
    tri = m["tri"].copy()
    tri[5, 1] = len(m["pos"]) + 1000
    from common import desc_from_arrays
    bad = dict(m, tri=tri)
    d, keep = desc_from_arrays(t["prims"], t["root"], [bad])
    with pytest.raises(McptError) as e:
        ctx.upload_scene(d)
    assert "out of range" in str(e.value)
    # the context is still usable
    ro, rd = random_rays(o.prim_data(), 10_000, seed=2, clip=150.0)
    got = _trace(ctx, s.desc(), ro, rd, 1)
    assert len(compare_hits(got, o.trace(ro, rd))) == 0


def test_edge_meshes_in_the_unified_tree(ctx):
    """A mesh of ONE triangle, a mesh of two, and a mesh that is added but never placed: the builder's edge cases
    (single-leaf subtree, single node, an instance list shorter than the mesh list)."""
    rng = np.random.default_rng(17)
    s = Scene()
    ids = []
    for n_tris in (1, 2, 5):
        pos = rng.uniform(-1, 1, (3 * n_tris, 3)).astype(np.float32)
        tri = np.arange(3 * n_tris, dtype=np.uint32).reshape(n_tris, 3)
        ids.append(s.add_mesh(pos, tri))
    s.place_mesh(ids[0], capi.transfo("tZs", -40, 0, 10, 30, 25, 25, 25), (0.8, 0.7, 0.6))
    s.place_mesh(ids[1], capi.transfo("tZs", 40, 0, 10, 60, 25, 25, 25), (0.2, 0.7, 0.6))
    s.place_mesh(ids[1], capi.transfo("tXs", 0, 50, 10, 45, 20, 30, 20), (0.2, 0.2, 0.6))      # ids[2] stays unplaced
    s.add_orientedQuad(capi.transfo("tZs", 0, 0, -40, 0, 300, 300, 1), (0.5, 0.5, 0.5))
    raw = s.prim_data()
    s.finalize()
    o = Oracle().load_unsorted(raw, s.tables()["meshes"]).finalize()
    ro, rd = random_rays(o.prim_data(), 100_000, seed=6, clip=120.0)
    rng2 = np.random.default_rng(3)
    k = 20_000
    for j, c in enumerate([(-40, 0, 10), (40, 0, 10), (0, 50, 10)]):
        tgt = np.float32(c) + rng2.uniform(-20, 20, (k, 3)).astype(np.float32)
        org = np.float32(c) + rng2.uniform(-150, 150, (k, 3)).astype(np.float32)
        ro[j * k:(j + 1) * k], rd[j * k:(j + 1) * k] = org, tgt - org
    ref = o.trace(ro, rd)
    for unified in (1, 0):
        for desc in (s.desc(), s.desc(mesh_trees=False)):
            got = _trace(ctx, desc, ro, rd, unified)
            assert len(compare_hits(got, ref)) == 0, unified
    assert (ref["tri_index"] >= 0).sum() > 500


def _trace_pool(ctx, desc, ro, rd, trigger):
    ctx.set_option(capi.OPT_LEAF_POOL, trigger)
    try:
        ctx.upload_scene(desc)
        return ctx.trace_rays(ro, rd)
    finally:
        ctx.set_option(capi.OPT_LEAF_POOL, LEAF_POOL_DEFAULT)


LEAF_POOL_DEFAULT = 0


@pytest.mark.parametrize("trigger", [1, 8, 16, 32])
def test_pooled_leaf_tests_find_the_same_records(ctx, trigger):
    """MCPT_OPT_LEAF_POOL: a lane that reaches a leaf of the unified tree queues a (ray, leaf) item and walks on, the warp
    runs the queued exact tests 32 at a time, the closest hit of a ray is kept by atomicMin on a key that encodes
    offer()'s order. Same tests in another order with the same ranking: the records are the same bits as with one test
    per lane, on triangles (with and without uploaded trees, instanced meshes) and on analytic primitives (repo scene 8,
    whose tree is large enough for the unified walk)."""
    cases = []
    s, o = both("mesh", n_meshes=4, tris_log2=12)
    ro, rd = random_rays(o.prim_data(), 300_000, seed=61, clip=150.0)
    cases.append(("meshes", o.scene_desc(), o, ro, rd))
    cases.append(("meshes without host trees", s.desc(mesh_trees=False), None, ro, rd))
    s2, o2 = _instanced_scene()
    ro2, rd2 = random_rays(o2.prim_data(), 200_000, seed=62, clip=1200.0)
    cases.append(("instances", o2.scene_desc(), o2, ro2, rd2))
    s8, o8 = both(8)
    ro8, rd8 = random_rays(o8.prim_data(), 200_000, seed=63)
    iv, ipv = capi.default_camera(320, 180)
    o8.set_uniforms(8, 1.0, DATE, 0, iv, ipv)
    co, cd = camera_rays(o8, 320, 180)
    cases.append(("scene 8", o8.scene_desc(), o8, np.concatenate([ro8, co]), np.concatenate([rd8, cd])))
    for name, desc, orc, a, b in cases:
        plain = _trace_pool(ctx, desc, a, b, 0)
        pooled = _trace_pool(ctx, desc, a, b, trigger)
        assert pooled.tobytes() == plain.tobytes(), name
        assert (plain["index"] >= 0).sum() > 10_000, name
        if orc is not None and trigger == 16:
            ref = orc.trace(a, b)
            assert len(compare_hits(pooled, ref)) == 0, name
            assert np.array_equal(pooled["dist"].view(np.uint32), ref["dist"].view(np.uint32)), name


@pytest.mark.parametrize("kind,kw,bounces,ior", [("mesh", dict(n_meshes=4, tris_log2=10), 8, 1.5), (8, {}, 9, 1.0), (3, {}, 9, 1.5)])
def test_pooled_leaf_tests_render_the_same_sums(kind, kw, bounces, ior):
    """... and so are the accumulated sums, ray and segment counts of a render (interior rays and, on analytic scenes, the
    fused exit included). Scene 3 has fewer than 256 leaves: the option changes nothing there."""
    s, o = both(kind, **kw)
    iv, ipv = capi.default_camera(240, 135)
    out = {}
    for name, trigger in (("plain", 0), ("pooled", 16), ("pooled at once", 1)):
        c = Context(0)
        try:
            c.set_option(capi.OPT_LEAF_POOL, trigger)
            c.set_option(capi.OPT_PASSES_PER_BATCH, 3)
            c.upload_scene(s)
            c.resize(240, 135)
            c.set_camera(iv, ipv)
            c.set_params(nb_bounces=bounces, refract_ind=ior, date=float(DATE))
            c.render(0, 5)
            out[name] = (c.read_accum(), c.stats())
        finally:
            c.close()
    base, st = out["plain"]
    assert base[..., :3].mean() > 0.005 and st["rays"] > 1.2 * st["paths"]
    for name in ("pooled", "pooled at once"):
        assert np.array_equal(base, out[name][0]), name
        assert out[name][1]["rays"] == st["rays"] and out[name][1]["segments"] == st["segments"], name
