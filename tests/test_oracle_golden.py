"""CPU: the hand-written oracle (oracle/liboracle.so) and the host mirror (libmcpt_host.so) against
the golden fixtures of tests/golden/, which were produced by the reference's own code compiled as
C++ (tests/golden/make_golden.py). This is what pins the oracle: integer / index / table results
bit-exact, hit records bit-exact, images bit-exact on the same libm and within 1e-5 otherwise."""
from pathlib import Path

import numpy as np
import pytest

from common import DATE, host_scene
from montecarlo_pathtracing_b200 import capi
from oracle.binding import Oracle, have_reference

GOLD = Path(__file__).resolve().parent / "golden"
SCENES = list(range(1, 9))


def gold(name):
    return np.load(GOLD / f"{name}.npz")


def mesh_inputs(g):
    return [{"pos": g[f"m{m}_pos"], "tri": g[f"m{m}_tri"], "nrm": g[f"m{m}_nrm"], "mesh_line": int(g[f"m{m}_line"])}
            for m in range(2)]


def assert_image(img, ref, what=""):
    assert np.array_equal(np.isfinite(img), np.isfinite(ref)), what
    # bit-exact on the glibc that produced the fixtures; sin/cos/log/pow may move by an ulp on another one
    assert np.allclose(img, ref, rtol=1e-5, atol=1e-6, equal_nan=True), (what, np.nanmax(np.abs(img - ref)))


def replay(g, meshes=()):
    o = Oracle().load_unsorted(g["raw_prims"], meshes).finalize()
    return o


@pytest.mark.parametrize("sid", SCENES)
def test_tables_sort_and_bvh_match_the_reference(sid):
    """sortEmissiveFirst + BVH_KDtree::compute (bvh_gpu/scene.cpp:70-88, bvh.cpp:34-93)."""
    g = gold(f"scene{sid}")
    o = replay(g)
    assert o.nb_emissives == int(g["nb_emissives"])
    assert np.array_equal(o.prim_data(), g["prims"])
    b = o.bvh()
    assert b["depth"] == int(g["bvh_depth"])
    assert np.array_equal(b["ind"], g["bvh_ind"])
    assert np.array_equal(b["bb"], g["bvh_bb"])


@pytest.mark.parametrize("sid", SCENES)
def test_host_mirror_builds_the_reference_scenes(sid):
    """RTViewer::scene_* through the product's host mirror: same records in the same order; the
    inverse differs from Eigen's only in rounding."""
    g = gold(f"scene{sid}")
    s = host_scene(sid, 0.4)
    raw = s.prim_data()
    ref = g["raw_prims"]
    assert raw.shape == ref.shape
    assert np.array_equal(raw[:, 0:16], ref[:, 0:16])            # transfo_
    assert np.array_equal(raw[:, 32:64], ref[:, 32:64])          # mesh transfo, type, colour, material + area
    assert np.allclose(raw[:, 16:32], ref[:, 16:32], rtol=1e-5, atol=1e-4 * np.abs(ref[:, 16:32]).max())
    s.finalize()
    t = s.tables()
    assert t["nb_emissives"] == int(g["nb_emissives"])
    assert t["root"]["depth"] == int(g["bvh_depth"])
    assert np.array_equal(t["root"]["ind"], g["bvh_ind"])
    assert np.array_equal(t["root"]["bb"], g["bvh_bb"])
    assert np.array_equal(t["prims"][:, 0:16], g["prims"][:, 0:16])


@pytest.mark.parametrize("sid", SCENES)
def test_traversal_matches_the_reference_shader(sid):
    g = gold(f"scene{sid}")
    o = replay(g)
    got = o.trace(g["ray_o"], g["ray_d"])
    assert got.tobytes() == g["hits"].tobytes()                   # ids, face, dist, pl, pg: bit for bit
    assert np.array_equal(o.trace(g["ray_o"], g["ray_d"], True)["index"] >= 0, g["any_hit"])


@pytest.mark.parametrize("sid", SCENES)
def test_image_matches_the_reference_shader(sid):
    g = gold(f"scene{sid}")
    o = replay(g)
    w, h, spp, bounces = (int(v) for v in g["size"])
    o.set_uniforms(bounces, float(g["ior"]), DATE, 0, g["inv_v"], g["inv_pv"])
    assert_image(o.render(w, h, 0, spp), g["image"], f"scene {sid}")
    prim = np.stack([np.concatenate(o.primary_ray(x, y, w, h)) for y in (0, 17, 47) for x in (0, 31, 63)])
    assert np.array_equal(prim, g["primary"])                     # raytracer.vert + interpolation
    if sid == 2:                                                  # the two debug shaders: no transcendental
        assert np.array_equal(o.render(w, h, 0, 2, 1), g["image_mode1"])
        assert np.array_equal(o.render(w, h, 0, 2, 2), g["image_mode2"])
    if sid == 6:
        o.set_uniforms(bounces, 2.0, DATE, 0, g["inv_v"], g["inv_pv"])
        assert_image(o.render(w, h, 0, 8), g["image_ior2"], "scene 6 ior 2")


def test_mesh_scene_matches_the_reference():
    """add_mesh / place_mesh / send_mesh tables, Mesh_intersect + Triangle_intersect, smooth and flat normals."""
    g = gold("mesh")
    o = replay(g, mesh_inputs(g))
    assert np.array_equal(o.prim_data(), g["prims"])
    b = o.bvh()
    assert b["depth"] == int(g["bvh_depth"]) and np.array_equal(b["ind"], g["bvh_ind"]) and np.array_equal(b["bb"], g["bvh_bb"])
    for m in range(2):
        mb = o.bvh(m)
        assert mb["depth"] == int(g[f"m{m}_depth"])
        assert np.array_equal(mb["ind"], g[f"m{m}_ind"]) and np.array_equal(mb["bb"], g[f"m{m}_bb"])
    got = o.trace(g["ray_o"], g["ray_d"])
    assert got.tobytes() == g["hits"].tobytes()
    assert (got["tri_index"] >= 0).sum() > 1000
    w, h, spp, bounces = (int(v) for v in g["size"])
    o.set_uniforms(bounces, float(g["ior"]), DATE, 0, g["inv_v"], g["inv_pv"])
    assert_image(o.render(w, h, 0, spp), g["image"], "mesh")
    o.set_uniforms(bounces, float(g["ior"]), DATE, 1, g["inv_v"], g["inv_pv"])
    assert_image(o.render(w, h, 0, 4), g["image_flat"], "mesh flat_face")


def test_rng_sampler_camera_transfo_units():
    g = gold("units")
    o = Oracle()
    for k, (tx, ty, p, d) in enumerate(g["rng_cases"]):
        s, seed = o.rng_stream(float(tx), float(ty), int(p), float(d), 256)
        assert np.array_equal(seed, g[f"rng_seed_{k}"]) and np.array_equal(s, g[f"rng_{k}"])
        assert (s >= 0).all() and (s < 1).all()
    for k in range(4):
        r = o.random_rays(g[f"sampler_D_{k}"], float(g[f"sampler_rough_{k}"]), n=512)
        assert np.allclose(r, g[f"sampler_{k}"], rtol=1e-6, atol=1e-7)
    for k in range(4):
        w, h = (int(v) for v in g[f"cam_size_{k}"])
        iv, ipv = capi.default_camera(w, h)                       # host mirror vs the reference's Camera
        assert np.allclose(iv, g[f"cam_inv_v_{k}"], rtol=1e-5, atol=1e-4)
        assert np.allclose(ipv, g[f"cam_inv_pv_{k}"], rtol=1e-5, atol=1e-6)
    assert np.array_equal(capi.transfo("tZs", 70, 20, -40, 20, 20, 20, 60), g["transfo_0"])
    assert np.array_equal(capi.transfo("tYs", 200, 0, 100, -110, 20, 20, 1), g["transfo_1"])
    assert np.array_equal(capi.transfo("tXu", 1, 2, 3, 45, 2.5), g["transfo_2"])


@pytest.mark.skipif(not have_reference(), reason="oracle/_ref/libref.so not built (needs /root/reference)")
@pytest.mark.parametrize("sid,bounces,ior", [(2, 9, 1.5), (5, 6, 1.0), (7, 9, 2.0)])
def test_live_reference_build_agrees_with_the_restatement(sid, bounces, ior):
    """Where the reference build is present: fresh rays and a fresh image, restatement vs reference."""
    from common import random_rays
    r = Oracle("reference").build_repo_scene(sid, 1.2)
    raw = r.prim_data()
    r.finalize()
    o = Oracle().load_unsorted(raw).finalize()
    ro, rd = random_rays(o.prim_data(), 20000, seed=900 + sid)
    assert r.trace(ro, rd).tobytes() == o.trace(ro, rd).tobytes()
    iv, ipv = r.default_camera(80, 50)
    for x in (r, o):
        x.set_uniforms(bounces, ior, np.float32(0.031), 0, iv, ipv)
    assert np.array_equal(r.render(80, 50, 3, 5), o.render(80, 50, 3, 5), equal_nan=True)


def test_scene8_waiver_is_the_noise_floor_of_the_transcendentals():
    """tests/test_gpu_golden.py lets 10 % of scene 8's 16-spp pixels differ from the reference fixture and 1.5 % on
    the other scenes. Evidence that this is amplification of last-ulp differences in sin / cos / log / pow (which
    GLSL, glibc and CUDA are all free to round differently), not a defect: the oracle against ITSELF with every
    transcendental result of the bounce sampler moved by ONE ulp. Scenes 1 and 6: no 16-spp pixel moves by more than
    the 2e-3 tolerance. Scene 8 (hall of columns on a near-mirror floor): identical up to 2 bounces, and at the
    fixture's 9 bounces 3-5 % of the pixels move -- from one ulp; CUDA's libm differs from glibc by up to 2 ulps
    and the MUFU sampler by more."""
    from common import both
    w, h, spp = 64, 48, 16
    drift = {}
    for sid in (1, 6, 8):
        s, o = both(sid)
        iv, ipv = capi.default_camera(w, h)
        for nb in (2, 9):
            o.set_uniforms(nb, 1.0, DATE, 0, iv, ipv)
            a = o.render(w, h, 0, spp) / spp
            try:
                o.set_ulp_perturb(1)
                b = o.render(w, h, 0, spp) / spp
            finally:
                o.set_ulp_perturb(0)
            drift[sid, nb] = 1.0 - np.isclose(a, b, rtol=2e-3, atol=2e-3).all(axis=2).mean()
    print({k: round(float(v), 4) for k, v in drift.items()})
    assert drift[1, 9] <= 0.002 and drift[6, 9] <= 0.002
    assert drift[8, 2] <= 0.002
    assert 0.01 <= drift[8, 9] <= 0.10


@pytest.mark.parametrize("name", ["c2", "c5", "scene3"])
def test_bench_reference_arm_builds_the_same_scene_by_itself(name):
    """bench.py --impl reference builds repo scenes (and the two configs derived from them) with the reference's own
    scene builders, BVH builder and camera; the tables must equal the host mirror's that the GPU arm uploads
    (records bit for bit except the inverse transform, which every stand-in for Eigen's inverse rounds differently)."""
    from oracle.binding import have_reference
    if not have_reference():
        pytest.skip("oracle/_ref/libref.so not built")
    import bench
    from common import host_scene
    from montecarlo_pathtracing_b200 import capi
    kind = bench.WORKLOADS[name][0]
    o, iv, ipv, how = bench.reference_scene(name)
    assert o.cpu_kind == "reference" and "own code" in how
    s = host_scene(kind, 0.4)
    s.finalize()
    a, b = o.prim_data(), s.prim_data()
    assert a.shape == b.shape
    keep = np.r_[0:16, 32:64]
    assert np.array_equal(a[:, keep].view(np.uint32), b[:, keep].view(np.uint32))
    assert np.abs(a[:, 16:32] - b[:, 16:32]).max() <= 1e-4
    w, h = bench.WORKLOADS[name][2:4]
    hv, hpv = capi.default_camera(w, h)
    # (the camera matrices are inverses too: equal to the rounding of the two inverse routines)
    assert np.allclose(np.float32(iv), np.float32(hv).ravel(), rtol=1e-5, atol=1e-5)
    assert np.allclose(np.float32(ipv), np.float32(hpv).ravel(), rtol=1e-5, atol=1e-5)
    t = s.tables()["root"]
    rb = o.bvh()
    assert int(rb["depth"]) == int(t["depth"]) and np.array_equal(rb["ind"], t["ind"])
