"""GPU parity against the golden fixtures of tests/golden/ — outputs of the reference's own code
(shader text + bvh_gpu/*.cpp compiled as C++, tests/golden/make_golden.py). The CUDA path is fed the
reference's tables through the C ABI; no oracle is involved in these tests.

Bar: hit ids / face ids exact and distance within 1e-5 relative (BASELINE.json); the library does
better (bit-identical hit records), which is asserted. Images: fixed seeds, per-pixel agreement
within 2e-3 on >= 98.5 % of the pixels (sin/cos/log/pow differ in the last ulps between CUDA and
glibc and a few chaotic paths flip a hit) and mean luminance within 0.5 %."""
from pathlib import Path

import numpy as np
import pytest

from common import DATE, desc_from_arrays
from montecarlo_pathtracing_b200 import Context, capi

pytestmark = pytest.mark.gpu
GOLD = Path(__file__).resolve().parent / "golden"


@pytest.fixture(scope="module")
def ctx():
    c = Context(0)
    yield c
    c.close()


def _upload(ctx, g, meshes=()):
    d, keep = desc_from_arrays(g["prims"], {"depth": g["bvh_depth"], "bb": g["bvh_bb"], "ind": g["bvh_ind"]}, meshes)
    ctx.upload_scene(d)
    return keep


def _image_ok(img, ref_sum, spp, frac=0.985):
    ref = ref_sum / np.float32(spp)
    ok = np.isclose(img, ref, rtol=2e-3, atol=2e-3).all(axis=2)
    assert ok.mean() >= frac, f"only {ok.mean():.4f} of the pixels agree"
    assert abs(img.mean() - ref.mean()) <= 5e-3 * abs(ref.mean())          # mean luminance within 0.5 %
    assert np.sqrt(np.mean((img - ref) ** 2)) <= 0.02                      # RMSE, radiance units (white = 1)


# Scene 8 (colonnes) is a hall of 891 columns on a near-mirror floor (roughness 0.9999): a last-ulp
# difference in sin/cos is amplified over 9 bounces, so more 16-spp pixels contain one diverged path.
# Measured, not assumed: the CPU oracle against itself with the sampler's transcendentals moved by ONE ulp loses
# 3-5 % of scene 8's pixels at 9 bounces and none on the other scenes (tests/test_oracle_golden.py::
# test_scene8_waiver_is_the_noise_floor_of_the_transcendentals); test_scene8_disagreement_grows_with_the_bounce_count
# below shows the GPU agreeing on >= 99.9 % of the pixels at 1-2 bounces with the same hit records everywhere.
PIXEL_FRACTION = {8: 0.90}


# Every walker the library can select goes through the SAME kernel a render launches (k_traverse; the ray batch
# rides through ray / hit sectors, mcpt_trace_rays with MCPT_OPT_TRACE_WALKER = 0): the warp-coherent flat list
# (scenes of <= 16 primitives), persistent warps over the SAH-built top2 tree (the default of scenes 3, 5, 7, 8 --
# the kernel bench.py times there), persistent warps over the reference's topology, the stack-free heap walk, and
# the reference-order walk of k_trace. Options in the order (flat, persistent, top_sah, trace_walker).
WALKERS = {
    "default": (1, 1, 1, 0),
    "persistent-sah": (0, 1, 1, 0),
    "persistent-ref-topology": (0, 1, 0, 0),
    "heap-walk": (0, 0, 1, 0),
    "k_trace": (1, 1, 1, 1),
}


def _set_walker(ctx, walker):
    flat, persistent, top_sah, trace_walker = WALKERS[walker]
    ctx.set_option(capi.OPT_FLAT_WALK, flat)
    ctx.set_option(capi.OPT_PERSISTENT, persistent)
    ctx.set_option(capi.OPT_TOP_SAH, top_sah)
    ctx.set_option(capi.OPT_TRACE_WALKER, trace_walker)


@pytest.mark.parametrize("sid", range(1, 9))
@pytest.mark.parametrize("walker", list(WALKERS))
def test_traversal_vs_reference_fixture(ctx, sid, walker):
    g = np.load(GOLD / f"scene{sid}.npz")
    _set_walker(ctx, walker)
    try:
        _upload(ctx, g)
        got = ctx.trace_rays(g["ray_o"], g["ray_d"])
        any_hit = ctx.trace_rays(g["ray_o"], g["ray_d"], any_hit=True)
    finally:
        _set_walker(ctx, "default")
    ref = g["hits"]
    for f in ("index", "tri_index", "shape", "dir"):
        assert np.array_equal(got[f], ref[f]), f
    hit = ref["index"] >= 0
    assert np.all(np.abs(got["dist"][hit] - ref["dist"][hit]) <= 1e-5 * ref["dist"][hit])   # the stated bar
    assert got.tobytes() == ref.tobytes()                                                    # what the library delivers
    assert np.array_equal(any_hit["index"] >= 0, g["any_hit"])


@pytest.mark.parametrize("sid", range(1, 9))
def test_image_vs_reference_fixture(ctx, sid):
    g = np.load(GOLD / f"scene{sid}.npz")
    _upload(ctx, g)
    w, h, spp, bounces = (int(v) for v in g["size"])
    ctx.resize(w, h)
    ctx.set_camera(g["inv_v"], g["inv_pv"])
    ctx.set_params(nb_bounces=bounces, refract_ind=float(g["ior"]), date=float(DATE))
    for fm in (0, 1):
        ctx.set_option(capi.OPT_FAST_MATH, fm)
        ctx.clear()
        ctx.render(0, spp)
        _image_ok(ctx.resolve(spp), g["image"], spp, frac=PIXEL_FRACTION.get(sid, 0.985 if fm == 0 else 0.98))
    ctx.set_option(capi.OPT_FAST_MATH, 0)
    if sid == 2:
        for mode, key in ((1, "image_mode1"), (2, "image_mode2")):
            ctx.set_params(nb_bounces=bounces, refract_ind=1.0, date=float(DATE), mode=mode)
            ctx.clear()
            ctx.render(0, 2)
            assert np.allclose(ctx.read_accum()[..., :3], g[key], rtol=1e-6, atol=1e-7)
    if sid == 6:
        ctx.set_params(nb_bounces=bounces, refract_ind=2.0, date=float(DATE))
        ctx.clear()
        ctx.render(0, 8)
        _image_ok(ctx.resolve(8), g["image_ior2"], 8)


def test_scene8_disagreement_grows_with_the_bounce_count(ctx):
    """The scene-8 waiver (PIXEL_FRACTION) backed by a measurement: against the oracle at the fixture's size and
    seeds, the first bounces agree on (nearly) every pixel -- same hit records (the traversal tests), same shading --
    and the share of 16-spp pixels holding a diverged path grows with the bounce count, as chaotic amplification of
    last-ulp differences in the sampler's sin / cos / log / pow does."""
    from common import both
    s, o = both(8)                                   # light 0.4, as the fixture (tests/golden/make_golden.py)
    g = np.load(GOLD / "scene8.npz")
    w, h, spp, _ = (int(v) for v in g["size"])
    ctx.upload_scene(o.scene_desc())
    ctx.resize(w, h)
    ctx.set_camera(g["inv_v"], g["inv_pv"])
    frac = {}
    for nb in (1, 2, 4, 9):
        o.set_uniforms(nb, float(g["ior"]), DATE, 0, g["inv_v"], g["inv_pv"])
        ref = o.render(w, h, 0, spp) / np.float32(spp)
        ctx.set_params(nb_bounces=nb, refract_ind=float(g["ior"]), date=float(DATE))
        ctx.clear()
        ctx.render(0, spp)
        frac[nb] = float(np.isclose(ctx.resolve(spp), ref, rtol=2e-3, atol=2e-3).all(axis=2).mean())
    print("scene 8, share of pixels within 2e-3 of the oracle by bounce count:", frac)
    assert frac[1] >= 0.999 and frac[2] >= 0.995
    assert frac[9] >= PIXEL_FRACTION[8]
    assert frac[9] <= frac[2]


def _mesh_fixture():
    g = np.load(GOLD / "mesh.npz")
    meshes = [{"bvh": {"depth": g[f"m{m}_depth"], "bb": g[f"m{m}_bb"], "ind": g[f"m{m}_ind"]}, "tri": g[f"m{m}_tri"],
               "pos": g[f"m{m}_pos"], "nrm": g[f"m{m}_nrm"], "mesh_line": g[f"m{m}_line"]} for m in range(2)]
    return g, meshes


# (persistent, mesh_bvh, top_sah, trace_walker): four-children persistent walk over the GPU-built trees (default,
# what bench.py times on configs 3 / 4), the same over the reference's top-level topology, one ray per lane over
# the GPU-built two-children trees, the same over the uploaded median-split trees, and k_trace.
MESH_WALKERS = {
    "default-wide-persistent": (1, 1, 1, 0),
    "wide-persistent-ref-topology": (1, 1, 0, 0),
    "one-ray-per-lane-lbvh": (0, 1, 1, 0),
    "one-ray-per-lane-reference-trees": (0, 0, 1, 0),
    "k_trace": (1, 1, 1, 1),
    "k_trace-reference-trees": (1, 0, 1, 1),
}


@pytest.mark.parametrize("walker", list(MESH_WALKERS))
def test_mesh_traversal_vs_reference_fixture(ctx, walker):
    g, meshes = _mesh_fixture()
    persistent, mesh_bvh, top_sah, trace_walker = MESH_WALKERS[walker]
    opts = ((capi.OPT_PERSISTENT, persistent, 1), (capi.OPT_MESH_BVH, mesh_bvh, 1), (capi.OPT_TOP_SAH, top_sah, 1),
            (capi.OPT_TRACE_WALKER, trace_walker, 0))
    for o, v, _ in opts:
        ctx.set_option(o, v)
    try:
        _upload(ctx, g, meshes)
        got = ctx.trace_rays(g["ray_o"], g["ray_d"])
    finally:
        for o, _, dflt in opts:
            ctx.set_option(o, dflt)
    assert got.tobytes() == g["hits"].tobytes()


def test_mesh_scene_vs_reference_fixture(ctx):
    g, meshes = _mesh_fixture()
    _upload(ctx, g, meshes)
    got = ctx.trace_rays(g["ray_o"], g["ray_d"])
    assert got.tobytes() == g["hits"].tobytes()
    w, h, spp, bounces = (int(v) for v in g["size"])
    ctx.resize(w, h)
    ctx.set_camera(g["inv_v"], g["inv_pv"])
    ctx.set_params(nb_bounces=bounces, refract_ind=float(g["ior"]), date=float(DATE))
    ctx.render(0, spp)
    _image_ok(ctx.resolve(spp), g["image"], spp, frac=0.975)
    ctx.set_params(nb_bounces=bounces, refract_ind=float(g["ior"]), date=float(DATE), flat_face=1)
    ctx.clear()
    ctx.render(0, 4)
    _image_ok(ctx.resolve(4), g["image_flat"], 4, frac=0.975)


def test_rng_is_bit_exact_through_the_debug_shader(ctx):
    """mode 2 (tp/montecarlo_mat_tr.frag) returns col.rgb * random_float(): one draw per pixel, so the
    image is the RNG stream itself; it is compared exactly above (image_mode2). Here: determinism."""
    g = np.load(GOLD / "scene2.npz")
    _upload(ctx, g)
    ctx.resize(64, 48)
    ctx.set_camera(g["inv_v"], g["inv_pv"])
    ctx.set_params(nb_bounces=3, refract_ind=1.0, date=float(DATE), mode=2)
    ctx.render(0, 2)
    a = ctx.read_accum()
    ctx.clear()
    ctx.render(0, 2)
    assert np.array_equal(a, ctx.read_accum())


@pytest.mark.parametrize("sid", [1, 6, "c2", "c5"])
def test_converges_to_the_reference_16k_spp_image(ctx, sid):
    """BASELINE.json: RMSE against a 16k-spp converged image of the reference under a stated threshold
    and mean luminance within 0.5 %. The converged image comes from the reference shader text
    (tests/golden/make_converged.py). (a) the same 16384 passes on the GPU: only chaotic paths and the
    summation order differ -> RMSE <= 2e-3; (b) 16384 OTHER passes: two independent estimates, Monte
    Carlo noise only -> RMSE <= 1e-2 (radiance units, white = 1)."""
    cv = np.load(GOLD / f"converged{sid}.npz")
    # BASELINE configs 2 and 5 ("c2": scene 6 at roughness 0.5, the headline workload; "c5": divergence stress, 16 bounces)
    # carry the tables the reference built in the converged fixture itself
    g = cv if isinstance(sid, str) else np.load(GOLD / f"scene{sid}.npz")
    _upload(ctx, g)
    w, h, bounces = (int(v) for v in cv["size"])
    spp = int(cv["spp"])
    ctx.resize(w, h)
    ctx.set_camera(cv["inv_v"], cv["inv_pv"])
    ctx.set_params(nb_bounces=bounces, refract_ind=float(cv["ior"]), date=float(DATE))
    ctx.set_option(capi.OPT_FAST_MATH, 1)
    try:
        # measured on a B200 (same passes / other passes): scene 1 1.3e-5 / 6.1e-3, scene 6 4.3e-5 / 3.5e-3,
        # c2 7.0e-6 / 2.9e-3, c5 1.0e-5 / 1.6e-3
        for first, limit in ((0, 2e-3), (spp, 1e-2)):
            acc = np.zeros((h, w, 3), np.float64)
            for k in range(0, spp, 1024):                 # float32 sums of 1024 passes, combined in float64
                ctx.clear()
                ctx.render(first + k, 1024)
                acc += ctx.read_accum()[..., :3]
            img = acc / spp
            ref = cv["image"].astype(np.float64)
            rmse = float(np.sqrt(np.mean((img - ref) ** 2)))
            dlum = abs(img.mean() - ref.mean()) / ref.mean()
            print(f"scene {sid}, {spp} passes from {first}: RMSE {rmse:.3e}, mean luminance off by {dlum:.3e}")
            assert rmse <= limit, (first, rmse)
            assert dlum <= 5e-3, (first, dlum)
    finally:
        ctx.set_option(capi.OPT_FAST_MATH, 0)


# ---- the bounce sampler on its own (tp/montecarlo.frag:49-89 = tp/hsphere.vert:1-41, the reference's DrawSampling)
@pytest.mark.parametrize("k", [0, 1, 2, 3])
def test_sampler_draws_vs_reference_fixture(ctx, k):
    """512 consecutive random_ray draws of one RNG stream against the reference's own (units.npz): the RNG is integer
    and exact, the transcendentals differ in the last ulps between CUDA and glibc -> 1e-5 absolute on unit vectors."""
    u = np.load(GOLD / "units.npz")
    D, rough, ref = u[f"sampler_D_{k}"], float(u[f"sampler_rough_{k}"]), u[f"sampler_{k}"]
    for fm in (0, 1):
        ctx.set_option(capi.OPT_FAST_MATH, fm)
        try:
            got = ctx.sample_directions(D, rough, len(ref))
        finally:
            ctx.set_option(capi.OPT_FAST_MATH, 0)
        assert np.abs(got - ref).max() <= (1e-5 if fm == 0 else 2e-4), (fm, np.abs(got - ref).max())
        assert np.abs(np.linalg.norm(got, axis=1) - 1.0).max() <= 1e-5


@pytest.mark.parametrize("rough", [1.0, 0.5, 0.1])
def test_sampling_cone_matches_the_oracle_statistics(ctx, rough, tmp_path):
    """10^5 draws at the three roughness values of the reference's README pictures: the GPU sampler and the CPU oracle
    draw the same stream (element-wise to 1e-5), so cone statistics and histogram agree; the cone narrows with the
    roughness as README.md:5-14 shows (roughness 1: the whole hemisphere, 0.1: a few degrees around the normal)."""
    from montecarlo_pathtracing_b200 import imageio
    from oracle.binding import Oracle
    n, D = 100_000, np.float32([0.0, 0.6, 0.8])
    got = ctx.sample_directions(D, rough, n)
    ref = Oracle().random_rays(D, rough, n=n)
    assert np.abs(got - ref).max() <= 1e-5
    sg, sr = imageio.cone_stats(got, D), imageio.cone_stats(ref, D)
    for key in ("mean_cos", "angle_p50", "angle_p90", "angle_p99"):
        assert abs(sg[key] - sr[key]) <= 5e-3 * max(abs(sr[key]), 1.0), (key, sg[key], sr[key])
    assert sg["below_plane"] == 0.0
    _, hg = imageio.cone_histogram(got, D)
    _, hr = imageio.cone_histogram(ref, D)
    assert np.abs(hg - hr).max() <= 1e-4
    # the cone of README.md:5-14: median angle to the normal
    lo, hi = {1.0: (35.0, 45.0), 0.5: (10.0, 13.5), 0.1: (0.3, 0.7)}[rough]
    assert lo <= sg["angle_p50"] <= hi, sg
    img = imageio.sampling_plot(got, D)
    p = imageio.write_png(tmp_path / f"sampling_{rough}.png", img, bottom_up=False)
    assert imageio.read_png(p).shape[:2] == (256, 256)
    print(f"roughness {rough}: {sg}")
