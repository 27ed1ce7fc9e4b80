"""Next-event estimation (MCPT_OPT_NEE): a NEW estimator with its own parity class -- SURVEY 8(f) n3.

The reference has no working light sampler (shaders/raytracer_func.frag:908-956 is commented out), so there are no
records to be equal to. What NEE must do instead:
  * converge to the SAME image: 16 384 spp against the reference's 16 384-spp image of the same scene
    (tests/golden/converged*.npz), inside the bar the plain estimator is held to for independent passes
    (RMSE <= 1e-2, mean luminance within 0.5 %);
  * be less noisy at equal spp on a scene lit by a quad (RMSE of 64-spp images against the converged one);
  * leave the bounce rays alone: same path segments as without it, only shadow rays on top; with the option off
    again the image is the default one bit for bit.
"""
import numpy as np
import pytest

from common import DATE, desc_from_arrays
from montecarlo_pathtracing_b200 import Context, capi

pytestmark = pytest.mark.gpu
GOLD = __import__("pathlib").Path(__file__).resolve().parent / "golden"


@pytest.fixture(scope="module")
def ctx():
    c = Context(0)
    c.set_option(capi.OPT_FAST_MATH, 1)
    yield c
    c.close()


def _setup(ctx, sid):
    cv = np.load(GOLD / f"converged{sid}.npz")
    g = cv if isinstance(sid, str) else np.load(GOLD / f"scene{sid}.npz")
    d, keep = desc_from_arrays(g["prims"], {"depth": g["bvh_depth"], "bb": g["bvh_bb"], "ind": g["bvh_ind"]}, ())
    d.nb_emissives = int(g["nb_emissives"])
    ctx.upload_scene(d)
    w, h, bounces = (int(v) for v in cv["size"])
    ctx.resize(w, h)
    ctx.set_camera(cv["inv_v"], cv["inv_pv"])
    ctx.set_params(nb_bounces=bounces, refract_ind=float(cv["ior"]), date=float(DATE))
    return cv, keep


def _mean_image(ctx, first, spp, chunk=1024):
    h, w = ctx.height, ctx.width
    acc = np.zeros((h, w, 3), np.float64)
    for k in range(0, spp, chunk):
        n = min(chunk, spp - k)
        ctx.clear()
        ctx.render(first + k, n)
        acc += ctx.read_accum()[..., :3]
    return acc / spp


def _rmse(a, b):
    return float(np.sqrt(np.mean((np.asarray(a, np.float64) - np.asarray(b, np.float64)) ** 2)))


@pytest.mark.parametrize("sid", [1, "c2"])
def test_nee_converges_to_the_reference_image(ctx, sid):
    cv, keep = _setup(ctx, sid)
    ref = cv["image"].astype(np.float64)
    spp = int(cv["spp"])
    ctx.set_option(capi.OPT_NEE, 1)
    try:
        img = _mean_image(ctx, spp, spp)           # 16 384 passes the reference image did not use
    finally:
        ctx.set_option(capi.OPT_NEE, 0)
    rmse, dlum = _rmse(img, ref), abs(img.mean() - ref.mean()) / ref.mean()
    print(f"scene {sid}, NEE, {spp} other passes: RMSE {rmse:.3e}, mean luminance off by {dlum:.3e}")
    assert rmse <= 1e-2 and dlum <= 5e-3


def test_nee_is_less_noisy_at_equal_spp_and_keeps_the_bounce_rays(ctx):
    cv, keep = _setup(ctx, 1)
    ref = cv["image"].astype(np.float64)
    off, on, stats = [], [], {}
    for nee in (0, 1):
        ctx.set_option(capi.OPT_NEE, nee)
        try:
            ctx.reset_stats()
            errs = [_rmse(_mean_image(ctx, 40_000 + 64 * k, 64), ref) for k in range(12)]
            stats[nee] = ctx.stats()
        finally:
            ctx.set_option(capi.OPT_NEE, 0)
        (on if nee else off).extend(errs)
    print(f"64 spp RMSE against the converged image: plain {np.mean(off):.4e}, NEE {np.mean(on):.4e}; "
          f"rays {stats[0]['rays']} -> {stats[1]['rays']}, segments {stats[0]['segments']} / {stats[1]['segments']}")
    assert np.mean(on) < 0.9 * np.mean(off)
    assert stats[1]["segments"] == stats[0]["segments"]          # the same paths were followed ...
    assert stats[1]["rays"] > stats[0]["rays"]                   # ... plus the shadow rays


def test_nee_off_is_the_default_image_bit_for_bit(ctx):
    cv, keep = _setup(ctx, 1)
    ctx.clear()
    ctx.render(0, 16)
    a = ctx.read_accum()
    ctx.set_option(capi.OPT_NEE, 1)
    try:
        ctx.clear()
        ctx.render(0, 16)
        b = ctx.read_accum()
    finally:
        ctx.set_option(capi.OPT_NEE, 0)
    ctx.clear()
    ctx.render(0, 16)
    c = ctx.read_accum()
    assert np.array_equal(a, c) and not np.array_equal(a, b)
    assert abs(b[..., :3].mean() - a[..., :3].mean()) < 0.05 * a[..., :3].mean()


def test_nee_agrees_with_the_plain_estimator_on_a_mesh_scene(ctx):
    """Meshes (one of them transparent: interior rays), the unified tree, rough surfaces (wide cones) and a small quad
    light that paths find rarely by chance: NEE and the plain estimator have the same mean image, and two disjoint
    2048-spp renders differ LESS from each other with NEE than without (that difference is pure noise)."""
    from test_gpu_unified import _instanced_scene
    from montecarlo_pathtracing_b200 import Scene
    base, _ = _instanced_scene()
    t = base.tables()
    s = Scene()
    m = t["meshes"][0]
    mid = s.add_mesh(m["pos"], m["tri"])
    s.place_mesh(mid, capi.transfo("tZs", -45, 10, 0, 30, 25, 25, 25), (0.8, 0.7, 0.6, 1.0))
    s.place_mesh(mid, capi.transfo("tXs", 45, -10, 5, 75, 20, 25, 18), (0.3, 0.7, 0.9, 0.3))                  # transparent
    s.add_orientedQuad(capi.transfo("tZs", 0, 0, -40, 0, 400, 400, 1), (0.7, 0.7, 0.7, 1.0))                   # floor, roughness 0: cone = hemisphere
    s.add_orientedQuad(capi.transfo("tXs", 0, 0, 90, 180, 12, 12, 1), (0.9, 0.9, 0.9, 1.0), emissivity=30.0)   # small light, facing down
    s.finalize()
    ctx.upload_scene(s)
    ctx.resize(96, 54)
    ctx.set_camera(*capi.default_camera(96, 54))
    ctx.set_params(nb_bounces=8, refract_ind=1.5, date=float(DATE))
    imgs, rays = {}, {}
    for nee in (0, 1):
        ctx.set_option(capi.OPT_NEE, nee)
        try:
            ctx.reset_stats()
            imgs[nee] = [_mean_image(ctx, first, 2048) for first in (0, 2048)]
            rays[nee] = ctx.stats()["rays"]
        finally:
            ctx.set_option(capi.OPT_NEE, 0)
    noise = {k: _rmse(v[0], v[1]) for k, v in imgs.items()}
    mean = {k: float((v[0].mean() + v[1].mean()) / 2) for k, v in imgs.items()}
    print(f"mesh scene, 2 x 2048 spp: noise plain {noise[0]:.3e}, NEE {noise[1]:.3e}; mean luminance {mean[0]:.4f} / {mean[1]:.4f}; rays {rays}")
    assert rays[1] > rays[0]                                  # shadow rays were traced
    assert abs(mean[0] - mean[1]) <= 2e-2 * mean[0]
    assert noise[1] < 0.8 * noise[0]


def test_nee_without_a_sampled_light_changes_nothing(ctx):
    """Scenes whose lights are not quads (or that have none), and a descriptor whose nb_emissives is 0 or too large:
    the option is a no-op, or clamps -- never an error and never a different image than the plain estimator."""
    cv, keep = _setup(ctx, 1)
    g = np.load(GOLD / "scene1.npz")
    imgs = []
    for nb_emi in (int(g["nb_emissives"]), 0, 10_000):
        d, keep2 = desc_from_arrays(g["prims"], {"depth": g["bvh_depth"], "bb": g["bvh_bb"], "ind": g["bvh_ind"]}, ())
        d.nb_emissives = nb_emi
        ctx.set_option(capi.OPT_NEE, 1 if nb_emi != int(g["nb_emissives"]) else 0)
        try:
            ctx.upload_scene(d)
            ctx.clear()
            ctx.render(0, 8)
            imgs.append(ctx.read_accum())
        finally:
            ctx.set_option(capi.OPT_NEE, 0)
    assert np.array_equal(imgs[0], imgs[1])                 # nb_emissives = 0: no light table, plain estimator
    assert np.isfinite(imgs[2]).all() and not np.array_equal(imgs[0], imgs[2])   # clamped to n_prims: NEE on the real light
