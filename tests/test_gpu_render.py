"""GPU parity: fixed-seed images through the C ABI vs the CPU oracle, plus the
accumulate / resolve contract."""
import numpy as np
import pytest

from common import DATE, both
from montecarlo_pathtracing_b200 import Context, capi

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    c = Context(0)
    yield c
    c.close()


def _render_both(ctx, kind, w, h, spp, bounces, ior, mode=0, **kw):
    s, o = both(kind, **kw)
    iv, ipv = capi.default_camera(w, h)
    o.set_uniforms(bounces, ior, DATE, 0, iv, ipv)
    ctx.upload_scene(o.scene_desc())
    ctx.resize(w, h)
    ctx.set_camera(iv, ipv)
    ctx.set_params(nb_bounces=bounces, refract_ind=ior, date=float(DATE), mode=mode)
    ctx.render(0, spp)
    img = ctx.resolve(spp)
    ref = o.render(w, h, 0, spp, mode) / np.float32(spp)
    return img, ref


def _assert_image_parity(img, ref, frac=0.985, lum=5e-3):
    """Per-pixel: RNG and hit ids are bit-exact, only sin/cos/log/pow differ in the last
    ulps, so all but a few chaotic paths (a last-ulp change flipping a hit) must agree."""
    assert np.isfinite(img).all() == np.isfinite(ref).all()
    ok = np.isclose(img, ref, rtol=2e-3, atol=2e-3).all(axis=2)
    assert ok.mean() >= frac, f"only {ok.mean():.4f} of the pixels agree"
    m_img, m_ref = np.nanmean(img), np.nanmean(ref)
    assert abs(m_img - m_ref) <= lum * abs(m_ref), (m_img, m_ref)   # mean luminance within 0.5 %


@pytest.mark.parametrize("kind,bounces,ior", [(1, 9, 1.0), (2, 8, 1.5), (4, 5, 2.0), ("c2", 8, 1.5), (7, 8, 1.0)])
def test_fixed_seed_image_matches_oracle(ctx, kind, bounces, ior):
    img, ref = _render_both(ctx, kind, 160, 120, 8, bounces, ior)
    _assert_image_parity(img, ref)


@pytest.mark.parametrize("kind,bounces,ior", [(2, 8, 1.5), ("c2", 8, 1.5), (8, 6, 1.0)])
def test_fast_math_image_parity(ctx, kind, bounces, ior):
    """MCPT_OPT_FAST_MATH (MUFU sin/cos/log/pow/rsqrt in the bounce sampler) and the register-capped
    traversal variants stay inside the same image-parity tolerance."""
    ctx.set_option(capi.OPT_FAST_MATH, 1)
    ctx.set_option(capi.OPT_TRAVERSE_OCC, 4)
    try:
        img, ref = _render_both(ctx, kind, 160, 120, 8, bounces, ior)
    finally:
        ctx.set_option(capi.OPT_FAST_MATH, 0)
        ctx.set_option(capi.OPT_TRAVERSE_OCC, 0)
    _assert_image_parity(img, ref, frac=0.98)


def test_divergence_stress_image(ctx):
    img, ref = _render_both(ctx, "c5", 160, 90, 4, 16, 1.5)
    _assert_image_parity(img, ref, frac=0.97)


def test_mesh_scene_image(ctx):
    img, ref = _render_both(ctx, "mesh", 128, 72, 4, 8, 1.5, n_meshes=4, tris_log2=9)
    _assert_image_parity(img, ref, frac=0.97)


@pytest.mark.parametrize("mode", [1, 2])
def test_debug_shaders(ctx, mode):
    """tp/montecarlo_mat.frag and tp/montecarlo_mat_tr.frag: no transcendental -> exact."""
    img, ref = _render_both(ctx, 2, 96, 64, 2, 3, 1.0, mode=mode)
    assert np.allclose(img, ref, rtol=1e-6, atol=1e-7)


def test_zero_bounces_is_black(ctx):
    img, ref = _render_both(ctx, 1, 32, 32, 2, 0, 1.0)
    assert not img.any() and not ref.any()


def test_accumulation_is_additive_in_pass_order(ctx):
    s, o = both(6)
    iv, ipv = capi.default_camera(96, 54)
    ctx.upload_scene(s)
    ctx.resize(96, 54)
    ctx.set_camera(iv, ipv)
    ctx.set_params(nb_bounces=6, refract_ind=1.5, date=float(DATE))
    ctx.render(0, 8)
    a = ctx.read_accum()
    ctx.clear()
    ctx.render(0, 3)
    ctx.render(3, 5)
    b = ctx.read_accum()
    assert np.array_equal(a, b)                       # bit-exact: blend order = pass order
    ctx.clear()
    ctx.set_option(capi.OPT_PASSES_PER_BATCH, 2)      # batch size must not change the result
    ctx.render(0, 8)
    assert np.array_equal(a, ctx.read_accum())
    ctx.set_option(capi.OPT_PASSES_PER_BATCH, 0)
    # resume from a checkpoint: write the accumulation back and continue
    ctx.clear()
    ctx.render(0, 3)
    saved = ctx.read_accum()
    ctx.clear()
    ctx.write_accum(saved)
    ctx.render(3, 5)
    assert np.array_equal(a, ctx.read_accum())
    st = ctx.stats()
    assert st["paths"] > 0 and st["rays"] >= st["paths"] and st["launches"] > 0


def test_resolve_variants(ctx):
    s, o = both(1)
    iv, ipv = capi.default_camera(40, 30)
    ctx.upload_scene(s)
    ctx.resize(40, 30)
    ctx.set_camera(iv, ipv)
    ctx.set_params(nb_bounces=4, refract_ind=1.0, date=float(DATE))
    ctx.render(0, 4)
    acc = ctx.read_accum()
    f = ctx.resolve(4)
    assert np.array_equal(f, acc[..., :3] / np.float32(4))        # fs_frag: tex / nb
    u8 = ctx.resolve(4, rgba8=True)
    want = np.rint(np.clip(f, 0, 1) * 255).astype(np.uint8)
    assert np.abs(u8[..., :3].astype(int) - want.astype(int)).max() <= 0 and (u8[..., 3] == 255).all()
    up = ctx.resolve(4, subsampling=3)                            # average.frag: texelFetch(xy / subsampling)
    assert up.shape == (90, 120, 3)
    assert np.array_equal(up, np.repeat(np.repeat(f, 3, axis=0), 3, axis=1))


def test_state_errors(ctx):
    from montecarlo_pathtracing_b200 import McptError
    c = Context(0)
    with pytest.raises(McptError):
        c.render(0, 1)          # no scene / size / camera yet
    with pytest.raises(McptError):
        c.clear()
    c.close()


@pytest.mark.parametrize("kind,bounces", [("c2", 8), (2, 6)])
def test_queue_order_does_not_change_the_image(ctx, kind, bounces):
    """The hit / survivor queues are filled by warp-aggregated atomics, so their order differs from
    run to run; per-path results, and therefore the pass-ordered sums, must be bit-identical, and
    the work counters too."""
    s, o = both(kind)
    iv, ipv = capi.default_camera(200, 120)
    ctx.upload_scene(s)
    ctx.resize(200, 120)
    ctx.set_camera(iv, ipv)
    ctx.set_params(nb_bounces=bounces, refract_ind=1.5, date=float(DATE))
    runs = []
    for ppb in (0, 0, 1):          # auto batch twice, then one pass per batch
        ctx.set_option(capi.OPT_PASSES_PER_BATCH, ppb)
        ctx.clear()
        ctx.reset_stats()
        ctx.render(0, 6)
        runs.append((ctx.read_accum(), ctx.stats()))
    ctx.set_option(capi.OPT_PASSES_PER_BATCH, 0)
    for img, st in runs[1:]:
        assert np.array_equal(runs[0][0], img)
        assert st["rays"] == runs[0][1]["rays"] and st["segments"] == runs[0][1]["segments"]
    assert runs[0][1]["rays"] > runs[0][1]["segments"] > 0


@pytest.mark.parametrize("kind,kw", [(3, {}), (7, {}), ("c5", {}), ("mesh", {"n_meshes": 3, "tris_log2": 10})])
def test_persistent_warps_do_not_change_the_image(kind, kw):
    """MCPT_OPT_PERSISTENT changes which lane traces which ray and when, nothing else: sums, ray and
    segment counts are bit-identical to the one-ray-per-lane kernel."""
    s, o = both(kind, **kw)
    iv, ipv = capi.default_camera(192, 108)
    out = []
    for persistent in (1, 0):
        c = Context(0)
        c.set_option(capi.OPT_PERSISTENT, persistent)
        c.upload_scene(s)
        c.resize(192, 108)
        c.set_camera(iv, ipv)
        c.set_params(nb_bounces=7, refract_ind=1.5, date=float(DATE))
        c.render(0, 5)
        out.append((c.read_accum(), c.stats()))
        c.close()
    assert np.array_equal(out[0][0], out[1][0])
    assert out[0][1]["rays"] == out[1][1]["rays"] and out[0][1]["segments"] == out[1][1]["segments"]
    assert out[0][0][..., :3].mean() > 0.01


def test_reupload_and_resize_give_the_same_bits():
    """A context that has held other scenes and sizes (parked and reused allocations, regrown wavefront,
    stale path tags in the reused buffers) renders exactly what a fresh context renders."""
    s2, _ = both(2)
    sm, _ = both("mesh", n_meshes=2, tris_log2=8)
    s7, _ = both(7)

    def render(c, s, w, h, spp):
        iv, ipv = capi.default_camera(w, h)
        c.upload_scene(s)
        c.resize(w, h)
        c.set_camera(iv, ipv)
        c.set_params(nb_bounces=6, refract_ind=1.5, date=float(DATE))
        c.clear()
        c.render(0, spp)
        return c.read_accum()

    fresh = Context(0)
    want = render(fresh, s2, 64, 48, 5)
    fresh.close()
    c = Context(0)
    render(c, sm, 200, 100, 3)
    render(c, s7, 96, 54, 2)
    render(c, sm, 64, 48, 4)
    got = render(c, s2, 64, 48, 5)
    again = render(c, s2, 64, 48, 5)
    c.close()
    assert np.array_equal(want, got) and np.array_equal(want, again)


def test_two_contexts_on_their_own_streams_do_not_interfere():
    import torch
    s2, _ = both(2)
    s6, _ = both("c2")
    iv, ipv = capi.default_camera(128, 72)
    ctxs, streams = [], []
    for s in (s2, s6):
        c = Context(0)
        st = torch.cuda.Stream()
        c.set_stream(st.cuda_stream)
        c.upload_scene(s)
        c.resize(128, 72)
        c.set_camera(iv, ipv)
        c.set_params(nb_bounces=6, refract_ind=1.5, date=float(DATE))
        ctxs.append(c)
        streams.append(st)
    for k in range(4):                       # interleaved, asynchronous launches
        for c in ctxs:
            c.render(2 * k, 2)
    both_imgs = [c.read_accum() for c in ctxs]
    for c in ctxs:
        c.clear()
        c.render(0, 8)                       # the same eight passes in one call, one context at a time
        c.synchronize()
    for c, img in zip(ctxs, both_imgs):
        assert np.array_equal(c.read_accum(), img)
        c.close()


@pytest.mark.gpu
def test_headless_render_tool_writes_the_back_buffer_and_reports_convergence(tmp_path):
    """montecarlo_pathtracing_b200/render.py: the viewer's call sequence replayed headless,
    PNG of the clamped 8-bit back buffer + float PFM, RMSE / mean luminance against the reference's
    16k-spp image of scene 6 (tests/golden/converged6.npz). 1024 passes: Monte Carlo noise only."""
    import json
    from pathlib import Path
    from montecarlo_pathtracing_b200 import imageio, render
    gold = Path(__file__).parent / "golden" / "converged6.npz"
    png, pfm = tmp_path / "s6.png", tmp_path / "s6.pfm"
    import contextlib
    import io
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf):
        render.main(["--scene", "6", "--width", "64", "--height", "48", "--spp", "1024", "--bounces", "8", "--ior", "1.5",
                     "--ref", str(gold), "--out", str(png), "--pfm", str(pfm)])
    rep = json.loads(buf.getvalue().strip().splitlines()[-1])
    assert rep["non_finite_pixels"] == 0
    assert rep["vs_ref"]["rmse"] <= 3e-2 and rep["vs_ref"]["luminance_rel_err"] <= 5e-3, rep["vs_ref"]
    img = imageio.read_pfm(pfm)
    assert img.shape == (48, 64, 3)
    assert np.array_equal(imageio.read_png(png), imageio.to_rgba8(img)[::-1, :, :3])
    # the progressive curve falls as passes accumulate
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf):
        render.main(["--scene", "6", "--width", "64", "--height", "48", "--bounces", "8", "--ior", "1.5", "--ref", str(gold),
                     "--curve", "4,64,1024"])
    curve = json.loads(buf.getvalue().strip().splitlines()[-1])["curve"]
    assert [c["spp"] for c in curve] == [4, 64, 1024]
    assert curve[0]["rmse"] > curve[1]["rmse"] > curve[2]["rmse"]


@pytest.mark.parametrize("kind,kw", [(1, {}), (2, {}), (6, {}), (3, {}), (8, {}), ("c5", {}), ("mesh", {"n_meshes": 3, "tris_log2": 9})])
def test_top_level_tree_choices_do_not_change_the_image(kind, kw):
    """The persistent warps walk a tree of the library's own (full-sweep SAH, rotated elongated primitives
    entered as slab boxes; MCPT_OPT_TOP_SAH) or the reference's median-split topology, small scenes the flat
    list (MCPT_OPT_FLAT_WALK): candidates are ranked by the reference's distance and visit order, so sums,
    ray and segment counts are the same bit for bit whichever structure found them."""
    s, o = both(kind, **kw)
    iv, ipv = capi.default_camera(160, 90)
    out = {}
    for flat_walk, top_sah in ((1, 1), (0, 1), (0, 0)):
        c = Context(0)
        try:
            c.set_option(capi.OPT_FLAT_WALK, flat_walk)
            c.set_option(capi.OPT_TOP_SAH, top_sah)
            c.upload_scene(s)
            c.resize(160, 90)
            c.set_camera(iv, ipv)
            c.set_params(nb_bounces=8, refract_ind=1.5, date=float(DATE))
            c.render(0, 4)
            out[(flat_walk, top_sah)] = (c.read_accum(), c.stats())
        finally:
            c.close()
    base, st = out[(1, 1)]
    assert base[..., :3].mean() > 0.01
    for key in ((0, 1), (0, 0)):
        assert np.array_equal(base, out[key][0]), key
        assert out[key][1]["rays"] == st["rays"] and out[key][1]["segments"] == st["segments"], key


def test_rotated_elongated_primitives_of_every_type():
    """A soup of 70 rotated, anisotropically scaled spheres, cubes, cylinders, cones and quads (none of the
    repo scenes rotates anything but cubes): the SAH tree with loose primitives entered as slab boxes, the
    reference's topology under persistent warps, and the stack-free walk over the reference's own tree give
    the same bits; the image agrees with the oracle."""
    from montecarlo_pathtracing_b200 import Scene
    from oracle.binding import Oracle
    T = capi.transfo
    rng = np.random.default_rng(11)
    s = Scene()
    s.add_orientedQuad(T("ts", 0, 0, -100, 400, 400, 1), [0.8, 0.8, 0.8, 1], roughness=0.8)
    s.add_orientedQuad(T("tXs", 0, 0, 160, 180, 60, 60, 1), [1, 1, 1, 1], emissivity=6.0)
    adders = [s.add_sphere, s.add_cube, s.add_cylinder, s.add_cone, s.add_orientedQuad]
    for i in range(70):
        pos = rng.uniform(-120, 120, 3)
        rot = rng.uniform(0, 360, 3)
        sc = rng.uniform(3, 12, 3)
        sc[rng.integers(0, 3)] *= rng.uniform(3, 8)                      # elongated along one local axis
        col = list(rng.uniform(0.2, 0.9, 3)) + [1.0 if i % 3 else 0.4]   # every third one transparent
        adders[i % 5](T("tZXYs", *pos, rot[0], rot[1], rot[2], *sc), col, shininess=0.5 * (i % 2), roughness=0.4)
    raw = s.prim_data()
    s.finalize()
    o = Oracle().load_unsorted(raw).finalize()
    w, h, spp, bounces = 160, 90, 3, 6
    iv, ipv = capi.default_camera(w, h)
    o.set_uniforms(bounces, 1.5, DATE, 0, iv, ipv)
    out = {}
    for persistent, top_sah in ((1, 1), (1, 0), (0, 1)):
        c = Context(0)
        try:
            c.set_option(capi.OPT_PERSISTENT, persistent)
            c.set_option(capi.OPT_TOP_SAH, top_sah)
            c.upload_scene(s)
            c.resize(w, h)
            c.set_camera(iv, ipv)
            c.set_params(nb_bounces=bounces, refract_ind=1.5, date=float(DATE))
            c.render(0, spp)
            out[(persistent, top_sah)] = (c.read_accum(), c.stats())
        finally:
            c.close()
    base, st = out[(1, 1)]
    for key in ((1, 0), (0, 1)):
        assert np.array_equal(base, out[key][0]), key
        assert out[key][1]["rays"] == st["rays"] and out[key][1]["segments"] == st["segments"], key
    ref = o.render(w, h, 0, spp) / np.float32(spp)
    _assert_image_parity(base[..., :3] / np.float32(spp), ref, frac=0.97)


@pytest.mark.parametrize("kind,bounces,ior", [(1, 9, 1.0), (2, 9, 1.0), (4, 8, 1.0), (6, 8, 1.5), ("c2", 8, 1.5)])
def test_pooled_flat_walk_finds_the_same_records(kind, bounces, ior):
    """MCPT_OPT_FLAT_POOL: after the first bounce a warp pools the exact tests of 64 rays per primitive type instead of
    walking the list with 32 rays in lock step. The candidates are ranked by the same total order (distance, leaf slot,
    offer order) through 64-bit keys, so accumulated sums, ray and segment counts are the same bit for bit -- also
    against the tree walk, which shares no code with either."""
    s, o = both(kind)
    iv, ipv = capi.default_camera(240, 135)
    out = {}
    for name, opts in (("pooled", {capi.OPT_FLAT_POOL: 1}), ("pooled, async ray copies", {capi.OPT_FLAT_POOL: 2}), ("lockstep", {capi.OPT_FLAT_POOL: 0}),
                       ("tree", {capi.OPT_FLAT_WALK: 0})):
        c = Context(0)
        try:
            for k, v in opts.items():
                c.set_option(k, v)
            c.set_option(capi.OPT_PASSES_PER_BATCH, 3)      # ragged batches: 7 passes = 3 + 3 + 1
            c.upload_scene(s)
            c.resize(240, 135)
            c.set_camera(iv, ipv)
            c.set_params(nb_bounces=bounces, refract_ind=ior, date=float(DATE))
            c.render(0, 7)
            out[name] = (c.read_accum(), c.stats())
        finally:
            c.close()
    base, st = out["pooled"]
    assert base[..., :3].mean() > 0.01 and st["rays"] > 2 * st["paths"]
    for name in ("pooled, async ray copies", "lockstep", "tree"):
        assert np.array_equal(base, out[name][0]), name
        assert out[name][1]["rays"] == st["rays"] and out[name][1]["segments"] == st["segments"], name


@pytest.mark.parametrize("kind,kw,mode,fast", [(1, {}, 0, 1), ("c2", {}, 0, 0), (5, {}, 0, 1), (8, {}, 0, 1), ("mesh", dict(n_meshes=3, tris_log2=9), 0, 1),
                                               (3, {}, 1, 1), (6, {}, 2, 0)])
def test_camera_rays_made_by_the_first_traversal(kind, kw, mode, fast):
    """MCPT_OPT_FUSED_RAYGEN: the first traversal of a batch computes (and stores) the camera ray of every path it takes
    instead of reading what k_raygen wrote -- the same arithmetic, so sums, ray and segment counts are the same bits, on
    the flat, the tree and the unified walk, in the debug modes and with libm shading (which reads the ray origin back)."""
    s, o = both(kind, **kw)
    w, h = 200, 117                      # not a multiple of the tile
    iv, ipv = capi.default_camera(w, h)
    out = {}
    for fused in (1, 0):
        c = Context(0)
        try:
            c.set_option(capi.OPT_FUSED_RAYGEN, fused)
            c.set_option(capi.OPT_FAST_MATH, fast)
            c.set_option(capi.OPT_PASSES_PER_BATCH, 3)
            c.upload_scene(s)
            c.resize(w, h)
            c.set_camera(iv, ipv)
            c.set_params(nb_bounces=6, refract_ind=1.5, date=float(DATE), mode=mode)
            c.render(0, 5)
            out[fused] = (c.read_accum(), c.stats())
        finally:
            c.close()
    assert out[1][0][..., :3].mean() > 0.005
    assert np.array_equal(out[1][0], out[0][0])
    assert out[1][1]["rays"] == out[0][1]["rays"] and out[1][1]["segments"] == out[0][1]["segments"]
    assert out[1][1]["launches"] < out[0][1]["launches"]
