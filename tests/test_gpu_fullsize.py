"""GPU, at the FULL sizes of BASELINE.json's configs, through properties that do not need a full-size oracle
run: bit-exact additivity of pass ranges and batch sizes, determinism, bands of rows against the oracle at the
full resolution (pixel indexing, seeds and camera rays of a 1920x1080 / 3840x2160 frame), equality of the
walkers (persistent / one ray per lane, both mesh trees) on the 2^20-triangle scene, and conservation of the
work counters. The small-size parity tests are in test_gpu_render.py / test_gpu_traversal.py / test_gpu_golden.py."""
import numpy as np
import pytest

from common import DATE, both, compare_hits, random_rays
from montecarlo_pathtracing_b200 import Context, capi

pytestmark = pytest.mark.gpu

W, H = 1920, 1080


def _setup(c, scene, w, h, bounces, ior):
    iv, ipv = capi.default_camera(w, h)
    c.upload_scene(scene)
    c.resize(w, h)
    c.set_camera(iv, ipv)
    c.set_params(nb_bounces=bounces, refract_ind=ior, date=float(DATE))
    return iv, ipv


def _band_parity(c, o, w, h, iv, ipv, bounces, ior, rows, spp, frac):
    """`rows` full-width rows in the middle of the frame, passes 0..spp-1, against the oracle."""
    o.set_uniforms(bounces, ior, DATE, 0, iv, ipv)
    y0 = (h - rows) // 2
    ref = np.zeros((h, w, 3), np.float32)
    o.render(w, h, 0, spp, 0, ref, y0, y0 + rows)
    c.clear()
    c.render(0, spp)
    got = c.read_accum()[y0:y0 + rows, :, :3]
    ref = ref[y0:y0 + rows]
    ok = np.isclose(got, ref, rtol=2e-3, atol=2e-3).all(axis=2)
    assert ok.mean() >= frac, f"only {ok.mean():.4f} of the band's pixels agree"
    assert abs(got.mean() - ref.mean()) <= 5e-3 * ref.mean()           # mean luminance within 0.5 %


def test_config2_full_size_additivity_determinism_and_band_parity():
    """BASELINE config 2: repo scene 6, 1920x1080, 64 spp, IOR 1.5, roughness 0.5, 8 bounces."""
    s, o = both("c2")
    c = Context(0)
    try:
        c.set_option(capi.OPT_FAST_MATH, 1)                             # what bench.py measures
        iv, ipv = _setup(c, s, W, H, 8, 1.5)
        c.render(0, 64)
        a = c.read_accum()
        st = c.stats()
        assert st["paths"] == W * H * 64 and st["rays"] > st["paths"] and st["segments"] > 0
        assert np.isfinite(a).all() and a[..., :3].min() >= 0.0
        c.clear()
        c.render(0, 32)                                                 # two calls = one call, bit for bit
        c.render(32, 32)
        assert np.array_equal(a, c.read_accum())
        c.clear()
        c.set_option(capi.OPT_PASSES_PER_BATCH, 5)                      # ragged batches (13 batches, the last of 4)
        c.render(0, 64)
        assert np.array_equal(a, c.read_accum())
        c.set_option(capi.OPT_PASSES_PER_BATCH, 0)
        # size-independent: the mean radiance of the frame does not depend on the resolution (same aspect)
        img = a[..., :3] / 64.0
        c.resize(480, 270)
        iv2, ipv2 = capi.default_camera(480, 270)
        c.set_camera(iv2, ipv2)
        c.render(0, 256)
        small = c.read_accum()[..., :3] / 256.0
        assert abs(img.mean() - small.mean()) <= 1e-2 * small.mean()
        # a band of the full-size frame against the oracle: pixel / seed / camera-ray indexing at 1920x1080
        c.resize(W, H)
        c.set_camera(iv, ipv)
        _band_parity(c, o, W, H, iv, ipv, 8, 1.5, rows=6, spp=3, frac=0.985)
    finally:
        c.close()


def test_config5_full_size_walkers_agree():
    """BASELINE config 5 (all transparent+reflective, 16 bounces, 1080p): the persistent walk and the
    one-ray-per-lane walk give the same bits, ray and segment counts; a band agrees with the oracle."""
    s, o = both("c5")
    out = []
    for persistent in (1, 0):
        c = Context(0)
        try:
            c.set_option(capi.OPT_PERSISTENT, persistent)
            c.set_option(capi.OPT_FAST_MATH, 1)
            iv, ipv = _setup(c, s, W, H, 16, 1.5)
            c.render(0, 2)
            out.append((c.read_accum(), c.stats()))
            if persistent:
                _band_parity(c, o, W, H, iv, ipv, 16, 1.5, rows=4, spp=2, frac=0.97)
        finally:
            c.close()
    assert np.array_equal(out[0][0], out[1][0])
    assert out[0][1]["rays"] == out[1][1]["rays"] and out[0][1]["segments"] == out[1][1]["segments"]
    assert out[0][1]["rays"] > 2 * W * H * 2                           # interior rays: more than one ray per bounce


@pytest.fixture(scope="module")
def million():
    """BASELINE configs 3 / 4: 16 meshes x 65 536 triangles (2^20), floor and light."""
    return both("mesh", n_meshes=16, tris_log2=16)


def test_config3_million_triangles_traversal(million):
    """Hit records on the 2^20-triangle scene, 2^20 random rays through the PRODUCTION kernel (mcpt_trace_rays
    runs k_traverse itself): the four-children persistent walk over the GPU-built trees (what bench.py times),
    the one-ray-per-lane walk over the two-children trees, the walk over the uploaded median-split trees and
    k_trace return the same records bit for bit, and equal the oracle on a subset."""
    s, o = million
    ro, rd = random_rays(s.prim_data(), 1 << 20, seed=7, clip=150.0)
    hits = {}
    for name, (persistent, mesh_bvh, trace_walker) in {"wide persistent": (1, 1, 0), "one ray per lane": (0, 1, 0),
                                                       "reference trees": (0, 0, 0), "k_trace": (1, 1, 1)}.items():
        c = Context(0)
        try:
            c.set_option(capi.OPT_PERSISTENT, persistent)
            c.set_option(capi.OPT_MESH_BVH, mesh_bvh)
            c.set_option(capi.OPT_TRACE_WALKER, trace_walker)
            c.upload_scene(s)
            hits[name] = c.trace_rays(ro, rd)
        finally:
            c.close()
    base = hits["wide persistent"]
    for name, h in hits.items():
        assert h.tobytes() == base.tobytes(), name
    assert (base["tri_index"] >= 0).sum() > 10000
    n = 20000
    ref = o.trace(ro[:n], rd[:n])
    assert len(compare_hits(base[:n], ref)) == 0
    assert np.array_equal(base["dist"][:n], ref["dist"])               # distances bit-exact, not just 1e-5
    assert np.array_equal(base["pl"][:n].view(np.uint32), ref["pl"].view(np.uint32))


def test_config3_full_size_band_parity(million):
    """BASELINE config 3 at its full resolution: a 4-row band of the 1920x1080 frame of the 16 x 65 536-triangle
    scene against the oracle (as for configs 2 and 5): pixel / seed / camera-ray indexing and the default
    (four-children persistent) walker on the real workload."""
    s, o = million
    c = Context(0)
    try:
        c.set_option(capi.OPT_FAST_MATH, 1)                             # what bench.py measures
        iv, ipv = _setup(c, s, W, H, 8, 1.5)
        _band_parity(c, o, W, H, iv, ipv, 8, 1.5, rows=4, spp=2, frac=0.975)
    finally:
        c.close()


def test_config3_config4_full_size_render_walkers_agree(million):
    """1080p (config 3) and 3840x2160 (config 4), 8 bounces: four-children persistent walk = one-ray-per-lane
    walk over the two-children trees = walk over the reference's own trees, bit for bit; pass ranges add up."""
    s, o = million
    variants = {"wide persistent": (1, 1), "one ray per lane": (0, 1), "reference trees": (0, 0)}
    for (w, h, spp) in ((W, H, 2), (3840, 2160, 1)):
        acc = {}
        for name, (persistent, mesh_bvh) in variants.items():
            c = Context(0)
            try:
                c.set_option(capi.OPT_PERSISTENT, persistent)
                c.set_option(capi.OPT_MESH_BVH, mesh_bvh)
                c.set_option(capi.OPT_FAST_MATH, 1)
                _setup(c, s, w, h, 8, 1.5)
                c.render(0, spp)
                acc[name] = (c.read_accum(), c.stats())
                if name == "wide persistent" and spp > 1:
                    c.clear()
                    c.render(0, 1)
                    c.render(1, spp - 1)
                    assert np.array_equal(acc[name][0], c.read_accum())
            finally:
                c.close()
        base, st = acc["wide persistent"]
        assert np.isfinite(base).all() and base[..., :3].mean() > 0.01
        for name in ("one ray per lane", "reference trees"):
            assert np.array_equal(base, acc[name][0]), (w, h, name)
            assert acc[name][1]["rays"] == st["rays"] and acc[name][1]["segments"] == st["segments"]
