"""CPU: image artefacts and convergence metrics (montecarlo_pathtracing_b200/imageio.py)."""
import numpy as np
import pytest

from montecarlo_pathtracing_b200 import imageio


def test_png_round_trip_and_row_order(tmp_path):
    rng = np.random.default_rng(1)
    a = rng.integers(0, 256, (5, 7, 3), dtype=np.uint8)
    p = imageio.write_png(tmp_path / "a.png", a, bottom_up=True)
    back = imageio.read_png(p)
    assert np.array_equal(back, a[::-1])            # file rows run top-down, the buffer bottom-up
    p = imageio.write_png(tmp_path / "b.png", np.dstack([a, a[..., :1]]), bottom_up=False)
    assert imageio.read_png(p).shape == (5, 7, 4)


def test_back_buffer_conversion_matches_the_resolve_kernel_rule():
    img = np.array([[[-1.0, 0.0, 0.5], [1.0, 2.0, np.nan], [0.002, 0.998, 1 / 255 * 0.49]]], np.float32)
    q = imageio.to_rgba8(img)
    assert q[0, 0].tolist() == [0, 0, 128, 255]      # 127.5 rounds to even -> 128
    assert q[0, 1].tolist() == [255, 255, 0, 255]    # clamp, NaN -> 0
    assert q[0, 2].tolist() == [1, 254, 0, 255]


def test_pfm_round_trip(tmp_path):
    a = np.random.default_rng(2).normal(size=(4, 6, 3)).astype(np.float32) * 10
    assert np.array_equal(imageio.read_pfm(imageio.write_pfm(tmp_path / "a.pfm", a)), a)
    with pytest.raises(ValueError):
        imageio.write_pfm(tmp_path / "b.pfm", a[..., :2])


def test_metrics():
    ref = np.full((8, 8, 3), 0.5, np.float32)
    img = ref.copy()
    img[0, 0] = 0.5 + 0.8
    m = imageio.compare(img, ref)
    assert m["rmse"] == pytest.approx(np.sqrt(3 * 0.64 / 192))
    assert m["luminance_rel_err"] == pytest.approx((3 * 0.8 / 192) / 0.5)
    assert m["non_finite_pixels"] == 0
    img[1, 1, 0] = np.inf
    assert imageio.compare(img, ref)["non_finite_pixels"] == 1
    assert imageio.rmse(ref, ref) == 0.0
    with pytest.raises(ValueError):
        imageio.rmse(ref, ref[:4])


def test_golden_converged_image_writes_a_png(tmp_path):
    from pathlib import Path
    cv = np.load(Path(__file__).parent / "golden" / "converged6.npz")
    p = imageio.write_png(tmp_path / "c6.png", cv["image"])
    back = imageio.read_png(p)
    h, w, _ = cv["image"].shape
    assert back.shape == (h, w, 3) and back.max() > 100


def test_cone_statistics_and_sampling_plot_of_the_oracle_sampler(tmp_path):
    """The DrawSampling artefact on the CPU side (tp/hsphere.vert:1-41, README.md:5-14): the cone narrows with the
    roughness; statistics, histogram and plot come out of imageio without a GPU."""
    from oracle.binding import Oracle
    D = np.float32([0.0, 0.6, 0.8])
    med = {}
    for rough in (1.0, 0.5, 0.1):
        d = Oracle().random_rays(D, rough, n=20_000)
        s = imageio.cone_stats(d, D)
        med[rough] = s["angle_p50"]
        assert s["below_plane"] == 0.0 and abs(np.linalg.norm(d, axis=1) - 1).max() < 1e-5
        edges, h = imageio.cone_histogram(d, D, bins=18)
        assert len(edges) == 19 and abs(h.sum() - 1.0) < 1e-9
        img = imageio.sampling_plot(d, D, size=128)
        assert img.shape == (128, 128, 3) and img.max() == 255
        assert imageio.read_png(imageio.write_png(tmp_path / "s.png", img, bottom_up=False)).shape[:2] == (128, 128)
    assert 35 < med[1.0] < 45 and 10 < med[0.5] < 13.5 and 0.3 < med[0.1] < 0.7
