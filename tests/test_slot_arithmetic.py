"""The slot -> (pass, pixel) arithmetic of fetch_ray() (csrc/kernels.cu): the pass index of path slot p is ESTIMATED in
fp32, k = trunc(float_rz(p) * (1 / npix)), and corrected by one either way; the pixel is what remains. This is a numpy
restatement of those four lines (same roundings: uint -> float toward zero, one rounded multiply, truncation), checked
against integer division on the sizes the library accepts -- the claim that one correction step is always enough."""
import numpy as np
import pytest


def float_rz(p):
    f = p.astype(np.float32)                       # round to nearest
    over = f.astype(np.float64) > p.astype(np.float64)
    return np.where(over, np.nextafter(f, np.float32(0)), f).astype(np.float32)


def slot_to_pixel(p, npix):
    inv = np.float32(1.0) / np.float32(npix)       # host: 1.0f / float(npix)
    k = np.trunc(float_rz(p) * inv).astype(np.uint32)
    pix = (p - k * np.uint32(npix)).astype(np.uint32)          # modulo 2^32, as on the device
    neg = pix.view(np.int32) < 0
    big = ~neg & (pix >= np.uint32(npix))
    return np.where(neg, pix + np.uint32(npix), np.where(big, pix - np.uint32(npix), pix)).astype(np.uint32)


@pytest.mark.parametrize("npix", [1, 2, 3, 200 * 117, 512 * 512, 1280 * 1000, 1920 * 1080, 3840 * 2160, (1 << 24) + 1, 7919 * 7907, (1 << 28) - 1, 1 << 28])
def test_one_correction_step_is_enough(npix):
    max_passes = min(64, ((1 << 32) - 1) // npix)              # slots = npix * passes must fit 32 bits
    rng = np.random.default_rng(npix & 0xffff)
    m = np.arange(0, max_passes + 1, dtype=np.uint64) * np.uint64(npix)
    edge = np.concatenate([m, m + np.uint64(1), m[1:] - np.uint64(1), m[1:] - np.uint64(2)])
    rnd = rng.integers(0, max_passes * npix, 200_000, dtype=np.uint64)
    p64 = np.concatenate([edge, rnd])
    p64 = p64[p64 < np.uint64(max_passes) * np.uint64(npix)]
    p = p64.astype(np.uint32)
    with np.errstate(over="ignore"):
        got = slot_to_pixel(p, npix)
    assert np.array_equal(got.astype(np.uint64), p64 % np.uint64(npix))
