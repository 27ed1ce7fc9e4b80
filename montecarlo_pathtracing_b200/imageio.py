"""Image artefacts and convergence metrics.

The reference's only artefacts are window grabs (captures/*.png, README.md:5-46): the back buffer
after fs_frag (MontecarloGPU/montecarlo.cpp:59-70) -- accum / nb, clamped to [0,1], 8 bit, no gamma,
row 0 at the bottom. This module writes the same picture headless:

  write_png   8-bit RGB(A), rows flipped to top-down file order (pure Python + zlib, no dependency)
  write_pfm / read_pfm   the unclamped float image (PFM stores rows bottom-up, like the accumulation buffer)
  rmse, mean_luminance, compare   the figures BASELINE.json's parity bar is stated in
                                  (RMSE against a converged image, mean luminance within 0.5 %)

Nothing here touches the traced path; images come from Context.resolve().
"""
from __future__ import annotations

import struct
import zlib
from pathlib import Path

import numpy as np


def to_rgba8(img: np.ndarray) -> np.ndarray:
    """Back-buffer conversion of a float image (h, w, 3): clamp to [0,1], round to nearest, alpha 255;
    the arithmetic of k_resolve's RGBA8 branch (NaN -> 0)."""
    v = np.nan_to_num(np.asarray(img, np.float32), nan=0.0, posinf=1.0, neginf=0.0)
    v = np.clip(v, 0.0, 1.0)
    q = np.rint(v * np.float32(255.0)).astype(np.uint8)
    out = np.full(q.shape[:2] + (4,), 255, np.uint8)
    out[..., :3] = q[..., :3]
    return out


def _chunk(tag: bytes, data: bytes) -> bytes:
    return struct.pack(">I", len(data)) + tag + data + struct.pack(">I", zlib.crc32(tag + data) & 0xFFFFFFFF)


def write_png(path, img: np.ndarray, bottom_up: bool = True, level: int = 6) -> Path:
    """img: uint8 (h, w, 3|4), or float (h, w, 3) which is converted like the back buffer. bottom_up:
    row 0 of img is the bottom row (gl_FragCoord convention of the accumulation buffer)."""
    a = np.asarray(img)
    if a.dtype != np.uint8:
        a = to_rgba8(a)[..., :3]
    if a.ndim != 3 or a.shape[2] not in (3, 4):
        raise ValueError(f"expected (h, w, 3|4), got {a.shape}")
    if bottom_up:
        a = a[::-1]
    h, w, c = a.shape
    raw = np.empty((h, 1 + w * c), np.uint8)
    raw[:, 0] = 0                       # filter type 0 on every scanline
    raw[:, 1:] = a.reshape(h, w * c)
    ihdr = struct.pack(">IIBBBBB", w, h, 8, 2 if c == 3 else 6, 0, 0, 0)
    data = b"\x89PNG\r\n\x1a\n" + _chunk(b"IHDR", ihdr) + _chunk(b"IDAT", zlib.compress(raw.tobytes(), level)) + _chunk(b"IEND", b"")
    path = Path(path)
    path.write_bytes(data)
    return path


def read_png(path) -> np.ndarray:
    """Decoder for the files write_png produces (8-bit RGB/RGBA, filter 0, no interlace): tests only."""
    b = Path(path).read_bytes()
    if b[:8] != b"\x89PNG\r\n\x1a\n":
        raise ValueError("not a PNG file")
    pos, idat, w = 8, b"", None
    while pos < len(b):
        n, tag = struct.unpack(">I4s", b[pos:pos + 8])
        body = b[pos + 8:pos + 8 + n]
        (crc,) = struct.unpack(">I", b[pos + 8 + n:pos + 12 + n])
        if crc != (zlib.crc32(tag + body) & 0xFFFFFFFF):
            raise ValueError(f"bad CRC in chunk {tag!r}")
        if tag == b"IHDR":
            w, h, depth, ctype, _, _, interlace = struct.unpack(">IIBBBBB", body)
            if depth != 8 or ctype not in (2, 6) or interlace:
                raise ValueError("unsupported PNG flavour")
            c = 3 if ctype == 2 else 4
        elif tag == b"IDAT":
            idat += body
        pos += 12 + n
    raw = np.frombuffer(zlib.decompress(idat), np.uint8).reshape(h, 1 + w * c)
    if raw[:, 0].any():
        raise ValueError("only filter type 0 is supported")
    return raw[:, 1:].reshape(h, w, c).copy()


def write_pfm(path, img: np.ndarray) -> Path:
    """Unclamped float RGB; PFM rows run bottom-up, which is the accumulation buffer's own order."""
    a = np.ascontiguousarray(img, dtype="<f4")
    if a.ndim != 3 or a.shape[2] != 3:
        raise ValueError(f"expected (h, w, 3), got {a.shape}")
    h, w, _ = a.shape
    path = Path(path)
    with open(path, "wb") as f:
        f.write(f"PF\n{w} {h}\n-1.0\n".encode())
        f.write(a.tobytes())
    return path


def read_pfm(path) -> np.ndarray:
    with open(path, "rb") as f:
        if f.readline().strip() != b"PF":
            raise ValueError("not a colour PFM file")
        w, h = (int(x) for x in f.readline().split())
        scale = float(f.readline())
        a = np.frombuffer(f.read(w * h * 12), "<f4" if scale < 0 else ">f4").reshape(h, w, 3)
    return a.astype(np.float32)


def mean_luminance(img: np.ndarray) -> float:
    """The figure the parity bar uses (tests/test_gpu_golden.py): the plain mean over pixels and channels."""
    return float(np.asarray(img, np.float64).mean())


def rmse(img: np.ndarray, ref: np.ndarray) -> float:
    a, b = np.asarray(img, np.float64), np.asarray(ref, np.float64)
    if a.shape != b.shape:
        raise ValueError(f"shape mismatch {a.shape} vs {b.shape}")
    return float(np.sqrt(np.mean((a - b) ** 2)))


def compare(img: np.ndarray, ref: np.ndarray) -> dict:
    """RMSE, relative mean-luminance error, PSNR (peak 1.0) and the number of non-finite pixels."""
    a = np.asarray(img, np.float32)
    bad = int((~np.isfinite(a)).any(axis=-1).sum())
    e = rmse(np.nan_to_num(a), ref)
    lr = mean_luminance(ref)
    return {"rmse": e, "mean_luminance": mean_luminance(np.nan_to_num(a)), "ref_mean_luminance": lr,
            "luminance_rel_err": abs(mean_luminance(np.nan_to_num(a)) - lr) / abs(lr) if lr else float("inf"),
            "psnr_db": float(20.0 * np.log10(1.0 / e)) if e > 0 else float("inf"), "non_finite_pixels": bad}


def convergence_curve(ctx, ref: np.ndarray, spp_marks, first_pass: int = 0) -> list:
    """Progressive accumulation as the viewer does it when locked (montecarlo.cpp:420-467): render up to
    each mark of spp_marks (ascending), resolve, compare with ref. ctx: a Context with scene, size,
    camera and params set; its accumulation buffer is cleared first."""
    ctx.clear()
    done, out = 0, []
    for m in spp_marks:
        if m <= done:
            raise ValueError("spp_marks must be ascending")
        ctx.render(first_pass + done, m - done)
        done = m
        row = compare(ctx.resolve(float(done)), ref)
        row["spp"] = done
        out.append(row)
    return out


# ---- the sampling cone (the reference's DrawSampling artefact, README.md:5-14, tp/hsphere.vert:1-41)
def cone_stats(dirs: np.ndarray, D) -> dict:
    """Statistics of sampled directions around the axis D: the angle to D (degrees) at the median / 90 % / 99 % / max,
    the mean cosine, and the share of samples below the tangent plane (none for a hemisphere sampler)."""
    d = np.asarray(dirs, np.float64).reshape(-1, 3)
    a = np.asarray(D, np.float64).reshape(3)
    a = a / np.linalg.norm(a)
    cosv = np.clip(d @ a / np.maximum(np.linalg.norm(d, axis=1), 1e-30), -1.0, 1.0)
    ang = np.degrees(np.arccos(cosv))
    return {"n": int(len(d)), "mean_cos": float(cosv.mean()), "angle_p50": float(np.percentile(ang, 50)),
            "angle_p90": float(np.percentile(ang, 90)), "angle_p99": float(np.percentile(ang, 99)), "angle_max": float(ang.max()),
            "below_plane": float((cosv < 0).mean())}


def cone_histogram(dirs: np.ndarray, D, bins: int = 18):
    """Histogram of the angle to D over [0, 90] degrees: (bin edges, share of the samples per bin)."""
    d = np.asarray(dirs, np.float64).reshape(-1, 3)
    a = np.asarray(D, np.float64).reshape(3)
    a = a / np.linalg.norm(a)
    ang = np.degrees(np.arccos(np.clip(d @ a / np.maximum(np.linalg.norm(d, axis=1), 1e-30), -1.0, 1.0)))
    h, edges = np.histogram(ang, bins=bins, range=(0.0, 90.0))
    return edges, h / max(len(d), 1)


def sampling_plot(dirs: np.ndarray, D, size: int = 256) -> np.ndarray:
    """The DrawSampling picture, headless: the sampled unit directions seen from a point on the axis D, i.e. projected
    on the plane perpendicular to D (the unit circle is the horizon of the hemisphere). Returns a (size, size, 3) uint8
    image -- log-scaled point density in white on black, the horizon as a grey circle -- for write_png(bottom_up=False)."""
    d = np.asarray(dirs, np.float64).reshape(-1, 3)
    a = np.asarray(D, np.float64).reshape(3)
    a = a / np.linalg.norm(a)
    w = np.array([a[0], a[1] + 5.0, a[2] + 3.0])          # the frame random_ray builds (tp/montecarlo.frag:72-89)
    w /= np.linalg.norm(w)
    u = np.cross(a, w)
    u /= np.linalg.norm(u)
    v = np.cross(a, u)
    x, y = d @ u, d @ v
    ix = np.clip(((x * 0.48 + 0.5) * size).astype(int), 0, size - 1)
    iy = np.clip(((y * 0.48 + 0.5) * size).astype(int), 0, size - 1)
    dens = np.zeros((size, size), np.float64)
    np.add.at(dens, (iy, ix), 1.0)
    img = np.log1p(dens) / max(np.log1p(dens.max()), 1e-9)
    out = np.repeat((img * 255).astype(np.uint8)[..., None], 3, axis=2)
    t = np.linspace(0, 2 * np.pi, 4 * size)
    cx = np.clip(((np.cos(t) * 0.48 + 0.5) * size).astype(int), 0, size - 1)
    cy = np.clip(((np.sin(t) * 0.48 + 0.5) * size).astype(int), 0, size - 1)
    out[cy, cx] = np.maximum(out[cy, cx], 96)
    return out
