"""Shared helpers for the tests: scene construction on both sides, ray batches."""
from __future__ import annotations

import numpy as np

from montecarlo_pathtracing_b200 import Scene, capi
from oracle.binding import Oracle

DATE = np.float32(1.0 / 60.0)


def host_scene(kind, light=0.4, **kw) -> Scene:
    """kind: 1..8 repo scene, 'c2' (scene 6, roughness 0.5), 'c5' (divergence stress),
    'mesh' (synthetic meshes: n_meshes, tris_log2)."""
    s = Scene()
    if kind == "c2":
        s.build_scene6_roughness(light, kw.get("roughness", 0.5))
    elif kind == "c5":
        s.build_divergence_stress(light)
    elif kind == "mesh":
        s.build_synthetic_meshes(kw.get("n_meshes", 4), kw.get("tris_log2", 9), kw.get("seed", 12345))
    else:
        s.build_repo_scene(int(kind), light)
    return s


def both(kind, light=0.4, **kw):
    """Build the same scene in the host mirror and in the oracle; returns (scene, oracle),
    both finalized."""
    s = host_scene(kind, light, **kw)
    raw = s.prim_data()           # insertion order, before the emissive-first sort
    s.finalize()
    meshes = s.tables()["meshes"]
    o = Oracle().load_unsorted(raw, meshes).finalize()
    return s, o


def scene_bounds(prims64):
    """Loose world bounds of a PrimData table: transformed unit-cube corners."""
    lo, hi = np.full(3, np.inf), np.full(3, -np.inf)
    corners = np.array([[x, y, z, 1.0] for x in (-1, 1) for y in (-1, 1) for z in (-1, 1)], np.float64)
    for r in prims64:
        M = r[0:16].reshape(4, 4).T.astype(np.float64)
        w = (M @ corners.T).T[:, :3]
        lo, hi = np.minimum(lo, w.min(0)), np.maximum(hi, w.max(0))
    return lo, hi


def random_rays(prims64, n, seed=0, clip=400.0):
    """Origins uniform in the (clipped) scene box, directions uniform on the sphere, plus a
    share of axis-aligned and near-degenerate directions."""
    rng = np.random.default_rng(seed)
    lo, hi = scene_bounds(prims64)
    lo, hi = np.maximum(lo, -clip), np.minimum(hi, clip)
    o = rng.uniform(lo - 5, hi + 5, size=(n, 3)).astype(np.float32)
    d = rng.normal(size=(n, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    d = d.astype(np.float32)
    k = max(n // 50, 1)
    axes = np.eye(3, dtype=np.float32)
    d[:k] = axes[rng.integers(0, 3, k)] * rng.choice([-1.0, 1.0], (k, 1)).astype(np.float32)
    d[k:2 * k] *= np.float32(3.7)   # non-unit directions are legal
    return o, d


def camera_rays(oracle: Oracle, w, h, step=1):
    xs, ys = np.arange(0, w, step), np.arange(0, h, step)
    o = np.zeros((len(ys) * len(xs), 3), np.float32)
    d = np.zeros_like(o)
    i = 0
    for y in ys:
        for x in xs:
            o[i], d[i] = oracle.primary_ray(int(x), int(y), w, h)
            i += 1
    return o, d


def compare_hits(a, b, rtol=1e-5):
    """Return indices where two hit arrays disagree (ids exact, dist within rtol)."""
    bad = (a["index"] != b["index"]) | (a["tri_index"] != b["tri_index"]) | (a["shape"] != b["shape"]) | (a["dir"] != b["dir"])
    hit = a["index"] >= 0
    with np.errstate(invalid="ignore", over="ignore"):
        rel = np.abs(a["dist"] - b["dist"]) > rtol * np.abs(b["dist"])
    bad |= hit & rel
    return np.nonzero(bad)[0]


def desc_from_arrays(prims, root, meshes=()):
    """An include/mcpt_scene.h descriptor straight from numpy tables (golden fixtures): prims (n,64)
    float32 finalized records, root = {'depth','bb','ind'}, meshes = [{'bvh','tri','pos','nrm','mesh_line'}].
    Returns (desc, keepalive)."""
    import ctypes as C
    from montecarlo_pathtracing_b200.capi import BvhDesc, MeshDesc, SceneDesc
    keep = []

    def arr(a, dt):
        a = np.ascontiguousarray(a, dtype=dt)
        keep.append(a)
        return a.ctypes.data

    def bvh(b):
        return BvhDesc(arr(b["bb"], np.float32), arr(b["ind"], np.int32), int(b["depth"]))

    marr = (MeshDesc * max(len(meshes), 1))()
    for i, m in enumerate(meshes):
        marr[i] = MeshDesc(bvh(m["bvh"]), arr(m["tri"], np.uint32), arr(m["pos"], np.float32), arr(m["nrm"], np.float32),
                           len(m["tri"]), len(m["pos"]), int(m["mesh_line"]))
    keep.append(marr)
    d = SceneDesc(arr(prims, np.float32), len(prims), 0, bvh(root), C.cast(marr, C.POINTER(MeshDesc)), len(meshes))
    return d, keep
