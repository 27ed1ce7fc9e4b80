#!/usr/bin/env python
"""Generates tests/golden/*.npz from the REFERENCE'S OWN code (oracle/_ref/libref.so: bvh_gpu/*.cpp,
easycppogl/{gl_eigen,camera}.cpp compiled unmodified, the scene builders of montecarlo.cpp and the
shader text of shaders/ + tp/ compiled as C++; see oracle/build_ref.py). Run in the container that
has /root/reference:

    python oracle/build_ref.py && python tests/golden/make_golden.py

The fixtures are what pins the hand-written oracle (tests/test_oracle_golden.py, CPU) and the CUDA
path (tests/test_gpu_golden.py) to the reference: inputs AND outputs are stored, so the tests need
neither /root/reference nor libref.so. `date` is pinned to 1/60 (the reference feeds the wall clock).
"""
import sys
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
ROOT = HERE.parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))

from oracle.binding import Oracle  # noqa: E402

DATE = np.float32(1.0 / 60.0)
LIGHT = 0.4   # captures/N-04-*.png and BASELINE configs use light intensity 0.4


def rays_for(prims64, n, seed):
    from common import random_rays
    return random_rays(prims64, n, seed=seed)


def scene_fixture(sid):
    r = Oracle("reference").build_repo_scene(sid, LIGHT)
    raw = r.prim_data()                       # insertion order (before sortEmissiveFirst)
    r.finalize()
    out = {"raw_prims": raw, "prims": r.prim_data(), "nb_emissives": np.int32(r.nb_emissives)}
    b = r.bvh()
    out.update(bvh_depth=np.int32(b["depth"]), bvh_bb=b["bb"], bvh_ind=b["ind"])
    ro, rd = rays_for(r.prim_data(), 4000, seed=100 + sid)
    out.update(ray_o=ro, ray_d=rd, hits=r.trace(ro, rd), any_hit=(r.trace(ro, rd, True)["index"] >= 0))
    w, h, spp = 64, 48, 16
    bounces, ior = (9, 1.0) if sid != 6 else (8, 1.5)
    iv, ipv = r.default_camera(w, h)
    r.set_uniforms(bounces, ior, DATE, 0, iv, ipv)
    out.update(inv_v=iv, inv_pv=ipv, size=np.int32([w, h, spp, bounces]), ior=np.float32(ior),
               image=r.render(w, h, 0, spp))            # accumulated sum of passes 0..15
    if sid == 2:
        out.update(image_mode1=r.render(w, h, 0, 2, 1), image_mode2=r.render(w, h, 0, 2, 2))
    if sid == 6:                                        # second IOR: captures/6-04-2.png
        r.set_uniforms(bounces, 2.0, DATE, 0, iv, ipv)
        out.update(image_ior2=r.render(w, h, 0, 8))
    prim_rays = np.stack([np.concatenate(r.primary_ray(x, y, w, h)) for y in (0, 17, 47) for x in (0, 31, 63)])
    out.update(primary=prim_rays)
    return out


def mesh_fixture():
    """Two small meshes + a floor quad + a light: inputs come from a fixed-seed numpy generator
    (not from the product), the tables and outputs from the reference code."""
    rng = np.random.default_rng(7)
    meshes = []
    for m in range(2):
        n = 9
        g = np.stack(np.meshgrid(np.linspace(-1, 1, n), np.linspace(-1, 1, n), indexing="ij"), -1).reshape(-1, 2)
        z = 0.3 * np.sin(3 * g[:, 0] + m) * np.cos(2 * g[:, 1]) + 0.05 * rng.standard_normal(len(g))
        pos = np.concatenate([g, z[:, None]], 1).astype(np.float32)
        tri = []
        for i in range(n - 1):
            for j in range(n - 1):
                a = i * n + j
                tri += [[a, a + n, a + 1], [a + 1, a + n, a + n + 1]]
        tri = np.asarray(tri, np.uint32)
        nrm = np.zeros_like(pos)
        for t in tri:
            fn = np.cross(pos[t[1]] - pos[t[0]], pos[t[2]] - pos[t[0]])
            nrm[t] += fn
        nrm = (nrm / np.linalg.norm(nrm, axis=1, keepdims=True)).astype(np.float32)
        meshes.append({"pos": pos, "tri": tri, "nrm": nrm})
    r = Oracle("reference")
    ids = [r.add_mesh(m["pos"], m["tri"], m["nrm"]) for m in meshes]
    T = r.transfo
    place = [(ids[0], T("tZs", -45, 10, -20, 30, 50, 50, 40), [0.9, 0.2, 0.2, 1.0], 0.0, 0.5, 0.0),
             (ids[1], T("tXs", 50, 0, 0, 25, 45, 45, 45), [0.2, 0.9, 0.9, 0.4], 0.6, 0.3, 0.0)]
    for mid, trf, col, sh, ro_, em in place:
        r.place_mesh(mid, trf, col, sh, ro_, em)
    r.add_prim(5, T("ts", 0, 0, -60, 300, 300, 1), [0.9, 0.9, 0.9, 1.0])
    r.add_prim(5, T("tXs", 0, 0, 120, 180, 60, 60, 1), [0.9, 0.9, 0.9, 1.0], 0.0, 0.0, 4.0)
    r.add_prim(1, T("tu", 0, -70, -20, 25), [0.9, 0.9, 0.2, 0.3], 0.7, 0.2)
    raw = r.prim_data()
    r.finalize()
    out = {"raw_prims": raw, "prims": r.prim_data(), "nb_emissives": np.int32(r.nb_emissives)}
    b = r.bvh()
    out.update(bvh_depth=np.int32(b["depth"]), bvh_bb=b["bb"], bvh_ind=b["ind"])
    for m, md in enumerate(meshes):
        mb = r.bvh(m)
        out.update({f"m{m}_pos": md["pos"], f"m{m}_tri": md["tri"], f"m{m}_nrm": md["nrm"],
                    f"m{m}_line": np.int32(r.L.oracle_mesh_line(r.h, m)), f"m{m}_depth": np.int32(mb["depth"]),
                    f"m{m}_bb": mb["bb"], f"m{m}_ind": mb["ind"]})
    ro, rd = rays_for(r.prim_data(), 6000, seed=77)
    # two thirds of the batch is aimed at the meshes (the floor quad inflates the scene box)
    centres = np.float32([[-45, 10, -20], [50, 0, 0]])
    tgt = centres[rng.integers(0, 2, 4000)] + rng.uniform(-35, 35, (4000, 3)).astype(np.float32)
    ro[:4000] = np.clip(ro[:4000], -200, 200)
    rd[:4000] = tgt - ro[:4000]
    out.update(ray_o=ro, ray_d=rd, hits=r.trace(ro, rd))
    w, h, spp = 64, 48, 8
    iv, ipv = r.default_camera(w, h)
    r.set_uniforms(8, 1.5, DATE, 0, iv, ipv)
    out.update(inv_v=iv, inv_pv=ipv, size=np.int32([w, h, spp, 8]), ior=np.float32(1.5), image=r.render(w, h, 0, spp))
    r.set_uniforms(8, 1.5, DATE, 1, iv, ipv)
    out.update(image_flat=r.render(w, h, 0, 4))
    return out


def unit_fixture():
    r = Oracle("reference")
    out = {}
    cases = [(0.3, 0.7, 0, 0.0), (0.0078125, 0.99, 5, float(DATE)), (0.5, 0.5, 1023, 0.25)]
    out["rng_cases"] = np.float32(cases)
    for k, (tx, ty, p, d) in enumerate(cases):
        s, seed = r.rng_stream(tx, ty, p, d, 256)
        out[f"rng_{k}"], out[f"rng_seed_{k}"] = s, seed
    for k, (D, rough) in enumerate([([0, 0, 1], 1.0), ([0, 0, 1], 0.5), ([0.6, 0.0, 0.8], 0.1), ([0, -1, 0], 0.0)]):
        out[f"sampler_D_{k}"], out[f"sampler_rough_{k}"] = np.float32(D), np.float32(rough)
        out[f"sampler_{k}"] = r.random_rays(D, rough, n=512)
    for k, (w, h) in enumerate([(64, 48), (512, 512), (1920, 1080), (1280, 1000)]):
        iv, ipv = r.default_camera(w, h)
        out[f"cam_size_{k}"], out[f"cam_inv_v_{k}"], out[f"cam_inv_pv_{k}"] = np.int32([w, h]), iv, ipv
    out["transfo_0"] = r.transfo("tZs", 70, 20, -40, 20, 20, 20, 60)
    out["transfo_1"] = r.transfo("tYs", 200, 0, 100, -110, 20, 20, 1)
    out["transfo_2"] = r.transfo("tXu", 1, 2, 3, 45, 2.5)
    return out


def main():
    from oracle.binding import have_reference
    if not have_reference():
        raise SystemExit("oracle/_ref/libref.so is missing: run python oracle/build_ref.py (needs /root/reference)")
    for sid in range(1, 9):
        np.savez_compressed(HERE / f"scene{sid}.npz", **scene_fixture(sid))
    np.savez_compressed(HERE / "mesh.npz", **mesh_fixture())
    np.savez_compressed(HERE / "units.npz", **unit_fixture())
    for f in sorted(HERE.glob("*.npz")):
        print(f.name, f.stat().st_size)


if __name__ == "__main__":
    main()
