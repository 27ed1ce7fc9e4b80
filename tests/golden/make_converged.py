#!/usr/bin/env python
"""16384-spp converged images of repo scenes 1 and 6 and of BASELINE configs 2 (scene 6 at roughness 0.5) and 5
(divergence stress: scene 5 with every material transparent + reflective, roughness 0.1, 16 bounces) from the
REFERENCE'S OWN shader text compiled as C++ (oracle/_ref/libref.so), 64x48, passes 0..16383, date = 1/60, light 0.4 —
the "16k-spp converged image" BASELINE.json measures RMSE and mean luminance against. The two configs also store the
scene tables the reference built (they have no scene*.npz fixture). Takes a few minutes on 8 cores:

    python oracle/build_ref.py && python tests/golden/make_converged.py
"""
import sys
import time
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
ROOT = HERE.parent.parent
sys.path.insert(0, str(ROOT))
from oracle.binding import Oracle  # noqa: E402

DATE = np.float32(1.0 / 60.0)
W, H, SPP = 64, 48, 16384
CASES = {1: (9, 1.0), 6: (8, 1.5),     # scene: (bounces, ior) as in tests/golden/scene*.npz
         "c2": (8, 1.5), "c5": (16, 1.5)}  # BASELINE configs 2 and 5 (bench.py WORKLOADS)
LIGHT = 0.4


def build(sid):
    r = Oracle("reference")
    if sid == "c2":                      # scene 6, roughness 0.5 on everything that is not a light
        r.build_repo_scene(6, LIGHT)
        raw = r.prim_data()
        for i in range(r.n_prims):
            if not raw[i, 58] > 0.0:
                r.poke_prim(i, 57, 0.5)
    elif sid == "c5":                    # scene 5: opacity 0.3, shininess 0.8, roughness 0.1 everywhere, one more light
        r.build_repo_scene(5, LIGHT)
        for i in range(r.n_prims):
            r.poke_prim(i, 55, 0.3)
            r.poke_prim(i, 56, 0.8)
            r.poke_prim(i, 57, 0.1)
        r.add_prim(5, r.transfo("tYs", 200, 0, 100, -110, 20, 20, 1), (0.9, 0.9, 0.9, 1.0), 0.0, 0.0, 20 * LIGHT)
    else:
        r.build_repo_scene(sid, LIGHT)
    return r.finalize()


only = sys.argv[1:]
for sid, (bounces, ior) in CASES.items():
    if only and str(sid) not in only:
        continue
    r = build(sid)
    iv, ipv = r.default_camera(W, H)
    r.set_uniforms(bounces, ior, DATE, 0, iv, ipv)
    r.set_threads(0)
    t0 = time.time()
    acc = np.zeros((H, W, 3), np.float64)
    for first in range(0, SPP, 256):          # float32 sums of 256 passes, combined in float64
        acc += r.render(W, H, first, 256)
    img = (acc / SPP).astype(np.float32)
    extra = {}
    if isinstance(sid, str):
        b = r.bvh()
        extra = dict(prims=r.prim_data(), nb_emissives=np.int32(r.nb_emissives), bvh_depth=np.int32(b["depth"]), bvh_bb=b["bb"], bvh_ind=b["ind"])
    np.savez_compressed(HERE / f"converged{sid}.npz", image=img, spp=np.int32(SPP), size=np.int32([W, H, bounces]),
                        ior=np.float32(ior), inv_v=iv, inv_pv=ipv, **extra)
    print(sid, "mean", img.mean(), "time", round(time.time() - t0, 1), "s", flush=True)
