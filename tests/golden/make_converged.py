#!/usr/bin/env python
"""16384-spp converged images of repo scenes 1 and 6 from the REFERENCE'S OWN shader text compiled as
C++ (oracle/_ref/libref.so), 64x48, passes 0..16383, date = 1/60, light 0.4 — the "16k-spp converged
image" BASELINE.json measures RMSE and mean luminance against. Takes a few minutes on 8 cores:

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
CASES = {1: (9, 1.0), 6: (8, 1.5)}     # scene: (bounces, ior) as in tests/golden/scene*.npz

for sid, (bounces, ior) in CASES.items():
    r = Oracle("reference").build_repo_scene(sid, 0.4).finalize()
    iv, ipv = r.default_camera(W, H)
    r.set_uniforms(bounces, ior, DATE, 0, iv, ipv)
    r.set_threads(0)
    t0 = time.time()
    acc = np.zeros((H, W, 3), np.float64)
    for first in range(0, SPP, 256):          # float32 sums of 256 passes, combined in float64
        acc += r.render(W, H, first, 256)
    img = (acc / SPP).astype(np.float32)
    np.savez_compressed(HERE / f"converged{sid}.npz", image=img, spp=np.int32(SPP), size=np.int32([W, H, bounces]),
                        ior=np.float32(ior), inv_v=iv, inv_pv=ipv)
    print(sid, "mean", img.mean(), "time", round(time.time() - t0, 1), "s", flush=True)
