#!/usr/bin/env python
"""Work counters and step time of the mesh workload for both mesh trees (diagnostic, GPU)."""
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))
import numpy as np  # noqa: E402
from common import DATE, host_scene  # noqa: E402
from montecarlo_pathtracing_b200 import Context, capi  # noqa: E402

n_meshes, tl = int(sys.argv[1]) if len(sys.argv) > 1 else 16, int(sys.argv[2]) if len(sys.argv) > 2 else 16
s = host_scene("mesh", 0.4, n_meshes=n_meshes, tris_log2=tl)
s.finalize()
W, H, SPP = 1920, 1080, 4
iv, ipv = capi.default_camera(W, H)
which = [int(sys.argv[3])] if len(sys.argv) > 3 else [0, 1]
for mesh_bvh in which:
    ctx = Context(0)
    ctx.set_option(capi.OPT_MESH_BVH, mesh_bvh)
    t0 = time.time()
    ctx.upload_scene(s)
    ctx.synchronize()
    t_up = time.time() - t0
    ctx.resize(W, H)
    ctx.set_camera(iv, ipv)
    ctx.set_params(nb_bounces=8, refract_ind=1.5, date=float(DATE))
    for count in (1, 0):
        ctx.set_option(capi.OPT_COUNT_WORK, count)
        ctx.clear()
        ctx.reset_stats()
        ctx.render(0, SPP)
        ctx.synchronize()
        t0 = time.time()
        ctx.render(SPP, SPP)
        ctx.synchronize()
        dt = time.time() - t0
        st = ctx.stats()
        if count:
            r = st["rays"]
            print(f"mesh_bvh={mesh_bvh} upload {t_up*1e3:.0f} ms; per ray: boxes {st['boxes']/r:.1f} prim tests {st['prim_tests']/r:.2f} "
                  f"tri tests {st['tri_tests']/r:.1f}; rays/path {r/st['paths']:.2f}")
        else:
            print(f"mesh_bvh={mesh_bvh} {SPP} spp in {dt*1e3:.1f} ms = {W*H*SPP/dt/1e6:.1f} Msamples/s")
    ctx.close()
