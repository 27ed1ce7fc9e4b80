#!/usr/bin/env python
"""Experiment: how much does the ORDER of a batch of secondary rays change the time of the production traversal
kernel (persistent warps over the unified tree)?

Camera rays are traced into the 1 M-triangle scene; from every hit a diffuse-like secondary ray is made (uniform
direction in the hemisphere facing the incoming ray). That batch is then traced through mcpt_trace_rays_device
  a) in pixel order (what the wavefront's queues roughly preserve),
  b) stable-sorted by direction octant,
  c) sorted by octant, then by a 30-bit Morton code of the origin,
  d) shuffled.
The kernel and the rays are the same; only the position of a ray in the batch differs. Wall clock around 5 calls
with a device synchronisation on both sides (the pack / unpack kernels of the entry point are included, ~5 %).

  python scripts/exp_ray_order.py [--workload mesh|scene8] [--refill 8]
"""
import argparse
import json
import sys
import time
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))
import bench  # noqa: E402
from montecarlo_pathtracing_b200 import Context, capi  # noqa: E402


def morton30(q):
    def spread(v):
        v = v.astype(np.uint64) & 0x3ff
        v = (v | (v << 16)) & 0x30000ff
        v = (v | (v << 8)) & 0x300f00f
        v = (v | (v << 4)) & 0x30c30c3
        v = (v | (v << 2)) & 0x9249249
        return v
    return spread(q[:, 0]) | (spread(q[:, 1]) << 1) | (spread(q[:, 2]) << 2)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="mesh")
    ap.add_argument("--refill", type=int, default=8)
    ap.add_argument("--reps", type=int, default=5)
    args = ap.parse_args()
    s, raw, (w, h, spp, bounces, ior, desc) = bench.build_scene(args.workload)
    ctx = Context(0)
    ctx.set_option(capi.OPT_REFILL_LANES, args.refill)
    ctx.upload_scene(s)
    iv, ipv = capi.default_camera(w, h)
    # camera rays: the four corner directions of raytracer.vert (csrc/api_render.inl camera_rays), bilinear in between --
    # close enough for an experiment about ray ORDER
    ivm, ipvm = iv.reshape(4, 4).T.astype(np.float64), ipv.reshape(4, 4).T.astype(np.float64)   # column-major
    P = (ivm @ np.array([0, 0, 0, 1.0]))[:3]
    c = {}
    for cx in (-1.0, 1.0):
        for cy in (-1.0, 1.0):
            Q = ipvm @ np.array([cx, cy, 1.0, 1.0])
            dd = Q[:3] / Q[3] - P
            c[(cx, cy)] = dd / np.linalg.norm(dd)
    u = (np.arange(w) / (w - 1))[None, :, None]
    v = (np.arange(h) / (h - 1))[:, None, None]
    d = c[(-1.0, -1.0)] * (1 - u) * (1 - v) + c[(1.0, -1.0)] * u * (1 - v) + c[(-1.0, 1.0)] * (1 - u) * v + c[(1.0, 1.0)] * u * v
    rd = (d / np.linalg.norm(d, axis=2, keepdims=True)).reshape(-1, 3).astype(np.float32)
    ro = np.tile(P.astype(np.float32), (h * w, 1))
    hits = ctx.trace_rays(ro, rd)
    ok = hits["index"] >= 0
    rng = np.random.default_rng(5)
    n = int(ok.sum())
    dirs = rng.normal(size=(n, 3)).astype(np.float32)
    dirs /= np.linalg.norm(dirs, axis=1, keepdims=True)
    flip = np.sum(dirs * rd[ok], axis=1) > 0          # leave on the side the ray came from
    dirs[flip] *= -1
    org = hits["pg"][ok] + 1e-2 * dirs
    octant = (dirs[:, 0] < 0).astype(np.uint64) | ((dirs[:, 1] < 0).astype(np.uint64) << 1) | ((dirs[:, 2] < 0).astype(np.uint64) << 2)
    lo, hi = org.min(0), org.max(0)
    q = np.clip(((org - lo) / (hi - lo + 1e-9) * 1023).astype(np.int64), 0, 1023)
    orders = {
        "pixel order": np.arange(n),
        "by octant (stable)": np.argsort(octant, kind="stable"),
        "by octant, then Morton(origin)": np.argsort((octant << np.uint64(30)) | morton30(q), kind="stable"),
        "by Morton(origin) only": np.argsort(morton30(q), kind="stable"),
        "shuffled": rng.permutation(n),
    }
    out = {"workload": args.workload, "camera_rays": int(len(ro)), "rays": n, "refill_lanes": args.refill, "ms": {}}
    res = torch.empty((n, 11), dtype=torch.float32, device="cuda")
    ref_bytes = None
    for name, perm in orders.items():
        o4 = torch.zeros((n, 4), dtype=torch.float32, device="cuda")
        d4 = torch.zeros((n, 4), dtype=torch.float32, device="cuda")
        o4[:, :3] = torch.from_numpy(org[perm]).cuda()
        d4[:, :3] = torch.from_numpy(dirs[perm]).cuda()
        ctx.trace_rays_device(n, o4.data_ptr(), d4.data_ptr(), res.data_ptr())
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(args.reps):
            ctx.trace_rays_device(n, o4.data_ptr(), d4.data_ptr(), res.data_ptr())
        torch.cuda.synchronize()
        ms = (time.perf_counter() - t0) * 1e3 / args.reps
        out["ms"][name] = round(ms, 3)
        # same records whatever the order
        back = np.empty(n, capi.HIT_DTYPE)
        back[perm] = res.cpu().numpy().view(np.uint8).reshape(n, 44).copy().view(capi.HIT_DTYPE).reshape(n)
        if ref_bytes is None:
            ref_bytes = back.tobytes()
        out.setdefault("same_records", True)
        out["same_records"] = out["same_records"] and back.tobytes() == ref_bytes
    print(json.dumps(out))


if __name__ == "__main__":
    main()
