#!/usr/bin/env python
"""bench.py — Msamples/s of the path-tracing hot path on N B200s (BASELINE.json metric).

  python bench.py [--gpus N --steps K --warmup W] [--impl reference] [--workload c2|c1|c5|mesh|c4|scene1..scene8]

One step = mcpt_render of SPP passes of the workload on every rank (rank r traces its own
seed range, passes [r*SPP, (r+1)*SPP)) followed, for N > 1, by ONE NCCL sum-reduce of the
accumulation buffers to rank 0: spp-split weak scaling (`--scaling strong`: a FIXED spp budget --
BASELINE config 4's 1024 spp for `--workload c4` -- is split over the ranks instead). `value` counts
paths (pixels x spp) of all ranks per second with the scene already resident in HBM; `e2e` repeats
the step through the C ABI from host buffers (scene tables in, resolved RGB32F image out).
`--impl reference` times the reference's algorithm on the host cores (oracle/_ref/libref.so: the
reference's own host code and shader text compiled as C++; its GLSL cannot run on this box).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))

DATE = float(np.float32(1.0 / 60.0))
WORKLOADS = {
    # name: (scene kind, kwargs, width, height, spp, bounces, ior, description)
    "c1": (1, {}, 512, 512, 16, 8, 1.0, "BASELINE config 1: repo scene 1, 512x512, 16 spp, light 0.4, IOR 1.0, 8 bounces"),
    "c2": ("c2", {}, 1920, 1080, 64, 8, 1.5, "BASELINE config 2: repo scene 6, 1920x1080, 64 spp, IOR 1.5, roughness 0.5, light 0.4, 8 bounces"),
    "c5": ("c5", {}, 1920, 1080, 64, 16, 1.5, "BASELINE config 5 (divergence stress) at 64 spp per step: all transparent+reflective, roughness 0.1, 16 bounces, 1080p"),
    "c5_256": ("c5", {}, 1920, 1080, 256, 16, 1.5, "BASELINE config 5: divergence stress, all transparent+reflective, roughness 0.1, 16 bounces, 1920x1080, 256 spp"),
    "mesh": ("mesh", {"n_meshes": 16, "tris_log2": 16}, 1920, 1080, 16, 8, 1.5, "BASELINE config 3 at 16 spp per step: 16 meshes x 65536 triangles, 1080p, 8 bounces"),
    "c3": ("mesh", {"n_meshes": 16, "tris_log2": 16}, 1920, 1080, 256, 8, 1.5, "BASELINE config 3: synthetic 1M-triangle scene (16 meshes x 65536 triangles), 1920x1080, 256 spp, 8 bounces"),
    # the repo's eight scenes (hot keys Q..I, MontecarloGPU/montecarlo.cpp:236-290) at the metric's settings
    **{f"scene{k}": (k, {}, 1920, 1080, 32, 8, 1.0, f"repo scene {k}, 1920x1080, 32 spp per step, light 0.4, IOR 1.0, 8 bounces")
       for k in range(1, 9)},
    # the reference's own screenshots: window 1280x1000, Bounces 9, Paths 1 per frame --
    # one step = one frame of the locked viewer; FPS = 1000 / ms_per_step
    **{f"fps{k}": (k, {"light": 0.443}, 1280, 1000, 1, 9, 1.0, f"viewer frame of the reference's screenshots: repo scene {k}, 1280x1000, 1 path per frame, 9 bounces, IOR 1.0")
       for k in range(1, 9)},
    "c4": ("mesh", {"n_meshes": 16, "tris_log2": 16}, 3840, 2160, 4, 8, 1.5, "BASELINE config 4 at 4 spp per step: 16 meshes x 65536 triangles, 3840x2160, 8 bounces"),
}


# Msamples/s derived from the FPS read-outs of the reference's screenshots (FPS x 1.28 Mpixel
# x 1 path per frame; unknown desktop GPU through OpenGL): the published numbers of the fpsK workloads
REFERENCE_SCREENSHOT_MSAMPLES = {"fps1": 34.5, "fps2": 20.0, "fps3": 24.0, "fps4": 27.6, "fps5": 83.6, "fps7": 13.0, "fps8": 6.6}


def peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)", float(d.get("sm_max_mhz", 1965.0))
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)", 1965.0


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.rows, self.proc, self.idx = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200",
                                          "-i", str(self.idx)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[1]))
                mx.append(float(r[2]))
            except Exception:
                continue
            for name, col in (("hw_slowdown", 5), ("hw_thermal_slowdown", 6), ("sw_thermal_slowdown", 7), ("sw_power_cap", 8)):
                if len(r) > col and r[col].lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def build_scene(name):
    from common import host_scene
    kind, kw, w, h, spp, bounces, ior, desc = WORKLOADS[name]
    kw = dict(kw)
    s = host_scene(kind, kw.pop("light", 0.4), **kw)
    raw = s.prim_data()
    s.finalize()
    return s, raw, (w, h, spp, bounces, ior, desc)


def pin_desc_arrays(d):
    """Page-lock the host tables behind an mcpt_scene_desc (cudaHostRegister), so that the library's host -> device copies
    of the end-to-end arm read PINNED host memory, as the bench contract asks: the copies then run at link speed instead
    of through the driver's staging buffer. Returns the registered (pointer, bytes) list; failures are skipped."""
    import torch
    rt = torch.cuda.cudart()
    regs = []

    def pin(ptr, nbytes):
        if ptr and nbytes > 0:
            try:
                if int(rt.cudaHostRegister(int(ptr), int(nbytes), 0)) == 0:
                    regs.append((int(ptr), int(nbytes)))
            except Exception:
                pass

    pin(d.prim_data, d.n_prims * 256)
    pin(d.root.bb, ((2 << d.root.depth) - 1) * 24)
    pin(d.root.ind, (1 << d.root.depth) * 4)
    for m in range(d.n_meshes):
        md = d.meshes[m]
        if md.bvh.bb:
            pin(md.bvh.bb, ((2 << md.bvh.depth) - 1) * 24)
            pin(md.bvh.ind, (1 << md.bvh.depth) * 4)
        pin(md.tri_idx, md.n_tris * 12)
        pin(md.pos, md.n_verts * 12)
        pin(md.nrm, md.n_verts * 12)
    return regs


def desc_bytes(d):
    """Bytes of host tables behind an mcpt_scene_desc (what upload copies host -> device)."""
    n = d.n_prims * 256 + ((2 << d.root.depth) - 1) * 24 + (1 << d.root.depth) * 4
    for m in range(d.n_meshes):
        md = d.meshes[m]
        if md.bvh.bb:
            n += ((2 << md.bvh.depth) - 1) * 24 + (1 << md.bvh.depth) * 4
        n += md.n_tris * 12 + md.n_verts * 24
    return int(n)


def reference_scene(name):
    """The workload's scene for the CPU arm, built WITHOUT the product where the reference can build it itself:
    repo scenes (and the two configs derived from them, c2 = scene 6 at roughness 0.5, c5 = scene 5 with every
    material made transparent + reflective plus one more light) come from libref.so's own scene builders
    (RTViewer::scene_*, cut out of MontecarloGPU/montecarlo.cpp) and its own camera; the synthetic mesh workloads
    have no reference builder, their tables are replayed from the repo's host mirror. Returns (oracle, inv_v, inv_pv, how)."""
    from oracle.binding import Oracle, have_reference
    kind, kw, w, h, spp, bounces, ior, desc = WORKLOADS[name]
    light = dict(kw).get("light", 0.4)
    if have_reference() and kind != "mesh":
        o = Oracle("reference")
        sid = {"c2": 6, "c5": 5}.get(kind, kind)
        o.build_repo_scene(int(sid), light)
        n = o.n_prims
        raw = o.prim_data()
        if kind == "c2":            # build_scene6_roughness: roughness 0.5 on everything that is not a light
            for i in range(n):
                if not raw[i, 58] > 0.0:
                    o.poke_prim(i, 57, 0.5)
        elif kind == "c5":          # build_divergence_stress: opacity 0.3, shininess 0.8, roughness 0.1, one more light
            for i in range(n):
                o.poke_prim(i, 55, 0.3)
                o.poke_prim(i, 56, 0.8)
                o.poke_prim(i, 57, 0.1)
            o.add_prim(5, o.transfo("tYs", 200, 0, 100, -110, 20, 20, 1), (0.9, 0.9, 0.9, 1.0), 0.0, 0.0, 20 * light)
        o.finalize()
        o.cpu_kind = "reference"
        iv, ipv = o.default_camera(w, h)
        return o, iv, ipv, "scene, BVH and camera built by the reference's own code (libref.so)"
    from montecarlo_pathtracing_b200 import capi
    s, raw, _ = build_scene(name)
    o = oracle_for(s, raw)
    iv, ipv = capi.default_camera(w, h)
    return o, iv, ipv, "scene tables replayed from the repo's host mirror (no reference builder for this workload)"


def workload_config(args, world):
    """The `config` object of the JSON line: identical for the GPU arm and the reference arm."""
    kind, kw, w, h, spp, bounces, ior, desc = WORKLOADS[args.workload]
    strong = args.scaling == "strong"
    return {"workload": desc, "name": args.workload, "width": w, "height": h,
            "spp_per_step_per_gpu": (strong_spp(args.workload) / world) if strong else spp, "bounces": bounces, "ior": ior,
            "parallelism": (f"spp-split x{world}, one ncclReduce(sum) per step" if world > 1 else "single GPU"),
            "l2": "wavefront state of a batch (>1 GB) exceeds the 126 MB L2; no flush needed", "fast_math": args.fast_math,
            "host_tree": getattr(args, "host_tree", 0), "unified": getattr(args, "unified", 1)}


def strong_spp(name):
    """Fixed spp budget of one step under --scaling strong: BASELINE config 4 as written (1024 spp) for c4, eight
    default steps' worth for the other workloads (so that every rank of an 8-GPU split traces one default step)."""
    return 1024 if name == "c4" else 8 * WORKLOADS[name][4]


def oracle_for(s, raw):
    """The CPU arm: oracle/_ref/libref.so (the reference's own sources compiled as C++, kind
    'reference') when it travelled with the repo, else the hand restatement (kind 'port')."""
    from oracle.binding import Oracle, have_reference
    meshes = s.tables()["meshes"]
    kind = "reference" if have_reference() else "port"
    o = Oracle(kind).load_unsorted(raw, meshes).finalize()
    o.cpu_kind = kind
    return o


def run_reference(args, result_out):
    """Reference arm: the reference's own code (libref.so; the hand port when it did not travel) on all host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    kind, kw, w, h, spp, bounces, ior, desc = WORKLOADS[args.workload]
    o, iv, ipv, how = reference_scene(args.workload)
    o.set_uniforms(bounces, ior, DATE, 0, iv, ipv)
    # torchrun exports OMP_NUM_THREADS=1: ask for every core this process may run on explicitly
    cores = o.set_threads(len(os.sched_getaffinity(0)))
    # bounded sample per step: the full frame at 1 spp (mesh workloads: every 8th of the rows)
    rows = h if args.workload not in ("mesh", "c3", "c4") else max(8, h // 8)
    y0 = (h - rows) // 2
    acc = np.zeros((h, w, 3), np.float32)
    for i in range(args.warmup):
        o.render(w, h, i, 1, 0, acc, y0, y0 + max(rows // 8, 1))
    t0 = time.perf_counter()
    for i in range(args.steps):
        o.render(w, h, i, 1, 0, acc, y0, y0 + rows)
    dt = time.perf_counter() - t0
    paths = rows * w * args.steps
    val = paths / dt / 1e6
    sample = f"{rows} rows x {w} px x 1 spp per step (centre band of the {w}x{h} frame), {args.steps} steps"
    cfg = workload_config(args, args.gpus)       # the same object as the GPU arm's line
    note = ("CPU arm: " + how + "; shader text compiled as C++ on the host cores (GLSL through a graphics API is unavailable on this box). "
            "Each timed step is a bounded sample of the workload -- the frame at 1 spp -- in the same unit (paths per second) "
            "as the GPU arm's 64-spp steps")
    result_out.write(json.dumps({
        "impl": "reference", "metric": "Msamples/s", "value": val, "unit": "Msamples/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True,
        "scaling": args.scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": cfg, "note": note,
        "cpu_baseline": {"value": val, "unit": "Msamples/s", "cores": cores, "kind": o.cpu_kind, "sample": sample},
        "e2e": {"value": val, "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }) + "\n")
    result_out.flush()


def quick_measure(name, local, rank, world, steps, warmup, strong=False, fast_math=1):
    """A second workload measured inside the default run, so that the other BASELINE configs are in the driver's own
    record next to the headline one: same timing rules (warm-up, barrier + synchronize, CUDA events on the launching
    stream, max over ranks, clocks sampled over the region), library defaults for every option, no host tree for meshes.
    Returns the sub-record (rank 0) or None."""
    import torch
    import torch.distributed as dist
    from montecarlo_pathtracing_b200 import Context, capi
    from montecarlo_pathtracing_b200.distributed import pass_range, reduce_accum
    s, raw, (w, h, spp, bounces, ior, desc) = build_scene(name)
    iv, ipv = capi.default_camera(w, h)
    ctx = Context(local)
    stream = torch.cuda.current_stream()
    ctx.set_stream(stream.cuda_stream)
    ctx.set_option(capi.OPT_FAST_MATH, fast_math)
    ctx.upload_scene(s.desc(mesh_trees=False) if s.nb_meshes else s.desc())
    ctx.resize(w, h)
    accum = torch.zeros(h * w * 4, dtype=torch.float32, device="cuda")
    ctx.bind_accum(accum.data_ptr(), accum.numel() * 4)
    ctx.set_camera(iv, ipv)
    ctx.set_params(nb_bounces=bounces, refract_ind=ior, date=DATE)
    spp_total = strong_spp(name) if strong else spp * world
    first, count = pass_range(rank, world, spp_total)

    def step():
        accum.zero_()
        ctx.render(first, count)
        reduce_accum(accum, root=0)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(warmup):
        step()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record(stream)
    for _ in range(steps):
        step()
    ev1.record(stream)
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    ms = ev0.elapsed_time(ev1)
    if world > 1:
        t = torch.tensor([ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    st = ctx.stats()
    ctx.close()
    del accum
    torch.cuda.empty_cache()
    if rank != 0:
        return None
    return {"workload": desc, "value": w * h * spp_total * steps / (ms * 1e-3) / 1e6, "unit": "Msamples/s", "ms_per_step": ms / steps,
            "steps": steps, "warmup": warmup, "n_gpus": world, "scaling": "strong" if strong else "weak", "spp_per_step_total": spp_total,
            "gpu_launches": int(st["launches"]), "clocks": clocks}


def claim_stdout():
    """Keep stdout for the ONE JSON line: everything else a library writes to fd 1 (NCCL prints its version
    banner there) goes to stderr. Returns the file to print the result line to."""
    sys.stdout.flush()
    result = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    return result


def main():
    result_out = claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100, help="timed steps (default 100: a timed region of >= 2 s on every workload)")
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"], help="N > 1: weak = spp per rank fixed (default), strong = a fixed spp budget per step split over the ranks")
    ap.add_argument("--reduce", default="torch", choices=["torch", "capi"], help="N > 1: the per-step sum-reduce through torch.distributed.reduce (default) or through the C ABI (mcpt_reduce_nccl on a communicator of its own)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="timed region only (no per-stage pass, no work counters, no end-to-end arm, no CPU baseline): what the ncu captures run")
    ap.add_argument("--fast-math", type=int, default=1, help="1 = MUFU sin/cos/log/pow/rsqrt in the bounce sampler (same parity tolerance, tests/test_gpu_golden.py)")
    ap.add_argument("--ppb", type=int, default=0, help="passes per wavefront batch (0 = auto)")
    ap.add_argument("--refill", type=int, default=0, help="persistent warps: idle lanes that trigger a refill (0 = auto)")
    ap.add_argument("--persistent", type=int, default=1, help="1 = persistent warps with lane refill on tree scenes")
    ap.add_argument("--mesh-bvh", type=int, default=1, help="1 = GPU-built LBVH per mesh, 0 = the uploaded median-split trees")
    ap.add_argument("--top-sah", type=int, default=1, help="1 = SAH-built top-level tree for the persistent walk, 0 = the reference's median-split topology")
    ap.add_argument("--flat-walk", type=int, default=1, help="1 = warp-coherent flat walk for scenes of <= 16 primitives, 0 = tree walk (persistent warps)")
    ap.add_argument("--unified", type=int, default=1, help="scenes with meshes: 1 = one world-space tree over the top level and every mesh instance, 0 = two-level walk")
    ap.add_argument("--fused-raygen", type=int, default=-1, help="1 = the first traversal makes its camera rays itself, 0 = k_raygen, -1 = library default")
    ap.add_argument("--leaf-pool", type=int, default=-1, help="unified walk: N > 0 = leaf tests queued per warp and run 32 at a time once N items wait, 0 = per-lane leaf tests, -1 = library default")
    ap.add_argument("--flat-pool", type=int, default=2, help="scenes of <= 16 primitives: 1 = exact tests pooled per warp and primitive type after the first bounce, 0 = warp-coherent walk everywhere")
    ap.add_argument("--host-tree", type=int, default=0, help="mesh workloads: 1 = upload the reference's median-split mesh trees too (tie-break ranks = its leaf slots), 0 = no host tree at all")
    ap.add_argument("--occ", type=int, default=0, help="traversal kernel register cap: CTAs per SM (0 = library default)")
    ap.add_argument("--other-configs", type=int, default=1, help="default workload only: 1 = also measure BASELINE configs 1, 3, 5 (N = 1) and the strong split of config 4 (every N) in short runs and report them under `other_configs`")
    ap.add_argument("--sustain-s", type=float, default=2.0, help="after the timed steps, keep stepping for this long and report the throughput and clocks of that region too (0 = off)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args, result_out)

    import torch
    import torch.distributed as dist
    from montecarlo_pathtracing_b200 import Context, capi

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the path has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    s, raw, (w, h, spp, bounces, ior, desc) = build_scene(args.workload)
    iv, ipv = capi.default_camera(w, h)
    ctx = Context(local)
    # a non-default torch stream: the library launches on it, torch zeroes / reduces on it,
    # and the CUDA events that time the region are recorded on it
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    assert stream.cuda_stream != 0
    ctx.set_stream(stream.cuda_stream)
    if args.ppb:
        ctx.set_option(capi.OPT_PASSES_PER_BATCH, args.ppb)
    ctx.set_option(capi.OPT_FAST_MATH, args.fast_math)
    if args.occ:
        ctx.set_option(capi.OPT_TRAVERSE_OCC, args.occ)
    ctx.set_option(capi.OPT_TOP_SAH, args.top_sah)
    ctx.set_option(capi.OPT_FLAT_WALK, args.flat_walk)
    ctx.set_option(capi.OPT_MESH_BVH, args.mesh_bvh)
    ctx.set_option(capi.OPT_UNIFIED, args.unified)
    ctx.set_option(capi.OPT_FLAT_POOL, args.flat_pool)
    if args.leaf_pool >= 0:
        ctx.set_option(capi.OPT_LEAF_POOL, args.leaf_pool)
    if args.fused_raygen >= 0:
        ctx.set_option(capi.OPT_FUSED_RAYGEN, args.fused_raygen)
    ctx.set_option(capi.OPT_PERSISTENT, args.persistent)
    ctx.set_option(capi.OPT_REFILL_LANES, args.refill)
    # what upload_scene is given: with --host-tree 0 the mesh descriptors carry no BVH (the library builds its own
    # trees on the GPU and needs none); the reference's median-split trees are then never built for the GPU arm
    scene_desc = s.desc()
    if s.nb_meshes and not args.host_tree and args.mesh_bvh:
        scene_desc = s.desc(mesh_trees=False)
    ctx.upload_scene(scene_desc)
    ctx.resize(w, h)
    # accumulate straight into a torch tensor: it is the NCCL send/recv buffer, no staging copy
    accum = torch.zeros(h * w * 4, dtype=torch.float32, device="cuda")
    ctx.bind_accum(accum.data_ptr(), accum.numel() * 4)
    ctx.set_camera(iv, ipv)
    ctx.set_params(nb_bounces=bounces, refract_ind=ior, date=DATE)
    from montecarlo_pathtracing_b200.distributed import make_nccl_comm, pass_range, reduce_accum
    strong = args.scaling == "strong"
    spp_total = strong_spp(args.workload) if strong else spp * world
    first, count = pass_range(rank, world, spp_total)     # weak: spp passes per rank; strong: a share of the fixed budget
    comm = make_nccl_comm(rank, world) if (world > 1 and args.reduce == "capi") else None

    def reduce_step():
        if comm is not None:
            ctx.reduce_nccl(comm.value, 0)       # ncclReduce issued by the library on its stream (C ABI)
        else:
            reduce_accum(accum, root=0)

    def step():
        accum.zero_()
        ctx.render(first, count)
        reduce_step()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(n_steps):
        """n_steps steps bracketed by barrier + synchronize, CUDA events on the launching stream, max over ranks."""
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        ev0.record(stream)
        for _ in range(n_steps):
            step()
        ev1.record(stream)
        barrier()
        t_ms = ev0.elapsed_time(ev1)
        if world > 1:
            t = torch.tensor([t_ms], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            t_ms = float(t.item())
        return t_ms

    for _ in range(args.warmup):
        step()
    barrier()
    ctx.reset_stats()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ms = timed(args.steps)
    clocks = sampler.stop() if rank == 0 else None
    st = ctx.stats()
    paths_per_step = w * h * spp_total
    value = paths_per_step * args.steps / (ms * 1e-3) / 1e6
    rays = st["rays"]

    # ---- a longer region right behind the timed one: is the number a sustained one? (clocks sampled over it)
    sustained = None
    if args.sustain_s > 0 and not args.no_extras:
        n_sus = max(int(np.ceil(args.sustain_s * 1e3 / max(ms / args.steps, 1e-3))), 1)
        sampler2 = ClockSampler(local)
        if rank == 0:
            sampler2.start()
        ms_sus = timed(n_sus)
        clocks2 = sampler2.stop() if rank == 0 else None
        sustained = {"steps": n_sus, "seconds": ms_sus * 1e-3, "value": paths_per_step * n_sus / (ms_sus * 1e-3) / 1e6,
                     "unit": "Msamples/s", "ms_per_step": ms_sus / n_sus, "clocks": clocks2}

    config = workload_config(args, world)
    if args.no_extras:
        if rank == 0:
            result_out.write(json.dumps({"metric": "Msamples/s", "value": value, "unit": "Msamples/s", "n_gpus": world, "steps": args.steps,
                                         "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
                                         "scaling": args.scaling, "dtype": "f32", "data": "synthetic", "config": config,
                                         "gpu_launches": int(st["launches"]), "clocks": clocks}) + "\n")
            result_out.flush()
        if world > 1:
            dist.destroy_process_group()
        return

    # per-stage device times: the same steps again with every launch bracketed by CUDA events on
    # the launching stream (kept out of the timed region above: the events cost a few percent);
    # the reduce is bracketed too, so the collective that a strong split leans on has a number
    ctx.set_option(capi.OPT_PROFILE_STAGES, 1)
    ctx.stage_times()
    n_prof = min(args.steps, 10)
    red_ev = []
    tiny = torch.zeros(1, device="cuda")
    for _ in range(n_prof):
        accum.zero_()
        ctx.render(first, count)
        if world > 1:
            dist.all_reduce(tiny)        # device-side rendezvous: the reduce below is timed without the ranks' skew
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        reduce_step()
        b.record(stream)
        red_ev.append((a, b))
    barrier()
    stages = ctx.stage_times()
    ctx.set_option(capi.OPT_PROFILE_STAGES, 0)
    reduce_ms = float(np.mean([a.elapsed_time(b) for a, b in red_ev])) if world > 1 else 0.0
    if world > 1:
        t = torch.tensor([reduce_ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        reduce_ms = float(t.item())
    # algorithmic work of one step: the counting instantiation of the traversal kernels (untimed)
    ctx.set_option(capi.OPT_COUNT_WORK, 1)
    ctx.reset_stats()
    step()
    barrier()
    work = ctx.stats()
    ctx.set_option(capi.OPT_COUNT_WORK, 0)
    ctx.reset_stats()

    # ---- end to end through the C ABI with host buffers (scene tables in, image out)
    ctx.bind_accum(accum.data_ptr(), accum.numel() * 4)
    h2d = desc_bytes(scene_desc) + 2 * 64 + 28
    img_t = torch.empty((h, w, 3), dtype=torch.float32, pin_memory=True)   # pinned host destination
    img = img_t.numpy()
    d2h = img.nbytes if rank == 0 else 0
    upload_s = []
    pinned = pin_desc_arrays(scene_desc)      # the scene tables the library copies from: page-locked, like the image

    def e2e_step(n_passes=count):
        t_up = time.perf_counter()
        ctx.upload_scene(scene_desc)              # host tables -> device, re-pack, tree builds (returns when they are done)
        upload_s.append(time.perf_counter() - t_up)
        ctx.set_camera(iv, ipv)
        ctx.set_params(nb_bounces=bounces, refract_ind=ior, date=DATE)
        accum.zero_()
        ctx.render(first, n_passes)
        reduce_step()
        if rank == 0:
            ctx._ck(ctx._L.mcpt_resolve(ctx._c, float(spp_total), 1, img.ctypes.data, capi.DST_HOST_RGB32F))

    e2e_step()
    e2e_step()
    barrier()
    upload_s.clear()
    t0 = time.perf_counter()
    marks = []
    n_e2e = min(args.steps, 20)
    for _ in range(n_e2e):
        e2e_step()
        marks.append(time.perf_counter())
    barrier()
    e2e_s = time.perf_counter() - t0
    upload_ms = 1e3 * float(np.median(upload_s))
    # time to first image after a scene change: upload + ONE pass + resolve to the host
    barrier()
    t1 = time.perf_counter()
    e2e_step(1 if count >= 1 else 0)
    barrier()
    first_image_ms = 1e3 * (time.perf_counter() - t1)
    if rank == 0:
        sys.stderr.write("e2e ms per step: " + " ".join(f"{1e3 * (b - a):.1f}" for a, b in zip([t0] + marks[:-1], marks)) + "\n")
    if world > 1:
        t = torch.tensor([e2e_s], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    e2e_val = paths_per_step * n_e2e / e2e_s / 1e6

    accum_bytes = int(accum.numel() * 4)
    # ---- the other BASELINE configs, in short runs of the same kind (only next to the default headline workload)
    other = None
    if args.other_configs and args.workload == "c2" and args.scaling == "weak":
        other = {}
        ctx.close()                      # free the headline context's wavefront buffers first
        accum = None
        torch.cuda.empty_cache()
        torch.cuda.set_stream(stream)
        if world == 1:
            other["c1"] = quick_measure("c1", local, rank, world, 50, 5, fast_math=args.fast_math)
            other["c3"] = quick_measure("c3", local, rank, world, 4, 1, fast_math=args.fast_math)          # 256 spp per step, as BASELINE states it
            other["c3_16spp"] = quick_measure("mesh", local, rank, world, 10, 3, fast_math=args.fast_math)
            other["c5"] = quick_measure("c5_256", local, rank, world, 3, 1, fast_math=args.fast_math)     # 256 spp per step, as BASELINE states it
            other["c5_64spp"] = quick_measure("c5", local, rank, world, 10, 3, fast_math=args.fast_math)
        # BASELINE config 4 as written: 3840x2160, a FIXED 1024 spp split over the N ranks, one reduce, rank 0 holds the sum
        other["c4_strong"] = quick_measure("c4", local, rank, world, 1 if world == 1 else 2, 1, strong=True, fast_math=args.fast_math)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel (the traversal kernel serves both traverse stages)
    peak, peak_src, sm_max = peaks()
    trav_ms = stages["traverse"][0] + stages["traverse_interior"][0]
    trav_launches = stages["traverse"][1] + stages["traverse_interior"][1]
    kernel_ms = {k: v[0] for k, v in stages.items()}
    total_kernel_ms = sum(kernel_ms.values())
    dominant = max(kernel_ms, key=kernel_ms.get)
    trav_ms_step = trav_ms / n_prof
    rays_step = work["rays"]
    sm_count = torch.cuda.get_device_properties(local).multi_processor_count
    clk = (clocks or {}).get("sm_mhz") or sm_max or 1965.0
    # What binds the traversal kernel is the ISSUE rate (ncu: 58..77 % of the issue slots busy, DRAM at 5..30 %): the
    # roof is SMs x 4 schedulers x one warp instruction per cycle at the clock sampled under load. `achieved` = the
    # warp instructions the traversal launches of ONE step execute -- deterministic for a fixed workload, counted by
    # the committed ncu pass of this same command (profiles/r02_step_metrics.json, scripts/ncu_step_metrics.py) --
    # divided by the traversal time of a step measured live here with CUDA events.
    issue_peak = sm_count * 4 * clk * 1e6 / 1e9          # G warp instructions / s
    sm_file = ROOT / "profiles" / "r02_step_metrics.json"
    m = json.loads(sm_file.read_text()).get(args.workload) if sm_file.exists() else None
    mt = (m or {}).get("by_stage", {}).get("traverse")
    # the committed counters belong to the library's default option set; any tuning switch on the command line voids them
    default_opts = (args.unified == 1 and args.persistent == 1 and args.mesh_bvh == 1 and args.top_sah == 1 and args.flat_walk == 1 and
                    args.flat_pool == 2 and args.leaf_pool == -1 and args.fused_raygen == -1 and args.refill == 0 and args.occ == 0 and args.ppb == 0 and args.fast_math == 1 and not os.environ.get("MCPT_LIBRARY"))
    same_launches = bool(mt) and default_opts and mt["launches"] * n_prof == trav_launches
    roofline = {"kernel": "k_traverse_unified" if (s.nb_meshes and args.unified and args.persistent and args.mesh_bvh) else "k_traverse",
                "bound": "issue", "unit": "Gwarp-inst/s", "peak": issue_peak,
                "peak_source": f"{sm_count} SMs x 4 schedulers x {clk:.0f} MHz (SM clock sampled under load)"}
    if mt and same_launches:
        achieved = mt["warp_inst"] / (trav_ms_step * 1e-3) / 1e9
        roofline.update({"achieved": achieved, "frac": achieved / issue_peak,
                         "warp_inst_per_step": mt["warp_inst"], "warp_inst_per_ray": mt["warp_inst"] / max(rays_step, 1),
                         "active_lanes": mt["active_lanes"], "useful_issue_frac": achieved / issue_peak * mt["active_lanes"] / 32.0,
                         "issue_active_pct_under_ncu": mt["issue_active_pct_time_weighted"],
                         "active_lanes_first_bounce": (m.get("traverse_first_bounce") or {}).get("active_lanes"),
                         "active_lanes_later_launches": (m.get("traverse_later_launches") or {}).get("active_lanes"),
                         "traffic": mt["dram_bytes"] / max(mt["launches"], 1), "traffic_unit": "DRAM bytes per launch (ncu dram__bytes_read + write)",
                         "l2_bytes_per_ray": mt["lts_bytes"] / max(rays_step, 1),
                         "l2_gbs": mt["lts_bytes"] / (trav_ms_step * 1e-3) / 1e9,
                         "counters_source": f"profiles/r02_step_metrics.json[{args.workload}] (ncu pass at commit {m.get('commit')})"})
    else:
        roofline.update({"achieved": None, "frac": None, "traffic": None,
                         "counters_source": "no ncu pass committed for this workload / option set (launch counts differ)" if mt else "no ncu pass committed for this workload"})
    if s.nb_meshes and roofline.get("l2_gbs"):
        # the mesh trees are served from L2 -- its bandwidth measured here with a resident-set copy
        # (32 MB -> 32 MB inside the 126 MB L2, read + write bytes, best of 5 runs of 20 copies)
        a_l2 = torch.empty(8 << 20, dtype=torch.float32, device="cuda")
        b_l2 = torch.empty_like(a_l2)
        best = 0.0
        for _ in range(5):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for _ in range(20):
                b_l2.copy_(a_l2)
            e1.record(stream)
            torch.cuda.synchronize()
            best = max(best, 20 * 2 * a_l2.numel() * 4 / (e0.elapsed_time(e1) * 1e-3) / 1e9)
        roofline["l2_peak_gbs"] = best
        roofline["l2_frac"] = roofline["l2_gbs"] / best if best > 0 else None
        del a_l2, b_l2
    roofline.update({"launches": trav_launches // n_prof, "avg_launch_ms": trav_ms / max(trav_launches, 1),
                     "rays_per_launch": rays_step / max(trav_launches // n_prof, 1),
                     "share_of_kernel_time": trav_ms / total_kernel_ms if total_kernel_ms else None,
                     "stage_ms_per_step": {k: v / n_prof for k, v in kernel_ms.items()}, "dominant_stage": dominant})
    # secondary: the HBM figure. Algorithmic bytes per traversed ray: 4 B queue index read, 32 B ray sector
    # (origin | direction) read, 32 B hit sector (local hit point, primitive | face, direction) written; scene data of
    # the repo scenes is served from shared memory (0 B), the trees of the mesh workloads from L1 / L2 (l2_* above).
    bytes_per_ray = 68.0
    hbm_achieved = rays_step * bytes_per_ray / (trav_ms_step * 1e-3) / 1e9 if trav_ms_step > 0 else 0.0
    roofline["hbm"] = {"achieved": hbm_achieved, "peak": peak, "unit": "GB/s", "frac": hbm_achieved / peak, "bytes_per_ray": bytes_per_ray,
                       "peak_source": peak_src, "note": "not the binding roof: the path state streams through HBM once per bounce, the scene does not"}

    # FP32 figure: flops = 24 N_box + (42 + F_type) N_record + 46 N_tri
    # + 60 N_hit + 230 N_segment, FMA = 2, counted by the kernels themselves in the untimed counting
    # step above. F_type is averaged over the scene's analytic records (the counters are not split
    # by type). Peak = SMs x 128 FP32 lanes x 2 x the SM clock sampled under load.
    f_type = {capi.PRIM_SPHERE: 34.0, capi.PRIM_CUBE: 48.0, capi.PRIM_CYLINDER: 58.0, capi.PRIM_CONE: 66.0,
              capi.PRIM_ORIENTED_QUAD: 14.0, capi.PRIM_MESH: 0.0}
    types = [int(t) for t in s.prim_data()[:, 48]]   # PrimData texel 12 .x = type code
    f_avg = sum(f_type.get(t, 0.0) for t in types) / max(len(types), 1)
    trav_flops = 24.0 * work["boxes"] + (42.0 + f_avg) * work["prim_tests"] + 46.0 * work["tri_tests"]
    step_flops = trav_flops + (60.0 + 230.0) * work["segments"]
    fp32_peak = sm_count * 128 * 2 * clk * 1e6 / 1e12
    roofline["fp32"] = {
        "unit": "TFLOP/s", "peak": fp32_peak, "peak_source": f"{sm_count} SMs x 128 lanes x 2 x {clk:.0f} MHz (clock under load)",
        "traverse": {"achieved": trav_flops / (trav_ms_step * 1e-3) / 1e12 if trav_ms_step > 0 else None,
                     "flops_per_ray": trav_flops / max(work["rays"], 1)},
        "step": {"achieved": step_flops / (ms / args.steps * 1e-3) / 1e12, "flops_per_path": step_flops / max(work["paths"], 1)},
        "work_per_step": {k: int(work[k]) for k in ("paths", "rays", "segments", "boxes", "prim_tests", "tri_tests")},
        "f_type_avg": f_avg,
    }
    roofline["fp32"]["traverse"]["frac"] = (roofline["fp32"]["traverse"]["achieved"] or 0.0) / fp32_peak
    roofline["fp32"]["step"]["frac"] = roofline["fp32"]["step"]["achieved"] / fp32_peak
    # SURVEY 8(d) byte formula for the whole step (every stage, wavefront state included):
    # 32 N_node + 36 N_tri + 96 N_ray + 32 N_path (primitive records are served from shared memory)
    step_bytes = 32.0 * work["boxes"] * (1.0 if s.nb_meshes else 0.0) + 36.0 * work["tri_tests"] + 96.0 * work["rays"] + 32.0 * work["paths"]
    roofline["step_bytes_survey"] = {"bytes_per_step": step_bytes, "achieved": step_bytes / (ms / args.steps * 1e-3) / 1e9,
                                     "unit": "GB/s", "frac": step_bytes / (ms / args.steps * 1e-3) / 1e9 / peak}

    cpu = None
    if not args.no_cpu_baseline:
        o, r_iv, r_ipv, how = reference_scene(args.workload)
        o.set_uniforms(bounces, ior, DATE, 0, r_iv, r_ipv)
        cores = o.set_threads(len(os.sched_getaffinity(0)))
        rows = h if args.workload not in ("mesh", "c3", "c4") else max(8, h // 8)
        y0 = (h - rows) // 2
        acc = np.zeros((h, w, 3), np.float32)
        t0 = time.perf_counter()
        o.render(w, h, 0, 1, 0, acc, y0, y0 + rows)          # 1 spp: warm-up and calibration
        t1 = time.perf_counter() - t0
        n_spp = int(min(32, max(1, round(12.0 / max(t1, 1e-3)))))   # ~12 s of CPU work
        t0 = time.perf_counter()
        o.render(w, h, 1, n_spp, 0, acc, y0, y0 + rows)
        dt = time.perf_counter() - t0
        cpu = {"value": rows * w * n_spp / dt / 1e6, "unit": "Msamples/s", "cores": cores, "kind": o.cpu_kind,
               "sample": f"{rows} rows x {w} px x {n_spp} spp (centre band of the {w}x{h} frame), {dt:.1f} s; {how}",
               **({"counters": o.last_counters} if any((o.last_counters or {}).values()) else {})}

    out = {
        "metric": "Msamples/s", "value": value, "unit": "Msamples/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": args.scaling,
        "vs_baseline": (value / REFERENCE_SCREENSHOT_MSAMPLES[args.workload]) if args.workload in REFERENCE_SCREENSHOT_MSAMPLES else None,
        "dtype": "f32", "data": "synthetic",
        "config": config,
        "mrays_per_s": rays / (ms * 1e-3) / 1e6 * world,
        "rays_per_path": rays / max(st["paths"], 1),
        "gpu_launches": int(st["launches"]),
        "e2e": {"value": e2e_val, "unit": "Msamples/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                "steps": n_e2e, "upload_ms": upload_ms, "first_image_ms": first_image_ms,
                "host_buffers": f"scene tables page-locked with cudaHostRegister ({len(pinned)} arrays), image in pinned memory"},
        "reduce": {"ms": reduce_ms, "bytes": accum_bytes, "through": args.reduce} if world > 1 else None,
        "sustained": sustained, "other_configs": other,
        "roofline": roofline, "cpu_baseline": cpu, "clocks": clocks,
    }
    result_out.write(json.dumps(out) + "\n")
    result_out.flush()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
