"""Headless replay of the viewer's render loop: scene -> image file + metrics.

  python -m montecarlo_pathtracing_b200.render --scene 6 --width 1920 --height 1080 --spp 64 \
      --bounces 8 --ior 1.5 --roughness 0.5 --out scene6.png [--pfm scene6.pfm] [--ref converged.npz|.pfm] \
      [--curve 1,4,16,64]

It issues the call sequence of RTViewer::draw_ogl's RT branch (MontecarloGPU/montecarlo.cpp:408-477)
through the C ABI -- upload, resize, camera, params, nb passes, resolve -- with the default camera of
montecarlo.cpp:388-405, and writes what the window would show (8-bit, clamped, no gamma) and/or the
float image. Under torchrun the passes are split over the ranks and reduced once (distributed.py).
Needs a CUDA device: there is no CPU path.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

import numpy as np

from . import Context, Scene, capi
from . import imageio

SCENES = {"1": "scene_box_diffuse", "2": "scene_box_balls", "3": "scene_menger", "4": "scene_box_no_top",
          "5": "scene_materials", "6": "scene_4boules", "7": "scene_menger_lights", "8": "scene_colonnes",
          "c5": "divergence stress (scene_materials, all transparent+reflective)", "mesh": "synthetic displaced-sphere meshes"}


def build_scene(args) -> Scene:
    s = Scene()
    if args.scene == "c5":
        s.build_divergence_stress(args.light)
    elif args.scene == "mesh":
        s.build_synthetic_meshes(args.meshes, args.tris_log2, args.seed)
    elif args.scene == "6" and args.roughness is not None:
        s.build_scene6_roughness(args.light, args.roughness)
    else:
        s.build_repo_scene(int(args.scene), args.light)
    s.finalize()
    return s


def load_ref(path):
    if path.endswith(".npz"):
        return np.load(path)["image"].astype(np.float32)
    return imageio.read_pfm(path)


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--scene", default="6", choices=sorted(SCENES))
    ap.add_argument("--width", type=int, default=1280)
    ap.add_argument("--height", type=int, default=1000)     # the reference's window, montecarlo.cpp:801
    ap.add_argument("--spp", type=int, default=64, help="passes in total (nb_paths x frames while locked)")
    ap.add_argument("--first-pass", type=int, default=0)
    ap.add_argument("--bounces", type=int, default=9)
    ap.add_argument("--ior", type=float, default=1.0)
    ap.add_argument("--light", type=float, default=0.4, help="light_intensity_ slider")
    ap.add_argument("--roughness", type=float, default=None, help="scene 6 only: roughness override (BASELINE config 2)")
    ap.add_argument("--mode", type=int, default=0, help="0 montecarlo.frag, 1 montecarlo_mat.frag, 2 montecarlo_mat_tr.frag")
    ap.add_argument("--date", type=float, default=1.0 / 60.0)
    ap.add_argument("--subsampling", type=int, default=1)
    ap.add_argument("--meshes", type=int, default=16)
    ap.add_argument("--tris-log2", type=int, default=16)
    ap.add_argument("--seed", type=int, default=12345)
    ap.add_argument("--fast-math", type=int, default=1)
    ap.add_argument("--out", default=None, help="PNG of the back buffer (clamped 8 bit)")
    ap.add_argument("--pfm", default=None, help="float image, unclamped")
    ap.add_argument("--ref", default=None, help="converged image (.npz with 'image', or .pfm) to compare with")
    ap.add_argument("--curve", default=None, help="comma-separated spp marks: RMSE / luminance after each (needs --ref)")
    args = ap.parse_args(argv)
    sys.stdout.flush()
    report_out = os.fdopen(os.dup(1), "w") if argv is None else sys.stdout   # CLI: keep fd 1 for the report line
    if argv is None:
        os.dup2(2, 1)                                                         # (NCCL prints its banner on stdout)

    world, rank, local = (int(os.environ.get(k, d)) for k, d in (("WORLD_SIZE", "1"), ("RANK", "0"), ("LOCAL_RANK", "0")))
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    s = build_scene(args)
    w, h = args.width, args.height
    iv, ipv = capi.default_camera(w, h)
    ctx = Context(local)
    ctx.set_option(capi.OPT_FAST_MATH, args.fast_math)
    ctx.upload_scene(s)
    ctx.resize(w, h)
    ctx.set_camera(iv, ipv)
    ctx.set_params(nb_bounces=args.bounces, refract_ind=args.ior, date=args.date, mode=args.mode,
                   nb_prims=s.nb_prim, nb_emissives=s.nb_emissives)
    report = {"scene": SCENES[args.scene], "width": w, "height": h, "spp": args.spp, "bounces": args.bounces, "ior": args.ior,
              "n_gpus": world}
    ref = load_ref(args.ref) if args.ref else None

    if args.curve:
        if ref is None or world > 1:
            raise SystemExit("--curve needs --ref and a single process")
        marks = sorted({int(x) for x in args.curve.split(",")})
        report["curve"] = imageio.convergence_curve(ctx, ref, marks, args.first_pass)
        args.spp = marks[-1]
        img = ctx.resolve(float(args.spp), args.subsampling)
    else:
        t0 = time.perf_counter()
        if world > 1:
            import torch
            from .distributed import SplitRenderer
            stream = torch.cuda.Stream()
            torch.cuda.set_stream(stream)
            ctx.set_stream(stream.cuda_stream)          # library launches, zeroing and the reduce share one stream
            SplitRenderer(ctx, w, h, rank, world).render(args.spp, args.first_pass, root=0)
            torch.cuda.synchronize()
        else:
            ctx.render(args.first_pass, args.spp)
        img = ctx.resolve(float(args.spp), args.subsampling)
        dt = time.perf_counter() - t0
        report["seconds"] = dt
        report["msamples_per_s"] = w * h * args.spp / dt / 1e6
    st = ctx.stats()
    report["rays"] = int(st["rays"])
    if rank == 0:
        if args.out:
            imageio.write_png(args.out, img)
            report["png"] = args.out
        if args.pfm:
            imageio.write_pfm(args.pfm, img)
            report["pfm"] = args.pfm
        if ref is not None and not args.curve:
            report["vs_ref"] = imageio.compare(img, ref)
        report["non_finite_pixels"] = int((~np.isfinite(img)).any(axis=-1).sum())
        report_out.write(json.dumps(report) + "\n")
        report_out.flush()
    ctx.close()
    if dist is not None:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
