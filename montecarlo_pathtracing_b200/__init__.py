"""B200-native Monte Carlo path-tracing core: a drop-in for the fragment-shader hot path of
ksaintmarc/Montecarlo-PathTracing (bounce loop, BVH traversal, progressive accumulation).

  capi      ctypes bindings of include/mcpt.h (CUDA, sm_100a) and include/mcpt_host.h
  renderer  Scene (host mirror of BVH_GPU_Scene) and Context (CUDA path tracer)
  build     in-tree nvcc / g++ builds
  imageio   PNG / PFM writers, RMSE / luminance / convergence curve (no GPU needed)
  render    headless replay of the viewer's render loop (CLI: python -m montecarlo_pathtracing_b200.render)
"""
from . import capi  # noqa: F401
from .capi import McptError  # noqa: F401
from .renderer import Context, Scene  # noqa: F401

__all__ = ["capi", "Context", "Scene", "McptError"]
