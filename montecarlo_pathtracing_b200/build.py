"""In-tree builds: nvcc for the sm_100a library, g++ for the host mirror and the oracle.

Everything is compiled by explicit compiler invocations (no JIT cache), outputs
land next to the sources so they travel to the GPU box with the repo snapshot:

  montecarlo_pathtracing_b200/lib/libmcpt.so       CUDA kernels + C ABI (include/mcpt.h)
  montecarlo_pathtracing_b200/lib/libmcpt_host.so  host-side mirror     (include/mcpt_host.h)
  oracle/liboracle.so                              CPU oracle (test infrastructure)
  oracle/_ref/libref.so                            reference sources compiled as C++ (when /root/reference exists)
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
PKG = ROOT / "montecarlo_pathtracing_b200"
LIB = PKG / "lib"
INCLUDE = ROOT / "include"
ORACLE = ROOT / "oracle"

NVCC_FLAGS = [
    "-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
    # hit records must be reproducible bit for bit: no FMA contraction (see csrc/vecmath.cuh)
    "-fmad=false", "-Xcompiler", "-fPIC,-ffp-contract=off", "-shared",
]
CXX_FLAGS = ["-O2", "-std=c++17", "-fPIC", "-shared", "-ffp-contract=off", "-Wall"]


def _newer(target: Path, sources) -> bool:
    if not target.exists():
        return True
    t = target.stat().st_mtime
    return any(Path(s).stat().st_mtime > t for s in sources)


def _run(cmd, verbose):
    if verbose:
        print("+", " ".join(str(c) for c in cmd), flush=True)
    r = subprocess.run([str(c) for c in cmd], capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError(f"build step failed: {' '.join(str(c) for c in cmd[:3])} ...")
    return r.stdout + r.stderr


def nvcc_path() -> str:
    p = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(p):
        raise RuntimeError("nvcc not found; the CUDA library cannot be built")
    return p


def build_cuda(force=False, verbose=False, extra=()):
    """libmcpt.so: every CUDA kernel of the path, for sm_100a only."""
    LIB.mkdir(exist_ok=True)
    out = LIB / "libmcpt.so"
    csrc = PKG / "csrc"
    deps = list(csrc.glob("*")) + list(INCLUDE.glob("*.h"))
    if force or _newer(out, deps):
        _run([nvcc_path(), *NVCC_FLAGS, *extra, "-I", INCLUDE, "-o", out, csrc / "kernels.cu", csrc / "api.cu", csrc / "bvh_build.cu", "-ldl"], verbose)
    return out


def build_host(force=False, verbose=False):
    """libmcpt_host.so: Eigen/GL-free mirror of the reference's scene/BVH/camera host code."""
    LIB.mkdir(exist_ok=True)
    out = LIB / "libmcpt_host.so"
    host = PKG / "host"
    srcs = sorted(host.glob("*.cpp"))
    deps = list(host.glob("*")) + list(INCLUDE.glob("*.h")) + [LIB / "libmcpt.so"]
    if force or _newer(out, deps):
        _run(["g++", *CXX_FLAGS, "-I", INCLUDE, "-I", host, "-o", out, *srcs, "-L", LIB, "-lmcpt", "-Wl,-rpath,$ORIGIN"], verbose)
    return out


def build_oracle(force=False, verbose=False):
    """liboracle.so: the CPU restatement of the reference shaders (checker only)."""
    out = ORACLE / "liboracle.so"
    srcs = [ORACLE / n for n in ("ref_tables.cpp", "ref_shader.cpp", "ref_shader_path.cpp", "oracle_capi.cpp")]
    deps = srcs + list(ORACLE.glob("*.h"))
    if force or _newer(out, deps):
        _run(["g++", "-O3", "-march=x86-64-v2", "-fopenmp", *CXX_FLAGS[1:], "-I", ORACLE, "-o", out, *srcs], verbose)
    return out


def build_ref(force=False, verbose=False):
    """oracle/_ref: the reference's own sources compiled as C++ (only where /root/reference exists)."""
    script = ORACLE / "build_ref.py"
    if not script.exists() or not Path("/root/reference").exists():
        return None
    r = subprocess.run([sys.executable, str(script)] + (["--force"] if force else []), capture_output=True, text=True)
    if verbose or r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
    if r.returncode != 0:
        raise RuntimeError("oracle/_ref build failed")
    # ... and the same sources with the INTEGRATION.md patch compiled in, linked against libmcpt.so
    integ = ORACLE / "build_integration.py"
    if integ.exists():
        r = subprocess.run([sys.executable, str(integ)] + (["--force"] if force else []), capture_output=True, text=True)
        if verbose or r.returncode != 0:
            sys.stderr.write(r.stdout + r.stderr)
        if r.returncode != 0:
            raise RuntimeError("oracle/_ref integration build failed")
    out = ORACLE / "_ref" / "libref.so"
    return out if out.exists() else None


def build_all(force=False, verbose=False):
    outs = [build_cuda(force, verbose), build_host(force, verbose), build_oracle(force, verbose)]
    ref = build_ref(force, verbose)
    if ref:
        outs.append(ref)
    return outs


if __name__ == "__main__":
    for o in build_all(force="--force" in sys.argv, verbose=True):
        print("built", o)
