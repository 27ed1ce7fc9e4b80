"""Tuning builds: libmcpt.so compiled with extra -D switches into montecarlo_pathtracing_b200/lib/variants/<name>/,
selected at run time with MCPT_LIBRARY=<path> (capi.lib()). Usage: build_variant.py name -DMACRO=1 ..."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from montecarlo_pathtracing_b200 import build as b

name, extra = sys.argv[1], sys.argv[2:]
out = b.LIB / "variants" / name
out.mkdir(parents=True, exist_ok=True)
csrc = b.PKG / "csrc"
b._run([b.nvcc_path(), *b.NVCC_FLAGS, *extra, "-I", b.INCLUDE, "-o", out / "libmcpt.so", csrc / "kernels.cu", csrc / "api.cu",
        csrc / "bvh_build.cu", "-ldl"], True)
print(out / "libmcpt.so")
