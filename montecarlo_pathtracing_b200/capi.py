"""ctypes bindings of the two in-tree shared libraries.

libmcpt.so       include/mcpt.h       the CUDA path tracer (product)
libmcpt_host.so  include/mcpt_host.h  host-side mirror of the reference's scene API

There is no CPU fallback: if libmcpt.so is missing, loading raises; if there is no
CUDA device, Context() raises McptError(MCPT_E_NODEVICE).
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

import numpy as np

LIB_DIR = Path(__file__).resolve().parent / "lib"

MCPT_OK = 0
MCPT_E_INVALID, MCPT_E_CUDA, MCPT_E_NODEVICE, MCPT_E_NOMEM, MCPT_E_NCCL, MCPT_E_UNSUPPORTED = -1, -2, -3, -4, -5, -6
PRIM_MESH, PRIM_SPHERE, PRIM_CUBE, PRIM_CYLINDER, PRIM_CONE, PRIM_ORIENTED_QUAD = range(6)
DST_HOST_RGB32F, DST_HOST_RGBA8, DST_DEVICE_RGB32F, DST_DEVICE_RGBA8 = range(4)
OPT_PASSES_PER_BATCH, OPT_SMEM_STAGING, OPT_FAST_MATH, OPT_MESH_BVH, OPT_PROFILE_STAGES, OPT_COUNT_WORK, OPT_FLAT_WALK, OPT_TRAVERSE_OCC, OPT_PERSISTENT, OPT_REFILL_LANES, OPT_TOP_SAH, OPT_TRACE_WALKER, OPT_UNIFIED, OPT_NEE, OPT_FLAT_POOL, OPT_LEAF_POOL, OPT_FUSED_RAYGEN = range(17)
STAGE_NAMES = ("raygen", "traverse", "shade", "bvh_build", "traverse_interior", "shade_exit", "accumulate", "resolve")


class PrimData(C.Structure):  # include/mcpt_scene.h mcpt_prim_data
    _fields_ = [("transfo", C.c_float * 16), ("inv_transfo", C.c_float * 16), ("mesh_transfo", C.c_float * 16),
                ("type", C.c_float * 4), ("color", C.c_float * 4), ("mat_info", C.c_float * 4), ("padding", C.c_float * 4)]


class BvhDesc(C.Structure):
    _fields_ = [("bb", C.c_void_p), ("ind", C.c_void_p), ("depth", C.c_int32)]


class MeshDesc(C.Structure):
    _fields_ = [("bvh", BvhDesc), ("tri_idx", C.c_void_p), ("pos", C.c_void_p), ("nrm", C.c_void_p),
                ("n_tris", C.c_uint32), ("n_verts", C.c_uint32), ("mesh_line", C.c_int32)]


class SceneDesc(C.Structure):
    _fields_ = [("prim_data", C.c_void_p), ("n_prims", C.c_int32), ("nb_emissives", C.c_int32), ("root", BvhDesc),
                ("meshes", C.POINTER(MeshDesc)), ("n_meshes", C.c_int32)]


class Params(C.Structure):  # mcpt_params
    _fields_ = [("nb_bounces", C.c_int32), ("refract_ind", C.c_float), ("date", C.c_float), ("flat_face", C.c_int32),
                ("mode", C.c_int32), ("nb_prims", C.c_int32), ("nb_emissives", C.c_int32)]


class Stats(C.Structure):
    _fields_ = [("paths", C.c_uint64), ("rays", C.c_uint64), ("segments", C.c_uint64), ("launches", C.c_uint64),
                ("boxes", C.c_uint64), ("prim_tests", C.c_uint64), ("tri_tests", C.c_uint64)]


class TreeStats(C.Structure):  # mcpt_tree_stats
    _fields_ = [("nodes", C.c_int32), ("leaves", C.c_int32), ("prims_reached", C.c_int32), ("prims_expected", C.c_int32),
                ("depth", C.c_int32), ("depth4", C.c_int32), ("leaves4", C.c_int32), ("violations", C.c_int32),
                ("sah_cost", C.c_double)]


class StageTimes(C.Structure):
    _fields_ = [("ms", C.c_double * 8), ("launches", C.c_uint64 * 8)]


TRACE_STATIC_MAX_RAYS = 65536   # MCPT_TRACE_STATIC_MAX_RAYS (include/mcpt.h)

HIT_DTYPE = np.dtype([("index", "<i4"), ("tri_index", "<i4"), ("shape", "<i4"), ("dir", "<i4"), ("dist", "<f4"),
                      ("pl", "<f4", 3), ("pg", "<f4", 3)])
assert HIT_DTYPE.itemsize == 44

# every symbol include/mcpt.h declares
MCPT_SYMBOLS = [
    "mcpt_abi_version", "mcpt_last_error", "mcpt_error_string", "mcpt_create", "mcpt_destroy", "mcpt_set_stream",
    "mcpt_synchronize", "mcpt_set_option", "mcpt_upload_scene", "mcpt_resize", "mcpt_set_camera", "mcpt_set_params",
    "mcpt_clear", "mcpt_render", "mcpt_resolve", "mcpt_get_accum", "mcpt_bind_accum", "mcpt_read_accum",
    "mcpt_write_accum", "mcpt_reduce_nccl", "mcpt_trace_rays", "mcpt_trace_rays_device", "mcpt_get_stats",
    "mcpt_reset_stats", "mcpt_get_stage_times", "mcpt_inspect_walk_tree", "mcpt_trace_rays_static", "mcpt_sample_directions",
]
# every symbol include/mcpt_host.h declares
MCPTH_SYMBOLS = [
    "mcpth_scene_new", "mcpth_scene_free", "mcpth_scene_clear", "mcpth_add_prim", "mcpth_add_mesh", "mcpth_place_mesh",
    "mcpth_build_repo_scene", "mcpth_build_scene6_roughness", "mcpth_build_divergence_stress",
    "mcpth_build_synthetic_meshes", "mcpth_finalize", "mcpth_desc", "mcpth_upload", "mcpth_nb_prim",
    "mcpth_nb_emissives", "mcpth_depth", "mcpth_prim_data", "mcpth_nb_meshes", "mcpth_default_camera", "mcpth_transfo",
]


class McptError(RuntimeError):
    def __init__(self, code, msg=""):
        super().__init__(f"mcpt error {code}: {msg}")
        self.code = code


_lib = None
_host = None


def lib():
    """libmcpt.so; raises if the CUDA extension has not been built (no fallback)."""
    global _lib
    if _lib is None:
        path = Path(os.environ.get("MCPT_LIBRARY", LIB_DIR / "libmcpt.so"))  # override: tuning builds only
        if not path.exists():
            raise ImportError(f"{path} is missing: run __graft_entry__.build(); there is no CPU fallback")
        L = C.CDLL(str(path), mode=C.RTLD_GLOBAL)
        L.mcpt_last_error.restype = C.c_char_p
        L.mcpt_last_error.argtypes = [C.c_void_p]
        L.mcpt_error_string.restype = C.c_char_p
        L.mcpt_create.argtypes = [C.c_int, C.POINTER(C.c_void_p)]
        L.mcpt_destroy.argtypes = [C.c_void_p]
        L.mcpt_set_stream.argtypes = [C.c_void_p, C.c_void_p]
        L.mcpt_synchronize.argtypes = [C.c_void_p]
        L.mcpt_set_option.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.mcpt_upload_scene.argtypes = [C.c_void_p, C.POINTER(SceneDesc)]
        L.mcpt_resize.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.mcpt_set_camera.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.mcpt_set_params.argtypes = [C.c_void_p, C.POINTER(Params)]
        L.mcpt_clear.argtypes = [C.c_void_p]
        L.mcpt_render.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.mcpt_resolve.argtypes = [C.c_void_p, C.c_float, C.c_int, C.c_void_p, C.c_int]
        L.mcpt_get_accum.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_size_t)]
        L.mcpt_bind_accum.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
        L.mcpt_read_accum.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
        L.mcpt_write_accum.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
        L.mcpt_reduce_nccl.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
        L.mcpt_trace_rays.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        L.mcpt_trace_rays_device.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        L.mcpt_trace_rays_static.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]
        L.mcpt_sample_directions.argtypes = [C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_float, C.c_int, C.c_float, C.c_int64, C.c_void_p]
        L.mcpt_get_stats.argtypes = [C.c_void_p, C.POINTER(Stats)]
        L.mcpt_reset_stats.argtypes = [C.c_void_p]
        L.mcpt_get_stage_times.argtypes = [C.c_void_p, C.POINTER(StageTimes)]
        L.mcpt_inspect_walk_tree.argtypes = [C.POINTER(SceneDesc), C.c_int, C.POINTER(TreeStats)]
        _lib = L
    return _lib


def host():
    """libmcpt_host.so (pure CPU; links libmcpt.so for mcpth_upload)."""
    global _host
    if _host is None:
        lib()
        path = LIB_DIR / "libmcpt_host.so"
        if not path.exists():
            raise ImportError(f"{path} is missing: run __graft_entry__.build()")
        H = C.CDLL(str(path))
        H.mcpth_scene_new.restype = C.c_void_p
        H.mcpth_scene_free.argtypes = [C.c_void_p]
        H.mcpth_scene_clear.argtypes = [C.c_void_p]
        fp = C.c_void_p
        H.mcpth_add_prim.argtypes = [C.c_void_p, C.c_int, fp, fp, C.c_float, C.c_float, C.c_float]
        H.mcpth_add_mesh.argtypes = [C.c_void_p, fp, C.c_int, fp, fp, C.c_int]
        H.mcpth_place_mesh.argtypes = [C.c_void_p, C.c_int, fp, fp, C.c_float, C.c_float, C.c_float]
        H.mcpth_build_repo_scene.argtypes = [C.c_void_p, C.c_int, C.c_float]
        H.mcpth_build_scene6_roughness.argtypes = [C.c_void_p, C.c_float, C.c_float]
        H.mcpth_build_divergence_stress.argtypes = [C.c_void_p, C.c_float]
        H.mcpth_build_synthetic_meshes.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_uint32]
        H.mcpth_finalize.argtypes = [C.c_void_p]
        H.mcpth_desc.argtypes = [C.c_void_p, C.POINTER(SceneDesc)]
        H.mcpth_upload.argtypes = [C.c_void_p, C.c_void_p]
        for f in ("mcpth_nb_prim", "mcpth_nb_emissives", "mcpth_nb_meshes"):
            getattr(H, f).argtypes = [C.c_void_p]
        H.mcpth_depth.argtypes = [C.c_void_p, C.c_int]
        H.mcpth_prim_data.argtypes = [C.c_void_p]
        H.mcpth_prim_data.restype = C.POINTER(PrimData)
        H.mcpth_default_camera.argtypes = [C.c_int, C.c_int, fp, fp]
        H.mcpth_transfo.argtypes = [C.c_char_p, fp, fp]
        _host = H
    return _host


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def transfo(ops: str, *args) -> np.ndarray:
    """Product of Transfo::translate/scale/rotate matrices, e.g. transfo('tZs', 70,20,-40, 20, 20,20,60)."""
    a = _f32(args)
    out = np.zeros(16, np.float32)
    host().mcpth_transfo(ops.encode(), a.ctypes.data, out.ctypes.data)
    return out


def default_camera(width: int, height: int):
    iv, ipv = np.zeros(16, np.float32), np.zeros(16, np.float32)
    host().mcpth_default_camera(width, height, iv.ctypes.data, ipv.ctypes.data)
    return iv, ipv
