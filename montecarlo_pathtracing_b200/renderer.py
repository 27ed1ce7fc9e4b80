"""Thin Python objects over the C ABI: Scene (host mirror) and Context (CUDA path tracer).

These add no behaviour of their own: every method is one call into
include/mcpt_host.h / include/mcpt.h, so tests written against them exercise the
drop-in boundary itself.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import capi
from .capi import McptError, Params, SceneDesc, StageTimes, Stats


class Scene:
    """ScenePrimitives + BVH_GPU_Scene of the reference (bvh_gpu/gpu_bvh_scene.h:35-121)."""

    def __init__(self):
        self._h = capi.host()
        self._s = C.c_void_p(self._h.mcpth_scene_new())
        if not self._s:
            raise MemoryError("mcpth_scene_new")

    def __del__(self):
        try:
            if self._s:
                self._h.mcpth_scene_free(self._s)
                self._s = None
        except Exception:
            pass

    def clear(self):
        self._h.mcpth_scene_clear(self._s)

    def add_prim(self, kind, trf, color, shininess=0.0, roughness=0.0, emissivity=0.0):
        t, c = capi._f32(trf).ravel(), capi._f32(color)
        r = self._h.mcpth_add_prim(self._s, int(kind), t.ctypes.data, c.ctypes.data, shininess, roughness, emissivity)
        if r < 0:
            raise McptError(capi.MCPT_E_INVALID, f"add_prim(kind={kind})")
        return r

    def add_sphere(self, trf, color, **kw):
        return self.add_prim(capi.PRIM_SPHERE, trf, color, **kw)

    def add_cube(self, trf, color, **kw):
        return self.add_prim(capi.PRIM_CUBE, trf, color, **kw)

    def add_cylinder(self, trf, color, **kw):
        return self.add_prim(capi.PRIM_CYLINDER, trf, color, **kw)

    def add_cone(self, trf, color, **kw):
        return self.add_prim(capi.PRIM_CONE, trf, color, **kw)

    def add_orientedQuad(self, trf, color, **kw):
        return self.add_prim(capi.PRIM_ORIENTED_QUAD, trf, color, **kw)

    def add_mesh(self, pos, tri, nrm=None):
        p = capi._f32(pos).reshape(-1, 3)
        t = np.ascontiguousarray(tri, dtype=np.uint32).reshape(-1, 3)
        n = None if nrm is None else capi._f32(nrm).reshape(-1, 3)
        r = self._h.mcpth_add_mesh(self._s, p.ctypes.data, len(p), None if n is None else n.ctypes.data, t.ctypes.data, len(t))
        if r < 0:
            raise McptError(capi.MCPT_E_INVALID, "add_mesh")
        return r

    def place_mesh(self, mesh_id, trf, color, shininess=0.0, roughness=0.0, emissivity=0.0):
        t, c = capi._f32(trf).ravel(), capi._f32(color)
        r = self._h.mcpth_place_mesh(self._s, mesh_id, t.ctypes.data, c.ctypes.data, shininess, roughness, emissivity)
        if r < 0:
            raise McptError(capi.MCPT_E_INVALID, "place_mesh")
        return r

    def _ck(self, rc, what):
        if rc != 0:
            raise McptError(rc, what)

    def build_repo_scene(self, scene_id, light_intensity=1.2):
        self._ck(self._h.mcpth_build_repo_scene(self._s, scene_id, light_intensity), f"scene {scene_id}")
        return self

    def build_scene6_roughness(self, light_intensity, roughness):
        self._ck(self._h.mcpth_build_scene6_roughness(self._s, light_intensity, roughness), "scene 6")
        return self

    def build_divergence_stress(self, light_intensity):
        self._ck(self._h.mcpth_build_divergence_stress(self._s, light_intensity), "divergence stress")
        return self

    def build_synthetic_meshes(self, n_meshes, tris_log2, seed=12345):
        self._ck(self._h.mcpth_build_synthetic_meshes(self._s, n_meshes, tris_log2, seed), "synthetic meshes")
        return self

    def finalize(self):
        self._ck(self._h.mcpth_finalize(self._s), "finalize")
        return self

    def desc(self, mesh_trees=True) -> SceneDesc:
        """The include/mcpt_scene.h descriptor of the finalized scene. mesh_trees=False: the mesh descriptors carry no
        BVH (bb = ind = NULL) -- what a host that never ran BVH_KDtree::compute on its meshes would upload; the
        library then builds every mesh tree on the GPU and ranks equal-distance triangles by index."""
        d = SceneDesc()
        self._ck(self._h.mcpth_desc(self._s, C.byref(d)), "desc (finalize first)")
        if not mesh_trees and d.n_meshes > 0:
            from .capi import BvhDesc, MeshDesc
            arr = (MeshDesc * d.n_meshes)()
            for i in range(d.n_meshes):
                m = d.meshes[i]
                arr[i] = MeshDesc(BvhDesc(None, None, 0), m.tri_idx, m.pos, m.nrm, m.n_tris, m.n_verts, m.mesh_line)
            d.meshes = C.cast(arr, C.POINTER(MeshDesc))
            d._keep = arr
        return d

    @property
    def nb_prim(self):
        return self._h.mcpth_nb_prim(self._s)

    @property
    def nb_emissives(self):
        return self._h.mcpth_nb_emissives(self._s)

    @property
    def nb_meshes(self):
        return self._h.mcpth_nb_meshes(self._s)

    def depth(self, i=0):
        return self._h.mcpth_depth(self._s, i)

    def prim_data(self) -> np.ndarray:
        """(n, 64) float32 copy of the PrimData table."""
        n = self.nb_prim
        p = self._h.mcpth_prim_data(self._s)
        if n == 0:
            return np.zeros((0, 64), np.float32)
        return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_float)), shape=(n, 64)).copy()

    def tables(self):
        """Numpy copies of everything finalize() produced (root tree, meshes)."""
        d = self.desc()

        def bvh(b):
            nl = 1 << b.depth
            bb = np.ctypeslib.as_array(C.cast(b.bb, C.POINTER(C.c_float)), shape=(2 * nl - 1, 6)).copy()
            ind = np.ctypeslib.as_array(C.cast(b.ind, C.POINTER(C.c_int32)), shape=(nl,)).copy()
            return {"depth": b.depth, "bb": bb, "ind": ind}

        meshes = []
        for m in range(d.n_meshes):
            md = d.meshes[m]
            meshes.append({
                "bvh": bvh(md.bvh), "mesh_line": md.mesh_line,
                "tri": np.ctypeslib.as_array(C.cast(md.tri_idx, C.POINTER(C.c_uint32)), shape=(md.n_tris, 3)).copy(),
                "pos": np.ctypeslib.as_array(C.cast(md.pos, C.POINTER(C.c_float)), shape=(md.n_verts, 3)).copy(),
                "nrm": np.ctypeslib.as_array(C.cast(md.nrm, C.POINTER(C.c_float)), shape=(md.n_verts, 3)).copy(),
            })
        return {"prims": self.prim_data(), "nb_emissives": d.nb_emissives, "root": bvh(d.root), "meshes": meshes}


class Context:
    """One CUDA path-tracing context (include/mcpt.h). Fails loudly without a GPU."""

    def __init__(self, device=0):
        self._L = capi.lib()
        self._c = C.c_void_p()
        rc = self._L.mcpt_create(device, C.byref(self._c))
        if rc != 0:
            self._c = None
            raise McptError(rc, self._L.mcpt_error_string(rc).decode())
        self.width = self.height = 0

    def close(self):
        if getattr(self, "_c", None):
            self._L.mcpt_destroy(self._c)
            self._c = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc):
        if rc != 0:
            raise McptError(rc, self._L.mcpt_last_error(self._c).decode())

    def set_stream(self, cuda_stream_ptr):
        self._ck(self._L.mcpt_set_stream(self._c, C.c_void_p(cuda_stream_ptr)))

    def synchronize(self):
        self._ck(self._L.mcpt_synchronize(self._c))

    def set_option(self, opt, value):
        self._ck(self._L.mcpt_set_option(self._c, opt, value))

    def upload_scene(self, scene_or_desc):
        d = scene_or_desc.desc() if hasattr(scene_or_desc, "desc") else scene_or_desc
        self._scene_keepalive = scene_or_desc
        self._ck(self._L.mcpt_upload_scene(self._c, C.byref(d)))

    def resize(self, w, h):
        self._ck(self._L.mcpt_resize(self._c, w, h))
        self.width, self.height = w, h

    def set_camera(self, inv_v, inv_pv):
        a, b = capi._f32(inv_v).ravel(), capi._f32(inv_pv).ravel()
        self._ck(self._L.mcpt_set_camera(self._c, a.ctypes.data, b.ctypes.data))

    def set_params(self, nb_bounces=3, refract_ind=1.0, date=0.0, flat_face=0, mode=0, nb_prims=0, nb_emissives=0):
        p = Params(nb_bounces, refract_ind, date, flat_face, mode, nb_prims, nb_emissives)
        self._ck(self._L.mcpt_set_params(self._c, C.byref(p)))

    def clear(self):
        self._ck(self._L.mcpt_clear(self._c))

    def render(self, first_pass, n_passes):
        self._ck(self._L.mcpt_render(self._c, first_pass, n_passes))

    def resolve(self, divisor, subsampling=1, rgba8=False) -> np.ndarray:
        W, H = self.width * subsampling, self.height * subsampling
        out = np.empty((H, W, 4), np.uint8) if rgba8 else np.empty((H, W, 3), np.float32)
        kind = capi.DST_HOST_RGBA8 if rgba8 else capi.DST_HOST_RGB32F
        self._ck(self._L.mcpt_resolve(self._c, float(divisor), subsampling, out.ctypes.data, kind))
        return out

    def resolve_device(self, divisor, dst_ptr, subsampling=1, rgba8=False):
        kind = capi.DST_DEVICE_RGBA8 if rgba8 else capi.DST_DEVICE_RGB32F
        self._ck(self._L.mcpt_resolve(self._c, float(divisor), subsampling, C.c_void_p(dst_ptr), kind))

    def accum_ptr(self):
        p, n = C.c_void_p(), C.c_size_t()
        self._ck(self._L.mcpt_get_accum(self._c, C.byref(p), C.byref(n)))
        return p.value, n.value

    def bind_accum(self, dev_ptr, nbytes):
        self._ck(self._L.mcpt_bind_accum(self._c, C.c_void_p(dev_ptr), nbytes))

    def read_accum(self) -> np.ndarray:
        out = np.empty((self.height, self.width, 4), np.float32)
        self._ck(self._L.mcpt_read_accum(self._c, out.ctypes.data, out.nbytes))
        return out

    def write_accum(self, a):
        a = capi._f32(a)
        self._ck(self._L.mcpt_write_accum(self._c, a.ctypes.data, a.nbytes))

    def reduce_nccl(self, comm_ptr, root=0):
        self._ck(self._L.mcpt_reduce_nccl(self._c, C.c_void_p(comm_ptr), root))

    def trace_rays(self, o, d, any_hit=False) -> np.ndarray:
        o, d = capi._f32(o).reshape(-1, 3), capi._f32(d).reshape(-1, 3)
        assert o.shape == d.shape
        out = np.zeros(len(o), capi.HIT_DTYPE)
        self._ck(self._L.mcpt_trace_rays(self._c, len(o), o.ctypes.data, d.ctypes.data, out.ctypes.data, int(any_hit)))
        return out

    def trace_rays_device(self, n, o4_ptr, d4_ptr, out_ptr, any_hit=False):
        self._ck(self._L.mcpt_trace_rays_device(self._c, n, C.c_void_p(o4_ptr), C.c_void_p(d4_ptr), C.c_void_p(out_ptr), int(any_hit)))

    def trace_rays_static(self, n, o4_ptr, d4_ptr, out_ptr):
        """mcpt_trace_rays_static: device buffers in and out, fixed scratch inside the library, no allocation."""
        self._ck(self._L.mcpt_trace_rays_static(self._c, n, C.c_void_p(o4_ptr), C.c_void_p(d4_ptr), C.c_void_p(out_ptr)))

    def sample_directions(self, D, roughness, n, tcx=0.3, tcy=0.7, numero_pass=0, date=0.25) -> np.ndarray:
        """mcpt_sample_directions: (n, 3) consecutive random_ray(D, roughness) draws of one RNG stream."""
        d = capi._f32(D).reshape(3)
        out = np.zeros((int(n), 3), np.float32)
        self._ck(self._L.mcpt_sample_directions(self._c, d.ctypes.data, float(roughness), float(tcx), float(tcy), int(numero_pass),
                                                float(date), int(n), out.ctypes.data))
        return out

    def stats(self) -> dict:
        s = Stats()
        self._ck(self._L.mcpt_get_stats(self._c, C.byref(s)))
        return {"paths": s.paths, "rays": s.rays, "segments": s.segments, "launches": s.launches,
                "boxes": s.boxes, "prim_tests": s.prim_tests, "tri_tests": s.tri_tests}

    def stage_times(self) -> dict:
        """{stage: (ms, launches)} since the last call (needs OPT_PROFILE_STAGES)."""
        t = StageTimes()
        self._ck(self._L.mcpt_get_stage_times(self._c, C.byref(t)))
        return {n: (t.ms[i], int(t.launches[i])) for i, n in enumerate(capi.STAGE_NAMES)}

    def reset_stats(self):
        self._ck(self._L.mcpt_reset_stats(self._c))
