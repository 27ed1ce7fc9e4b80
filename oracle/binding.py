"""TEST INFRASTRUCTURE — ctypes binding of oracle/liboracle.so (the CPU oracle).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module. The product package never does.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
HIT_DTYPE = np.dtype([("index", "<i4"), ("tri_index", "<i4"), ("shape", "<i4"), ("dir", "<i4"), ("dist", "<f4"),
                      ("pl", "<f4", 3), ("pg", "<f4", 3)])

_libs = {}
REF_LIB = HERE / "_ref" / "libref.so"
INTEGRATION_LIB = HERE / "_ref" / "libref_integration.so"   # the same + the INTEGRATION.md patch compiled in (oracle/build_integration.py)


def have_reference():
    """True when oracle/_ref/libref.so (the reference's own sources compiled as C++) is present."""
    return REF_LIB.exists()


def lib(kind="port"):
    """kind 'port': liboracle.so, the hand restatement. kind 'reference': oracle/_ref/libref.so, the
    reference's own .cpp / shader text compiled by oracle/build_ref.py. Same entry points."""
    if kind not in _libs:
        path = REF_LIB if kind == "reference" else INTEGRATION_LIB if kind == "integration" else HERE / "liboracle.so"
        if not path.exists():
            raise ImportError(f"{path} missing: run __graft_entry__.build()")
        L = C.CDLL(str(path))
        vp, fp = C.c_void_p, C.c_void_p
        L.oracle_new.restype = vp
        L.oracle_free.argtypes = [vp]
        L.oracle_clear.argtypes = [vp]
        L.oracle_add_prim.argtypes = [vp, C.c_int, fp, fp, C.c_float, C.c_float, C.c_float]
        L.oracle_add_mesh.argtypes = [vp, fp, C.c_int, fp, fp, C.c_int]
        L.oracle_place_mesh.argtypes = [vp, C.c_int, fp, fp, C.c_float, C.c_float, C.c_float]
        L.oracle_poke_prim.argtypes = [vp, C.c_int, C.c_int, C.c_float]
        L.oracle_finalize.argtypes = [vp]
        for f in ("oracle_n_prims", "oracle_nb_emissives", "oracle_n_meshes"):
            getattr(L, f).argtypes = [vp]
        L.oracle_prim_data.argtypes = [vp]
        L.oracle_prim_data.restype = vp
        L.oracle_bvh_depth.argtypes = [vp, C.c_int]
        L.oracle_bvh_bb.argtypes = [vp, C.c_int]
        L.oracle_bvh_bb.restype = vp
        L.oracle_bvh_ind.argtypes = [vp, C.c_int]
        L.oracle_bvh_ind.restype = vp
        for f in ("oracle_mesh_line", "oracle_mesh_n_tris", "oracle_mesh_n_verts"):
            getattr(L, f).argtypes = [vp, C.c_int]
        for f in ("oracle_mesh_tri", "oracle_mesh_pos", "oracle_mesh_nrm"):
            getattr(L, f).argtypes = [vp, C.c_int]
            getattr(L, f).restype = vp
        L.oracle_tex_height.argtypes = [vp, C.c_int]
        L.oracle_tex_comps.argtypes = [vp, C.c_int]
        L.oracle_tex_data.argtypes = [vp, C.c_int]
        L.oracle_tex_data.restype = vp
        L.oracle_set_uniforms.argtypes = [vp, C.c_int, C.c_float, C.c_float, C.c_int, fp, fp]
        L.oracle_set_threads.argtypes = [C.c_int]
        if hasattr(L, "oracle_set_ulp_perturb"):
            L.oracle_set_ulp_perturb.argtypes = [C.c_int]
        L.oracle_trace.argtypes = [vp, C.c_int64, fp, fp, vp, C.c_int, vp]
        L.oracle_render.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, fp, C.c_int, C.c_int, vp]
        L.oracle_rng_stream.argtypes = [C.c_float, C.c_float, C.c_int, C.c_float, C.c_int, fp, vp]
        L.oracle_random_rays.argtypes = [fp, C.c_float, C.c_float, C.c_float, C.c_int, fp]
        L.oracle_primary_ray.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int, fp, fp]
        if kind == "integration":
            L.integration_set_ctx.argtypes = [vp]
        if kind in ("reference", "integration"):
            L.oracle_build_repo_scene.argtypes = [vp, C.c_int, C.c_float]
            L.oracle_default_camera.argtypes = [C.c_int, C.c_int, fp, fp]
            L.oracle_transfo.argtypes = [C.c_char_p, fp, fp]
            L.oracle_tex_width.argtypes = [vp, C.c_int]
        _libs[kind] = L
    return _libs[kind]


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


class _quiet_stdout:
    """The reference's BVH builder prints the root box to std::cout (bvh_gpu/bvh.cpp:92) and add_mesh
    the mesh box (gpu_bvh_scene.cpp:53); callers such as bench.py own stdout (one JSON line)."""

    def __enter__(self):
        import os
        import sys
        sys.stdout.flush()
        self._saved = os.dup(1)
        self._null = os.open(os.devnull, os.O_WRONLY)
        os.dup2(self._null, 1)

    def __exit__(self, *exc):
        import os
        C.CDLL(None).fflush(None)
        os.dup2(self._saved, 1)
        os.close(self._saved)
        os.close(self._null)


class Oracle:
    """CPU restatement of the reference scene tables + shaders."""

    def __init__(self, kind="port"):
        self.kind = kind
        self.L = lib(kind)
        self.h = C.c_void_p(self.L.oracle_new())
        self._keep = []

    # ---- only on the 'reference' library: the reference's own scene builders / camera / Transfo
    def build_repo_scene(self, scene_id, light_intensity=1.2):
        assert self.kind in ("reference", "integration")
        rc = self.L.oracle_build_repo_scene(self.h, int(scene_id), light_intensity)
        assert rc == 0, scene_id
        return self

    def default_camera(self, w, h):
        assert self.kind in ("reference", "integration")
        iv, ipv = np.zeros(16, np.float32), np.zeros(16, np.float32)
        self.L.oracle_default_camera(w, h, iv.ctypes.data, ipv.ctypes.data)
        return iv, ipv

    def transfo(self, ops, *args):
        assert self.kind in ("reference", "integration")
        a, out = _f32(args), np.zeros(16, np.float32)
        self.L.oracle_transfo(ops.encode(), a.ctypes.data, out.ctypes.data)
        return out

    def __del__(self):
        try:
            if self.h:
                self.L.oracle_free(self.h)
                self.h = None
        except Exception:
            pass

    def clear(self):
        self.L.oracle_clear(self.h)

    def add_prim(self, kind, trf, color, shininess=0.0, roughness=0.0, emissivity=0.0):
        t, c = _f32(trf).ravel(), _f32(color)
        return self.L.oracle_add_prim(self.h, int(kind), t.ctypes.data, c.ctypes.data, shininess, roughness, emissivity)

    def add_mesh(self, pos, tri, nrm):
        p, n = _f32(pos).reshape(-1, 3), _f32(nrm).reshape(-1, 3)
        t = np.ascontiguousarray(tri, dtype=np.uint32).reshape(-1, 3)
        if self.kind in ("reference", "integration"):
            with _quiet_stdout():
                return self.L.oracle_add_mesh(self.h, p.ctypes.data, len(p), n.ctypes.data, t.ctypes.data, len(t))
        return self.L.oracle_add_mesh(self.h, p.ctypes.data, len(p), n.ctypes.data, t.ctypes.data, len(t))

    def place_mesh(self, mesh_id, trf, color, shininess=0.0, roughness=0.0, emissivity=0.0):
        t, c = _f32(trf).ravel(), _f32(color)
        return self.L.oracle_place_mesh(self.h, mesh_id, t.ctypes.data, c.ctypes.data, shininess, roughness, emissivity)

    def poke_prim(self, prim, word, value):
        self.L.oracle_poke_prim(self.h, prim, word, value)

    def load_unsorted(self, prims64, meshes=()):
        """Replay a pre-finalize scene: prims64 = (n,64) PrimData rows in insertion order,
        meshes = [{'pos','tri','nrm','mesh_line'}] from the host mirror."""
        self.clear()
        line_to_id = {}
        for m in meshes:
            line_to_id[int(m["mesh_line"])] = self.add_mesh(m["pos"], m["tri"], m["nrm"])
        for r in np.asarray(prims64, np.float32).reshape(-1, 64):
            kind = int(r[48])
            if kind == 0:
                i = self.place_mesh(line_to_id[int(r[49])], r[32:48], r[52:56], r[56], r[57], r[58])
            else:
                i = self.add_prim(kind, r[0:16], r[52:56], r[56], r[57], r[58])
            # keep the inverse the table's author computed (trf.inverse() is Eigen's in the reference;
            # every stand-in rounds it a little differently), so both sides trace the SAME table
            for w in range(16, 32):
                self.poke_prim(i, w, float(r[w]))
        return self

    def finalize(self):
        if self.kind in ("reference", "integration"):
            with _quiet_stdout():
                self.L.oracle_finalize(self.h)
        else:
            self.L.oracle_finalize(self.h)
        return self

    @property
    def n_prims(self):
        return self.L.oracle_n_prims(self.h)

    @property
    def nb_emissives(self):
        return self.L.oracle_nb_emissives(self.h)

    @property
    def n_meshes(self):
        return self.L.oracle_n_meshes(self.h)

    def prim_data(self):
        n = self.n_prims
        if n == 0:
            return np.zeros((0, 64), np.float32)
        p = C.cast(self.L.oracle_prim_data(self.h), C.POINTER(C.c_float))
        return np.ctypeslib.as_array(p, shape=(n, 64)).copy()

    def bvh(self, which=-1):
        depth = self.L.oracle_bvh_depth(self.h, which)
        nl = 1 << depth
        bb = np.ctypeslib.as_array(C.cast(self.L.oracle_bvh_bb(self.h, which), C.POINTER(C.c_float)), shape=(2 * nl - 1, 6)).copy()
        ind = np.ctypeslib.as_array(C.cast(self.L.oracle_bvh_ind(self.h, which), C.POINTER(C.c_int32)), shape=(nl,)).copy()
        return {"depth": depth, "bb": bb, "ind": ind}

    def texture(self, k):
        """k: 0 prim, 1 bb, 2 ind, 3 tri, 4 pos, 5 nrm -> (height, 1024, comps) array."""
        hgt, comps = self.L.oracle_tex_height(self.h, k), self.L.oracle_tex_comps(self.h, k)
        wid = self.L.oracle_tex_width(self.h, k) if self.kind in ("reference", "integration") else 1024
        ctype = C.c_int32 if k in (2, 3) else C.c_float
        if hgt == 0:
            return np.zeros((0, 1024, comps), np.int32 if k in (2, 3) else np.float32)
        p = C.cast(self.L.oracle_tex_data(self.h, k), C.POINTER(ctype))
        # the reference allocates its primitive texture 1204 wide (scene.h:77) and fills 1024 columns
        return np.ctypeslib.as_array(p, shape=(hgt, wid, comps))[:, :1024].copy()

    def scene_desc(self):
        """The oracle's tables as an include/mcpt_scene.h descriptor (borrowed pointers)."""
        from montecarlo_pathtracing_b200.capi import BvhDesc, MeshDesc, SceneDesc
        L, h = self.L, self.h
        nm = self.n_meshes
        arr = (MeshDesc * max(nm, 1))()
        for m in range(nm):
            arr[m] = MeshDesc(BvhDesc(L.oracle_bvh_bb(h, m), L.oracle_bvh_ind(h, m), L.oracle_bvh_depth(h, m)),
                              L.oracle_mesh_tri(h, m), L.oracle_mesh_pos(h, m), L.oracle_mesh_nrm(h, m),
                              L.oracle_mesh_n_tris(h, m), L.oracle_mesh_n_verts(h, m), L.oracle_mesh_line(h, m))
        self._keep.append(arr)
        return SceneDesc(L.oracle_prim_data(h), self.n_prims, self.nb_emissives,
                         BvhDesc(L.oracle_bvh_bb(h, -1), L.oracle_bvh_ind(h, -1), L.oracle_bvh_depth(h, -1)),
                         C.cast(arr, C.POINTER(MeshDesc)), nm)

    def set_uniforms(self, nb_bounces=3, refract_ind=1.0, date=0.0, flat_face=0, inv_v=None, inv_pv=None):
        a = None if inv_v is None else _f32(inv_v).ravel()
        b = None if inv_pv is None else _f32(inv_pv).ravel()
        self.L.oracle_set_uniforms(self.h, nb_bounces, refract_ind, date, flat_face,
                                   None if a is None else a.ctypes.data, None if b is None else b.ctypes.data)

    def set_threads(self, n=0):
        return self.L.oracle_set_threads(n)

    def set_ulp_perturb(self, n=0):
        """Port oracle only: shift every sin / cos / log / pow result of the bounce sampler by n ulps."""
        self.L.oracle_set_ulp_perturb(int(n))

    def trace(self, o, d, any_hit=False):
        o, d = _f32(o).reshape(-1, 3), _f32(d).reshape(-1, 3)
        out = np.zeros(len(o), HIT_DTYPE)
        cnt = np.zeros(5, np.uint64)
        self.L.oracle_trace(self.h, len(o), o.ctypes.data, d.ctypes.data, out.ctypes.data, int(any_hit), cnt.ctypes.data)
        self.last_counters = dict(zip(("traversals", "box_tests", "prim_tests", "tri_tests", "segments"), map(int, cnt)))
        return out

    def render(self, w, h, first_pass, n_passes, mode=0, accum=None, y0=-1, y1=-1):
        """accum (h,w,3) float32 += passes first_pass..first_pass+n_passes-1 (blend order = pass order)."""
        if accum is None:
            accum = np.zeros((h, w, 3), np.float32)
        assert accum.dtype == np.float32 and accum.flags.c_contiguous and accum.shape == (h, w, 3)
        cnt = np.zeros(5, np.uint64)
        self.L.oracle_render(self.h, w, h, first_pass, n_passes, mode, accum.ctypes.data, y0, y1, cnt.ctypes.data)
        self.last_counters = dict(zip(("traversals", "box_tests", "prim_tests", "tri_tests", "segments"), map(int, cnt)))
        return accum

    def rng_stream(self, tcx, tcy, numero_pass, date, n):
        out, seed = np.zeros(n, np.float32), np.zeros(3, np.uint32)
        self.L.oracle_rng_stream(tcx, tcy, numero_pass, date, n, out.ctypes.data, seed.ctypes.data)
        return out, seed

    def random_rays(self, D, roughness, tcx=0.3, tcy=0.7, n=1000):
        d, out = _f32(D), np.zeros((n, 3), np.float32)
        self.L.oracle_random_rays(d.ctypes.data, roughness, tcx, tcy, n, out.ctypes.data)
        return out

    def primary_ray(self, x, y, w, h):
        o, d = np.zeros(3, np.float32), np.zeros(3, np.float32)
        self.L.oracle_primary_ray(self.h, x, y, w, h, o.ctypes.data, d.ctypes.data)
        return o, d
