"""CPU-only checks of the drop-in boundary: the C-ABI libraries load, export every symbol the
headers declare, agree with the ctypes mirrors on struct sizes, and fail loudly without a GPU."""
import ctypes as C
import re
from pathlib import Path

import numpy as np
import pytest

from montecarlo_pathtracing_b200 import Context, McptError, capi

ROOT = Path(__file__).resolve().parent.parent


def _declared(header, prefix):
    txt = (ROOT / "include" / header).read_text()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(" + prefix + r"[a-z0-9_]+)\s*\(", txt)))


def test_every_declared_symbol_is_exported():
    L, H = capi.lib(), capi.host()
    names = _declared("mcpt.h", "mcpt_")
    assert len(names) >= 25
    assert sorted(capi.MCPT_SYMBOLS) == names            # the ctypes list tracks the header
    for n in names:
        assert hasattr(L, n), f"libmcpt.so does not export {n}"
    hnames = _declared("mcpt_host.h", "mcpth_")
    assert sorted(capi.MCPTH_SYMBOLS) == hnames
    for n in hnames:
        assert hasattr(H, n), f"libmcpt_host.so does not export {n}"


def test_no_torch_or_cxx_types_in_the_abi():
    for h in ("mcpt.h", "mcpt_scene.h", "mcpt_host.h"):
        txt = (ROOT / "include" / h).read_text()
        assert 'extern "C"' in txt
        code = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)   # comments may mention torch's stream
        assert "torch" not in code and "at::" not in code and "std::" not in code and "#include <c" not in code


def test_struct_sizes_match_the_reference_layout():
    assert C.sizeof(capi.PrimData) == 256                # PrimData, bvh_gpu/scene.h:64-73 (16 RGBA32F texels)
    assert capi.HIT_DTYPE.itemsize == 44
    assert C.sizeof(capi.Params) == 28
    assert C.sizeof(capi.BvhDesc) == 24
    assert capi.lib().mcpt_abi_version() == 1


def test_error_strings():
    L = capi.lib()
    assert L.mcpt_error_string(0) == b"ok"
    for code in range(-6, 0):
        assert L.mcpt_error_string(code) not in (b"ok", b"unknown error")
    assert L.mcpt_error_string(-99) == b"unknown error"
    assert L.mcpt_last_error(None) == b"null context"


def test_null_arguments_are_rejected_not_dereferenced():
    L = capi.lib()
    assert L.mcpt_create(0, None) == capi.MCPT_E_INVALID
    assert L.mcpt_destroy(None) == capi.MCPT_E_INVALID
    assert L.mcpt_render(None, 0, 1) == capi.MCPT_E_INVALID
    assert L.mcpt_upload_scene(None, None) == capi.MCPT_E_INVALID
    assert L.mcpt_resolve(None, 1.0, 1, None, 0) == capi.MCPT_E_INVALID


def test_no_cpu_fallback_without_a_device():
    """On a box without a GPU the product path must refuse to run, not fall back."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    with pytest.raises(McptError) as e:
        Context(0)
    assert e.value.code == capi.MCPT_E_NODEVICE


def test_product_package_does_not_import_the_oracle():
    pkg = ROOT / "montecarlo_pathtracing_b200"
    for f in list(pkg.rglob("*.py")) + list(pkg.rglob("*.c*")) + list(pkg.rglob("*.h")) + list(pkg.rglob("*.inl")):
        txt = f.read_text(errors="ignore")
        assert "oracle" not in txt.lower() or f.name == "build.py", f"{f} mentions the oracle"


def test_default_camera_matches_the_viewer_maths():
    """mcpth_default_camera: view = T(0,0,-focal) * rotateX(-80), fov 0.78, radius 145
    (montecarlo.cpp:388-389,404-405; camera.cpp:28-70), evaluated independently in float64."""
    w, h = 1920, 1080
    iv, ipv = capi.default_camera(w, h)
    iv, ipv = iv.reshape(4, 4).T.astype(np.float64), ipv.reshape(4, 4).T.astype(np.float64)
    fov, r = 0.78, 145.0
    focal = r / np.tan(fov / 2)
    a = np.deg2rad(-80.0)
    R = np.array([[1, 0, 0, 0], [0, np.cos(a), -np.sin(a), 0], [0, np.sin(a), np.cos(a), 0], [0, 0, 0, 1]])
    T = np.eye(4)
    T[2, 3] = -focal
    V = T @ R
    f = 1 / np.tan(fov / 2)
    asp = w / h
    zn, zf = max(0.01, focal - r), focal + r
    P = np.zeros((4, 4))
    P[0, 0], P[1, 1] = f / asp, f
    P[2, 2], P[2, 3], P[3, 2] = (zn + zf) / (zn - zf), 2 * zn * zf / (zn - zf), -1
    assert np.allclose(iv, np.linalg.inv(V), rtol=1e-5, atol=1e-4)
    assert np.allclose(ipv, np.linalg.inv(P @ V), rtol=1e-4, atol=1e-3)


def test_python_constants_mirror_the_header_enums():
    """capi.OPT_* / STAGE_NAMES are hand-written mirrors of the enums in include/mcpt.h."""
    txt = (ROOT / "include" / "mcpt.h").read_text()
    opts = dict(re.findall(r"MCPT_OPT_(\w+)\s*=\s*(\d+)", txt))
    assert len(opts) >= 10
    for name, value in opts.items():
        assert getattr(capi, "OPT_" + name) == int(value), name
    stages = {int(v): n.lower() for n, v in re.findall(r"MCPT_STAGE_(\w+)\s*=\s*(\d+)", txt) if n != "COUNT"}
    assert tuple(stages[i] for i in range(len(stages))) == tuple(capi.STAGE_NAMES)


@pytest.mark.parametrize("kind,kw", [(1, {}), (3, {}), (5, {}), (7, {}), (8, {}), ("c5", {}), ("mesh", {"n_meshes": 5, "tris_log2": 6})])
def test_walk_tree_invariants_on_the_host(kind, kw):
    """The top-level tree the persistent warps walk (full-sweep SAH, loose primitives as slab boxes, the
    four-children fold) is host code: mcpt_inspect_walk_tree builds it as mcpt_upload_scene does -- no CUDA
    device -- and reports the invariants the traversal relies on."""
    import ctypes as C
    from common import host_scene
    s = host_scene(kind, 0.4, **kw)
    s.finalize()
    d = s.desc()
    L = capi.lib()
    st = {}
    for top_sah in (1, 0):
        t = capi.TreeStats()
        assert L.mcpt_inspect_walk_tree(C.byref(d), top_sah, C.byref(t)) == 0
        st[top_sah] = t
        assert t.violations == 0                                   # boxes contain what is below them; slabs cover their primitive
        assert t.prims_reached == t.prims_expected == s.nb_prim    # every primitive owns a leaf
        assert t.leaves4 == t.leaves >= t.prims_reached            # the fold reaches the same leaves
        assert t.depth <= 30                                       # one stack entry per level (kWalkStack)
        assert t.depth4 <= (t.depth + 1) // 2 + 1
    sah, ref = st[1], st[0]
    assert sah.nodes == sah.leaves - 1                             # one primitive (or slab) per leaf, two children per node
    assert sah.sah_cost <= ref.sah_cost * 1.001                    # never worse than the reference's median split
    if kind == 8:
        assert sah.leaves > sah.prims_reached                      # the rotated beams entered as slab boxes
        assert sah.sah_cost < 0.7 * ref.sah_cost
    assert L.mcpt_inspect_walk_tree(None, 1, C.byref(capi.TreeStats())) == capi.MCPT_E_INVALID
