"""The integration patch of INTEGRATION.md section 2, COMPILED into the reference's own host code and run.

oracle/build_integration.py inserts oracle/integration_patch.inc (the documented lines) at the end of
BVH_GPU_Scene::finalize() in a generated copy of bvh_gpu/gpu_bvh_scene.cpp and links the reference's sources with
libmcpt.so. Here a scene is built through the reference's BVH_GPU_Scene API (its own scene builders, its own mesh
path); finalize() then uploads it into an mcpt context through the patch; the context must return the records of the
reference-generated fixtures bit for bit. No graphics is involved. Skipped when the prebuilt library did not travel."""
import numpy as np
import pytest

from montecarlo_pathtracing_b200 import Context
from oracle.binding import INTEGRATION_LIB, Oracle

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not INTEGRATION_LIB.exists(), reason="oracle/_ref/libref_integration.so not built")]
GOLD = __import__("pathlib").Path(__file__).resolve().parent / "golden"
LIGHT = 0.4


@pytest.fixture()
def patched():
    ctx = Context(0)
    r = Oracle("integration")
    r.L.integration_set_ctx(ctx._c)
    yield ctx, r
    r.L.integration_set_ctx(None)
    ctx.close()


@pytest.mark.parametrize("sid", [6, 3, 8])
def test_repo_scene_through_the_compiled_patch(patched, sid):
    ctx, r = patched
    g = np.load(GOLD / f"scene{sid}.npz")
    r.build_repo_scene(sid, LIGHT)
    r.finalize()                      # the reference's finalize() + the patch: mcpt_upload_scene(ctx_, &d)
    got = ctx.trace_rays(g["ray_o"], g["ray_d"])
    assert got.tobytes() == g["hits"].tobytes()


def test_mesh_scene_through_the_compiled_patch(patched):
    """The mesh half of the patch: mesh_data_ / bvh_data_ indexing, Eigen vertex arrays as float*, tri_begins_ as
    mesh_line. The scene of tests/golden/mesh.npz is rebuilt from the fixture's inputs through the reference API."""
    ctx, r = patched
    g = np.load(GOLD / "mesh.npz")
    ids = {int(g[f"m{m}_line"]): r.add_mesh(g[f"m{m}_pos"], g[f"m{m}_tri"], g[f"m{m}_nrm"]) for m in range(2)}
    raw = g["raw_prims"]
    for row in raw:
        if int(row[48]) == 0:
            r.place_mesh(ids[int(row[49])], row[32:48], row[52:56], row[56], row[57], row[58])
        else:
            r.add_prim(int(row[48]), row[0:16], row[52:56], row[56], row[57], row[58])
    r.finalize()
    got = ctx.trace_rays(g["ray_o"], g["ray_d"])
    assert got.tobytes() == g["hits"].tobytes()
