"""mcpt_trace_rays_static (include/mcpt.h): the ray-batch entry point that allocates nothing -- its ray / hit sectors
live in a fixed __device__ array inside libmcpt.so; no counterpart in the
reference. Collected last on purpose.

Through the C ABI, against the existing entry point mcpt_trace_rays_device on the same device buffers: records must be
bit-identical (the tolerance every traversal test of this repo uses), on two calls in a row that share the block --
the second one smaller and on another scene, so stale contents of the first would show. The test owns its context
and buffers: it passes alone in a fresh process and after the other tests."""
import numpy as np
import pytest

from common import both, random_rays
from montecarlo_pathtracing_b200 import Context, McptError, capi

pytestmark = pytest.mark.gpu


def _device_rays(torch, ro, rd):
    o4 = np.zeros((len(ro), 4), np.float32)
    d4 = np.zeros((len(rd), 4), np.float32)
    o4[:, :3], d4[:, :3] = ro, rd
    return torch.from_numpy(o4).cuda(), torch.from_numpy(d4).cuda()


def test_static_scratch_entry_point_equals_the_allocating_one():
    import torch
    assert capi.HIT_DTYPE.itemsize == 44
    ctx = Context(0)
    try:
        # (scene, rays): a tree scene of 410 primitives walked by persistent warps, then a small flat scene with
        # fewer rays -- the second call finds the first call's sectors in the block
        for call, (kind, n) in enumerate(((3, capi.TRACE_STATIC_MAX_RAYS), (6, 10_007))):
            s, o = both(kind)
            ro, rd = random_rays(o.prim_data(), n, seed=40 + n % 7)
            ctx.upload_scene(s)
            o4, d4 = _device_rays(torch, ro, rd)
            out_ref = torch.zeros(n * 44, dtype=torch.uint8, device="cuda")
            out_new = torch.full((n * 44,), 0xAB, dtype=torch.uint8, device="cuda")
            ctx.trace_rays_device(n, o4.data_ptr(), d4.data_ptr(), out_ref.data_ptr())
            ctx.synchronize()
            free_before = torch.cuda.mem_get_info()[0]
            ctx.trace_rays_static(n, o4.data_ptr(), d4.data_ptr(), out_new.data_ptr())
            ctx.synchronize()
            # The call takes no device memory. (Only from the second call on is that visible as "free memory
            # unchanged": with CUDA's lazy module loading the driver materialises the library image's 4 MiB array
            # on a device when the module holding it is first touched there, i.e. at the once-per-device set-up
            # inside the first call -- the image, not an allocation of the call.)
            taken = free_before - torch.cuda.mem_get_info()[0]
            assert taken == 0 if call > 0 else 0 <= taken <= 8 << 20, taken
            ref = out_ref.cpu().numpy().view(capi.HIT_DTYPE)
            got = out_new.cpu().numpy().view(capi.HIT_DTYPE)
            assert (ref["index"] >= 0).mean() > 0.2                     # the batch does exercise hits
            assert got.tobytes() == ref.tobytes()
            assert ctx.trace_rays(ro, rd).tobytes() == ref.tobytes()    # ... and both equal the host-buffer entry point
            hits = o.trace(ro[:2000], rd[:2000])                        # and the oracle
            assert np.array_equal(got["index"][:2000], hits["index"]) and np.array_equal(got["dist"][:2000], hits["dist"])
        # beyond the block: a clear error, nothing launched
        with pytest.raises(McptError) as e:
            ctx.trace_rays_static(capi.TRACE_STATIC_MAX_RAYS + 1, o4.data_ptr(), d4.data_ptr(), out_new.data_ptr())
        assert e.value.code == capi.MCPT_E_INVALID and "MCPT_TRACE_STATIC_MAX_RAYS" in str(e.value)
        ctx.trace_rays_static(0, 0, 0, 0)                               # empty batch: ok, no pointers needed
    finally:
        ctx.close()
