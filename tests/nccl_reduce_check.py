#!/usr/bin/env python
"""N-GPU check of the spp split through the C ABI's own collective (mcpt_reduce_nccl), run under
torchrun on a multi-GPU box (not collected by pytest: it needs >= 2 GPUs):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29511 tests/nccl_reduce_check.py

Every rank renders its pass range of scene 6 into its own accumulation buffer, the buffers are
combined by ONE ncclReduce issued by the library on its stream, and rank 0 compares the sum with
its own single-GPU render of all passes (fp32 reassociation tolerance) and with the torch path
(bind_accum + torch.distributed.reduce), which must give the same bits as the C path.
"""
import ctypes as C
import glob
import os
import site
import sys
from pathlib import Path

import numpy as np
import torch
import torch.distributed as dist

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))

from common import DATE, host_scene  # noqa: E402
from montecarlo_pathtracing_b200 import Context, capi  # noqa: E402
from montecarlo_pathtracing_b200.distributed import SplitRenderer, pass_range  # noqa: E402

W, H, SPP = 640, 360, 13     # 13 passes: an uneven split for every world size > 1


def nccl_lib():
    paths = [p for sp in site.getsitepackages() for p in glob.glob(sp + "/nvidia/nccl/lib/libnccl.so.2")]
    # the copy torch itself uses, promoted to the global namespace so the library's dlsym finds it
    return C.CDLL(paths[0] if paths else "libnccl.so.2", mode=C.RTLD_GLOBAL)


class NcclUniqueId(C.Structure):   # ncclUniqueId: passed BY VALUE to ncclCommInitRank
    _fields_ = [("internal", C.c_byte * 128)]


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    nccl = nccl_lib()
    uid = NcclUniqueId()
    if rank == 0:
        assert nccl.ncclGetUniqueId(C.byref(uid)) == 0
    t = torch.tensor(list(bytes(uid)), dtype=torch.uint8, device="cuda")
    dist.broadcast(t, 0)
    uid = NcclUniqueId.from_buffer_copy(t.cpu().numpy().tobytes())
    comm = C.c_void_p()
    nccl.ncclCommInitRank.argtypes = [C.POINTER(C.c_void_p), C.c_int, NcclUniqueId, C.c_int]
    nccl.ncclCommInitRank.restype = C.c_int
    rc = nccl.ncclCommInitRank(C.byref(comm), world, uid, rank)
    assert rc == 0, f"ncclCommInitRank -> {rc}"

    s = host_scene("c2", 0.4)
    s.finalize()
    iv, ipv = capi.default_camera(W, H)

    def make_ctx():
        c = Context(local)
        c.upload_scene(s)
        c.set_camera(iv, ipv)
        c.set_params(nb_bounces=8, refract_ind=1.5, date=float(DATE))
        return c

    # ---- C path: library-owned buffer + mcpt_reduce_nccl
    ctx = make_ctx()
    ctx.resize(W, H)
    first, count = pass_range(rank, world, SPP)
    ctx.render(first, count)
    ctx.reduce_nccl(comm.value, 0)
    ctx.synchronize()
    got_c = ctx.read_accum()

    # ---- torch path: accumulation buffer bound to a tensor + torch.distributed.reduce
    ctx2 = make_ctx()
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    ctx2.set_stream(stream.cuda_stream)
    sr = SplitRenderer(ctx2, W, H, rank, world)
    acc = sr.render(SPP)
    torch.cuda.synchronize()
    got_t = acc.cpu().numpy().reshape(H, W, 4)

    ok = True
    if rank == 0:
        ref_ctx = make_ctx()
        ref_ctx.resize(W, H)
        ref_ctx.render(0, SPP)
        ref = ref_ctx.read_accum()
        same_bits = np.array_equal(got_c, got_t)
        close = np.allclose(got_c, ref, rtol=1e-5, atol=1e-5)
        rel = float(np.abs(got_c - ref).max() / max(np.abs(ref).max(), 1e-9))
        print(f"world {world}: C-path == torch-path bits: {same_bits}; vs single-GPU sum: allclose {close}, "
              f"max rel diff {rel:.2e}; mean {ref[..., :3].mean():.5f}")
        # with two ranks a + b is order-free; beyond that NCCL may pick another reduction order per communicator
        ok = (same_bits or world > 2) and close and ref[..., :3].mean() > 0.01
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.broadcast(flag, 0)
    nccl.ncclCommDestroy.argtypes = [C.c_void_p]
    nccl.ncclCommDestroy(comm)
    dist.destroy_process_group()
    sys.exit(0 if int(flag.item()) == 1 else 1)


if __name__ == "__main__":
    main()
