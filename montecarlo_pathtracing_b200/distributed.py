"""spp-split rendering across the GPUs of one box: one process per GPU (torchrun), every rank
traces its own seed range of the SAME frame into its own accumulation buffer, and the buffers are
combined by ONE sum-reduce to the root (NCCL over NVLink on GPUs; gloo in the CPU tests).

The path shards without any data-path exchange: passes are independent given numero_pass
(shaders/raytracer_func.frag:105-110) and accumulation is a plain sum (additive blending,
MontecarloGPU/montecarlo.cpp:450-466). The accumulation buffer IS the collective's send/recv buffer
(mcpt_bind_accum), so the accumulate kernel writes straight into it and no staging copy exists."""
from __future__ import annotations


def pass_range(rank: int, world: int, n_passes: int, first_pass: int = 0):
    """Contiguous, disjoint, exhaustive split of passes [first_pass, first_pass + n_passes) over
    `world` ranks; the first n_passes % world ranks take one extra pass. Returns (first, count)."""
    if world <= 0 or not (0 <= rank < world) or n_passes < 0:
        raise ValueError(f"bad split: rank {rank} of {world}, {n_passes} passes")
    base, rem = divmod(n_passes, world)
    return first_pass + rank * base + min(rank, rem), base + (1 if rank < rem else 0)


def reduce_accum(accum, root: int = 0, group=None):
    """One sum-reduce of the per-rank accumulation buffers to `root` (in place)."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.reduce(accum, dst=root, op=dist.ReduceOp.SUM, group=group)
    return accum


def make_nccl_comm(rank: int, world: int):
    """An ncclComm_t of our own over the ranks of the (already initialised) torch process group, for the C ABI's
    mcpt_reduce_nccl: rank 0 draws the unique id, torch broadcasts it, every rank joins. The library resolves NCCL's
    symbols from the copy already loaded in the process (torch's), which is promoted to the global namespace here.
    Returns a ctypes.c_void_p (pass .value); the caller ends it with destroy_nccl_comm."""
    import ctypes as C
    import glob
    import site

    import torch
    import torch.distributed as dist

    paths = [p for sp in site.getsitepackages() for p in glob.glob(sp + "/nvidia/nccl/lib/libnccl.so.2")]
    nccl = C.CDLL(paths[0] if paths else "libnccl.so.2", mode=C.RTLD_GLOBAL)

    class NcclUniqueId(C.Structure):   # passed BY VALUE to ncclCommInitRank
        _fields_ = [("internal", C.c_byte * 128)]

    uid = NcclUniqueId()
    if rank == 0 and nccl.ncclGetUniqueId(C.byref(uid)) != 0:
        raise RuntimeError("ncclGetUniqueId failed")
    t = torch.tensor(list(bytes(uid)), dtype=torch.uint8, device="cuda")
    dist.broadcast(t, 0)
    uid = NcclUniqueId.from_buffer_copy(t.cpu().numpy().tobytes())
    comm = C.c_void_p()
    nccl.ncclCommInitRank.argtypes = [C.POINTER(C.c_void_p), C.c_int, NcclUniqueId, C.c_int]
    nccl.ncclCommInitRank.restype = C.c_int
    rc = nccl.ncclCommInitRank(C.byref(comm), world, uid, rank)
    if rc != 0:
        raise RuntimeError(f"ncclCommInitRank -> {rc}")
    comm._nccl = nccl
    return comm


def destroy_nccl_comm(comm):
    import ctypes as C
    comm._nccl.ncclCommDestroy.argtypes = [C.c_void_p]
    comm._nccl.ncclCommDestroy(comm)


class SplitRenderer:
    """Drives one Context per rank. `accum` is a float32 torch tensor of h*w*4 elements on the
    context's device, bound as the library's accumulation target."""

    def __init__(self, ctx, width, height, rank=0, world=1, group=None):
        import torch
        self.ctx, self.rank, self.world, self.group = ctx, rank, world, group
        # zero_(), k_accumulate and the reduce must be ordered on ONE stream: the context launches on torch's
        # current stream from here on (its own stream is non-blocking and would not be ordered against torch's)
        # (handle 0 is torch's legacy default stream; mcpt_set_stream reads NULL as "own stream", so it is passed
        # as cudaStreamLegacy = 1)
        ctx.set_stream(torch.cuda.current_stream().cuda_stream or 1)
        ctx.resize(width, height)
        self.accum = torch.zeros(height * width * 4, dtype=torch.float32, device="cuda")
        ctx.bind_accum(self.accum.data_ptr(), self.accum.numel() * 4)

    def render(self, n_passes, first_pass=0, root=0):
        """Clear, trace this rank's share of the passes, reduce to root. Asynchronous on the torch
        stream that was current when the renderer was made (the context launches on it)."""
        first, count = pass_range(self.rank, self.world, n_passes, first_pass)
        self.accum.zero_()
        self.ctx.render(first, count)
        reduce_accum(self.accum, root, self.group)
        return self.accum
