"""CPU, world_size 2 over gloo: the spp-split host logic (pass partition + one sum-reduce to the
root). Each rank's "device" is the CPU oracle here; the partition and the reduce call are the
product's own (montecarlo_pathtracing_b200/distributed.py)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from montecarlo_pathtracing_b200.distributed import pass_range, reduce_accum

W, H, SPP = 48, 32, 7          # 7 passes over 2 ranks: an uneven split (4 + 3)


def test_pass_range_partitions_exactly():
    for world in (1, 2, 3, 4, 8):
        for n in (0, 1, 7, 64, 1023):
            got = [pass_range(r, world, n, first_pass=5) for r in range(world)]
            passes = [p for f, c in got for p in range(f, f + c)]
            assert passes == list(range(5, 5 + n))                       # disjoint, ordered, exhaustive
            assert max(c for _, c in got) - min(c for _, c in got) <= 1  # balanced
    with pytest.raises(ValueError):
        pass_range(2, 2, 8)


def _oracle_for_scene6():
    from common import DATE, both
    from montecarlo_pathtracing_b200 import capi
    s, o = both("c2")
    iv, ipv = capi.default_camera(W, H)
    o.set_uniforms(8, 1.5, DATE, 0, iv, ipv)
    o.set_threads(2)
    return o


def _worker(rank, world, port, out_path):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        o = _oracle_for_scene6()
        first, count = pass_range(rank, world, SPP)
        acc = torch.from_numpy(o.render(W, H, first, count)).reshape(-1)    # this rank's partial sums
        reduce_accum(acc, root=0)
        if rank == 0:
            np.save(out_path, acc.numpy().reshape(H, W, 3))
    finally:
        dist.destroy_process_group()


def test_two_ranks_equal_one(tmp_path):
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    out = str(tmp_path / "sum.npy")
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    got = np.load(out)
    ref = _oracle_for_scene6().render(W, H, 0, SPP)                         # one rank, passes in order
    assert np.isfinite(got).all()
    assert np.allclose(got, ref, rtol=1e-5, atol=1e-6)                      # fp32 reassociation only
    assert got.mean() > 0.05


def test_strong_split_of_the_bench_covers_the_budget_exactly():
    """bench.py --scaling strong: BASELINE config 4's 1024 spp (8 default steps' worth elsewhere) split over N ranks --
    disjoint, ordered, exhaustive for every N the driver uses, and the per-rank shares differ by at most one pass."""
    import sys
    sys.path.insert(0, str(__import__("pathlib").Path(__file__).resolve().parent.parent))
    import bench
    assert bench.strong_spp("c4") == 1024 and bench.strong_spp("c2") == 8 * bench.WORKLOADS["c2"][4]
    for name in ("c4", "c2"):
        total = bench.strong_spp(name)
        for world in (1, 2, 3, 4, 8):
            parts = [pass_range(r, world, total) for r in range(world)]
            assert parts[0][0] == 0 and sum(c for _, c in parts) == total
            assert all(parts[r][0] + parts[r][1] == parts[r + 1][0] for r in range(world - 1))
            assert max(c for _, c in parts) - min(c for _, c in parts) <= 1
