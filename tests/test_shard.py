"""Multi-GPU host logic on CPU: frame sharding + final result gather with world_size 2 over gloo."""
import os

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from triton_client_b200 import shard


def test_shard_frames_partitions_everything():
    for n in (0, 1, 7, 64, 65):
        for world in (1, 2, 4, 8):
            for mode in ("interleave", "contiguous"):
                parts = [shard.shard_frames(n, r, world, mode) for r in range(world)]
                assert sorted(i for p in parts for i in p) == list(range(n))
                assert max(len(p) for p in parts) - min(len(p) for p in parts) <= 1


def _worker(rank, world, port, n_frames, mode, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        ids = shard.shard_frames(n_frames, rank, world, mode)
        # frame i has (i % 5) detections whose rows carry the frame index
        rows = torch.zeros((len(ids), 6, 3), dtype=torch.float64)
        counts = torch.zeros((len(ids),), dtype=torch.int32)
        for k, i in enumerate(ids):
            counts[k] = i % 5
            rows[k, :i % 5] = float(i)
        out_r, out_c = shard.gather_detections(rows, counts, ids, n_frames)
        ok = all(int(out_c[i]) == i % 5 and bool((out_r[i, :i % 5] == float(i)).all()) and bool((out_r[i, i % 5:] == 0).all())
                 for i in range(n_frames))
        ret[rank] = ok
    finally:
        dist.destroy_process_group()


def test_gather_detections_world2_gloo():
    for n_frames, mode, port in [(9, "interleave", 29611), (10, "contiguous", 29612)]:
        ctx = mp.get_context("spawn")
        ret = ctx.Manager().dict()
        procs = [ctx.Process(target=_worker, args=(r, 2, port, n_frames, mode, ret)) for r in range(2)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(120)
            assert p.exitcode == 0
        assert ret[0] and ret[1]


def test_gather_single_process():
    rows = torch.arange(12, dtype=torch.float64).view(2, 2, 3)
    out_r, out_c = shard.gather_detections(rows, torch.tensor([1, 2], dtype=torch.int32), [2, 0], 3)
    assert out_c.tolist() == [2, 0, 1] and torch.equal(out_r[2], rows[0])


def test_bind_to_gpu_numa_node_never_raises():
    """Affinity is an optimisation: without a GPU (or on a one-node box) the call reports and changes nothing."""
    import os
    from triton_client_b200.shard import bind_to_gpu_numa_node
    before = os.sched_getaffinity(0)
    info = bind_to_gpu_numa_node(0)
    assert set(info) >= {"numa_node", "cpus", "bound"} and info["bound"] in (False, True)
    if not info["bound"]:
        assert os.sched_getaffinity(0) == before
