#!/usr/bin/env python
"""Host-link ceiling of the end-to-end path: N ranks (one per GPU, torchrun) each copy pinned host memory up and down
with large cudaMemcpyAsync calls, alone per direction and both directions at once, all ranks at the same time.
Prints one JSON line: GB/s per rank and direction, summed over the box.

  python profiles/probe_pcie_nranks.py                                                   # N = 1
  python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 profiles/probe_pcie_nranks.py
"""
import json
import os
import sys
import time

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from triton_client_b200.shard import bind_to_gpu_numa_node  # noqa: E402

world = int(os.environ.get("WORLD_SIZE", "1"))
rank = int(os.environ.get("RANK", "0"))
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
numa = bind_to_gpu_numa_node(local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
MB = 256
up_h = torch.empty(MB << 20, dtype=torch.uint8).pin_memory()
dn_h = torch.empty(MB << 20, dtype=torch.uint8).pin_memory()
up_d = torch.empty(MB << 20, dtype=torch.uint8, device=dev)
dn_d = torch.zeros(MB << 20, dtype=torch.uint8, device=dev)
s_up, s_dn = torch.cuda.Stream(dev), torch.cuda.Stream(dev)


def run(up, dn, reps=12):
    def once(n):
        for _ in range(n):
            if up:
                with torch.cuda.stream(s_up):
                    up_d.copy_(up_h, non_blocking=True)
            if dn:
                with torch.cuda.stream(s_dn):
                    dn_h.copy_(dn_d, non_blocking=True)
    once(2)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    once(reps)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    t = torch.tensor([dt], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return reps * (MB << 20) / float(t.item()) / 1e9      # GB/s per rank and direction, at the slowest rank's time


res = {"n_ranks": world, "copy_mb": MB, "h2d_alone_gbs_per_rank": run(True, False), "d2h_alone_gbs_per_rank": run(False, True),
       "both_gbs_per_rank_per_direction": run(True, True), "host_numa_rank0": numa, "cpus": os.cpu_count()}
res["both_total_gbs_per_direction"] = res["both_gbs_per_rank_per_direction"] * world
if rank == 0:
    print(json.dumps(res), flush=True)
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
