"""Frame-stream sharding across the GPUs of one box and the final result gather.

Frames are independent (communicator/ros_inference.py:117-175 overwrites all per-frame state), so the stream is
partitioned, one process per GPU, with no collective on the data path. The only communication is one gather of
the per-frame detection counts and padded rows at the end (KBs): torch.distributed all_gather over NCCL on the
GPU box, gloo in the CPU tests."""
import os

import torch
import torch.distributed as dist


def bind_to_gpu_numa_node(device_index):
    """Host side of one process per GPU: run this process (and therefore first-touch its pinned staging buffers) on the
    NUMA node the GPU hangs off, so that the per-step host<->device copies of every rank stay on their own socket's
    memory instead of crossing the inter-socket link. Reads /sys (PCI numa_node, node cpulist); does nothing when the
    box has one node or the files are missing. Returns a dict describing what was done."""
    info = {"numa_node": None, "cpus": None, "bound": False}
    try:
        p = torch.cuda.get_device_properties(device_index)
        bdf = "%04x:%02x:%02x.0" % (p.pci_domain_id, p.pci_bus_id, p.pci_device_id)
        node_path = f"/sys/bus/pci/devices/{str(bdf).lower()}/numa_node"
        node = int(open(node_path).read().strip())
        info["numa_node"] = node
        nodes = [d for d in os.listdir("/sys/devices/system/node") if d.startswith("node") and d[4:].isdigit()]
        if node < 0 or len(nodes) < 2:
            return info
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        allowed = cpus & set(os.sched_getaffinity(0))
        if allowed:
            os.sched_setaffinity(0, allowed)
            info["cpus"], info["bound"] = len(allowed), True
    except Exception as e:      # affinity is an optimisation, never a reason to fail
        info["error"] = str(e)[:120]
    return info


def shard_frames(n_frames, rank, world, mode="interleave"):
    """Indices of the frames rank `rank` owns: 'interleave' = i % world == rank (live stream), 'contiguous' = a
    block of the bag file (the last ranks get one frame less when world does not divide n_frames)."""
    if mode == "interleave":
        return list(range(rank, n_frames, world))
    base, extra = divmod(n_frames, world)
    start = rank * base + min(rank, extra)
    return list(range(start, start + base + (1 if rank < extra else 0)))


def gather_detections(rows, counts, frame_ids, n_frames, group=None):
    """rows [f, max_det, C] and counts [f] of this rank's frames (frame_ids: their global indices) -> on every rank
    (rows_all [n_frames, max_det, C], counts_all [n_frames]) in global frame order. Ranks may own different numbers
    of frames; shards are padded to the largest for the all_gather."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if world == 1:
        out_r = rows.new_zeros((n_frames,) + tuple(rows.shape[1:]))
        out_c = counts.new_zeros((n_frames,))
        idx = torch.as_tensor(frame_ids, dtype=torch.long, device=rows.device)
        out_r[idx], out_c[idx] = rows, counts
        return out_r, out_c
    dev = rows.device
    f_max = (n_frames + world - 1) // world
    pad_r = rows.new_zeros((f_max,) + tuple(rows.shape[1:]))
    pad_c = counts.new_zeros((f_max,))
    pad_i = torch.full((f_max,), -1, dtype=torch.int64, device=dev)
    f = rows.shape[0]
    pad_r[:f], pad_c[:f] = rows, counts
    pad_i[:f] = torch.as_tensor(frame_ids, dtype=torch.int64, device=dev)
    all_r = [torch.empty_like(pad_r) for _ in range(world)]
    all_c = [torch.empty_like(pad_c) for _ in range(world)]
    all_i = [torch.empty_like(pad_i) for _ in range(world)]
    dist.all_gather(all_i, pad_i, group=group)
    dist.all_gather(all_c, pad_c, group=group)
    dist.all_gather(all_r, pad_r, group=group)
    out_r = rows.new_zeros((n_frames,) + tuple(rows.shape[1:]))
    out_c = counts.new_zeros((n_frames,))
    for r, c, i in zip(all_r, all_c, all_i):
        ok = i >= 0
        out_r[i[ok]], out_c[i[ok]] = r[ok], c[ok]
    return out_r, out_c
