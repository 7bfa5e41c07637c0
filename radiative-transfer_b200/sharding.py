"""Energy-group sharding across ranks (one process per GPU).

Groups never couple in the reference's solve (no scattering, no temperature update), so the path shards by
contiguous blocks of groups with NO data-path collective (the material-temperature extension, rt_params::cv, adds the
one exchange step: an all-reduce of three N-vectors per time step, `make_allreduce_hook`); every rank keeps the G-length tables for all groups (the
edge stencils of correction.cpp:125-249 and the "remainder into the last group" of Planck.cpp:58-76 need them).
The only communication is the final gather of per-group results, done with torch.distributed (NCCL on GPUs, gloo in the
CPU tests)."""
from __future__ import annotations

from typing import List, Tuple

import numpy as np


def group_range(G: int, world: int, rank: int) -> Tuple[int, int]:
    """(first group, number of groups) of `rank`: contiguous blocks, the first G % world ranks get one more."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError("bad world/rank")
    if G < world:
        raise ValueError(f"{G} groups cannot be sharded over {world} ranks")
    base, extra = divmod(G, world)
    begin = rank * base + min(rank, extra)
    return begin, base + (1 if rank < extra else 0)


def all_ranges(G: int, world: int) -> List[Tuple[int, int]]:
    return [group_range(G, world, r) for r in range(world)]


def batches(group_count: int, groups_per_batch: int) -> List[Tuple[int, int]]:
    """(local first group, size) of the batches a rank processes one after the other when its state exceeds HBM."""
    if groups_per_batch < 1:
        raise ValueError("groups_per_batch must be >= 1")
    return [(b, min(groups_per_batch, group_count - b)) for b in range(0, group_count, groups_per_batch)]


def gather_group_rows(local: np.ndarray, G: int, dist=None, device=None) -> np.ndarray:
    """All-gather arrays whose FIRST axis is the rank's group block, e.g. phi in group-major order [gl, c], into the
    full [G, ...] array on every rank.  `dist` is torch.distributed (initialised) or None for a single process."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        assert local.shape[0] == G
        return local
    import torch
    world, rank = dist.get_world_size(), dist.get_rank()
    ranges = all_ranges(G, world)
    assert local.shape[0] == ranges[rank][1]
    dev = device if device is not None else ("cuda" if dist.get_backend() == "nccl" else "cpu")
    tail = local.shape[1:]
    nmax = max(n for _, n in ranges)
    pad = np.zeros((nmax,) + tail, dtype=local.dtype)
    pad[: local.shape[0]] = local
    mine = torch.from_numpy(pad).to(dev)
    outs = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(outs, mine)
    full = np.empty((G,) + tail, dtype=local.dtype)
    for (b, n), t in zip(ranges, outs):
        full[b:b + n] = t[:n].cpu().numpy()
    return full


def reduce_scalar(value: float, op: str = "max", dist=None, device=None) -> float:
    """max / sum of one fp64 scalar over the ranks (device time of a step = max over ranks; units = sum over ranks)."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return float(value)
    import torch
    dev = device if device is not None else ("cuda" if dist.get_backend() == "nccl" else "cpu")
    t = torch.tensor([float(value)], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX if op == "max" else dist.ReduceOp.SUM)
    return float(t.item())


def reduce_vector_sum(local: np.ndarray, dist=None, device=None) -> np.ndarray:
    """Element-wise fp64 sum over the ranks of a small vector -- the validation sums of correction.cpp:39-63,100-122
    (sum_g B_g, sum_g dB/dT_g, sum_g kappa_g B_g) when every rank only contributes the groups it owns."""
    local = np.ascontiguousarray(local, dtype=np.float64)
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return local.copy()
    import torch
    dev = device if device is not None else ("cuda" if dist.get_backend() == "nccl" else "cpu")
    t = torch.from_numpy(local.copy()).to(dev)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return t.cpu().numpy()


def allreduce_tensor_(t, dist):
    """In-place element-wise sum of a torch tensor over the ranks (NCCL for device tensors, gloo for CPU tensors)."""
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return t


class _DevicePointer:
    """A raw device buffer of `count` float64 as something torch.as_tensor can wrap without copying."""

    def __init__(self, ptr: int, count: int):
        self.__cuda_array_interface__ = {"shape": (count,), "typestr": "<f8", "data": (ptr, False), "version": 3,
                                         "strides": None}


def make_allreduce_hook(dist, device: int):
    """The callback for rt_set_allreduce (include/rt_cuda.h): sums the 3*N group sums of the material-temperature update
    (absorbed power, emitted power, emission derivative per cell) over the ranks with torch.distributed -- NCCL over
    NVLink/NVSwitch on a multi-GPU box -- in place on the library's own device buffer, ordered on the library's stream.
    Returns a ctypes function object; the caller must keep it alive as long as the handle uses it."""
    import torch
    from . import capi

    def hook(buf, count, stream, user):
        try:
            dev = torch.device("cuda", device)
            ext = torch.cuda.ExternalStream(int(stream), device=dev)
            with torch.cuda.device(dev), torch.cuda.stream(ext):
                t = torch.as_tensor(_DevicePointer(int(buf), int(count)), device=dev)
                allreduce_tensor_(t, dist)
            return 0
        except Exception as exc:  # an exception must not unwind through the C frames
            import sys
            print(f"rt all-reduce hook: {exc!r}", file=sys.stderr)
            return 1

    return capi.ALLREDUCE_FN(hook)


def broadcast_nccl_id(dist, device: int):
    """The 128-byte ncclUniqueId for rt_set_nccl (include/rt_cuda.h): made by rank 0 inside the library, handed to every
    rank through torch.distributed (the host's own transport, used once at set-up -- never in the time loop).
    Returns a ctypes buffer."""
    import ctypes as C
    import torch
    from . import capi
    buf = C.create_string_buffer(128)
    if dist.get_rank() == 0:
        capi.check(capi.lib().rt_nccl_unique_id(buf))
    dev = torch.device("cuda", device) if dist.get_backend() == "nccl" else torch.device("cpu")
    t = torch.tensor(list(buf.raw), dtype=torch.uint8, device=dev)
    dist.broadcast(t, src=0)
    C.memmove(buf, bytes(t.cpu().tolist()), 128)
    return buf


def bind_to_gpu_numa_node(device: int):
    """Pin this process to the CPU cores of the NUMA node its GPU hangs off, so that pinned host buffers allocated
    afterwards (first touch) and the copy engine's writes stay on that node.  With 8 ranks draining phi to the host at
    once, buffers on a remote node cost more than half of the PCIe rate (profiles/r02_d2h_ceiling.json).
    Returns a dict describing what was done (nothing, when the topology cannot be read)."""
    import os
    info = {"device": device, "numa_node": None, "cpus": None, "bound": False}
    try:
        import torch
        pr = torch.cuda.get_device_properties(device)
        bus = None
        if all(hasattr(pr, k) for k in ("pci_domain_id", "pci_bus_id", "pci_device_id")) and isinstance(pr.pci_bus_id, int):
            bus = f"{pr.pci_domain_id:04x}:{pr.pci_bus_id:02x}:{pr.pci_device_id:02x}.0"
        if bus is None:
            import subprocess
            bus = subprocess.run(["nvidia-smi", "-i", str(device), "--query-gpu=pci.bus_id", "--format=csv,noheader"],
                                 capture_output=True, text=True, timeout=10).stdout.strip()
        bus = str(bus).lower()
        if bus.count(":") == 2 and len(bus.split(":")[0]) == 8:
            bus = bus[4:]  # nvidia-smi prints an 8-digit domain, sysfs uses 4
        node = int(open(f"/sys/bus/pci/devices/{bus}/numa_node").read())
        info["numa_node"] = node
        if node < 0:
            return info
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        cpus &= set(os.sched_getaffinity(0))
        if cpus:
            os.sched_setaffinity(0, cpus)
            info["cpus"], info["bound"] = len(cpus), True
    except Exception as exc:  # best effort: an unreadable topology must not stop a run
        info["error"] = repr(exc)
    return info
