#!/usr/bin/env python
"""The host-drain ceiling of the end-to-end path: what N ranks copying device -> pinned host AT ONCE can sustain, per
GPU, with plain asynchronous copies -- no kernels of ours involved.  bench.py's `e2e` ships phi of every step to the host
(G_local*N*8 bytes per rank and step); when that takes longer than the step itself the copy, not the solve, is the clock.
Run under torchrun like bench.py; rank 0 prints one JSON line.  RT_NUMA_BIND=0 skips the NUMA binding (A/B)."""
import importlib, json, os, sys, time
import torch
import torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sharding = importlib.import_module("radiative-transfer_b200.sharding")
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
bind = sharding.bind_to_gpu_numa_node(local) if os.environ.get("RT_NUMA_BIND", "1") != "0" else {"bound": False}
nbytes = int(float(os.environ.get("D2H_GB", "2")) * (1 << 30))
src = torch.empty(nbytes, dtype=torch.uint8, device="cuda").fill_(1)
dst = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
dst.fill_(0)  # touch
def timed(reps):
    if world > 1: dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): dst.copy_(src, non_blocking=True)
    e1.record(); torch.cuda.synchronize()
    return nbytes * reps / (e0.elapsed_time(e1) * 1e-3) / 1e9
timed(2)
gbs = timed(5)
t = torch.tensor([gbs], dtype=torch.float64, device="cuda")
lo, hi, tot = t.clone(), t.clone(), t.clone()
if world > 1:
    dist.all_reduce(lo, op=dist.ReduceOp.MIN); dist.all_reduce(hi, op=dist.ReduceOp.MAX); dist.all_reduce(tot, op=dist.ReduceOp.SUM)
if rank == 0:
    print(json.dumps({"what": "concurrent device->pinned-host copies, one per rank, plain cudaMemcpyAsync", "n_gpus": world, "bytes_per_copy": nbytes,
                      "GBps_per_gpu_min": lo.item(), "GBps_per_gpu_max": hi.item(), "GBps_total": tot.item(), "numa_bind_rank0": bind}))
if world > 1:
    dist.barrier(); dist.destroy_process_group()
