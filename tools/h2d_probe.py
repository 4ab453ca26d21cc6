"""Aggregate pinned host-to-device bandwidth with N ranks copying at once (one rank per GPU).
   python -m torch.distributed.run --nproc-per-node N tools/h2d_probe.py   (or plain python for N=1)"""
import json, os, sys, time
import torch
import torch.distributed as dist
rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1)); lr = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(lr)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
nbytes = 2_880_000_000
host = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
dev = torch.empty(nbytes, dtype=torch.uint8, device=f"cuda:{lr}")
best = 0.0
for rep in range(4):
    if world > 1: dist.barrier()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    dev.copy_(host, non_blocking=True); torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    if rep: best = max(best, nbytes / dt / 1e9)
t = torch.tensor([best], device=f"cuda:{lr}", dtype=torch.float64)
if world > 1:
    g = [torch.zeros_like(t) for _ in range(world)]; dist.all_gather(g, t); per = [float(x.item()) for x in g]
else:
    per = [best]
if rank == 0:
    print(json.dumps({"ranks": world, "h2d_gbs_per_rank": [round(x, 1) for x in per], "aggregate_gbs": round(sum(per), 1),
                      "bytes_per_copy": nbytes, "host_cores": os.cpu_count()}))
if world > 1: dist.destroy_process_group()
