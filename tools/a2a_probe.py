#!/usr/bin/env python
"""all_to_all_single cost vs size on this box (context for the sharded step)."""
import os, time, torch, torch.distributed as dist
rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
for mb in (0.001, 1, 16, 64, 128, 256, 430):
    n = int(mb * 1e6) // world * world
    n = max(n, world)
    s = torch.empty(n, dtype=torch.uint8, device="cuda"); r = torch.empty_like(s)
    for _ in range(3): dist.all_to_all_single(r, s)
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(10): dist.all_to_all_single(r, s)
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / 10
    if rank == 0: print(f"{mb:8.3f} MB per rank: {dt*1e3:.3f} ms  ({n*(world-1)/world/dt/1e9:.1f} GB/s off-GPU per rank)", flush=True)
dist.destroy_process_group()
