#!/usr/bin/env python
"""Phase timing of one rank of the sharded step (run under torchrun, 2+ GPUs)."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist

import chronos_b200  # noqa
from chronos_b200 import sharded, tracegen

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
n = int(sys.argv[1]) if len(sys.argv) > 1 else 12_500_000
dev = tracegen.generate_device(tracegen.WORKLOADS["sharded_100m_1m"], device=lr, first=rank * n, n=n)
mode = sys.argv[2] if len(sys.argv) > 2 else "auto"
r = sharded.ShardedRunner(dev.ptr, n, rank * n, lr, rank, world, mode=mode)
if rank == 0:
    print(f"mode {r.mode} {getattr(r, '_direct_error', '')}", flush=True)
for it in range(6):
    dist.barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    if r.mode == "direct":
        c = r.step()
        t1 = time.perf_counter()
        if rank == 0:
            print(f"it{it}: step {1e3*(t1-t0):.3f} ms wall, library events {c['gpu_ms']:.3f} ms, filter {c['filter_ms']:.3f} ms, "
                  f"received {c['n_survivors']} windows {c['n_windows']} launches {c['launches']}", flush=True)
        continue
    r.ctx.reset()
    ec, sc = r.filter_partition(r.rec_ptr, r.n, r.first)
    t1 = time.perf_counter()
    n_recv, n_slots = sharded.exchange(r.send_slots, r.send_hdr, ec, sc, sharded.SLOT_BYTES, r.recv_slots, r.recv_hdr)
    torch.cuda.current_stream().synchronize()
    t2 = time.perf_counter()
    r.ctx.submit_survivors_ptr(r.recv_slots.data_ptr(), r.recv_hdr.data_ptr(), n_recv, n_slots)
    c = r.ctx.fetch_counts()
    t3 = time.perf_counter()
    if rank == 0:
        print(f"it{it}: stage1 {1e3*(t1-t0):.3f} ms  exchange {1e3*(t2-t1):.3f} ms  stage2 {1e3*(t3-t2):.3f} ms  sent {int(ec.sum())} entries {int(sc.sum())*16/1e6:.0f} MB recv {n_recv} filter_ms {r.ctx.last_timing()[1]:.3f}", flush=True)
dist.barrier()
dist.destroy_process_group()
