#!/usr/bin/env python
"""Short driver for ncu: a few steps of the N=1 bench workload (device-resident input),
nothing else. Usage: python tools/profile_step.py [events] [steps] [workload]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import chronos_b200  # noqa: E402
from chronos_b200 import tracegen  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
name = sys.argv[3] if len(sys.argv) > 3 else "uniform_10m_4k"
dev = tracegen.generate_device(tracegen.WORKLOADS[name], device=0, first=0, n=n)
ctx = chronos_b200.Context(device=0, max_batch_events=n)
for i in range(steps):
    ctx.reset()
    ctx.submit_ptr(dev.ptr, n, 0)
    c = ctx.fetch_counts()
    t, f = ctx.last_timing()
    print(f"step {i}: {c} gpu_ms={t:.3f} filter_ms={f:.3f} launches={ctx.last_launches()}", flush=True)
