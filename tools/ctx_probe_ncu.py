"""Six contexts, two filtered batches each (for ncu: compare a slow context's filter launch with a fast one's)."""
import os, sys
sys.path.insert(0, os.getcwd())
import chronos_b200
from chronos_b200 import tracegen
n = 10_000_000
dev = tracegen.generate_device(tracegen.WORKLOADS["uniform_10m_4k"], device=0, first=0, n=n)
keep = []
for k in range(6):
    ctx = chronos_b200.Context(device=0, max_batch_events=n)
    for i in range(2):
        ctx.reset(); ctx.submit_ptr(dev.ptr, n, 0); ctx.fetch_counts()
    print(f"context {k}: filter {ctx.last_timing()[1]:.4f} ms", flush=True)
    if k % 2 == 0: keep.append(ctx)
    else: ctx.close()
    keep.append(tracegen.DeviceBuffer(0, (37 + 11 * k) << 20))
