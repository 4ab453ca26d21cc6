"""Does the filter time depend on WHICH context (= which allocation addresses) runs it?"""
import os, sys
sys.path.insert(0, os.getcwd())
import numpy as np
import chronos_b200
from chronos_b200 import tracegen
n = 10_000_000
dev = tracegen.generate_device(tracegen.WORKLOADS["uniform_10m_4k"], device=0, first=0, n=n)
def measure(ctx, reset=True, steps=15):
    f = []
    for i in range(steps + 5):
        if reset: ctx.reset()
        ctx.submit_ptr(dev.ptr, n, i * n if not reset else 0); ctx.fetch_counts()
        if i >= 5: f.append(ctx.last_timing()[1])
    return float(np.mean(f)), float(np.min(f)), float(np.max(f))
keep = []
for k in range(6):
    ctx = chronos_b200.Context(device=0, max_batch_events=n)
    a = measure(ctx, True); b = measure(ctx, False); c = measure(ctx, True)
    print(f"context {k}: reset each step {a[0]:.4f} [{a[1]:.4f}..{a[2]:.4f}]  state carried {b[0]:.4f} [{b[1]:.4f}..{b[2]:.4f}]  reset again {c[0]:.4f}", flush=True)
    if k % 2 == 0: keep.append(ctx)      # stays alive: the next context gets other addresses
    else: ctx.close()
    keep.append(tracegen.DeviceBuffer(0, (37 + 11 * k) << 20))
