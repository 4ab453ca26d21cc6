"""One context, many steps: per-step filter time against wall time (regime changes?)."""
import os, sys, time
sys.path.insert(0, os.getcwd())
import numpy as np
import chronos_b200
from chronos_b200 import tracegen
n = 10_000_000
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 3000
gb = lambda k: int(float(os.environ.get(k, "0")) * (1 << 30))
d0 = tracegen.DeviceBuffer(0, gb("DUMMY_FIRST")) if gb("DUMMY_FIRST") else None
dev = tracegen.generate_device(tracegen.WORKLOADS["uniform_10m_4k"], device=0, first=0, n=n)
d1 = tracegen.DeviceBuffer(0, gb("DUMMY_BEFORE")) if gb("DUMMY_BEFORE") else None
if os.environ.get("PENDING") is not None:
    ctx = chronos_b200.Context(device=0, max_batch_events=n, max_pending_events=int(os.environ["PENDING"]))
else:
    ctx = chronos_b200.Context(device=0, max_batch_events=n)
d2 = tracegen.DeviceBuffer(0, gb("DUMMY_AFTER")) if gb("DUMMY_AFTER") else None
sweep = os.environ.get("SWEEP")
if sweep:
    import torch
    big = torch.zeros(512 << 18, dtype=torch.float32, device="cuda:0")  # 512 MB
t0 = time.perf_counter(); f = []; ts = []
for i in range(steps):
    if sweep == "read":
        big.sum().item()
    elif sweep == "write":
        big.add_(1.0); torch.cuda.synchronize()
    ctx.reset(); ctx.submit_ptr(dev.ptr, n, 0); ctx.fetch_counts()
    f.append(ctx.last_timing()[1]); ts.append(time.perf_counter() - t0)
f = np.array(f); slow = f > 0.6
# run-length encode the regimes
runs = []; start = 0
for i in range(1, len(f) + 1):
    if i == len(f) or slow[i] != slow[start]:
        runs.append(("slow" if slow[start] else "fast", i - start, round(ts[start], 3), round(float(f[start:i].mean()), 3))); start = i
print({k: os.environ[k] for k in ("DUMMY_FIRST", "DUMMY_BEFORE", "DUMMY_AFTER", "PENDING", "SWEEP") if k in os.environ}, os.path.basename(os.getcwd()), "steps", steps, "slow fraction %.2f" % slow.mean(), "regimes (kind, steps, t_start_s, mean_ms):", runs[:40], flush=True)
