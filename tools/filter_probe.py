"""Per-step filter-kernel and step times of the N=1 workload, without bench.py around it.
  python tools/filter_probe.py [steps] [sampler 0/1]"""
import os, sys, threading, time
sys.path.insert(0, os.getcwd())
import numpy as np
import chronos_b200
from chronos_b200 import tracegen
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 60
sampler = int(sys.argv[2]) if len(sys.argv) > 2 else 0
n = 10_000_000
dev = tracegen.generate_device(tracegen.WORKLOADS["uniform_10m_4k"], device=0, first=0, n=n)
ctx = chronos_b200.Context(device=0, max_batch_events=n)
stop = threading.Event()
if sampler:
    import pynvml
    pynvml.nvmlInit(); h = pynvml.nvmlDeviceGetHandleByIndex(0)
    def run():
        while not stop.is_set():
            pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM); time.sleep(0.002)
    threading.Thread(target=run, daemon=True).start()
f, t = [], []
for i in range(steps + 10):
    ctx.reset(); ctx.submit_ptr(dev.ptr, n, 0); ctx.fetch_counts()
    a, b = ctx.last_timing()
    if i >= 10: t.append(a); f.append(b)
stop.set()
f = np.array(f); t = np.array(t)
print(f"{os.path.basename(os.getcwd()):8s} sampler={sampler} filter mean {f.mean():.4f} min {f.min():.4f} slow(>0.6) {int((f>0.6).sum())}/{len(f)}  step mean {t.mean():.4f} median {np.median(t):.4f}", flush=True)
