"""long_probe with an NVML sampler beside it: what does the GPU report in the fast and in the slow regime?"""
import os, sys, time, threading
sys.path.insert(0, os.getcwd())
import numpy as np, pynvml
import chronos_b200
from chronos_b200 import tracegen
pynvml.nvmlInit(); h = pynvml.nvmlDeviceGetHandleByIndex(0)
n = 10_000_000
dev = tracegen.generate_device(tracegen.WORKLOADS["uniform_10m_4k"], device=0, first=0, n=n)
ctx = chronos_b200.Context(device=0, max_batch_events=n)
samples = []; stop = threading.Event()
def run():
    while not stop.is_set():
        try:
            samples.append((time.perf_counter(), pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM), pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_MEM),
                            pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_GRAPHICS), pynvml.nvmlDeviceGetPowerUsage(h) // 1000,
                            pynvml.nvmlDeviceGetPerformanceState(h), pynvml.nvmlDeviceGetCurrentClocksEventReasons(h),
                            pynvml.nvmlDeviceGetUtilizationRates(h).memory))
        except Exception as e:
            samples.append((time.perf_counter(), str(e)))
        time.sleep(0.001)
time.sleep(0.5)
th = threading.Thread(target=run, daemon=True); th.start()
time.sleep(0.05)
f = []; ts = []
for i in range(400):
    ctx.reset(); ctx.submit_ptr(dev.ptr, n, 0); ctx.fetch_counts()
    f.append(ctx.last_timing()[1]); ts.append(time.perf_counter())
stop.set(); th.join()
f = np.array(f)
print(os.path.basename(os.getcwd()), "slow fraction %.2f" % (f > 0.6).mean(), "first slow step", int(np.argmax(f > 0.6)) if (f > 0.6).any() else None)
seen = []
for s in samples:
    key = s[1:]
    if not seen or seen[-1][1] != key: seen.append((round(s[0] - ts[0], 4), key))
print("nvml states (t rel. to first step, (sm, mem, gr, W, pstate, reasons, memutil)):", seen[:30])
