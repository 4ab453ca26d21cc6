"""Throughput with one vs two batches in flight (device-resident input, state carried)."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import chronos_b200
from chronos_b200 import tracegen, _ffi
n = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
K = int(sys.argv[2]) if len(sys.argv) > 2 else 20
name = sys.argv[3] if len(sys.argv) > 3 else "uniform_10m_4k"
spec = tracegen.WORKLOADS[name]
bufs = [tracegen.generate_device(spec, device=0, first=i * n, n=n) for i in range(2)]
L = _ffi.lib()
for mode in ("serial", "pipelined"):
    ctx = chronos_b200.Context(device=0, max_batch_events=n)
    for rep in range(2):
        ctx.reset()
        L.chronos_device_synchronize(0)
        t0 = time.perf_counter()
        if mode == "serial":
            for k in range(K):
                ctx.submit_ptr(bufs[k & 1].ptr, n, k * n)
                c = ctx.fetch_counts()
        else:
            ctx.submit_ptr(bufs[0].ptr, n, 0)
            for k in range(K):
                if k + 1 < K:
                    ctx.submit_ptr(bufs[(k + 1) & 1].ptr, n, (k + 1) * n)
                c = ctx.fetch_counts()
        L.chronos_device_synchronize(0)
        dt = (time.perf_counter() - t0) / K
    print(f"{name} n={n} {mode:10s} {1e3*dt:.4f} ms/batch  {n/dt/1e9:.2f} G ev/s  last: S={c['n_survivors']} W={c['n_windows']} P={c['n_pending']}", flush=True)
    ctx.close()
