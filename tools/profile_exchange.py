#!/usr/bin/env python
"""Direct exchange with W emulated ranks on ONE GPU (contexts of one process, windows connected
by plain pointers): kernel-inherent cost of stage 1 / stage 2 without NVLink.
Usage: python tools/profile_exchange.py [events_per_rank] [world] [steps]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import chronos_b200  # noqa: E402
from chronos_b200 import _ffi, tracegen  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 12_500_000
world = int(sys.argv[2]) if len(sys.argv) > 2 else 2
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 4
spec = tracegen.WORKLOADS["sharded_100m_1m"]
devs = [tracegen.generate_device(spec, device=0, first=r * n, n=n) for r in range(world)]
cap = max(n // (2 * world), 1024)
ctxs = [chronos_b200.Context(device=0, max_batch_events=max(n, world * cap)) for _ in range(world)]
wins = [c.exchange_create(world, r, cap)[0] for r, c in enumerate(ctxs)]
for c in ctxs:
    c.exchange_connect(world, windows=wins)
for it in range(steps):
    for c in ctxs:
        c.reset()
    for r, c in enumerate(ctxs):
        c.filter_exchange_ptr(devs[r].ptr, n, r * n)
    # ranks share ONE GPU here: all stage 1s must have finished before a stage 2 starts to spin
    # (see tests/test_gpu_exchange.py)
    assert _ffi.lib().chronos_device_synchronize(0) == 0
    for c in ctxs:
        c.submit_exchanged()
    for r, c in enumerate(ctxs):
        k = c.fetch_counts()
        t, f = c.last_timing()
        st = c.last_stage_times()
        print(f"it{it} rank{r}: {k['n_survivors']} received, {k['n_windows']} windows, library {t:.3f} ms, filter {f:.3f} ms, "
              f"pack {st['pack']:.3f} ms, launches {c.last_launches()}", flush=True)
