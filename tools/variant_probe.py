#!/usr/bin/env python
"""Step and stage times of the N=1 bench workload for the library named by CHRONOS_LIB (and
whatever CHRONOS_* knobs are set). Usage: python tools/variant_probe.py [steps] [events] [workload]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import chronos_b200  # noqa: E402
from chronos_b200 import tracegen  # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 60
n = int(sys.argv[2]) if len(sys.argv) > 2 else 10_000_000
name = sys.argv[3] if len(sys.argv) > 3 else "uniform_10m_4k"
dev = tracegen.generate_device(tracegen.WORKLOADS[name], device=0, first=0, n=n)
ctx = chronos_b200.Context(device=0, max_batch_events=n)
ctx.set_overlap(0)
tot, stg = [], []
for i in range(steps + 10):
    ctx.reset()
    ctx.submit_ptr(dev.ptr, n, 0)
    c = ctx.fetch_counts()
    if i >= 10:
        tot.append(ctx.last_timing()[0])
        stg.append(list(ctx.last_stage_times().values()))
tot = np.array(tot); stg = np.array(stg)
knobs = {k: v for k, v in os.environ.items() if k.startswith("CHRONOS_")}
print("%-28s step mean %.4f median %.4f min %.4f | filter %.4f compact %.4f sort %.4f windows+gather %.4f | %s %s" % (
    os.path.basename(knobs.pop("CHRONOS_LIB", "product")), tot.mean(), np.median(tot), tot.min(),
    stg[:, 0].mean(), stg[:, 1].mean(), stg[:, 4].mean(), stg[:, 5].mean(), knobs, c), flush=True)
