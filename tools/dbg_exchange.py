import sys, os
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import numpy as np
import chronos_b200
from chronos_b200 import tracegen
world = int(sys.argv[1]); n_rows = 120000
rows = tracegen.generate_host(dict(tracegen.WORKLOADS["skewed_zipf_chains"]), n=n_rows)
half = n_rows // 2
cap_pair = half // 2
ctxs = [chronos_b200.Context(device=0, max_batch_events=max(half, world * cap_pair)) for _ in range(world)]
wins = [c.exchange_create(world, r, cap_pair)[0] for r, c in enumerate(ctxs)]
for c in ctxs:
    c.exchange_connect(world, windows=wins); c.exchange_set_timeout(3.0)
for step, (lo, hi) in enumerate([(0, half), (half, n_rows)]):
    batch = rows[lo:hi]; n = len(batch)
    b = [r * n // world for r in range(world + 1)]
    order = list(range(world)) if step == 0 else list(reversed(range(world)))
    for r in order: ctxs[r].filter_exchange(batch[b[r]:b[r+1]], lo + b[r])
    for r in order: ctxs[r].submit_exchanged()
    for r in range(world):
        try:
            c = ctxs[r].fetch_counts(); print("step", step, "rank", r, c["n_survivors"], c["n_windows"], flush=True)
        except Exception as e:
            print("step", step, "rank", r, "ERR", e, flush=True)
