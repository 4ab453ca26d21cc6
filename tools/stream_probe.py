"""Where does a 64 K-event micro-batch spend its time? host-side submit / fetch and the library's stage times."""
import os, sys, time
sys.path.insert(0, os.getcwd())
import numpy as np
import chronos_b200
from chronos_b200 import tracegen
B = 65536; ring = 64
dev = tracegen.generate_device(tracegen.WORKLOADS["stream_64k"], device=0, first=0, n=ring * B)
ctx = chronos_b200.Context(device=0, max_batch_events=B)
ts, tf, tot, st = [], [], [], []
for i in range(600):
    t0 = time.perf_counter()
    ctx.submit_ptr(dev.ptr + (i % ring) * B * 288, B, i * B)
    t1 = time.perf_counter()
    c = ctx.fetch_counts()
    t2 = time.perf_counter()
    if i >= 100:
        ts.append(t1 - t0); tf.append(t2 - t1); tot.append(ctx.last_timing()[0]); st.append(ctx.last_stage_times())
med = lambda a: 1e3 * float(np.median(a))
print(f"fused_max={os.environ.get('CHRONOS_FUSED_MAX','default')}: submit {med(ts):.4f} ms  fetch {med(tf):.4f} ms  sum {med(np.array(ts)+np.array(tf)):.4f} ms | library total {float(np.median(tot)):.4f} ms, stages " +
      ", ".join(f"{k} {np.median([s[k] for s in st]):.4f}" for k in st[0]) + f" launches {ctx.last_launches()}")
