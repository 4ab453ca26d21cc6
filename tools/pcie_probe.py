#!/usr/bin/env python
"""H2D / D2H bandwidth of pinned memory on this box (context for the e2e number)."""
import time, torch
n = 2_880_000_000
h = torch.empty(n, dtype=torch.uint8).pin_memory()
d = torch.empty(n, dtype=torch.uint8, device="cuda")
for name, fn in (("h2d", lambda: d.copy_(h, non_blocking=True)), ("d2h", lambda: h.copy_(d, non_blocking=True))):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / 3
    print(f"{name}: {n/dt/1e9:.1f} GB/s ({dt*1e3:.1f} ms for 2.88 GB)")
