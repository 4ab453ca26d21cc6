#!/usr/bin/env python
"""Where the end-to-end (host buffers in, host results out) step spends its time."""
import ctypes, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import chronos_b200
from chronos_b200 import _ffi, tracegen
sys.path.insert(0, ROOT)
from bench import Pinned, REC

L = _ffi.lib()
n = 10_000_000
dev = tracegen.generate_device(tracegen.WORKLOADS["uniform_10m_4k"], device=0, n=n)
host = Pinned(n * REC)
assert L.chronos_memcpy_d2h(0, ctypes.c_void_p(host.ptr), ctypes.c_void_p(dev.ptr), n * REC) == 0
ctx = chronos_b200.Context(device=0, max_batch_events=n)
ctx.submit_ptr(host.ptr, n, 0); c = ctx.fetch_counts()
S, W = c["n_survivors"], c["n_windows"]
r_idx, r_meta, r_win, r_rec = Pinned(S * 8), Pinned(S), Pinned(W * 32), Pinned(S * REC)
h = ctx._h
for it in range(4):
    t = [time.perf_counter()]
    ctx.reset(); t.append(time.perf_counter())
    ctx.submit_ptr(host.ptr, n, 0); t.append(time.perf_counter())
    c = ctx.fetch_counts(); t.append(time.perf_counter())
    L.chronos_copy_survivors(h, ctypes.c_void_p(r_idx.ptr), ctypes.c_void_p(r_meta.ptr), S); t.append(time.perf_counter())
    L.chronos_copy_windows(h, ctypes.c_void_p(r_win.ptr), W); t.append(time.perf_counter())
    L.chronos_copy_window_records(h, ctypes.c_void_p(r_rec.ptr), S); t.append(time.perf_counter())
    names = ["reset", "submit", "fetch", "copy_surv", "copy_win", "copy_winrec"]
    print(" ".join(f"{nm}={1e3*(b-a):.2f}" for nm, a, b in zip(names, t[:-1], t[1:])), f"total={1e3*(t[-1]-t[0]):.2f} ms", flush=True)
