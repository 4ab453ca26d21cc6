#!/usr/bin/env python
"""bench.py -- events/sec through the CHRONOS hot path (filter + per-PID grouping +
window/flush logic + compaction) on N B200s, with the roofline of the dominant kernel
and the reference-style CPU path timed beside it.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

One "step" = one pass of the whole hot path over one batch of synthetic records:
  N=1  : BASELINE.json configs[1]  -- 10M events, 4K PIDs, 256-byte path fields
  N>1  : BASELINE.json configs[2]  -- 12.5M events per GPU (100M at N=8), 1M PIDs,
         contiguous slice per rank, survivors repartitioned by PID hash (one all-to-all)
`value` is measured with the batch already resident in HBM; `e2e` goes through the
public host API with pinned HOST buffers (H2D of the batch and D2H of the results inside
the timed region). The input (2.88 GB per step) is far larger than the 126 MB L2, so no
explicit L2 flush is needed between steps (stated in config.l2).
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

REC = 288
METRIC = "events/sec filtered+chain-matched"
UNIT = "events/s"


def measured_peak_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


# ------------------------------------------------------------------------------------------
# clocks sampling during the timed region
# ------------------------------------------------------------------------------------------
class ClockSampler:
    """SM clock + throttle reasons sampled every few ms DURING the timed region (NVML in a
    thread; the timed region of a default run is only tens of milliseconds long, far too
    short for an `nvidia-smi -lms` loop to land a sample in it)."""

    def __init__(self, index, period_s=0.002):
        self.index, self.period = index, period_s
        self.sm, self.mem, self.power, self.temp, self.reasons = [], [], [], [], set()
        self.sm_max = None
        self._stop = threading.Event()
        self._t = None
        self._h = None
        try:
            import pynvml

            pynvml.nvmlInit()
            self._nv = pynvml
            self._h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.sm_max = float(pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self._h = None

    def _sample(self):
        nv = self._nv
        try:
            self.sm.append(float(nv.nvmlDeviceGetClockInfo(self._h, nv.NVML_CLOCK_SM)))
            self.mem.append(float(nv.nvmlDeviceGetClockInfo(self._h, nv.NVML_CLOCK_MEM)))
            self.power.append(nv.nvmlDeviceGetPowerUsage(self._h) / 1e3)
            self.temp.append(float(nv.nvmlDeviceGetTemperature(self._h, nv.NVML_TEMPERATURE_GPU)))
            get = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or nv.nvmlDeviceGetCurrentClocksThrottleReasons
            r = get(self._h)
            for bit, name in ((nv.nvmlClocksEventReasonHwSlowdown, "hw_slowdown"),
                              (nv.nvmlClocksEventReasonHwThermalSlowdown, "hw_thermal_slowdown"),
                              (nv.nvmlClocksEventReasonSwThermalSlowdown, "sw_thermal_slowdown"),
                              (nv.nvmlClocksEventReasonSwPowerCap, "sw_power_cap")):
                if r & bit:
                    self.reasons.add(name)
        except Exception:
            pass

    def _run(self):
        while not self._stop.is_set():
            self._sample()
            time.sleep(self.period)

    def start(self):
        if self._h is None:
            return
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()

    def stop(self):
        self._stop.set()
        if self._t is not None:
            self._t.join(timeout=1)
        if self._h is not None and not self.sm:
            self._sample()
        return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": self.sm_max,
                "mem_mhz": float(np.median(self.mem)) if self.mem else None,
                "power_w_max": float(max(self.power)) if self.power else None,
                "temp_c_max": float(max(self.temp)) if self.temp else None,
                "reasons": sorted(self.reasons), "samples": len(self.sm), "source": "nvml, sampled inside the timed region"}


# ------------------------------------------------------------------------------------------
# CPU arm: the reference's own path on the host cores (oracle/_ref C filter + Python port
# of print_event, one worker per core, records partitioned by pid -- exact, PIDs are
# independent: chronos_sensor.py:105,138,142,157)
# ------------------------------------------------------------------------------------------
_CPU_PARTS = None


def _cpu_worker(k):
    from oracle import pyoracle

    part, drop = _CPU_PARTS[k]
    t0 = time.perf_counter()
    s = pyoracle.PySensor()
    s.feed(part, 0, drop=drop)
    return time.perf_counter() - t0, len(s.survivors), len(s.windows)


def cpu_reference_rate(rows, cores, repeats=1):
    """Returns (events/s, seconds, kind). rows: (n, 288) uint8 sample."""
    global _CPU_PARTS
    import multiprocessing as mp

    from oracle import build_oracle

    n = len(rows)
    ref_so = build_oracle.build_ref()
    pid = rows[:, :4].copy().view("<u4").reshape(-1)
    parts = []
    t_filter = 0.0
    for k in range(cores):
        sel = rows[(pid % cores) == k]
        buf = sel.tobytes()
        drop = None
        if ref_so is not None:  # stage A: the reference's own C helpers (chronos_sensor.py:28-47)
            lib = ctypes.CDLL(ref_so)
            lib.chronos_ref_filter_batch.argtypes = [ctypes.c_char_p, ctypes.c_uint64, ctypes.c_char_p]
            out = ctypes.create_string_buffer(max(len(sel), 1))
            t0 = time.perf_counter()
            lib.chronos_ref_filter_batch(buf, len(sel), out)
            t_filter = max(t_filter, time.perf_counter() - t0)
            drop = out.raw[:len(sel)]
        parts.append((buf, drop))
    _CPU_PARTS = parts
    best = None
    ctx = mp.get_context("fork")
    for _ in range(repeats):
        with ctx.Pool(cores) as pool:
            res = pool.map(_cpu_worker, range(cores))
        t = max(r[0] for r in res) + t_filter
        best = t if best is None else min(best, t)
    kind = "port"
    return n / best, best, kind, ("oracle/_ref C filter + " if ref_so else "") + "oracle.pyoracle.PySensor"


def cpu_sample_rows(spec, n):
    """The sample the CPU arm times, from the host-only generator (g++ build of the same
    source as the CUDA twin): the reference arm never maps libchronos_b200.so."""
    from chronos_b200 import tracegen

    return tracegen.generate_host_nocuda(spec, first=0, n=n)


CPU_CLOCK_NOTE = ("clock = slowest worker's feed() + the C filter pass; partitioning the sample by pid and starting "
                  "the worker pool are outside it (generous to the CPU arm)")


def run_reference_arm(args, rank, world):
    if rank != 0:
        return
    from chronos_b200 import tracegen

    spec = tracegen.WORKLOADS["uniform_10m_4k" if world == 1 else "sharded_100m_1m"]
    cores = os.cpu_count() or 1
    # bounded sample: ~0.5M raw events/s/core for this path -> about 3 s of work per step
    n_sample = int(min(spec["n"], 1_000_000 * cores))
    rows = cpu_sample_rows(spec, n_sample)
    times = []
    for i in range(args.warmup + args.steps):
        rate, secs, kind, what = cpu_reference_rate(rows, cores)
        if i >= args.warmup:
            times.append(secs)
    ms = 1e3 * float(np.mean(times))
    value = n_sample / (ms / 1e3)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8", "data": "synthetic",
        "config": {"workload": workload_name(world), "events_per_gpu": events_per_gpu(world), "record_bytes": REC,
                   "sample_events_per_step": n_sample},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind,
                         "sample": f"first {n_sample} events of the workload, partitioned by pid over {cores} processes; "
                                   f"{what}; {CPU_CLOCK_NOTE}"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def events_per_gpu(world):
    return 10_000_000 if world == 1 else 12_500_000


def workload_name(world):
    if world == 1:
        return "configs[1]: 10M synthetic events, 4K PIDs, 256-byte path fields, 1xB200 (uniform_10m_4k)"
    return (f"configs[2]: {12.5 * world:g}M events ({'100M' if world == 8 else 'scaled'}), 1M PIDs, PID-hash sharded "
            f"across {world}xB200, survivors repartitioned all-to-all over NVLink (sharded_100m_1m)")


# ------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------
class Pinned:
    def __init__(self, nbytes):
        from chronos_b200 import _ffi

        self.nbytes = int(nbytes)
        p = ctypes.c_void_p()
        rc = _ffi.lib().chronos_host_alloc_pinned(max(self.nbytes, 16), ctypes.byref(p))
        if rc != 0:
            raise RuntimeError(f"pinned allocation of {nbytes} bytes failed ({rc})")
        self.ptr = p.value
        self.array = np.ctypeslib.as_array((ctypes.c_uint8 * max(self.nbytes, 16)).from_address(self.ptr))

    def free(self):
        from chronos_b200 import _ffi

        if self.ptr:
            self.array = None
            _ffi.lib().chronos_host_free_pinned(ctypes.c_void_p(self.ptr))
            self.ptr = None


def _same_pairs(pid_a, arr_a, pid_b, arr_b):
    """Two (pid, arrival) lists hold the same pairs (order ignored)."""
    if len(pid_a) != len(pid_b):
        return False
    ia, ib = np.lexsort((arr_a, pid_a)), np.lexsort((arr_b, pid_b))
    return bool(np.array_equal(pid_a[ia], pid_b[ib]) and np.array_equal(arr_a[ia], arr_b[ib]))


def _gather_to_rank0(dist, group, obj, rank, world):
    out = [None] * world if rank == 0 else None
    dist.gather_object(obj, out, dst=0, group=group)
    return out


def verify_sharded(chronos_b200, tracegen, runner, spec, n, device, rank, world, dist, gloo):
    """ONE verification step on the real ranks before anything is timed (N > 1).

    Every rank runs the step the bench times (its contiguous slice, survivors delivered to their
    owners over NVLink, window logic on the owner) and sends rank 0 what it produced: windows
    (trigger index, pid, length, label), a GPU-computed digest of every window's records in
    canonical form, the surviving arrival indices and the pending (pid, arrival) pairs. Rank 0 runs
    the WHOLE world x n batch through one single-GPU Context -- the path the -m gpu tests hold
    bit-exact to the oracle -- and compares: merged by trigger index, the sharded output must be
    the single-GPU output (chronos_sensor.py:105,138,141-142,157: PIDs are independent, so the
    shard must be unobservable)."""
    from chronos_b200.engine import WINDOW_DTYPE

    t0 = time.perf_counter()
    c = runner.step()
    W, S, P = c["n_windows"], c["n_survivors"], c["n_pending"]
    ctx = runner.ctx
    win = np.zeros(W, dtype=WINDOW_DTYPE)
    ctx._check(ctx._L.chronos_copy_windows(ctx._h, win.ctypes.data, W))
    dig = ctx.window_digests(W)
    surv = np.zeros(S, dtype=np.uint64)
    ctx._check(ctx._L.chronos_copy_survivors(ctx._h, surv.ctypes.data, None, S))
    ppid, parr, _ = ctx.state_arrays(P, records=False)
    mine = {"trig": win["trigger_idx"].copy(), "pid": win["pid"].copy(), "len": win["len"].copy(),
            "flag": (win["flags"] & 1).astype(np.uint8), "dig": dig, "surv": surv, "ppid": ppid, "parr": parr}
    got = _gather_to_rank0(dist, gloo, mine, rank, world)
    res = None
    if rank == 0:
        total = n * world
        whole = tracegen.generate_device(spec, device=device, first=0, n=total)
        ref = chronos_b200.Context(device=device, max_batch_events=total, max_pending_events=0)
        ref.submit_ptr(whole.ptr, total, 0)
        rc = ref.fetch_counts()
        rw = np.zeros(rc["n_windows"], dtype=WINDOW_DTYPE)
        ref._check(ref._L.chronos_copy_windows(ref._h, rw.ctypes.data, rc["n_windows"]))
        rdig = ref.window_digests(rc["n_windows"])
        rs = np.zeros(rc["n_survivors"], dtype=np.uint64)
        ref._check(ref._L.chronos_copy_survivors(ref._h, rs.ctypes.data, None, rc["n_survivors"]))
        rppid, rparr, _ = ref.state_arrays(rc["n_pending"], records=False)
        ref.close()
        whole.free()
        cat = lambda k: np.concatenate([g[k] for g in got])  # noqa: E731
        order = np.argsort(cat("trig"), kind="stable")
        checks = {
            "window_count": int(len(order)) == rc["n_windows"],
            "trigger_order": np.array_equal(cat("trig")[order], rw["trigger_idx"]),
            "window_pid": np.array_equal(cat("pid")[order], rw["pid"]),
            "window_len": np.array_equal(cat("len")[order], rw["len"]),
            "window_label": np.array_equal(cat("flag")[order], (rw["flags"] & 1).astype(np.uint8)),
            "window_record_digests": np.array_equal(cat("dig")[order], rdig),
            "survivors": np.array_equal(np.sort(cat("surv")), rs),
            "pending": _same_pairs(cat("ppid"), cat("parr"), rppid, rparr),
            "owner_of_every_window": all(bool(np.all(np.array([chronos_b200.owner_of(int(p), world) for p in g["pid"][:2000]]) == r))
                                         for r, g in enumerate(got)),
        }
        res = {"checked": True, "ok": bool(all(checks.values())), "checks": {k: bool(v) for k, v in checks.items()},
               "windows": int(rc["n_windows"]), "survivors": int(rc["n_survivors"]), "pending": int(rc["n_pending"]),
               "events": int(total), "seconds": None,
               "reference": "one single-GPU Context over the whole batch on rank 0 (that path is held bit-exact to the "
                            "oracle by tests/test_gpu_full_size.py); window records compared through "
                            "chronos_window_digests (canonical bytes)",
               "received_per_rank": [int(len(g["surv"])) for g in got]}
        res["seconds"] = round(time.perf_counter() - t0, 2)
    box = [res]
    dist.broadcast_object_list(box, src=0, group=gloo)
    return box[0]


def same_slice_single_gpu(chronos_b200, dev_ptr, n, device, steps=5):
    """The rank's own slice through the single-GPU path (no exchange): the per-GPU time the
    sharded step is compared with. Mean library time per step in ms."""
    ctx = chronos_b200.Context(device=device, max_batch_events=n)
    ms = []
    for i in range(steps + 2):
        ctx.reset()
        ctx.submit_ptr(dev_ptr, n, 0)
        ctx.fetch_counts()
        if i >= 2:
            ms.append(ctx.last_timing()[0])
    ctx.close()
    return float(np.mean(ms))


def skew_leg(chronos_b200, tracegen, device, rank, world, dist, gloo, n, steps=5):
    """BASELINE.json configs[3] across the GPUs: 95 % noise, PIDs ~ Zipf(1.1) over 4096 PIDs, so a
    few owners receive most survivors. Per rank: survivors received and the time of its stage 2."""
    from chronos_b200.sharded import ShardedRunner

    spec = tracegen.WORKLOADS["skewed_zipf_chains"]
    first = rank * n
    dev = tracegen.generate_device(spec, device=device, first=first, n=n)
    r = ShardedRunner(dev.ptr, n, first, device, rank, world, cap=n // 4, cap_entries_per_peer=max(n // 8, 4096))
    tot, st2, recv, last = [], [], 0, None
    for i in range(steps + 2):
        dist.barrier()
        last = r.step()
        if i >= 2:
            tot.append(last["gpu_ms"])
            s = r.ctx.last_stage_times()
            st2.append(s["wait_unpack"] + s["sort"] + s["windows_gather"])
            recv = last["n_survivors"]
    mine = {"received": int(recv), "step_ms": float(np.mean(tot)), "stage2_ms": float(np.mean(st2)),
            "windows": int(last["n_windows"])}
    got = _gather_to_rank0(dist, gloo, mine, rank, world)
    dist.barrier()
    r.ctx.close()
    dev.free()
    if rank != 0:
        return None
    rec = [g["received"] for g in got]
    return {"workload": f"configs[3] sharded: skewed_zipf_chains, {n} events per GPU, PIDs ~ Zipf(1.1) over 4096 PIDs",
            "received_per_rank": rec, "imbalance_max_over_mean": float(max(rec) / max(np.mean(rec), 1.0)),
            "step_ms_per_rank": [round(g["step_ms"], 4) for g in got],
            "stage2_ms_per_rank": [round(g["stage2_ms"], 4) for g in got],
            "step_ms_max": float(max(g["step_ms"] for g in got)),
            "events_per_s": float(n * world / (max(g["step_ms"] for g in got) / 1e3)),
            "windows_total": int(sum(g["windows"] for g in got))}


def pipelined_leg(chronos_b200, dev_ptr, n, device, steps, serial_survivors):
    """Same batch size, TWO batches in flight: submit(k+1) is enqueued before fetch(k), so the
    filter of batch k+1 runs beside the window stage of batch k and the host never idles the
    GPU. Every batch's work is inside the timed region and every batch is fetched; the state
    carries from batch to batch (no reset), which only adds the carried-in pending entries."""
    from chronos_b200 import _ffi

    L = _ffi.lib()
    ctx = chronos_b200.Context(device=device, max_batch_events=n)
    out = {}
    for mode in ("one_in_flight", "two_in_flight"):
        ctx.reset()
        best = None
        for rep in range(2):
            L.chronos_device_synchronize(device)
            t0 = time.perf_counter()
            surv = []
            if mode == "one_in_flight":
                for k in range(steps):
                    ctx.submit_ptr(dev_ptr, n, k * n)
                    surv.append(ctx.fetch_counts()["n_survivors"])
            else:
                ctx.submit_ptr(dev_ptr, n, 0)
                for k in range(steps):
                    if k + 1 < steps:
                        ctx.submit_ptr(dev_ptr, n, (k + 1) * n)
                    surv.append(ctx.fetch_counts()["n_survivors"])
            L.chronos_device_synchronize(device)
            dt = (time.perf_counter() - t0) / steps
            best = dt if best is None else min(best, dt)
            assert all(s == serial_survivors for s in surv), "pipelined batches must report the same survivors"
        out[mode] = {"ms_per_batch": 1e3 * best, "events_per_s": n / best}
    ctx.close()
    out["clock"] = "host wall clock around the whole loop (device idle before and after), state carried, best of 2"
    return out


def run_gpu_arm(args, rank, world, local_rank):
    import torch

    import chronos_b200
    from chronos_b200 import _ffi, tracegen

    dist = None
    if world > 1:
        import torch.distributed as dist_mod

        dist = dist_mod
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        gloo = dist.new_group(backend="gloo")  # host-side objects (parity check, per-rank reports)
    device = local_rank
    torch.cuda.set_device(device)
    L = _ffi.lib()
    if L.chronos_device_ok(device) != 0:
        raise SystemExit("chronos_b200 needs a B200 (sm_100); there is no CPU fallback")

    if world == 1:
        spec = tracegen.WORKLOADS["uniform_10m_4k"]
        n = int(args.events or spec["n"])
        first = 0
    else:
        spec = tracegen.WORKLOADS["sharded_100m_1m"]
        n = int(args.events or 12_500_000)
        first = rank * n
    dev = tracegen.generate_device(spec, device=device, first=first, n=n)
    peak, peak_src = measured_peak_gbs()

    if world == 1:
        runner = SingleGpuRunner(chronos_b200, dev, n, device)
    else:
        from chronos_b200.sharded import ShardedRunner

        runner = ShardedRunner(dev.ptr, n, first, device, rank, world)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    parity = {"checked": False, "why": "N=1: the -m gpu tests compare this path with the oracle at the full 10 M events"}
    if world > 1 and not args.no_parity:
        parity = verify_sharded(chronos_b200, tracegen, runner, spec, n, device, rank, world, dist, gloo)
        barrier()

    for _ in range(args.warmup):
        runner.step()
    # The W warm-up steps of a default run last a few milliseconds, far less than the GPU needs to
    # reach its sustained state after sitting idle: keep stepping (untimed) until the step time
    # has settled -- the last 5 steps within 2 % of each other -- or 200 extra steps have run.
    extra, recent = 0, []
    settled = torch.zeros(1, dtype=torch.int32, device=f"cuda:{device}")
    while extra < 200:
        recent = (recent + [runner.step()["gpu_ms"]])[-5:]
        extra += 1
        ok = len(recent) == 5 and max(recent) <= 1.02 * min(recent)
        if dist is not None:  # every rank takes the same number of steps
            settled[0] = 1 if ok else 0
            dist.all_reduce(settled, op=dist.ReduceOp.MIN)
            ok = bool(settled.item())
        if ok:
            break
    barrier()
    sampler = ClockSampler(device)
    if rank == 0 and not os.environ.get("CHRONOS_BENCH_NO_SAMPLER"):
        sampler.start()
    ext = torch.cuda.ExternalStream(runner.stream_ptr(), device=torch.device("cuda", device))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    filt_ms, step_ms, launches, last = [], [], 0, None
    stage_ms = []
    e0.record(ext)
    t_wall0 = time.perf_counter()
    for last in runner.timed_steps(args.steps):
        filt_ms.append(last["filter_ms"])
        step_ms.append(last["gpu_ms"])
        stage_ms.append(last["stages"])
        launches += last["launches"]
    e1.record(ext)
    e1.synchronize()
    barrier()
    t_wall = time.perf_counter() - t_wall0
    dev_ms = e0.elapsed_time(e1)
    clocks = sampler.stop() if rank == 0 else None
    if dist is not None:
        t = torch.tensor([dev_ms], device=f"cuda:{device}", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dev_ms = float(t.item())
        lt = torch.tensor([launches], device=f"cuda:{device}", dtype=torch.int64)
        dist.all_reduce(lt)
        launches = int(lt.item())
    ms_per_step = dev_ms / args.steps
    total_events = n * world
    value = total_events / (ms_per_step / 1e3)

    # ---- end-to-end through the public host API with pinned host buffers -------------------
    e2e = runner.e2e(max(1, min(args.steps, 10)), barrier, dist)

    # ---- CPU baseline on a bounded sample (rank 0, N=1 only) -------------------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cores = os.cpu_count() or 1
        n_sample = int(min(n, 1_000_000 * cores))
        rows = dev.download(n_sample * REC).reshape(-1, REC)
        rate, secs, kind, what = cpu_reference_rate(rows, cores)
        n_one = int(min(n_sample, 500_000))
        rate1, secs1, _, _ = cpu_reference_rate(rows[:n_one], 1)
        cpu = {"value": rate, "unit": UNIT, "cores": cores, "kind": kind,
               "sample": f"first {n_sample} events of the workload in {secs:.2f}s, partitioned by pid over {cores} processes; "
                         f"{what}; {CPU_CLOCK_NOTE}",
               "one_core": {"value": rate1, "unit": UNIT, "sample": f"first {n_one} events in {secs1:.2f}s, one process"}}

    streaming = None
    pipelined = None
    if rank == 0 and world == 1 and not args.no_stream:
        streaming = stream_latency(chronos_b200, device)
        pipelined = pipelined_leg(chronos_b200, dev.ptr, n, device, max(args.steps, 10), last["n_survivors"])

    # ---- per-rank view of the sharded step, the same slice on one GPU, the skewed workload ----
    stages = {k: float(np.mean([s[k] for s in stage_ms])) for k in stage_ms[0]} if stage_ms else None
    per_rank = same_slice = skew = None
    if world > 1:
        same_slice = same_slice_single_gpu(chronos_b200, dev.ptr, n, device)
        mine = {"stages_ms": stages, "step_ms": float(np.mean(step_ms)), "received": int(last["n_survivors"]),
                "single_gpu_same_slice_ms": same_slice}
        per_rank = _gather_to_rank0(dist, gloo, mine, rank, world)
        barrier()
        if not args.no_skew:
            skew = skew_leg(chronos_b200, tracegen, device, rank, world, dist, gloo, n)
            barrier()

    if rank == 0:
        traffic, traffic_src = ncu_traffic()
        f_ms = float(np.mean(filt_ms))
        alg_bytes = REC * n  # the filter reads every raw record once (288 B per event)
        achieved = alg_bytes / (f_ms / 1e3) / 1e9
        S, W = last["n_survivors"], last["n_windows"]
        pipe_bytes = REC * n + REC * S + 16 * W
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u8", "data": "synthetic",
            "config": {"workload": workload_name(world), "events_per_gpu": n, "record_bytes": REC,
                       "survivors_rank0": S, "windows_rank0": W, "pending_rank0": last["n_pending"],
                       "single_gpu_same_slice_ms": same_slice,
                       "l2": "input per step (2.88 GB+) exceeds the 126 MB L2; no flush needed",
                       "timing": "torch CUDA events on the library stream around all K steps, max over ranks",
                       "steps_on_gpu": ("one batch at a time (a batch starts when the previous one has ended); the host "
                                        "enqueues one step ahead"),
                       "warmup_extra_steps": extra,
                       "exchange": (None if world == 1 else
                                    ("direct: stage 1 writes survivors into the owners' HBM over NVLink (CUDA IPC windows), "
                                     "stage 2 waits for the peers' flags on the GPU" if getattr(runner, "mode", "") == "direct"
                                     else "torch.distributed.all_to_all_single (NCCL) of the transport form"))},
            "roofline": {"bound": "hbm", "kernel": "filter_kernel", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "frac_of_8TBs_nominal": achieved / 8000.0,
                         "traffic": traffic if (world == 1 and n == 10_000_000) else None, "traffic_source": traffic_src,
                         "peak_source": peak_src,
                         "kernel_ms": f_ms, "algorithmic_bytes_per_launch": alg_bytes,
                         "pipeline": {"algorithmic_bytes_per_gpu": pipe_bytes, "ms_per_step": ms_per_step,
                                      "achieved": pipe_bytes / (ms_per_step / 1e3) / 1e9,
                                      "frac": pipe_bytes / (ms_per_step / 1e3) / 1e9 / peak,
                                      "frac_of_8TBs_nominal": pipe_bytes / (ms_per_step / 1e3) / 1e9 / 8000.0,
                                      "library_event_ms": float(np.mean(step_ms)),
                                      "library_event_ms_median": float(np.median(step_ms)),
                                      "library_event_ms_min": float(np.min(step_ms)),
                                      "filter_ms_per_step": [round(float(x), 4) for x in filt_ms[:64]]}},
            "cpu_baseline": cpu,
            "e2e": e2e,
            "parity": parity,
            "stages_ms_rank0": stages,
            "per_rank": per_rank,
            "skew": skew,
            "streaming": streaming,
            "pipelined": pipelined,
            "gpu_launches": launches,
            "clocks": clocks,
            "wall_ms_per_step": 1e3 * t_wall / args.steps,
        }
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    if parity.get("checked") and not parity.get("ok"):
        raise SystemExit("bench.py: the sharded output differs from the single-GPU output (see parity.checks)")


def ncu_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum of one filter_kernel launch on the N=1 workload,
    read from the newest committed `ncu --set full` raw page (profiles/r*_filter_ncu_raw.csv).
    Returns (bytes or None, file name)."""
    import csv
    import glob

    for path in sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_filter_ncu_raw.csv")), reverse=True):
        try:
            rows = list(csv.reader(open(path)))
            hdr = next(r for r in rows if "dram__bytes_read.sum" in r)
            units = rows[rows.index(hdr) + 1]
            val = rows[rows.index(hdr) + 2]
            scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
            tot = 0.0
            for name in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                i = hdr.index(name)
                tot += float(val[i].replace(",", "")) * scale[units[i]]
            return int(tot), os.path.basename(path)
        except Exception:
            continue
    return None, None


class SingleGpuRunner:
    def __init__(self, chronos_b200, dev, n, device):
        self.ctx = chronos_b200.Context(device=device, max_batch_events=n)
        self.dev, self.n, self.device = dev, n, device
        self.mod = chronos_b200

    def stream_ptr(self):
        from chronos_b200 import _ffi

        return int(_ffi.lib().chronos_stream(self.ctx._h))

    def step(self):
        self.ctx.reset()  # every step is the same one-shot batch on a fresh process_memory
        self.ctx.submit_ptr(self.dev.ptr, self.n, 0)
        c = self.ctx.fetch_counts()
        t, f = self.ctx.last_timing()
        c.update(gpu_ms=t, filter_ms=f, launches=self.ctx.last_launches())
        return c

    def timed_steps(self, steps):
        """K steps for the timed region. Every step is the same one-shot batch on a fresh
        process_memory (chronos_reset), and the GPU runs the batches strictly one after the other
        (chronos_set_overlap(0): a batch starts when the previous one has ended); only the HOST
        works ahead -- step k+1 is enqueued before step k is fetched, so the GPU does not sit idle
        while Python and the driver prepare the next launch sequence."""
        ctx = self.ctx
        ahead = not os.environ.get("CHRONOS_BENCH_NO_AHEAD")
        ctx.set_overlap(False)
        ctx.reset()
        ctx.submit_ptr(self.dev.ptr, self.n, 0)
        for k in range(steps):
            if not ahead and k > 0:
                ctx.reset()
                ctx.submit_ptr(self.dev.ptr, self.n, 0)
            if ahead and k + 1 < steps:
                ctx.reset()
                ctx.submit_ptr(self.dev.ptr, self.n, 0)
            c = ctx.fetch_counts()
            t, f = ctx.last_timing()
            c.update(gpu_ms=t, filter_ms=f, launches=ctx.last_launches(), stages=ctx.last_stage_times())
            yield c
        ctx.set_overlap(True)

    def e2e(self, steps, barrier, dist):
        """The same step through the public host API: the batch starts in pinned HOST memory
        (H2D inside submit) and survivors, windows and the windows' records end in pinned host
        memory. The download of step k runs on its own stream while step k+1's batch comes up
        (chronos_download_begin); every download has completed before the clock stops."""
        from chronos_b200 import _ffi

        L = _ffi.lib()
        n = self.n
        host = Pinned(n * REC)
        rc = L.chronos_memcpy_d2h(self.device, ctypes.c_void_p(host.ptr), ctypes.c_void_p(self.dev.ptr), n * REC)
        assert rc == 0
        # result buffers sized from a dry run
        self.ctx.reset()
        self.ctx.submit_ptr(host.ptr, n, 0)
        c = self.ctx.fetch_counts()
        cap = (c["n_survivors"] * (9 + REC) + c["n_windows"] * 32 + 4096 + 255) // 256 * 256
        out = [Pinned(cap), Pinned(cap)]
        def loop():
            barrier()
            t0 = time.perf_counter()
            lay = None
            for k in range(steps):
                self.ctx.reset()
                self.ctx.submit_ptr(host.ptr, n, 0)          # H2D of the whole batch inside
                c = self.ctx.fetch_counts()
                lay = self.ctx.download_begin(out[k & 1].ptr, cap)
            self.ctx.download_wait()
            barrier()
            return (time.perf_counter() - t0) / steps, lay, c, self.ctx.last_upload_bytes()

        # The host's memory system is shared with whatever else runs on the box, and these loops live
        # off it: each mode is timed twice and the better run counts (both are reported).
        runs_plain = [loop(), loop()]
        dt_plain, lay, c, up_plain = min(runs_plain, key=lambda r: r[0])
        # the same loop with the upload trimmed by the host's cores (chronos_set_host_trim): used when
        # there are cores to spare for it; both figures are reported, the better one is `value`
        threads = os.cpu_count() or 1
        dt_trim = up_trim = None
        runs_trim = []
        if threads >= 2 and not os.environ.get("CHRONOS_BENCH_NO_TRIM"):
            self.ctx.set_host_trim(threads)
            loop()
            runs_trim = [loop(), loop()]
            dt_trim, lay, c, up_trim = min(runs_trim, key=lambda r: r[0])
            self.ctx.set_host_trim(0)
        dt = min(dt_plain, dt_trim) if dt_trim else dt_plain
        # the last download must be what the blocking calls return
        idx, meta, win, rec = self.ctx.parse_download(out[(steps - 1) & 1].array, lay)
        assert len(idx) == c["n_survivors"] and len(win) == c["n_windows"] and int(win["len"].sum()) == len(rec)
        assert np.all(np.diff(win["trigger_idx"].astype(np.int64)) > 0) and np.all(np.diff(idx.astype(np.int64)) > 0)
        d2h = lay["total_bytes"]
        for p in [host] + out:
            p.free()
        trimmed = dt_trim is not None and dt_trim <= dt_plain
        return {"value": n / dt, "unit": UNIT, "h2d_bytes_per_step": int(up_trim if trimmed else up_plain),
                "d2h_bytes_per_step": int(d2h),
                "ms_per_step": 1e3 * dt, "steps": steps,
                "timing": "wall clock around K steps; each mode run twice, the better run reported",
                "plain_copy": {"value": n / dt_plain, "ms_per_step": 1e3 * dt_plain,
                               "ms_per_step_runs": [1e3 * r[0] for r in runs_plain]},
                "host_trim": (None if dt_trim is None else
                              {"value": n / dt_trim, "ms_per_step": 1e3 * dt_trim, "host_threads": threads,
                               "h2d_bytes_per_step": int(up_trim),
                               "ms_per_step_runs": [1e3 * r[0] for r in runs_trim],
                               "note": "host threads drop the NUL padding of the records they take (lossless, a kernel "
                                       "rebuilds them) while the copy engine moves the rest of the batch as it is"}),
                "upload": "host_trim" if trimmed else "plain_copy",
                "path": "Context.submit(pinned host records) -> fetch -> download_begin(pinned host results: survivors, "
                        "windows, window records); the download of step k overlaps the upload of step k+1, all "
                        "downloads complete inside the timed region"}


def _percentiles(lat):
    a = np.array(lat)
    return {"p50_ms": float(np.percentile(a, 50)), "p99_ms": float(np.percentile(a, 99)), "mean_ms": float(a.mean()),
            "min_ms": float(a.min()), "max_ms": float(a.max())}


def stream_latency(chronos_b200, device, batches=1000, batch_events=65_536, ring=64, warm=50):
    """BASELINE.json configs[4]: a stream of 64K-event micro-batches through ONE context whose
    per-PID state persists (process_memory, chronos_sensor.py:105); per-batch latency of
    submit -> fetch, wall clock around the blocking calls. Legs: batch already in HBM; batch in
    pinned host memory with the windows and their records read back (18.9 MB up); the HBM leg
    again with two batches in flight (throughput, not latency); and the same stream over the
    1 M-PID population of configs[2], where the pending state grows to several 100 K entries."""
    from chronos_b200 import _ffi, tracegen

    L = _ffi.lib()
    spec = tracegen.WORKLOADS["stream_64k"]
    dev = tracegen.generate_device(spec, device=device, first=0, n=ring * batch_events)
    host = Pinned(ring * batch_events * REC)
    assert L.chronos_memcpy_d2h(device, ctypes.c_void_p(host.ptr), ctypes.c_void_p(dev.ptr), ring * batch_events * REC) == 0
    r_win, r_rec = Pinned(batch_events * 32), Pinned(batch_events * REC)
    out = {"workload": "configs[4]: 64K-event micro-batches, persistent per-PID state (stream_64k)",
           "batch_events": batch_events, "batches": batches, "batch_bytes": batch_events * REC,
           "clock": "host wall clock around submit..fetch (blocking), one batch in flight; the first batches "
                    "are reported separately (first_batches) -- nothing is allocated by a submit"}
    for leg in ("device_resident", "host_buffers"):
        ctx = chronos_b200.Context(device=device, max_batch_events=batch_events)
        lat, first, windows, survivors = [], [], 0, 0
        for i in range(warm + batches):
            slot = i % ring
            src = (dev.ptr if leg == "device_resident" else host.ptr) + slot * batch_events * REC
            t0 = time.perf_counter()
            ctx.submit_ptr(src, batch_events, i * batch_events)
            c = ctx.fetch_counts()
            if leg == "host_buffers":
                W, S = c["n_windows"], c["n_survivors"]
                if W:
                    ctx._check(L.chronos_copy_windows(ctx._h, ctypes.c_void_p(r_win.ptr), W))
                    ctx._check(L.chronos_copy_window_records(ctx._h, ctypes.c_void_p(r_rec.ptr), batch_events))
            t1 = time.perf_counter()
            if i >= warm:
                lat.append(1e3 * (t1 - t0))
                windows += c["n_windows"]
                survivors += c["n_survivors"]
            else:
                first.append(1e3 * (t1 - t0))
        out[leg] = dict(_percentiles(lat), events_per_s=batch_events / (np.mean(lat) / 1e3), survivors=survivors,
                        windows=windows, pending_at_end=c["n_pending"],
                        first_batches={"n": warm, "max_ms": float(max(first)), "p50_ms": float(np.percentile(first, 50))})
        ctx.close()
    # two batches in flight: the host enqueues batch k+1 while batch k runs (throughput figure)
    ctx = chronos_b200.Context(device=device, max_batch_events=batch_events)
    for rep in range(2):
        ctx.reset()
        L.chronos_device_synchronize(device)
        t0 = time.perf_counter()
        ctx.submit_ptr(dev.ptr, batch_events, 0)
        for i in range(batches):
            if i + 1 < batches:
                ctx.submit_ptr(dev.ptr + ((i + 1) % ring) * batch_events * REC, batch_events, (i + 1) * batch_events)
            c = ctx.fetch_counts()
        L.chronos_device_synchronize(device)
        dt = (time.perf_counter() - t0) / batches
    out["device_resident_two_in_flight"] = {"ms_per_batch": 1e3 * dt, "events_per_s": batch_events / dt,
                                            "pending_at_end": c["n_pending"]}
    ctx.close()
    for b in (host, r_win, r_rec):
        b.free()
    dev.free()

    # the same stream over 1 M PIDs: every batch brings ~7 K new PIDs, the pending lists grow
    spec1m = tracegen.WORKLOADS["sharded_100m_1m"]
    nb = 600
    ctx = chronos_b200.Context(device=device, max_batch_events=batch_events, max_pending_events=4_000_000)
    lat, pend = [], []
    chunk = 100  # batches generated per refill of the device ring
    dev = tracegen.DeviceBuffer(device, chunk * batch_events * REC)
    cfg, _ = tracegen.make_cfg(spec1m)
    for i in range(nb):
        if i % chunk == 0:
            assert L.chronos_tracegen_device(device, ctypes.byref(cfg), None, i * batch_events, chunk * batch_events,
                                             ctypes.c_void_p(dev.ptr)) == 0
        t0 = time.perf_counter()
        ctx.submit_ptr(dev.ptr + (i % chunk) * batch_events * REC, batch_events, i * batch_events)
        c = ctx.fetch_counts()
        lat.append(1e3 * (time.perf_counter() - t0))
        pend.append(c["n_pending"])
    out["stream_64k_1m_pids"] = {
        "workload": "64K-event micro-batches of configs[2]'s trace (1 M PIDs), device resident, one batch in flight",
        "batches": nb,
        "first_100": dict(_percentiles(lat[10:100]), pending=int(pend[99])),
        "last_100": dict(_percentiles(lat[-100:]), pending=int(pend[-1])),
        "pids_seen_at_end": c["n_pids_seen"]}
    ctx.close()
    dev.free()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--events", type=int, default=0, help="override events per GPU (testing only)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-stream", action="store_true", help="skip the configs[4] micro-batch latency legs")
    ap.add_argument("--no-parity", action="store_true", help="N>1: skip the verification step against one GPU")
    ap.add_argument("--no-skew", action="store_true", help="N>1: skip the configs[3] (Zipf) leg")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world == 1 and args.gpus > 1:
        # launched without torchrun: re-exec under it so that one rank drives each GPU
        port = 29500 + (os.getpid() % 2000)
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.abspath(__file__)] + sys.argv[1:]
        raise SystemExit(subprocess.call(cmd))
    if args.impl == "reference":
        run_reference_arm(args, rank, world)
    else:
        run_gpu_arm(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
