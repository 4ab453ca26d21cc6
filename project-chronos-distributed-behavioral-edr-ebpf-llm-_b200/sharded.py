"""PID-hash sharding of the hot path over the GPUs of one box.

Each rank filters a contiguous slice of the arrival-ordered batch, survivors are
repartitioned with ONE all-to-all so that every PID's events meet on
`owner = hash32(pid) % world`, and the window logic runs there against that rank's
resident per-PID state (process_memory is independent per PID: chronos_sensor.py:105,
:138, :142, :157). Receivers concatenate what they get in source-rank order; since the
slices are contiguous and ordered, that is global arrival order, which is all the
reference's semantics need (there is no timestamp in struct data_t, :18-23).

`exchange()` is plain torch.distributed host logic and works on any backend/device (the
CPU tests run it over gloo); `ShardedRunner` binds it to the CUDA stages of the C ABI, or --
on one NVSwitch box -- skips it: stage 1 then writes the survivors directly into the owners'
HBM (csrc/exchange.cuh).
"""
import ctypes
from ctypes import c_void_p

import numpy as np
import torch
import torch.distributed as dist

from . import _ffi
from .engine import Context
from .records import RECORD_BYTES


def hash32_np(x):
    """Vectorised twin of chronos_owner_of's hash (csrc/common.cuh hash32)."""
    x = np.asarray(x, dtype=np.uint64) & np.uint64(0xFFFFFFFF)
    m = np.uint64(0xFFFFFFFF)
    x ^= x >> np.uint64(16)
    x = (x * np.uint64(0x7FEB352D)) & m
    x ^= x >> np.uint64(15)
    x = (x * np.uint64(0x846CA68B)) & m
    x ^= x >> np.uint64(16)
    return x.astype(np.uint32)


def owner_of_np(pids, world):
    return (hash32_np(pids) % np.uint32(world)).astype(np.int64)


SLOT_BYTES = 16
HDR_ARRIVAL_MASK = (1 << 48) - 1


def pack_hdr(arrival, meta, nslots):
    """Header word of the transport form: arrival (48 bits) | META bits << 48 | slot count << 56."""
    return (np.asarray(arrival, dtype=np.uint64) & np.uint64(HDR_ARRIVAL_MASK)) | \
        (np.asarray(meta, dtype=np.uint64) << np.uint64(48)) | (np.asarray(nslots, dtype=np.uint64) << np.uint64(56))


def exchange(send_payload, send_hdr, entry_counts, unit_counts, unit_bytes, recv_payload, recv_hdr, group=None):
    """One all-to-all-v of packed survivors.

    send_payload: uint8 tensor holding, destination after destination, `unit_counts[d]`
    units of `unit_bytes` bytes for rank d (16-byte transport slots on the GPU path);
    send_hdr: int64 tensor with `entry_counts[d]` header words for rank d. The recv_*
    tensors are filled in source-rank order, which -- the slices being contiguous and
    ordered -- is global arrival order. Counts travel first, then the two payloads.
    Returns (n_entries_received, n_units_received)."""
    world = dist.get_world_size(group)
    dev = send_hdr.device
    sc = torch.as_tensor(np.concatenate([np.asarray(entry_counts, dtype=np.int64), np.asarray(unit_counts, dtype=np.int64)]),
                         device=dev)
    rc = torch.empty(2 * world, dtype=torch.int64, device=dev)
    # counts of (entries, units) for every peer in one small all-to-all: layout [e_0..e_{w-1}, u_0..u_{w-1}]
    # is regrouped so that rank d receives (e_d, u_d) from everybody
    sc2 = torch.stack([sc[:world], sc[world:]], dim=1).contiguous().view(-1)
    dist.all_to_all_single(rc, sc2, group=group)
    rcn = rc.cpu().numpy().astype(np.int64).reshape(world, 2)
    re_, ru = rcn[:, 0], rcn[:, 1]
    se, su = np.asarray(entry_counts, dtype=np.int64), np.asarray(unit_counts, dtype=np.int64)
    n_re, n_ru, n_se, n_su = int(re_.sum()), int(ru.sum()), int(se.sum()), int(su.sum())
    if n_re > recv_hdr.numel() or n_ru * unit_bytes > recv_payload.numel():
        raise _ffi.ChronosError(_ffi.E_CAPACITY, f"{n_re} entries / {n_ru} units received exceed the receive buffers")
    dist.all_to_all_single(recv_payload.view(-1)[: n_ru * unit_bytes], send_payload.view(-1)[: n_su * unit_bytes],
                           [int(x) * unit_bytes for x in ru], [int(x) * unit_bytes for x in su], group=group)
    dist.all_to_all_single(recv_hdr[:n_re], send_hdr[:n_se], re_.tolist(), se.tolist(), group=group)
    return n_re, n_ru


class ShardedRunner:
    """One rank of the sharded pipeline.

    mode "direct": CUDA stage 1 writes the survivors straight into the owners' receive
    windows over NVLink (chronos_filter_exchange / chronos_submit_exchanged), nothing but
    kernel launches on the host. mode "nccl": stage 1 packs per owner into a send buffer,
    torch.distributed.all_to_all_single moves it, stage 2 runs on what arrived. "auto" takes
    "direct" when every rank could map every peer's window (CUDA IPC), else "nccl"."""

    def __init__(self, rec_ptr, n, first, device, rank, world, cap=None, mode="auto", cap_entries_per_peer=None):
        self.rec_ptr, self.n, self.first = int(rec_ptr), int(n), int(first)
        self.device, self.rank, self.world = int(device), int(rank), int(world)
        self.cap = int(cap or max(n // 2, 1024))   # survivors per rank, sent or received
        self.cap_slots = self.cap * 8              # 16-byte transport slots (3..18 per survivor)
        self.cap_pair = int(cap_entries_per_peer or max(self.cap // self.world, 1024))
        self.ctx = Context(device=self.device, max_batch_events=max(self.n, self.cap, self.cap_pair * self.world))
        self._L = _ffi.lib()
        self.mode = "nccl"
        if mode in ("auto", "direct"):
            ok = self._connect_direct()
            if ok:
                self.mode = "direct"
            elif mode == "direct":
                raise _ffi.ChronosError(_ffi.E_STATE, "direct exchange unavailable: " + self._direct_error)
        if self.mode == "nccl":
            dev = torch.device("cuda", self.device)
            self.send_slots = torch.empty(self.cap_slots * SLOT_BYTES, dtype=torch.uint8, device=dev)
            self.send_hdr = torch.empty(self.cap, dtype=torch.int64, device=dev)
            self.recv_slots = torch.empty(self.cap_slots * SLOT_BYTES, dtype=torch.uint8, device=dev)
            self.recv_hdr = torch.empty(self.cap, dtype=torch.int64, device=dev)
            self._ec = (ctypes.c_uint64 * self.world)()
            self._sc = (ctypes.c_uint64 * self.world)()

    def _connect_direct(self):
        """Create the receive window, all-gather the IPC handles, map the peers. Collective:
        every rank gets the same answer."""
        self._direct_error = ""
        dev = torch.device("cuda", self.device)
        ok = 1
        handle = bytes(_ffi.IPC_HANDLE_BYTES)
        try:
            _, handle = self.ctx.exchange_create(self.world, self.rank, self.cap_pair)
        except _ffi.ChronosError as e:
            ok, self._direct_error = 0, str(e)
        mine = torch.tensor(list(handle), dtype=torch.uint8, device=dev)
        everyone = torch.empty(self.world * _ffi.IPC_HANDLE_BYTES, dtype=torch.uint8, device=dev)
        dist.all_gather_into_tensor(everyone, mine)
        flag = torch.tensor([ok], dtype=torch.int32, device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if int(flag.item()) == 1:
            blob = bytes(everyone.cpu().numpy().tobytes())
            hs = [blob[i * _ffi.IPC_HANDLE_BYTES:(i + 1) * _ffi.IPC_HANDLE_BYTES] for i in range(self.world)]
            try:
                self.ctx.exchange_connect(self.world, handles=hs)
            except _ffi.ChronosError as e:
                ok, self._direct_error = 0, str(e)
            flag = torch.tensor([ok], dtype=torch.int32, device=dev)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        return int(flag.item()) == 1

    def stream_ptr(self):
        return int(self._L.chronos_stream(self.ctx._h))

    def filter_partition(self, rec_ptr, n, base):
        rc = self._L.chronos_filter_partition(self.ctx._h, c_void_p(int(rec_ptr)), int(n), int(base), self.world,
                                              c_void_p(self.send_slots.data_ptr()), self.cap_slots,
                                              c_void_p(self.send_hdr.data_ptr()), self.cap, self._ec, self._sc)
        self.ctx._check(rc)
        return np.array(list(self._ec), dtype=np.int64), np.array(list(self._sc), dtype=np.int64)

    def step(self, reset=True):
        if reset:
            self.ctx.reset()  # benchmark steps repeat the same batch on a fresh process_memory
        if self.mode == "direct":
            self.ctx.filter_exchange_ptr(self.rec_ptr, self.n, self.first)  # only enqueues
            self.ctx.submit_exchanged()                                     # only enqueues
            c = self.ctx.fetch_counts()
            t, f = self.ctx.last_timing()
            c.update(gpu_ms=t, filter_ms=f, launches=self.ctx.last_launches(), n_received=c["n_survivors"])
            return c
        ec, sc = self.filter_partition(self.rec_ptr, self.n, self.first)  # waits for stage 1
        launches = self.ctx.last_launches()
        n_recv, n_slots = exchange(self.send_slots, self.send_hdr, ec, sc, SLOT_BYTES, self.recv_slots, self.recv_hdr)
        torch.cuda.current_stream().synchronize()  # NCCL ran on torch's stream; stage 2 runs on the library's
        self.ctx.submit_survivors_ptr(self.recv_slots.data_ptr(), self.recv_hdr.data_ptr(), n_recv, n_slots)
        c = self.ctx.fetch_counts()
        t, f = self.ctx.last_timing()
        c.update(gpu_ms=t, filter_ms=f, launches=launches + self.ctx.last_launches(), n_sent=int(ec.sum()),
                 n_received=n_recv, slots_sent=int(sc.sum()))
        return c

    def timed_steps(self, steps):
        """K steps for bench.py's timed region. Direct mode: every step is a fresh process_memory
        and the GPU runs the steps strictly one after the other (everything of a step is queued on
        one stream); only the host works ahead -- step k+1 is enqueued before step k is fetched."""
        if self.mode != "direct":
            for _ in range(steps):
                c = self.step()
                c["stages"] = self.ctx.last_stage_times()
                yield c
            return

        def enqueue():
            self.ctx.reset()
            self.ctx.filter_exchange_ptr(self.rec_ptr, self.n, self.first)
            self.ctx.submit_exchanged()

        enqueue()
        for k in range(steps):
            if k + 1 < steps:
                enqueue()
            c = self.ctx.fetch_counts()
            t, f = self.ctx.last_timing()
            c.update(gpu_ms=t, filter_ms=f, launches=self.ctx.last_launches(), n_received=c["n_survivors"],
                     stages=self.ctx.last_stage_times())
            yield c

    def e2e(self, steps, barrier, dist_mod):
        """Same step, but the slice starts in pinned HOST memory (H2D inside the timed region)
        and survivors, windows and window records are read back to pinned host memory; the
        download of step k overlaps the upload of step k+1 (chronos_download_begin)."""
        import time

        L = self._L
        n = self.n
        p = c_void_p()
        assert L.chronos_host_alloc_pinned(n * RECORD_BYTES, ctypes.byref(p)) == 0
        assert L.chronos_memcpy_d2h(self.device, p, c_void_p(self.rec_ptr), n * RECORD_BYTES) == 0
        dev_ptr, self.rec_ptr = self.rec_ptr, p.value
        c = self.step()
        cap = ((c["n_survivors"] * 2 + 4096) * (9 + RECORD_BYTES + 32) + 255) // 256 * 256
        outs = []
        for _ in range(2):
            q = c_void_p()
            assert L.chronos_host_alloc_pinned(cap, ctypes.byref(q)) == 0
            outs.append(q)
        def loop():
            barrier()
            t0 = time.perf_counter()
            d2h = 0
            for k in range(steps):
                self.step()
                d2h = self.ctx.download_begin(outs[k & 1].value, cap)["total_bytes"]
            self.ctx.download_wait()
            barrier()
            return (time.perf_counter() - t0) / steps, d2h, self.ctx.last_upload_bytes()

        # (each mode is timed twice and the better run counts: the loops live off the host's memory
        #  system, which is shared with whatever else runs on the box)
        runs_plain = [loop(), loop()]
        dt_plain, d2h, up_plain = min(runs_plain, key=lambda r: r[0])
        # the upload is also tried with the host's cores trimming part of it (chronos_set_host_trim)
        import os

        threads = (os.cpu_count() or 1) // self.world
        dt_trim = up_trim = None
        if threads >= 2 and not os.environ.get("CHRONOS_BENCH_NO_TRIM"):
            self.ctx.set_host_trim(threads)
            loop()
            runs_trim = [loop(), loop()]
            dt_trim, d2h, up_trim = min(runs_trim, key=lambda r: r[0])
            self.ctx.set_host_trim(0)
        trimmed = dt_trim is not None and dt_trim <= dt_plain
        dt = dt_trim if trimmed else dt_plain
        self.rec_ptr = dev_ptr
        L.chronos_host_free_pinned(p)
        for q in outs:
            L.chronos_host_free_pinned(q)
        total = n * self.world
        how = ("chronos_filter_exchange(pinned host slice, survivors written into the owners' windows over NVLink) -> "
               "chronos_submit_exchanged" if self.mode == "direct" else
               "chronos_filter_partition(pinned host slice) -> all_to_all -> chronos_submit_survivors")
        return {"value": total / dt, "unit": "events/s",
                "h2d_bytes_per_step": int(up_trim if trimmed else up_plain) * self.world,
                "d2h_bytes_per_step": int(d2h) * self.world, "ms_per_step": 1e3 * dt, "steps": steps,
                "timing": "wall clock around K steps between barriers; each mode run twice, the better run reported",
                "plain_copy": {"value": total / dt_plain, "ms_per_step": 1e3 * dt_plain,
                               "ms_per_step_runs": [1e3 * r[0] for r in runs_plain]},
                "host_trim": None if dt_trim is None else {"value": total / dt_trim, "ms_per_step": 1e3 * dt_trim,
                                                           "ms_per_step_runs": [1e3 * r[0] for r in runs_trim],
                                                           "host_threads_per_rank": threads},
                "upload": "host_trim" if trimmed else "plain_copy",
                "path": "per rank: " + how + " -> fetch -> download_begin(pinned host results), overlapped with the next upload"}
