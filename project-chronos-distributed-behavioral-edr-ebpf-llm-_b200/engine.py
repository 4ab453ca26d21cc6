"""Context: one GPU's filter + per-PID window state behind the C ABI (include/chronos.h)."""
import ctypes
from ctypes import byref, c_float, c_uint64, c_void_p

import numpy as np

from . import _ffi
from .records import RECORD_BYTES, as_record_bytes, event_str

WINDOW_DTYPE = np.dtype([("trigger_idx", "<u8"), ("start", "<u8"), ("len", "<u4"), ("pid", "<u4"),
                         ("flags", "<u4"), ("reserved", "<u4")])
assert WINDOW_DTYPE.itemsize == ctypes.sizeof(_ffi.Window)


class BatchResult:
    """What one submitted batch produced. `windows` is sorted by trigger arrival index,
    the order the reference calls analyze_sequence in (chronos_sensor.py:143-147)."""

    def __init__(self, counts, survivors, survivor_meta, windows, window_records, gpu_ms, filter_ms, launches,
                 chain_pid=None, chain_idx=None):
        self.counts = counts
        self.survivors = survivors            # uint64 global arrival indices, ascending
        self.survivor_meta = survivor_meta    # uint8 META_* bits
        self.windows = windows                # WINDOW_DTYPE
        self.window_records = window_records  # (sum(len), 288) uint8, window after window, or None
        self.gpu_ms = gpu_ms
        self.filter_ms = filter_ms
        self.launches = launches
        # extension: completed curl -> chmod -> EXEC chains of this batch, in arrival order
        self.chain_pid = chain_pid
        self.chain_idx = chain_idx

    def histories(self):
        """[(pid, [event_str, ...]), ...]: exactly the arguments of analyze_sequence."""
        out = []
        o = 0
        recs = self.window_records
        for w in self.windows:
            n = int(w["len"])
            out.append((int(w["pid"]), [event_str(recs[o + k]) for k in range(n)]))
            o += n
        return out


class Context:
    """One GPU's state. All device memory is allocated here (chronos_create): room for batches
    of max_batch_events on top of max_pending_events pending in process_memory (default: as many
    as one batch holds). Two batches may be in flight: a second submit() before fetch() lets
    the filter of the new batch run beside the window stage of the previous one; fetch()
    returns batches in submit order."""

    def __init__(self, device=0, max_batch_events=1 << 20, max_pending_events=None):
        self._L = _ffi.lib()
        self.device = int(device)
        self._h = c_void_p()
        if max_pending_events is None:
            rc = self._L.chronos_create(self.device, int(max_batch_events), byref(self._h))
        else:
            rc = self._L.chronos_create_sized(self.device, int(max_batch_events), int(max_pending_events), byref(self._h))
        if rc != _ffi.OK:
            self._h = c_void_p()
            raise _ffi.ChronosError(rc, "chronos_create failed (no usable sm_100 GPU? there is no CPU fallback)")
        self.max_batch_events = int(max_batch_events)
        self._keep = []

    def _check(self, rc):
        if rc != _ffi.OK:
            msg = self._L.chronos_last_error(self._h)
            raise _ffi.ChronosError(rc, msg.decode("utf-8", "replace") if msg else "")

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            self._L.chronos_destroy(self._h)
            self._h = c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def reset(self):
        """Forget all per-PID state (a fresh process_memory, chronos_sensor.py:105). Takes effect in
        submit order: batches already submitted finish against the old state and stay fetchable."""
        self._check(self._L.chronos_reset(self._h))

    def set_host_trim(self, threads):
        """threads > 0: large submits from host memory are cut to their non-zero 16-byte chunks by that
        many host threads before they cross PCIe and rebuilt on the GPU (lossless); 0: off."""
        self._check(self._L.chronos_set_host_trim(self._h, int(threads)))

    def last_upload_bytes(self):
        n = c_uint64()
        self._check(self._L.chronos_last_upload_bytes(self._h, byref(n)))
        return int(n.value)

    def set_overlap(self, on):
        """False: batches run strictly one after the other on the GPU even when submitted ahead."""
        self._check(self._L.chronos_set_overlap(self._h, 1 if on else 0))

    # ---- submit ------------------------------------------------------------------------
    def submit(self, records, base_arrival_index=0):
        """records: bytes-like / numpy array of n*288 bytes on the host."""
        a = as_record_bytes(records)
        self._check(self._L.chronos_submit(self._h, a.ctypes.data, a.size // RECORD_BYTES, int(base_arrival_index)))
        self._keep = (self._keep + [a])[-2:]  # the async H2D copies of the (at most two) batches in flight read them

    def submit_ptr(self, ptr, n, base_arrival_index=0):
        """ptr: raw host or device address of n records (device pointers are used in place)."""
        self._check(self._L.chronos_submit(self._h, c_void_p(int(ptr)), int(n), int(base_arrival_index)))

    def submit_survivors_ptr(self, slots_ptr, hdr_ptr, n_entries, n_slots):
        """Stage 2 of the sharded path: already-filtered events in transport form (device pointers)."""
        self._check(self._L.chronos_submit_survivors(self._h, c_void_p(int(slots_ptr)), c_void_p(int(hdr_ptr)),
                                                     int(n_entries), int(n_slots)))

    # ---- checkpoint ------------------------------------------------------------------------
    def checkpoint_save(self):
        """The whole carried state (pending lists with records, PIDs seen, chain states) as bytes."""
        n = c_uint64()
        self._check(self._L.chronos_checkpoint_bytes(self._h, byref(n)))
        buf = np.zeros(max(int(n.value), 1), dtype=np.uint8)
        self._check(self._L.chronos_checkpoint_save(self._h, buf.ctypes.data, buf.size))
        return buf[:int(n.value)].tobytes()

    def checkpoint_load(self, blob):
        """Replace this context's state with a saved one; the next submit continues the stream."""
        a = np.frombuffer(blob, dtype=np.uint8)
        self._check(self._L.chronos_checkpoint_load(self._h, a.ctypes.data, a.size))

    # ---- asynchronous download ------------------------------------------------------------
    def download_begin(self, host_ptr, cap_bytes):
        """Start one device-to-host copy of the fetched batch's survivors, windows and window
        records into (pinned) host memory at host_ptr; returns the layout as a dict. The data
        is complete after download_wait(); the context can take the next submit meanwhile."""
        d = _ffi.Download()
        self._check(self._L.chronos_download_begin(self._h, c_void_p(int(host_ptr)), int(cap_bytes), byref(d)))
        return {k: int(getattr(d, k)) for k, _ in _ffi.Download._fields_}

    def download_wait(self):
        self._check(self._L.chronos_download_wait(self._h))

    @staticmethod
    def parse_download(buf, layout):
        """Views into a completed download (buf: uint8 array over the host buffer)."""
        S, W, R = layout["n_survivors"], layout["n_windows"], layout["n_window_records"]
        idx = buf[layout["off_survivor_idx"]:layout["off_survivor_idx"] + 8 * S].view("<u8")
        meta = buf[layout["off_survivor_meta"]:layout["off_survivor_meta"] + S]
        win = buf[layout["off_windows"]:layout["off_windows"] + 32 * W].view(WINDOW_DTYPE)
        rec = buf[layout["off_window_records"]:layout["off_window_records"] + RECORD_BYTES * R].reshape(-1, RECORD_BYTES)
        return idx, meta, win, rec

    # ---- direct exchange (include/chronos.h, "direct exchange") -----------------------------
    def exchange_create(self, world, rank, cap_entries_per_peer, cap_slots_per_peer=0):
        """Allocate this rank's receive window. Returns (device address, 64-byte IPC handle)."""
        win = c_void_p()
        handle = (ctypes.c_uint8 * _ffi.IPC_HANDLE_BYTES)()
        self._check(self._L.chronos_exchange_create(self._h, int(world), int(rank), int(cap_entries_per_peer),
                                                    int(cap_slots_per_peer), byref(win), handle))
        return int(win.value), bytes(handle)

    def exchange_connect(self, world, windows=None, handles=None):
        """windows: per-rank device addresses usable in this process (or None entries);
        handles: per-rank 64-byte IPC handles of windows owned by other processes."""
        wp = hp = None
        if windows is not None:
            wp = (c_void_p * world)(*[c_void_p(w) if w else c_void_p() for w in windows])
        if handles is not None:
            blob = b"".join(bytes(h) for h in handles)
            assert len(blob) == world * _ffi.IPC_HANDLE_BYTES
            hp = (ctypes.c_uint8 * len(blob)).from_buffer_copy(blob)
        self._check(self._L.chronos_exchange_connect(self._h, wp, hp))

    def filter_exchange_ptr(self, ptr, n, base_arrival_index=0):
        """Stage 1: filter n records at ptr (host or device) and deliver the survivors to their owners."""
        self._check(self._L.chronos_filter_exchange(self._h, c_void_p(int(ptr)), int(n), int(base_arrival_index)))

    def filter_exchange(self, records, base_arrival_index=0):
        a = as_record_bytes(records)
        self.filter_exchange_ptr(a.ctypes.data if a.size else 0, a.size // RECORD_BYTES, base_arrival_index)
        self._keep = (self._keep + [a])[-2:]

    def submit_exchanged(self):
        """Stage 2 on what the peers delivered for this step."""
        self._check(self._L.chronos_submit_exchanged(self._h))

    def exchange_set_timeout(self, seconds):
        self._check(self._L.chronos_exchange_set_timeout(self._h, float(seconds)))

    # ---- results -----------------------------------------------------------------------
    def fetch_counts(self):
        c = _ffi.Counts()
        self._check(self._L.chronos_fetch(self._h, byref(c)))
        return {k: int(getattr(c, k)) for k, _ in _ffi.Counts._fields_ if k != "reserved"}

    def fetch(self, records=True):
        counts = self.fetch_counts()
        S, W = counts["n_survivors"], counts["n_windows"]
        surv = np.zeros(S, dtype=np.uint64)
        meta = np.zeros(S, dtype=np.uint8)
        self._check(self._L.chronos_copy_survivors(self._h, surv.ctypes.data, meta.ctypes.data, S))
        win = np.zeros(W, dtype=WINDOW_DTYPE)
        self._check(self._L.chronos_copy_windows(self._h, win.ctypes.data, W))
        wrec = None
        if records:
            tot = int(win["len"].sum()) if W else 0
            wrec = np.zeros((tot, RECORD_BYTES), dtype=np.uint8)
            self._check(self._L.chronos_copy_window_records(self._h, wrec.ctypes.data, tot))
        K = counts["n_chains"]
        cpid = np.zeros(K, dtype=np.uint32)
        cidx = np.zeros(K, dtype=np.uint64)
        self._check(self._L.chronos_copy_chains(self._h, cpid.ctypes.data, cidx.ctypes.data, K))
        t, f = c_float(), c_float()
        self._check(self._L.chronos_last_timing(self._h, byref(t), byref(f)))
        ln = c_uint64()
        self._check(self._L.chronos_last_launches(self._h, byref(ln)))
        return BatchResult(counts, surv, meta, win, wrec, t.value, f.value, ln.value, cpid, cidx)

    def window_digests(self, n_windows):
        """uint64 digest of every window of the fetched batch (canonical record bytes), window order."""
        out = np.zeros(int(n_windows), dtype=np.uint64)
        self._check(self._L.chronos_window_digests(self._h, out.ctypes.data, out.size))
        return out

    def arena(self):
        c = _ffi.Counts()
        self._check(self._L.chronos_fetch(self._h, byref(c)))
        a = np.zeros((int(c.n_arena), RECORD_BYTES), dtype=np.uint8)
        self._check(self._L.chronos_copy_arena(self._h, a.ctypes.data, int(c.n_arena)))
        return a

    def last_timing(self):
        t, f = c_float(), c_float()
        self._check(self._L.chronos_last_timing(self._h, byref(t), byref(f)))
        return t.value, f.value

    STAGES = ("filter", "compact", "pack", "wait_unpack", "sort", "windows_gather")

    def last_stage_times(self):
        """{stage: ms} of the last fetched batch (see chronos_last_stage_times)."""
        a = (c_float * 6)()
        self._check(self._L.chronos_last_stage_times(self._h, a, 6))
        return dict(zip(self.STAGES, [float(x) for x in a]))

    def last_launches(self):
        ln = c_uint64()
        self._check(self._L.chronos_last_launches(self._h, byref(ln)))
        return int(ln.value)

    def state_arrays(self, n_pending, records=True):
        """Pending process_memory entries: (pid[P], global_idx[P], records[P,288] or None),
        ascending pid, arrival order inside a pid."""
        P = int(n_pending)
        pid = np.zeros(P, dtype=np.uint32)
        idx = np.zeros(P, dtype=np.uint64)
        rec = np.zeros((P, RECORD_BYTES), dtype=np.uint8) if records else None
        self._check(self._L.chronos_state(self._h, pid.ctypes.data, idx.ctypes.data,
                                          rec.ctypes.data if records else None, P))
        return pid, idx, rec

    def state_pids(self, n_pids_seen):
        """Keys of process_memory: every PID that ever had a surviving event (unordered)."""
        K = int(n_pids_seen)
        out = np.zeros(K, dtype=np.uint32)
        self._check(self._L.chronos_state_pids(self._h, out.ctypes.data, K))
        return out

    def process_memory(self, n_pending, n_pids_seen=None):
        """{pid: [event_str, ...]} as the reference's process_memory (chronos_sensor.py:105).
        With n_pids_seen the PIDs whose list is currently [] (:157) are included too."""
        pid, _, rec = self.state_arrays(n_pending, records=True)
        pm = {}
        if n_pids_seen is not None:
            pm = {int(p): [] for p in self.state_pids(n_pids_seen)}
        for k in range(pid.size):
            pm.setdefault(int(pid[k]), []).append(event_str(rec[k]))
        return pm
