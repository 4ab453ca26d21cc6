"""-m gpu: the corners of the context's bookkeeping, each against the oracle.

* a batch in which EVERY survivor flushes a window (one trigger event per PID whose list was
  filled by an earlier batch): the window buffers must hold as many windows as survivors;
* refused submits (bad alignment, too large) leave process_memory exactly as it was;
* submits allocate nothing once the context exists (cudaMemGetInfo before / after);
* two batches in flight give the same stream of results as one at a time;
* pending lists that outgrow the budget of chronos_create_sized re-size the context once;
* malformed checkpoints are refused before anything is touched.
"""
import ctypes
from ctypes import byref, c_uint64, c_void_p

import numpy as np
import pytest

from helpers import REC, Trace, as_rows, compare_batch
from oracle import coracle
from oracle.pyoracle import make_record as R

pytestmark = pytest.mark.gpu


def _free_bytes(dev=0):
    from chronos_b200 import _ffi

    f, t = c_uint64(), c_uint64()
    assert _ffi.lib().chronos_device_mem_info(dev, byref(f), byref(t)) == 0
    return int(f.value)


def test_every_survivor_flushes_a_window_over_carried_in_pids():
    """W == S == n: batch 1 leaves one non-trigger event pending for each of 300 PIDs, batch 2
    is one 'curl' event per PID (chronos_sensor.py:141-142: trigger and len >= 2 -> flush)."""
    import chronos_b200

    pids = [5000 + 7 * k for k in range(300)]
    first = b"".join(R(p, b"sh", b"/tmp/a%d" % p, b"OPEN") for p in pids)
    second = b"".join(R(p, b"sh", b"curl", b"EXEC") for p in reversed(pids))
    third = b"".join(R(p, b"sh", b"/tmp/b", b"OPEN") + R(p, b"sh", b"/tmp/cat", b"OPEN") for p in pids)
    for max_batch in (600, 1 << 16):
        with chronos_b200.Context(device=0, max_batch_events=max_batch, max_pending_events=300) as ctx:
            eng, tr = coracle.COracle(), Trace()
            compare_batch(ctx, eng, tr, first, 0)
            res, _ = compare_batch(ctx, eng, tr, second, 300)
            assert res.counts["n_windows"] == 300 == res.counts["n_survivors"]
            res, _ = compare_batch(ctx, eng, tr, third, 600)
            assert res.counts["n_windows"] == 300 and res.counts["n_survivors"] == 600


def test_refused_submits_leave_the_state_untouched():
    import chronos_b200
    from chronos_b200 import _ffi, tracegen

    rows = tracegen.generate_host(tracegen.WORKLOADS["skewed_zipf_chains"], n=60_000)
    dev = tracegen.DeviceBuffer(0, 20_001 * REC + 64)
    with chronos_b200.Context(device=0, max_batch_events=20_000) as ctx:
        eng, tr = coracle.COracle(), Trace()
        compare_batch(ctx, eng, tr, rows[:20_000], 0)
        dev.upload(rows[20_000:40_000])
        for _ in range(3):
            with pytest.raises(_ffi.ChronosError) as ei:       # misaligned device pointer
                ctx.submit_ptr(dev.ptr + 4, 100, 20_000)
            assert ei.value.code == _ffi.E_ARG
            with pytest.raises(_ffi.ChronosError) as ei:       # larger than the context was created for
                ctx.submit_ptr(dev.ptr, 20_001, 20_000)
            assert ei.value.code == _ffi.E_CAPACITY
            with pytest.raises(_ffi.ChronosError) as ei:       # arrival index beyond 2^48
                ctx.submit_ptr(dev.ptr, 100, 1 << 48)
            assert ei.value.code == _ffi.E_ARG
        # the stream goes on as if nothing had happened, state included
        compare_batch(ctx, eng, tr, rows[20_000:40_000], 20_000)
        compare_batch(ctx, eng, tr, rows[40_000:], 40_000)
    dev.free()


def test_submits_allocate_nothing_once_the_context_exists():
    """cudaMemGetInfo is the same before the first and after the 200th submit (device-resident
    batches; host-memory batches add the upload staging buffer on their first submit only)."""
    import chronos_b200
    from chronos_b200 import tracegen

    B = 16_384
    spec = tracegen.WORKLOADS["stream_64k"]
    dev = tracegen.generate_device(spec, device=0, first=0, n=8 * B)
    host = tracegen.generate_host(spec, first=0, n=8 * B)
    ctx = chronos_b200.Context(device=0, max_batch_events=B)
    free0 = _free_bytes()
    pending = []
    for i in range(200):
        ctx.submit_ptr(dev.ptr + (i % 8) * B * REC, B, i * B)
        pending.append(ctx.fetch_counts()["n_pending"])
    assert _free_bytes() == free0, "a device-pointer submit allocated (or freed) device memory"
    ctx.submit(host[:B], 200 * B)
    ctx.fetch_counts()
    ctx.submit(host[B:2 * B], 201 * B)   # second set's staging buffer
    ctx.fetch_counts()
    free1 = _free_bytes()
    assert free1 < free0
    for i in range(202, 260):
        ctx.submit(host[(i % 8) * B:(i % 8 + 1) * B], i * B)
        ctx.fetch_counts()
    assert _free_bytes() == free1, "a host-memory submit allocated after the staging buffers existed"
    ctx.close()
    dev.free()


def _check_result(res, ref, rows_of, check_records=True):
    np.testing.assert_array_equal(res.survivors, ref["survivors"])
    np.testing.assert_array_equal(res.windows["pid"], ref["win_pid"])
    np.testing.assert_array_equal(res.windows["trigger_idx"], ref["win_trigger"])
    np.testing.assert_array_equal(res.windows["len"], np.diff(ref["win_off"]).astype(np.uint32))
    np.testing.assert_array_equal(res.windows["flags"] & 1, ref["win_flags"] & 1)
    np.testing.assert_array_equal(res.chain_idx, ref["chain_idx"])
    if check_records and len(ref["members"]):
        np.testing.assert_array_equal(res.window_records, rows_of(ref["members"].astype(np.int64)))


@pytest.mark.parametrize("name,B,K", [("stream_64k", 65_536, 24), ("sharded_100m_1m", 50_000, 16),
                                      ("skewed_zipf_chains", 30_000, 20)])
def test_two_batches_in_flight_equal_one_at_a_time(name, B, K):
    """submit(k+1) before fetch(k): the filter of batch k+1 runs beside the window stage of
    batch k; results come back in submit order and equal the oracle's stream, and so does the
    final process_memory."""
    import chronos_b200
    from chronos_b200 import tracegen

    rows = tracegen.generate_host(tracegen.WORKLOADS[name], n=B * K)
    dev = tracegen.DeviceBuffer(0, B * K * REC)
    dev.upload(rows)
    eng = coracle.COracle()
    rows_of = lambda idx: rows[idx]  # noqa: E731
    with chronos_b200.Context(device=0, max_batch_events=B, max_pending_events=B * K) as ctx:
        ctx.submit_ptr(dev.ptr, B, 0)
        for k in range(K):
            if k + 1 < K:
                ctx.submit_ptr(dev.ptr + (k + 1) * B * REC, B, (k + 1) * B)   # two in flight from here
            res = ctx.fetch(records=True)
            ref = eng.feed(rows[k * B:(k + 1) * B], k * B)
            assert res.counts["n_events"] == B
            _check_result(res, ref, rows_of)
        pids, offs, mem = eng.state()
        c = ctx.fetch_counts()
        assert c["n_pending"] == len(mem) and c["n_pids_seen"] == len(pids)
        gp, gi, grec = ctx.state_arrays(c["n_pending"], records=True)
        np.testing.assert_array_equal(gi, mem)
        np.testing.assert_array_equal(grec, rows[mem.astype(np.int64)])
        np.testing.assert_array_equal(np.sort(ctx.state_pids(len(pids))), pids)
    dev.free()


def test_two_in_flight_from_host_memory_and_a_dropped_batch():
    """Host-memory submits in flight use one staging buffer per set; a third submit without a
    fetch drops the OLDEST batch's results but keeps its effect on the state."""
    import chronos_b200
    from chronos_b200 import _ffi, tracegen

    B = 20_000
    rows = tracegen.generate_host(tracegen.WORKLOADS["skewed_zipf_chains"], n=5 * B)
    eng = coracle.COracle()
    refs = [eng.feed(rows[k * B:(k + 1) * B], k * B) for k in range(5)]
    with chronos_b200.Context(device=0, max_batch_events=B, max_pending_events=5 * B) as ctx:
        ctx.submit(rows[:B].copy(), 0)
        ctx.submit(rows[B:2 * B].copy(), B)
        ctx.submit(rows[2 * B:3 * B].copy(), 2 * B)        # batch 0's results are gone now
        r1 = ctx.fetch(records=True)
        _check_result(r1, refs[1], lambda i: rows[i])
        with pytest.raises(_ffi.ChronosError):              # state is only readable when nothing is in flight
            ctx.state_arrays(1, records=False)
        r2 = ctx.fetch(records=True)
        _check_result(r2, refs[2], lambda i: rows[i])
        ctx.submit(rows[3 * B:4 * B].copy(), 3 * B)
        ctx.submit(rows[4 * B:].copy(), 4 * B)
        _check_result(ctx.fetch(records=True), refs[3], lambda i: rows[i])
        _check_result(ctx.fetch(records=True), refs[4], lambda i: rows[i])
        _, _, mem = eng.state()
        c = ctx.fetch_counts()
        assert c["n_pending"] == len(mem)
        np.testing.assert_array_equal(ctx.state_arrays(len(mem), records=False)[1], mem)


def test_pending_lists_outgrow_the_budget():
    """process_memory is unbounded (chronos_sensor.py:105): non-trigger events of ever new PIDs
    pile up far beyond max_pending_events; the context re-sizes itself and stays exact."""
    import chronos_b200

    eng, tr = coracle.COracle(), Trace()
    with chronos_b200.Context(device=0, max_batch_events=512, max_pending_events=64) as ctx:
        base = 0
        for step in range(40):
            recs = b"".join(R(100_000 + step * 512 + k, b"sh", b"/tmp/x%d" % k, b"OPEN") for k in range(512))
            res, _ = compare_batch(ctx, eng, tr, recs, base, check_state=(step % 13 == 0))
            base += 512
        assert res.counts["n_pending"] == 40 * 512
        # ... and now every one of them flushes
        for step in range(40):
            recs = b"".join(R(100_000 + step * 512 + k, b"sh", b"curl", b"EXEC") for k in range(512))
            res, _ = compare_batch(ctx, eng, tr, recs, base, check_state=(step == 39))
            base += 512
            assert res.counts["n_windows"] == 512
        assert res.counts["n_pending"] == 0


def test_malformed_checkpoints_are_refused_before_anything_changes():
    import chronos_b200
    from chronos_b200 import _ffi, tracegen

    rows = tracegen.generate_host(tracegen.WORKLOADS["skewed_zipf_chains"], n=60_000)
    eng, tr = coracle.COracle(), Trace()
    with chronos_b200.Context(device=0, max_batch_events=20_000) as ctx:
        compare_batch(ctx, eng, tr, rows[:20_000], 0)
        blob = bytearray(ctx.checkpoint_save())
        P, K = int.from_bytes(blob[16:24], "little"), int.from_bytes(blob[24:32], "little")
        assert P > 10 and K > 100
        tab, ppid = 64, 64 + K * 8

        def corrupt(edit):
            b = bytearray(blob)
            edit(b)
            return bytes(b)

        bad = [
            corrupt(lambda b: b.__setitem__(slice(tab + 8, tab + 16), b[tab:tab + 8])),            # duplicate pid entry
            corrupt(lambda b: b.__setitem__(tab + 4, 3)),                                           # DFA state 3
            corrupt(lambda b: b.__setitem__(tab + 7, 0x80)),                                        # bits above 33
            corrupt(lambda b: b.__setitem__(slice(ppid, ppid + 4), (1).to_bytes(4, "little"))),     # pending pid not a key
            corrupt(lambda b: b.__setitem__(slice(ppid + 4 * (P - 1), ppid + 4 * P), bytes(4))),    # pending out of order
        ]
        for b in bad:
            with pytest.raises(_ffi.ChronosError) as ei:
                ctx.checkpoint_load(b)
            assert ei.value.code == _ffi.E_ARG
        # nothing was touched: the stream continues, and the good blob still loads elsewhere
        compare_batch(ctx, eng, tr, rows[20_000:40_000], 20_000)
        with chronos_b200.Context(device=0, max_batch_events=20_000) as other:
            other.checkpoint_load(bytes(blob))
            assert other.checkpoint_save() == bytes(blob)


def test_create_rejects_sizes_the_entry_ids_cannot_address():
    import chronos_b200
    from chronos_b200 import _ffi

    with pytest.raises(_ffi.ChronosError):
        chronos_b200.Context(device=0, max_batch_events=1 << 27)


def test_window_digests_see_content_and_ignore_bytes_behind_the_nul():
    """chronos_window_digests (what bench.py --gpus N compares a sharded run with): equal for
    records that differ only behind a field's NUL -- the reference never reads those bytes
    (chronos_sensor.py:125-130) -- and different as soon as anything it does read differs."""
    import chronos_b200

    def batch(tail, path=b"/tmp/cat"):
        a = bytearray(R(7, b"sh", b"/tmp/a", b"OPEN") + R(9, b"zsh", b"/tmp/b", b"OPEN") + R(7, b"sh", path, b"OPEN") +
                      R(9, b"zsh", b"curl", b"EXEC"))
        for r in range(4):
            a[r * REC + 4 + 8:r * REC + 4 + 12] = tail      # comm bytes 8..11, behind the NUL
            a[r * REC + 20 + 200:r * REC + 20 + 204] = tail  # argv bytes 200..203, behind the NUL
            a[r * REC + 286:r * REC + 288] = tail[:2]        # tail padding
        return bytes(a)

    digs = []
    zero = bytes(4)
    for recs in (batch(zero), batch(b"junk"), batch(b"junk", path=b"/tmp/cab"), batch(zero, path=b"/tmp/catx")):
        with chronos_b200.Context(device=0, max_batch_events=16) as ctx:
            ctx.submit(recs, 0)
            res = ctx.fetch(records=True)
            digs.append((res.counts["n_windows"], ctx.window_digests(res.counts["n_windows"])))
    assert digs[0][0] == 2 and digs[1][0] == 2
    np.testing.assert_array_equal(digs[0][1], digs[1][1])          # garbage behind the NUL does not count
    assert digs[2][0] == 1                                        # "/tmp/cab" no longer triggers: one window less
    assert digs[3][0] == 2 and digs[3][1][0] != digs[0][1][0]     # one more byte in a path: another digest
    assert digs[3][1][1] == digs[0][1][1]                         # ... while the untouched window keeps its digest


def test_trimmed_host_upload_is_lossless():
    """chronos_set_host_trim: host threads cut every record to its non-zero 16-byte chunks, the GPU
    rebuilds it. The result -- survivors, windows, the window records BYTE FOR BYTE, the pending
    state -- must be what the plain upload and the oracle give, also for records with garbage
    behind a NUL, fields without a NUL and bytes in the very last argv chunk."""
    import chronos_b200
    from chronos_b200 import tracegen

    rows = tracegen.generate_host(tracegen.WORKLOADS["skewed_zipf_chains"], n=300_000).copy()
    rng = np.random.default_rng(7)
    for i in rng.integers(0, len(rows), 400):           # garbage everywhere behind the first bytes
        rows[i, 60:276] = rng.integers(1, 256, 216, dtype=np.uint8)
    for i in rng.integers(0, len(rows), 200):           # a lone byte in the last argv chunk
        rows[i, 20 + 250] = 0x41
    rows[5] = 0xFF
    rows[6] = 0
    eng, tr = coracle.COracle(), Trace()
    with chronos_b200.Context(device=0, max_batch_events=len(rows)) as ctx:
        for threads in (3, 0, 8):
            ctx.set_host_trim(threads)
            base = 0 if threads == 3 else (len(rows) if threads == 0 else 2 * len(rows))
            compare_batch(ctx, eng, tr, rows, base)
