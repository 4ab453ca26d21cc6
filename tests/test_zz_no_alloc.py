"""-m gpu, collected last on purpose: the allocation-free entry point.

chronos_filter_noalloc serves callers that may not allocate GPU memory at call time (a
sensor thread with a latency budget, a process whose GPU memory belongs to another
framework's pool, a plug-in host that forbids allocation after start-up). It runs the
filter stage with its scratch in a fixed array inside the library instead of buffers
owned by a context. Here it is called through the C ABI twice in a row on different
batches -- the second call finds the first call's leftovers in the shared scratch -- and
each result must equal, bit for bit, what the ordinary entry points (chronos_submit +
chronos_copy_survivors) report for the same batch. Oversized batches must be refused
with a readable message. The buffers this test itself allocates are its own inputs and
outputs, which the entry point expects from its caller.
"""
import ctypes
from ctypes import c_void_p

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _noalloc(L, dev, rows):
    from chronos_b200 import tracegen

    n = len(rows)
    d_rec = tracegen.DeviceBuffer(dev, max(n, 1) * 288)
    d_idx = tracegen.DeviceBuffer(dev, max(n, 1) * 4)
    d_pid = tracegen.DeviceBuffer(dev, max(n, 1) * 4)
    d_meta = tracegen.DeviceBuffer(dev, max(n, 1))
    d_cnt = tracegen.DeviceBuffer(dev, 4)
    if n:
        d_rec.upload(rows)
    err = ctypes.create_string_buffer(256)
    rc = L.chronos_filter_noalloc(dev, None, c_void_p(d_rec.ptr), n, c_void_p(d_idx.ptr), c_void_p(d_pid.ptr),
                                  c_void_p(d_meta.ptr), c_void_p(d_cnt.ptr), err, 256)
    assert rc == 0, err.value
    cnt = int(d_cnt.download(4).view(np.uint32)[0])  # synchronous copy: also waits for the queued work
    idx = d_idx.download(max(n, 1) * 4).view(np.uint32)[:cnt].copy()
    pid = d_pid.download(max(n, 1) * 4).view(np.uint32)[:cnt].copy()
    meta = d_meta.download(max(n, 1))[:cnt].copy()
    for b in (d_rec, d_idx, d_pid, d_meta, d_cnt):
        b.free()
    return cnt, idx, pid, meta


def _oracle(rows):
    """What the CPU oracle says survives, with pid and class bits (the checker of every other test)."""
    from oracle import coracle

    keep, meta = coracle.classify(rows)
    idx = np.nonzero(keep)[0].astype(np.uint32)
    pid = rows[idx.astype(np.int64), :4].copy().view("<u4").reshape(-1) if len(idx) else np.zeros(0, np.uint32)
    return idx, pid, meta[keep == 1] & 0x0F


def _existing(rows):
    import chronos_b200

    with chronos_b200.Context(device=0, max_batch_events=max(len(rows), 1)) as ctx:
        ctx.submit(rows, 0)
        res = ctx.fetch(records=False)
    return res.survivors.astype(np.uint32), res.survivor_meta


def test_noalloc_matches_existing_entry_point_twice_in_a_row():
    from chronos_b200 import _ffi, tracegen

    L = _ffi.lib()
    a = tracegen.generate_host(tracegen.WORKLOADS["uniform_10m_4k"], first=0, n=150_000)
    b = tracegen.generate_host(tracegen.WORKLOADS["skewed_zipf_chains"], first=777, n=64_000)
    for rows in (a, b):  # back to back: the second call reuses the scratch the first one dirtied
        cnt, idx, pid, meta = _noalloc(L, 0, rows)
        exp_idx, exp_meta = _existing(rows)
        assert cnt == len(exp_idx)
        np.testing.assert_array_equal(idx, exp_idx)
        np.testing.assert_array_equal(meta, exp_meta)
        np.testing.assert_array_equal(pid, rows[exp_idx.astype(np.int64), :4].copy().view("<u4").reshape(-1))
        o_idx, o_pid, o_meta = _oracle(rows)   # ... and both agree with the oracle
        np.testing.assert_array_equal(idx, o_idx)
        np.testing.assert_array_equal(pid, o_pid)
        np.testing.assert_array_equal(meta & 0x0F, o_meta)
    # an empty batch is legal and reports zero survivors
    cnt, idx, _, _ = _noalloc(L, 0, a[:0])
    assert cnt == 0 and idx.size == 0


def test_noalloc_rejects_oversized_batches():
    from chronos_b200 import _ffi

    L = _ffi.lib()
    err = ctypes.create_string_buffer(256)
    too_many = _ffi.NOALLOC_MAX_EVENTS + 1
    rc = L.chronos_filter_noalloc(0, None, c_void_p(16), too_many, c_void_p(16), c_void_p(16), c_void_p(16),
                                  c_void_p(16), err, 256)
    assert rc == _ffi.E_CAPACITY
    assert b"exceed" in err.value and str(_ffi.NOALLOC_MAX_EVENTS).encode() in err.value


@pytest.mark.parametrize("n", [1, 31, 33, 262_144])
def test_noalloc_edge_sizes_against_the_oracle(n):
    """One record, one short of a tile, one over a tile, and exactly the documented maximum."""
    from chronos_b200 import _ffi, tracegen

    assert _ffi.NOALLOC_MAX_EVENTS == 262_144
    rows = tracegen.generate_host(tracegen.WORKLOADS["skewed_zipf_chains"], first=4_321, n=n)
    for _ in range(2):   # twice: the second call starts from the first one's leftovers
        cnt, idx, pid, meta = _noalloc(_ffi.lib(), 0, rows)
        o_idx, o_pid, o_meta = _oracle(rows)
        assert cnt == len(o_idx)
        np.testing.assert_array_equal(idx, o_idx)
        np.testing.assert_array_equal(pid, o_pid)
        np.testing.assert_array_equal(meta & 0x0F, o_meta)


def test_noalloc_two_streams_ordered_by_an_event():
    """The documented way to share the one scratch array between streams: the second call's
    stream waits for an event recorded behind the first call. Both results must be right."""
    import torch

    from chronos_b200 import _ffi, tracegen

    L = _ffi.lib()
    a = tracegen.generate_host(tracegen.WORKLOADS["uniform_10m_4k"], first=0, n=100_000)
    b = tracegen.generate_host(tracegen.WORKLOADS["sharded_100m_1m"], first=5_000, n=70_000)
    bufs = []
    for rows in (a, b):
        n = len(rows)
        d = {k: tracegen.DeviceBuffer(0, sz) for k, sz in (("rec", n * 288), ("idx", n * 4), ("pid", n * 4), ("meta", n), ("cnt", 4))}
        d["rec"].upload(rows)
        bufs.append(d)
    s1, s2 = torch.cuda.Stream(device=0), torch.cuda.Stream(device=0)
    err = ctypes.create_string_buffer(256)

    def call(stream, d, n):
        rc = L.chronos_filter_noalloc(0, c_void_p(stream.cuda_stream), c_void_p(d["rec"].ptr), n, c_void_p(d["idx"].ptr),
                                      c_void_p(d["pid"].ptr), c_void_p(d["meta"].ptr), c_void_p(d["cnt"].ptr), err, 256)
        assert rc == 0, err.value

    call(s1, bufs[0], len(a))
    done = torch.cuda.Event()
    done.record(s1)
    s2.wait_event(done)          # without this the two calls would race on the shared scratch
    call(s2, bufs[1], len(b))
    s2.synchronize()
    for rows, d in ((a, bufs[0]), (b, bufs[1])):
        n = len(rows)
        cnt = int(d["cnt"].download(4).view(np.uint32)[0])
        o_idx, o_pid, o_meta = _oracle(rows)
        assert cnt == len(o_idx)
        np.testing.assert_array_equal(d["idx"].download(n * 4).view(np.uint32)[:cnt], o_idx)
        np.testing.assert_array_equal(d["pid"].download(n * 4).view(np.uint32)[:cnt], o_pid)
        np.testing.assert_array_equal(d["meta"].download(n)[:cnt] & 0x0F, o_meta)
        for v in d.values():
            v.free()
