"""Direct exchange (chronos_filter_exchange / chronos_submit_exchanged): stage 1 writes the
survivors into the owners' receive windows, stage 2 waits for the flags on the GPU.

Ranks are emulated by several contexts of ONE process on one GPU (windows connected by plain
device pointers instead of IPC handles); the kernels and the protocol are the ones the
multi-process runner uses. The merged output of all ranks must equal the oracle's on the
whole trace (PIDs are independent: chronos_sensor.py:105, :138, :142, :157).

Emulation caveat: a rank's stage 2 SPINS on the GPU until its peers have delivered. With every
rank on one GPU, a peer's stage 1 must therefore never queue up behind that spin: _step()
enqueues stage 1 of all ranks before any stage 2. Streams of different contexts may still share
a hardware queue (there are at most 32, CUDA_DEVICE_MAX_CONNECTIONS in conftest.py, and a long
pytest process creates hundreds of streams), and then a stage 1 sits behind a spinning stage 2
until that times out. Unless a test says otherwise, _step() therefore waits for all stage 1s
before it enqueues the first stage 2; the waiting path itself is covered by the time-out test
here and by the real ranks of test_gpu_multiprocess.py, which are one process per GPU and
cannot block each other this way."""
import numpy as np
import pytest

from chronos_b200 import _ffi
from helpers import canonical
from oracle import coracle

pytestmark = pytest.mark.gpu


def _make_ranks(world, n_max, cap_pair):
    import chronos_b200

    ctxs = [chronos_b200.Context(device=0, max_batch_events=max(n_max, world * cap_pair)) for _ in range(world)]
    wins = [c.exchange_create(world, r, cap_pair)[0] for r, c in enumerate(ctxs)]
    for c in ctxs:
        c.exchange_connect(world, windows=wins)
    return ctxs


def _step(ctxs, slices, bases, order=None, drain=True):
    """One step on every emulated rank; returns per-rank BatchResult."""
    world = len(ctxs)
    order = list(order or range(world))
    for r in order:
        ctxs[r].filter_exchange(slices[r], bases[r])
    if drain:
        assert _ffi.lib().chronos_device_synchronize(0) == 0
    for r in order:
        ctxs[r].submit_exchanged()
    return [ctxs[r].fetch(records=True) for r in range(world)]


def _merge(results):
    merged = []
    for res in results:
        o = 0
        for w in res.windows:
            ln = int(w["len"])
            merged.append((int(w["trigger_idx"]), int(w["pid"]), int(w["flags"]) & 1, res.window_records[o:o + ln].copy()))
            o += ln
    merged.sort(key=lambda t: t[0])
    return merged


@pytest.mark.parametrize("world,n_rows,name", [(1, 120_000, "skewed_zipf_chains"), (2, 120_000, "skewed_zipf_chains"),
                                               (3, 120_000, "skewed_zipf_chains"), (8, 120_000, "skewed_zipf_chains"),
                                               (2, 1_200_000, "skewed_zipf_chains"),
                                               (4, 1_400_000, "sharded_100m_1m")])   # the 8-GPU bench's 1 M-PID trace
def test_direct_exchange_matches_oracle_over_two_steps(world, n_rows, name):
    import chronos_b200
    from chronos_b200 import tracegen

    spec = dict(tracegen.WORKLOADS[name])
    rows = tracegen.generate_host(spec, n=n_rows)
    half = len(rows) // 2
    eng = coracle.COracle()
    ctxs = _make_ranks(world, half, cap_pair=half // 2)
    total_windows = 0
    for step, (lo, hi) in enumerate([(0, half), (half, len(rows))]):
        batch = rows[lo:hi]
        n = len(batch)
        bounds = [r * n // world for r in range(world + 1)]
        slices = [batch[bounds[r]:bounds[r + 1]] for r in range(world)]
        bases = [lo + bounds[r] for r in range(world)]
        # ranks enqueue in a different order on the second step: nobody may depend on it
        results = _step(ctxs, slices, bases, order=range(world) if step == 0 else reversed(range(world)))
        ref = eng.feed(batch, lo)
        keep, _ = coracle.classify(batch)
        got = np.sort(np.concatenate([r.survivors.astype(np.int64) for r in results]))
        np.testing.assert_array_equal(got, lo + np.nonzero(keep)[0])
        for d, res in enumerate(results):
            assert np.all(np.diff(res.survivors.astype(np.int64)) > 0), "owner's entries must arrive in arrival order"
            if len(res.window_records):
                pid = res.window_records[:, :4].copy().view("<u4").reshape(-1)
                assert all(chronos_b200.owner_of(int(p), world) == d for p in pid[:300])
        merged = _merge(results)
        assert [m[0] for m in merged] == ref["win_trigger"].tolist()
        assert [m[1] for m in merged] == ref["win_pid"].tolist()
        assert [m[2] for m in merged] == (ref["win_flags"] & 1).tolist()
        if len(ref["members"]):
            exp = canonical(rows[ref["members"].astype(np.int64)])
            np.testing.assert_array_equal(canonical(np.concatenate([m[3] for m in merged])), exp)
        total_windows += len(merged)
        chains = sorted((int(i), int(p)) for r in results for p, i in zip(r.chain_pid, r.chain_idx))
        assert chains == sorted(zip(ref["chain_idx"].tolist(), ref["chain_pid"].tolist()))
    assert total_windows > 100
    # per-PID state after both steps = the oracle's process_memory, split by owner
    _, _, members = eng.state()
    assert sum(r.counts["n_pending"] for r in results) == len(members)
    pend = np.sort(np.concatenate([c.state_arrays(r.counts["n_pending"], records=False)[1]
                                   for c, r in zip(ctxs, results)]).astype(np.uint64))
    np.testing.assert_array_equal(pend, np.sort(members))
    for c in ctxs:
        c.close()


def test_empty_slices_and_empty_steps():
    from chronos_b200 import tracegen

    rows = tracegen.generate_host(dict(tracegen.WORKLOADS["attack_replay_10k"]), n=3000)
    world = 3
    ctxs = _make_ranks(world, len(rows), cap_pair=len(rows))
    empty = rows[:0]
    # everything on rank 1, the others contribute empty slices
    results = _step(ctxs, [empty, rows, empty], [0, 0, len(rows)])
    ref = coracle.COracle().feed(rows, 0)
    merged = _merge(results)
    assert [m[0] for m in merged] == ref["win_trigger"].tolist()
    # a step in which nobody has anything
    results = _step(ctxs, [empty, empty, empty], [0, 0, 0])
    assert all(r.counts["n_survivors"] == 0 and r.counts["n_windows"] == 0 for r in results)
    for c in ctxs:
        c.close()


def test_window_overflow_is_reported_not_truncated():
    from chronos_b200 import _ffi, tracegen

    rows = tracegen.generate_host(dict(tracegen.WORKLOADS["uniform_10m_4k"]), n=60_000)
    keep, _ = coracle.classify(rows)
    assert keep.sum() > 2000
    world = 2
    ctxs = _make_ranks(world, len(rows), cap_pair=256)  # far too small
    for r in range(world):
        ctxs[r].filter_exchange(rows[r * 30_000:(r + 1) * 30_000], r * 30_000)
    assert _ffi.lib().chronos_device_synchronize(0) == 0
    for r in range(world):
        ctxs[r].submit_exchanged()
    for r in range(world):
        with pytest.raises(_ffi.ChronosError) as ei:
            ctxs[r].fetch_counts()
        assert ei.value.code == _ffi.E_CAPACITY
    for c in ctxs:
        c.close()


def test_missing_peer_times_out_instead_of_hanging():
    from chronos_b200 import _ffi, tracegen

    rows = tracegen.generate_host(dict(tracegen.WORKLOADS["attack_replay_10k"]), n=2000)
    ctxs = _make_ranks(2, len(rows), cap_pair=len(rows))
    ctxs[0].exchange_set_timeout(0.5)
    ctxs[0].filter_exchange(rows, 0)   # rank 1 never sends
    ctxs[0].submit_exchanged()
    with pytest.raises(_ffi.ChronosError) as ei:
        ctxs[0].fetch_counts()
    assert ei.value.code == _ffi.E_STATE
    for c in ctxs:
        c.close()


def test_protocol_misuse_is_rejected():
    import chronos_b200
    from chronos_b200 import _ffi, tracegen

    rows = tracegen.generate_host(dict(tracegen.WORKLOADS["attack_replay_10k"]), n=500)
    c = chronos_b200.Context(device=0, max_batch_events=4096)
    with pytest.raises(_ffi.ChronosError):
        c.filter_exchange(rows, 0)             # no window yet
    with pytest.raises(_ffi.ChronosError):
        c.exchange_create(17, 0, 64)           # world too large
    with pytest.raises(_ffi.ChronosError):
        c.exchange_create(2, 0, 1 << 20)       # 2 x capacity exceeds max_batch_events
    c.exchange_create(1, 0, 4096)
    with pytest.raises(_ffi.ChronosError):
        c.submit_exchanged()                   # nothing sent
    c.filter_exchange(rows, 0)
    with pytest.raises(_ffi.ChronosError):
        c.filter_exchange(rows, 0)             # previous step not consumed
    c.submit_exchanged()
    res = c.fetch(records=True)
    ref = coracle.COracle().feed(rows, 0)
    assert res.counts["n_windows"] == len(ref["win_pid"])
    c.close()
