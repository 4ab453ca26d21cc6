"""Parity at BASELINE.json's full sizes (SURVEY 8d "parity gate"): configs[1] (10 M events,
4 K PIDs), configs[3] (skewed, Zipf PIDs, planted chains) one-shot, and configs[4] (1000
micro-batches of 64 K events through one context) against the C oracle fed the same bytes.

The traces come from the device generator (bit-identical host twin: test_gpu_parity.py) and
are downloaded for the oracle; nothing here reads /root/reference."""
import numpy as np
import pytest

from oracle import coracle

pytestmark = pytest.mark.gpu
REC = 288


def _check_batch(res, ref, rows_of, base, n, check_records=True):
    """res: BatchResult of the CUDA path, ref: oracle output for the same batch.
    rows_of(idx array) -> the records with those arrival indices."""
    assert res.counts["n_events"] == n
    np.testing.assert_array_equal(res.survivors, ref["survivors"], err_msg="survivor set")
    assert res.counts["n_windows"] == len(ref["win_pid"])
    np.testing.assert_array_equal(res.windows["pid"], ref["win_pid"])
    np.testing.assert_array_equal(res.windows["trigger_idx"], ref["win_trigger"])
    np.testing.assert_array_equal(res.windows["len"], np.diff(ref["win_off"]).astype(np.uint32))
    np.testing.assert_array_equal(res.windows["flags"] & 1, ref["win_flags"] & 1)
    np.testing.assert_array_equal(res.chain_pid, ref["chain_pid"])
    np.testing.assert_array_equal(res.chain_idx, ref["chain_idx"])
    if check_records and len(ref["members"]):
        exp = rows_of(ref["members"].astype(np.int64))
        assert res.window_records.shape == exp.shape
        assert np.array_equal(res.window_records, exp), "window records differ"


@pytest.mark.parametrize("name", ["uniform_10m_4k", "skewed_zipf_chains"])
def test_one_shot_10m_events_bit_exact(name):
    import chronos_b200
    from chronos_b200 import tracegen

    spec = tracegen.WORKLOADS[name]
    n = 10_000_000
    dev = tracegen.generate_device(spec, device=0, first=0, n=n)
    rows = dev.download().reshape(-1, REC)
    ctx = chronos_b200.Context(device=0, max_batch_events=n)
    ctx.submit_ptr(dev.ptr, n, 0)
    res = ctx.fetch(records=True)
    eng = coracle.COracle()
    ref = eng.feed(rows, 0)
    keep, meta = coracle.classify(rows)
    np.testing.assert_array_equal(res.survivor_meta & 0x0F, meta[keep == 1] & 0x0F)
    _check_batch(res, ref, lambda idx: rows[idx], 0, n)
    # final process_memory (a13): pending entries and the PIDs ever seen
    pids, offs, members = eng.state()
    assert res.counts["n_pending"] == len(members)
    assert res.counts["n_pids_seen"] == len(pids)
    gp, gi, grec = ctx.state_arrays(res.counts["n_pending"], records=True)
    np.testing.assert_array_equal(gi, members)
    np.testing.assert_array_equal(gp, np.repeat(pids, np.diff(offs).astype(np.int64)))
    assert np.array_equal(grec, rows[members.astype(np.int64)])
    np.testing.assert_array_equal(np.sort(ctx.state_pids(res.counts["n_pids_seen"])), pids)
    if name == "skewed_zipf_chains":
        assert res.counts["n_chains"] > 0, "planted chains must be found"
    # size-independent properties: every survivor is in exactly one window or still pending
    assert int(res.windows["len"].sum()) + res.counts["n_pending"] == res.counts["n_survivors"]
    assert np.all(np.diff(res.windows["trigger_idx"].astype(np.int64)) > 0)
    ctx.close()
    dev.free()


def test_stream_of_1000_micro_batches_equals_the_oracle_stream():
    import chronos_b200
    from chronos_b200 import tracegen

    spec = tracegen.WORKLOADS["stream_64k"]
    B, n_batches, group = 65_536, 1000, 20
    ctx = chronos_b200.Context(device=0, max_batch_events=B)
    eng = coracle.COracle()
    total_w = total_s = 0
    for g0 in range(0, n_batches, group):
        dev = tracegen.generate_device(spec, device=0, first=g0 * B, n=group * B)
        rows = dev.download().reshape(-1, REC)
        for k in range(group):
            i = g0 + k
            batch = rows[k * B:(k + 1) * B]
            ctx.submit_ptr(dev.ptr + k * B * REC, B, i * B)
            # window records are compared on every 25th batch (members can be arbitrarily old:
            # they are regenerated from the deterministic trace)
            with_records = i % 25 == 0
            res = ctx.fetch(records=with_records)
            ref = eng.feed(batch, i * B)

            def rows_of(idx):
                out = np.empty((len(idx), REC), dtype=np.uint8)
                for j, a in enumerate(idx):
                    out[j] = tracegen.generate_host(spec, first=int(a), n=1)[0]
                return out

            _check_batch(res, ref, rows_of, i * B, B, check_records=with_records)
            total_w += res.counts["n_windows"]
            total_s += res.counts["n_survivors"]
        dev.free()
    assert total_w > 100_000 and total_s > total_w
    pids, offs, members = eng.state()
    c = res.counts
    assert c["n_pending"] == len(members) and c["n_pids_seen"] == len(pids)
    _, gi, _ = ctx.state_arrays(c["n_pending"], records=False)
    np.testing.assert_array_equal(gi, members)
    ctx.close()
