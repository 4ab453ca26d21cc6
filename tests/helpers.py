"""Shared comparison code: CUDA path (through the C ABI) vs the CPU oracle."""
import numpy as np

from oracle import coracle

REC = 288


def as_rows(records):
    a = np.frombuffer(records, dtype=np.uint8) if isinstance(records, (bytes, bytearray)) else np.asarray(records)
    return a.reshape(-1, REC)


class Trace:
    """Keeps every record fed so far so arrival indices can be mapped back to bytes."""

    def __init__(self):
        self.chunks = []
        self.bases = []

    def add(self, rows, base):
        self.chunks.append(rows)
        self.bases.append(base)

    def record(self, idx):
        for rows, base in zip(self.chunks, self.bases):
            if base <= idx < base + len(rows):
                return rows[idx - base]
        raise KeyError(idx)


def compare_batch(ctx, engine, trace, records, base=0, check_state=True):
    """Submit one batch to the GPU context and the oracle engine; assert bit-exact parity
    of survivor set, meta bits, ordered windows (incl. record bytes) and pending state."""
    rows = as_rows(records)
    trace.add(rows, base)
    ctx.submit(rows, base)
    res = ctx.fetch(records=True)
    ref = engine.feed(rows, base)

    assert res.counts["n_events"] == len(rows)
    np.testing.assert_array_equal(res.survivors, ref["survivors"], err_msg="survivor set differs")
    keep, meta = coracle.classify(rows)
    np.testing.assert_array_equal(res.survivor_meta & 0x0F, meta[keep == 1] & 0x0F, err_msg="meta bits differ")

    assert res.counts["n_windows"] == len(ref["win_pid"]), (res.counts, len(ref["win_pid"]))
    np.testing.assert_array_equal(res.windows["pid"], ref["win_pid"])
    np.testing.assert_array_equal(res.windows["trigger_idx"], ref["win_trigger"])
    ref_len = np.diff(ref["win_off"]).astype(np.uint32)
    np.testing.assert_array_equal(res.windows["len"], ref_len)
    # extension outputs (semantics defined by the oracle, "parity unpinned" vs the reference)
    np.testing.assert_array_equal(res.windows["flags"] & 1, ref["win_flags"] & 1, err_msg="t1105 window label")
    np.testing.assert_array_equal(res.chain_pid, ref["chain_pid"], err_msg="chain pids")
    np.testing.assert_array_equal(res.chain_idx, ref["chain_idx"], err_msg="chain arrival indices")
    # window contents, byte for byte, in window order
    if len(ref["members"]):
        exp = np.stack([trace.record(int(i)) for i in ref["members"]])
        np.testing.assert_array_equal(res.window_records, exp, err_msg="window records differ")
    else:
        assert res.window_records.shape[0] == 0

    if check_state:
        pids, offs, mem = engine.state()
        exp_pid = np.repeat(pids, np.diff(offs).astype(np.int64))
        assert res.counts["n_pending"] == len(mem)
        assert res.counts["n_pids_seen"] == len(pids)  # keys of process_memory, incl. PIDs whose list is []
        np.testing.assert_array_equal(np.sort(ctx.state_pids(len(pids))), pids)
        gp, gi, grec = ctx.state_arrays(res.counts["n_pending"], records=True)
        np.testing.assert_array_equal(gp, exp_pid, err_msg="pending pids differ")
        np.testing.assert_array_equal(gi, mem, err_msg="pending arrival indices differ")
        if len(mem):
            exp = np.stack([trace.record(int(i)) for i in mem])
            np.testing.assert_array_equal(grec, exp, err_msg="pending records differ")
    return res, ref


def canonical(rows):
    """Records as the transport form delivers them: bytes behind each field's NUL zeroed
    (a live sensor zero-fills them anyway, chronos_sensor.py:52/:69)."""
    out = np.array(rows, dtype=np.uint8, copy=True).reshape(-1, REC)
    for off, cap in ((4, 16), (20, 256), (276, 10)):
        f = out[:, off:off + cap]
        seen_nul = np.cumsum(f == 0, axis=1) > 0
        f[seen_nul] = 0
    out[:, 286:288] = 0
    return out
