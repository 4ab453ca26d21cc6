"""Checkpoint of the per-PID state (chronos_checkpoint_save/_load) and replay from a trace
file: a stream that is interrupted, saved and resumed in a NEW context must produce exactly
what the uninterrupted stream -- and the oracle -- produce."""
import numpy as np
import pytest

from helpers import Trace, compare_batch
from oracle import coracle

pytestmark = pytest.mark.gpu


def test_resume_from_checkpoint_equals_uninterrupted_stream(tmp_path):
    import chronos_b200
    from chronos_b200 import _ffi, tracefile, tracegen

    spec = dict(tracegen.WORKLOADS["skewed_zipf_chains"])
    rows = tracegen.generate_host(spec, n=160_000)
    path = tmp_path / "stream.chrn"
    tracefile.write_trace(path, rows, base_arrival_index=1_000_000, meta={"spec": "skewed_zipf_chains"})
    batches = list(tracefile.iter_batches(path, 20_000))
    assert len(batches) == 8
    eng, tr = coracle.COracle(), Trace()
    a = chronos_b200.Context(device=0, max_batch_events=20_000)
    for part, base in batches[:4]:
        compare_batch(a, eng, tr, part, base)
    blob = a.checkpoint_save()
    assert blob[:8] == b"CHRNCKP1"
    c_a = a.fetch_counts()
    assert c_a["n_pending"] > 0 and c_a["n_pids_seen"] > 100
    # a second save gives the same bytes (entries are sorted), and the state is still usable
    assert a.checkpoint_save() == blob
    b = chronos_b200.Context(device=0, max_batch_events=20_000)
    b.checkpoint_load(blob)
    assert b.checkpoint_save() == blob
    # both contexts continue; b (restored) must match the oracle that never stopped
    for part, base in batches[4:]:
        compare_batch(b, eng, tr, part, base)
    # final state equal to the oracle's
    pids, offs, members = eng.state()
    cb = b.fetch_counts()
    assert cb["n_pending"] == len(members) and cb["n_pids_seen"] == len(pids)
    np.testing.assert_array_equal(np.sort(b.state_pids(cb["n_pids_seen"])), pids)
    # loading into a context that already holds other state replaces it
    a.checkpoint_load(blob)
    assert a.checkpoint_save() == blob
    # malformed input is rejected
    with pytest.raises(_ffi.ChronosError):
        b.checkpoint_load(b"garbage" * 20)
    with pytest.raises(_ffi.ChronosError):
        b.checkpoint_load(blob[:-17])
    for c in (a, b):
        c.close()


def test_empty_state_round_trip():
    import chronos_b200

    c = chronos_b200.Context(device=0, max_batch_events=1024)
    blob = c.checkpoint_save()
    assert len(blob) == 64
    c.checkpoint_load(blob)
    c.submit(b"", 0)
    assert c.fetch_counts()["n_pending"] == 0
    c.close()
