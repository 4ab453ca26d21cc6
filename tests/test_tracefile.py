"""Trace file format (host-side, no GPU): round trip, batching, rejection of foreign files."""
import numpy as np
import pytest

from oracle import pyoracle


def _rows(n):
    recs = [pyoracle.make_record(100 + i % 7, b"sh", b"/tmp/f%d" % i, b"OPEN" if i % 3 else b"EXEC") for i in range(n)]
    return np.frombuffer(b"".join(recs), dtype=np.uint8).reshape(-1, 288)


def test_round_trip_and_batches(tmp_path):
    from chronos_b200 import tracefile

    rows = _rows(1000)
    p = tmp_path / "t.chrn"
    assert tracefile.write_trace(p, rows, base_arrival_index=5000, meta={"seed": 7, "workload": "unit"}) == 1000
    for mm in (True, False):
        got, base, meta = tracefile.read_trace(p, mmap=mm)
        assert base == 5000 and meta == {"seed": 7, "workload": "unit"}
        np.testing.assert_array_equal(np.asarray(got), rows)
    parts = list(tracefile.iter_batches(p, 300))
    assert [len(x) for x, _ in parts] == [300, 300, 300, 100]
    assert [b for _, b in parts] == [5000, 5300, 5600, 5900]
    np.testing.assert_array_equal(np.concatenate([x for x, _ in parts]), rows)
    assert tracefile.read_header(p)["data_offset"] % 64 == 0


def test_empty_and_foreign_files(tmp_path):
    from chronos_b200 import tracefile

    p = tmp_path / "e.chrn"
    tracefile.write_trace(p, b"")
    got, base, meta = tracefile.read_trace(p)
    assert got.shape == (0, 288) and base == 0 and meta == {}
    q = tmp_path / "x.bin"
    q.write_bytes(b"\0" * 200)
    with pytest.raises(ValueError):
        tracefile.read_trace(q)
    q.write_bytes(b"short")
    with pytest.raises(ValueError):
        tracefile.read_trace(q)
    # truncated payload
    rows = _rows(10)
    tracefile.write_trace(p, rows)
    raw = p.read_bytes()
    p.write_bytes(raw[:-100])
    with pytest.raises(ValueError):
        tracefile.read_trace(p)
