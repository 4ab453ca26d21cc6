"""On-disk replay format for raw sensor records (the reference has none:
it consumes the perf buffer live, chronos_sensor.py:160-162).

A trace file is what BCC would have handed `print_event` (:124-125), record after record:

  offset 0   8 bytes   magic  b"CHRNTRC1"
         8   u32       version (1)
        12   u32       record_bytes (288 = sizeof(struct data_t), :18-23)
        16   u64       n_records
        24   u64       base_arrival_index of the first record
        32   u32       meta_bytes (length of the JSON block that follows the header)
        36   28 bytes  zero
        64   meta_bytes of UTF-8 JSON (generator spec, seed, host, ...), zero-padded to a
             multiple of 64
        ...  n_records x 288 bytes, arrival order

All integers little-endian. Records are stored raw -- noise included -- so that a replay
exercises the filter exactly as a live capture would."""
import json
import struct

import numpy as np

from .records import RECORD_BYTES, as_record_bytes

MAGIC = b"CHRNTRC1"
VERSION = 1
_HDR = struct.Struct("<8sIIQQI28x")
assert _HDR.size == 64


def write_trace(path, records, base_arrival_index=0, meta=None):
    a = as_record_bytes(records)
    blob = json.dumps(meta or {}, sort_keys=True).encode("utf-8")
    pad = (-len(blob)) % 64
    with open(path, "wb") as f:
        f.write(_HDR.pack(MAGIC, VERSION, RECORD_BYTES, a.size // RECORD_BYTES, int(base_arrival_index), len(blob)))
        f.write(blob + b"\0" * pad)
        f.write(a.tobytes() if not a.flags.c_contiguous else memoryview(a))
    return a.size // RECORD_BYTES


def read_header(path):
    with open(path, "rb") as f:
        raw = f.read(_HDR.size)
        if len(raw) < _HDR.size:
            raise ValueError(f"{path}: not a trace file (too short)")
        magic, version, rec_bytes, n, base, meta_len = _HDR.unpack(raw)
        if magic != MAGIC or version != VERSION or rec_bytes != RECORD_BYTES:
            raise ValueError(f"{path}: not a version-{VERSION} trace of {RECORD_BYTES}-byte records")
        meta = json.loads(f.read(meta_len).decode("utf-8")) if meta_len else {}
    data_off = _HDR.size + meta_len + ((-meta_len) % 64)
    return {"n_records": n, "base_arrival_index": base, "meta": meta, "data_offset": data_off}


def read_trace(path, mmap=True):
    """-> (records [n, 288] uint8, base_arrival_index, meta). mmap=True maps the file instead
    of reading it (a 100 M-event trace is 28.8 GB)."""
    h = read_header(path)
    n, off = h["n_records"], h["data_offset"]
    if mmap:
        rows = np.memmap(path, dtype=np.uint8, mode="r", offset=off, shape=(n, RECORD_BYTES)) if n else \
            np.zeros((0, RECORD_BYTES), dtype=np.uint8)
    else:
        with open(path, "rb") as f:
            f.seek(off)
            rows = np.frombuffer(f.read(n * RECORD_BYTES), dtype=np.uint8).reshape(-1, RECORD_BYTES)
    if rows.shape[0] != n:
        raise ValueError(f"{path}: truncated ({rows.shape[0]} of {n} records)")
    return rows, h["base_arrival_index"], h["meta"]


def iter_batches(path, batch_events):
    """Yield (records, base_arrival_index) micro-batches in arrival order."""
    rows, base, _ = read_trace(path, mmap=True)
    for lo in range(0, len(rows), int(batch_events)):
        part = np.ascontiguousarray(rows[lo:lo + int(batch_events)])
        yield part, base + lo
