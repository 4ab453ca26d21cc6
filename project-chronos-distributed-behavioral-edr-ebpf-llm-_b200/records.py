"""Host-side view of `struct data_t` (chronos_sensor.py:18-23) and the strings the
reference derives from it. Only flushed windows and pending state are ever formatted
here; the per-event work happens on the GPU."""
import numpy as np

RECORD_BYTES = 288
OFF_PID, OFF_COMM, OFF_ARGV, OFF_TYPE = 0, 4, 20, 276
CAP_COMM, CAP_ARGV, CAP_TYPE = 16, 256, 10

RECORD_DTYPE = np.dtype(
    {"names": ["pid", "comm", "argv", "type"],
     "formats": ["<u4", "S16", "S256", "S10"],
     "offsets": [OFF_PID, OFF_COMM, OFF_ARGV, OFF_TYPE],
     "itemsize": RECORD_BYTES})


def _field(rec, off, cap):
    f = bytes(rec[off:off + cap])
    k = f.find(b"\0")
    return f if k < 0 else f[:k]


def pack_record(pid, comm=b"", argv=b"", etype=b"OPEN"):
    """One zero-filled record as the hooks build it (chronos_sensor.py:52-55, :69-72, :94)."""
    def fld(b, cap):
        b = bytes(b)[:cap]
        return b + b"\0" * (cap - len(b))
    return (int(pid) & 0xFFFFFFFF).to_bytes(4, "little") + fld(comm, CAP_COMM) + fld(argv, CAP_ARGV) + \
        fld(etype, CAP_TYPE) + b"\0\0"


def event_str(rec):
    """f"[{etype}] {cmd} -> {args}" (chronos_sensor.py:128-130, :137) for a surviving record."""
    cmd = _field(rec, OFF_COMM, CAP_COMM).decode("utf-8")
    args = _field(rec, OFF_ARGV, CAP_ARGV).decode("utf-8")
    etype = _field(rec, OFF_TYPE, CAP_TYPE).decode("utf-8")
    return f"[{etype}] {cmd} -> {args}"


def as_record_bytes(records):
    """Accept bytes / bytearray / memoryview / numpy array; return a C-contiguous uint8 array."""
    if isinstance(records, np.ndarray):
        a = np.ascontiguousarray(records).view(np.uint8).reshape(-1)
    else:
        a = np.frombuffer(records, dtype=np.uint8)
    if a.size % RECORD_BYTES:
        raise ValueError(f"record buffer size {a.size} is not a multiple of {RECORD_BYTES}")
    return a
