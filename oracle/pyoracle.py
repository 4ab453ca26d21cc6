"""Pure-Python restatement of the reference hot path.  TEST INFRASTRUCTURE ONLY.

Small-case oracle and the timed "port" CPU baseline: it performs the same interpreter
operations the reference's callback performs (bytes.decode, `in`, f-string, dict of
lists), so its speed is representative of chronos_sensor.py itself.  Pinned against
the unmodified reference by tests/test_oracle_pinning.py (build container) and against
tests/golden/*.json everywhere.  Citations are into /root/reference/chronos_sensor.py.
"""
from collections import defaultdict

REC_BYTES = 288
OFF_COMM, OFF_ARGV, OFF_TYPE = 4, 20, 276

PREFIX_A = (b"/lib", b"/usr/lib", b"/usr/share", b"/etc/ssl", b"/etc/fonts", b"/etc/host")  # :76-78
SUFFIX = (b".so", b".cache", b".mo", b".conf", b".crt", b".curlrc")  # :83-85
PREFIX_B = (b"/dev/", b"/proc/")  # :90
IGNORE = ["node", "code", "ollama", "python", "chrome", "vmtools", "git"]  # :134
TRIGGERS = ["curl", "chmod", "bash", "nc", "cat"]  # :141


def make_record(pid, comm=b"", argv=b"", etype=b"OPEN", junk=b""):
    """Build one 288-byte struct data_t (chronos_sensor.py:18-23). A field longer than
    its capacity is cut (it then has no NUL). `junk`, if given, is repeated into the
    bytes AFTER each field's NUL (a live sensor zero-fills, a replay need not)."""
    def fld(b, cap):
        b = bytes(b)[:cap]
        if len(b) == cap:
            return b
        pad = cap - len(b) - 1
        fill = (junk * (pad // len(junk) + 1))[:pad] if junk else b"\0" * pad
        return b + b"\0" + fill
    return (int(pid) & 0xFFFFFFFF).to_bytes(4, "little") + fld(comm, 16) + fld(argv, 256) + fld(etype, 10) + b"\0\0"


def c_char_field(rec, off, cap):
    """ctypes c_char[N] member read (:125, :128-130): bytes before the first NUL."""
    f = rec[off:off + cap]
    k = f.find(b"\0")
    return f if k < 0 else f[:k]


def openat_noise(argv_field):
    """The openat hook's drop decision (:74-92) on the raw 256-byte field."""
    k = argv_field.find(b"\0")
    s = argv_field if k < 0 else argv_field[:k]
    # starts_with (:41-47) compares bytes up to the prefix NUL; a string NUL mismatches.
    if any(s.startswith(p) for p in PREFIX_A):
        return True
    # ends_with (:28-38): length from first NUL (<=256), suffix must fit.
    if any(s.endswith(x) for x in SUFFIX):
        return True
    return any(s.startswith(p) for p in PREFIX_B)


def event_str_of(rec):
    """event_str (:137) of a record that decodes; raises UnicodeDecodeError otherwise."""
    cmd = c_char_field(rec, OFF_COMM, 16).decode("utf-8")
    args = c_char_field(rec, OFF_ARGV, 256).decode("utf-8")
    etype = c_char_field(rec, OFF_TYPE, 10).decode("utf-8")
    return f"[{etype}] {cmd} -> {args}"


class PySensor:
    """process_memory + print_event (:105, :124-157) over raw records."""

    def __init__(self):
        self.process_memory = defaultdict(list)  # :105
        self.windows = []     # (pid, [event_str]) as handed to analyze_sequence (:147)
        self.survivors = []   # arrival indices appended to process_memory
        self.trigger_idx = []

    def feed(self, records, base=0, drop=None):
        """drop: optional per-record 0/1 decisions of stage A already taken by the
        reference's own C helpers (oracle/_ref); without it the filter is restated here."""
        n = len(records) // REC_BYTES
        pm = self.process_memory
        for i in range(n):
            rec = records[i * REC_BYTES:(i + 1) * REC_BYTES]
            etype_b = c_char_field(rec, OFF_TYPE, 10)
            if drop is not None:
                if drop[i]:
                    continue
            elif etype_b == b"OPEN" and openat_noise(rec[OFF_ARGV:OFF_ARGV + 256]):
                continue  # never submitted by the kernel hook (:79, :86, :91)
            pid = int.from_bytes(rec[0:4], "little")
            try:  # :127-131
                cmd = c_char_field(rec, OFF_COMM, 16).decode("utf-8")
                args = c_char_field(rec, OFF_ARGV, 256).decode("utf-8")
                etype = etype_b.decode("utf-8")
            except UnicodeDecodeError:
                continue
            if any(x in cmd for x in IGNORE):  # :134-135
                continue
            event_str = f"[{etype}] {cmd} -> {args}"  # :137
            pm[pid].append(event_str)  # :138
            self.survivors.append(base + i)
            if any(x in event_str for x in TRIGGERS):  # :141
                if len(pm[pid]) >= 2:  # :142
                    self.windows.append((pid, pm[pid]))  # :147
                    self.trigger_idx.append(base + i)
                    pm[pid] = []  # :157
        return self


def prompt_of(history):
    """The prompt analyze_sequence builds (:109-114), byte for byte."""
    return f"""
    Analyze this sequence. Return JSON ONLY.
    Sequence: {history}
    Context: 'curl' -> 'chmod' -> 'exec' is a Dropper.
    Format: {{"risk_score": <0-10>, "verdict": "<SAFE/MALICIOUS>", "reason": "<Short Explanation>"}}
    """
