"""Run the UNMODIFIED reference sensor over a buffer of raw records.

TEST INFRASTRUCTURE ONLY.  Works only where /root/reference exists (the build
container); it is how the oracle is pinned and how tests/golden/*.json are produced.

What is faked is everything *around* the hot path, never the hot path itself:
  * `bcc`      -- not installable here (no eBPF).  The fake BPF object accepts the
                  program text, ignores kprobe attachment, and its perf_buffer_poll()
                  replays our records through the callback the script registered
                  (chronos_sensor.py:160-162).  event() is the same ctypes cast BCC does.
  * `requests` -- post() records the prompt and answers like an idle Ollama.
Stage A (the in-kernel openat filter, chronos_sensor.py:74-92) is the reference's own
C helpers compiled by oracle/build_oracle.py::build_ref; stage B is the script itself,
executed with runpy, byte for byte as it lies in /root/reference.
"""
import ctypes
import io
import os
import runpy
import sys
import types
from contextlib import redirect_stdout

REF_SENSOR = "/root/reference/chronos_sensor.py"
REC_BYTES = 288


class Data(ctypes.Structure):
    # struct data_t, chronos_sensor.py:18-23 (TASK_COMM_LEN == 16)
    _fields_ = [
        ("pid", ctypes.c_uint),
        ("comm", ctypes.c_char * 16),
        ("argv", ctypes.c_char * 256),
        ("type", ctypes.c_char * 10),
    ]


assert ctypes.sizeof(Data) == REC_BYTES


def available():
    return os.path.exists(REF_SENSOR)


def _ref_filter_lib():
    from . import build_oracle

    path = build_oracle.build_ref()
    if path is None:
        raise RuntimeError("reference filter library unavailable")
    lib = ctypes.CDLL(path)
    lib.chronos_ref_filter_batch.argtypes = [ctypes.c_char_p, ctypes.c_uint64, ctypes.c_char_p]
    lib.chronos_ref_filter_batch.restype = None
    return lib


def ref_filter_drop(records: bytes):
    """Stage A alone: list of 0/1 drop decisions from the reference's C helpers."""
    n = len(records) // REC_BYTES
    out = ctypes.create_string_buffer(max(n, 1))
    _ref_filter_lib().chronos_ref_filter_batch(records, n, out)
    return list(out.raw[:n])


def run_reference(records: bytes, llm_response='{"risk_score": 0, "verdict": "SAFE", "reason": "stub"}'):
    """Feed `records` (n*288 bytes, arrival order) through filter + unmodified script.

    Returns dict with
      survivors      : arrival indices whose event was appended to process_memory
      windows        : [(pid, [event_str, ...]), ...] in emission order (what
                       analyze_sequence received, chronos_sensor.py:147)
      prompts        : the prompt strings POSTed (chronos_sensor.py:109-119)
      process_memory : {pid: [event_str, ...]} after the last record (a13)
      stdout         : everything the script printed
    """
    if not available():
        raise RuntimeError("reference not present on this machine")
    n = len(records) // REC_BYTES
    drop = ref_filter_drop(records)
    buf = ctypes.create_string_buffer(records, len(records)) if n else ctypes.create_string_buffer(1)
    base = ctypes.addressof(buf)
    out = {"survivors": [], "windows": [], "prompts": []}
    holder = {}

    class _Table:
        def event(self, data):
            return ctypes.cast(data, ctypes.POINTER(Data)).contents

        def open_perf_buffer(self, callback, page_cnt=8, lost_cb=None):
            holder["cb"] = callback

    class _BPF:
        def __init__(self, text=None, **kw):
            self.text = text
            self._table = _Table()
            self._done = False

        def get_syscall_fnname(self, name):
            return "__x64_sys_" + name

        def attach_kprobe(self, event=None, fn_name=None):
            assert fn_name in ("syscall__execve", "syscall__openat")

        def __getitem__(self, key):
            assert key == "events"
            return self._table

        def perf_buffer_poll(self, timeout=-1):
            if self._done:
                raise KeyboardInterrupt
            self._done = True
            cb = holder["cb"]
            pm = cb.__globals__["process_memory"]
            for i in range(n):
                if drop[i]:
                    continue
                pid = ctypes.c_uint.from_address(base + i * REC_BYTES).value
                before = pm.get(pid)
                n0 = len(before) if before is not None else -1
                cb(0, ctypes.c_void_p(base + i * REC_BYTES), REC_BYTES)
                after = pm.get(pid)
                if after is not before or (after is not None and len(after) != n0):
                    out["survivors"].append(i)
            raise KeyboardInterrupt

    class _Resp:
        def json(self):
            return {"response": llm_response}

    def _post(url, json=None, timeout=None, **kw):
        frame = sys._getframe(1)  # analyze_sequence(pid, history)
        out["windows"].append((int(frame.f_locals["pid"]), list(frame.f_locals["history"])))
        out["prompts"].append(json["prompt"])
        return _Resp()

    fake_bcc = types.ModuleType("bcc")
    fake_bcc.BPF = _BPF
    fake_requests = types.ModuleType("requests")
    fake_requests.post = _post
    saved = {k: sys.modules.get(k) for k in ("bcc", "requests")}
    sys.modules["bcc"] = fake_bcc
    sys.modules["requests"] = fake_requests
    stdout = io.StringIO()
    saved_stdin = sys.stdin
    sys.stdin = io.StringIO()  # site's exit() closes sys.stdin; give it a throw-away
    try:
        with redirect_stdout(stdout):
            try:
                runpy.run_path(REF_SENSOR, run_name="__main__")
            except SystemExit:
                pass  # chronos_sensor.py:163 calls exit() on KeyboardInterrupt
    finally:
        sys.stdin = saved_stdin
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
    pm = holder["cb"].__globals__["process_memory"]
    out["process_memory"] = {int(k): list(v) for k, v in pm.items()}
    out["stdout"] = stdout.getvalue()
    return out
