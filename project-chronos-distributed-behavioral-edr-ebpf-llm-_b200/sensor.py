"""Drop-in for the user-space half of chronos_sensor.py.

`GpuSensor` offers the same surface the reference script wires together at module level:
a BCC-compatible `print_event(cpu, data, size)` callback (chronos_sensor.py:124, :160),
the `process_memory` mapping (:105) and the `analyze_sequence(pid, history)` hook (:107),
and prints the same lines (:143-155). The per-event work (openat noise filter, decode
gate, ignore-list, per-PID lists, trigger/flush) runs on the GPU in micro-batches; the
LLM call itself is unchanged and out of scope -- pass the reference's own function, or
use `default_analyze_sequence`, which builds the same prompt and POST body.
"""
import ctypes
import json

from .engine import Context
from .records import RECORD_BYTES


def build_prompt(history):
    """The prompt of chronos_sensor.py:109-114; `history` is interpolated with str(list)."""
    return f"""
    Analyze this sequence. Return JSON ONLY.
    Sequence: {history}
    Context: 'curl' -> 'chmod' -> 'exec' is a Dropper.
    Format: {{"risk_score": <0-10>, "verdict": "<SAFE/MALICIOUS>", "reason": "<Short Explanation>"}}
    """


def default_analyze_sequence(ai_url, post=None, out=print):
    """analyze_sequence as the reference defines it (chronos_sensor.py:107-122): same
    progress line, prompt, request body, 30 s timeout and error convention."""
    def analyze_sequence(pid, history):
        out(f"    [?] Analyzing Behavioral Chain for PID {pid} (Waiting for AI)...")
        prompt = build_prompt(history)
        try:
            poster = post
            if poster is None:
                import requests

                poster = requests.post
            response = poster(ai_url, json={"model": "llama3", "prompt": prompt, "stream": False, "format": "json"},
                              timeout=30)
            return json.loads(response.json()["response"])
        except Exception as e:  # noqa: BLE001  (the reference catches everything, :121)
            return {"risk_score": 0, "verdict": "ERROR", "reason": str(e)}
    return analyze_sequence


class GpuSensor:
    def __init__(self, analyze_sequence, device=0, batch_events=65536, out=print):
        self.analyze_sequence = analyze_sequence
        self.batch_events = int(batch_events)
        self.out = out
        self._ctx = Context(device=device, max_batch_events=self.batch_events)
        self._buf = (ctypes.c_char * (RECORD_BYTES * self.batch_events))()
        self._fill = 0
        self._arrival = 0
        self._n_pending = 0
        self._n_pids = 0
        self.results = []  # what analyze_sequence returned, in order

    # ---- ingest ------------------------------------------------------------------------
    def print_event(self, cpu, data, size):
        """Same signature BCC calls (chronos_sensor.py:124): `data` points at one struct data_t."""
        addr = data.value if isinstance(data, ctypes.c_void_p) else int(data)
        ctypes.memmove(ctypes.addressof(self._buf) + RECORD_BYTES * self._fill, addr, RECORD_BYTES)
        self._fill += 1
        if self._fill == self.batch_events:
            self.flush()

    def feed(self, records):
        """Replay a buffer of raw records (bytes-like, n*288) in arrival order."""
        mv = memoryview(records).cast("B")
        n = len(mv) // RECORD_BYTES
        done = 0
        while done < n:
            take = min(n - done, self.batch_events - self._fill)
            ctypes.memmove(ctypes.addressof(self._buf) + RECORD_BYTES * self._fill,
                           bytes(mv[done * RECORD_BYTES:(done + take) * RECORD_BYTES]), take * RECORD_BYTES)
            self._fill += take
            done += take
            if self._fill == self.batch_events:
                self.flush()

    def flush(self):
        """Process what has accumulated (call it after every perf_buffer_poll(), :162)."""
        if self._fill == 0:
            return
        self._ctx.submit(memoryview(self._buf)[: RECORD_BYTES * self._fill], self._arrival)
        res = self._ctx.fetch(records=True)
        self._arrival += self._fill
        self._fill = 0
        self._n_pending = res.counts["n_pending"]
        self._n_pids = res.counts["n_pids_seen"]
        out = self.out
        for pid, history in res.histories():
            out(f"\n[!] CAPTURED KILL CHAIN (PID {pid}):")           # :143
            for step in history:
                out(f"    {step}")                                    # :144-145
            result = self.analyze_sequence(pid, history)              # :147
            self.results.append(result)
            score = result.get("risk_score", 0)                       # :149
            if score > 5:
                out(f"\033[91m    ==> ALERT: {result.get('verdict')} (Risk {score})\033[0m")
                out(f"    ==> REASON: {result.get('reason')}")
            else:
                out(f"\033[92m    ==> CLEAN: {result.get('verdict')} (Risk {score})\033[0m")
                out(f"    ==> REASON: {result.get('reason')}")
        return res

    # ---- state -------------------------------------------------------------------------
    @property
    def process_memory(self):
        """{pid: [event_str, ...]} including the PIDs whose list is currently [] (:105, :157).
        Events still sitting in the ingest buffer are flushed first."""
        self.flush()
        return self._ctx.process_memory(self._n_pending, self._n_pids)

    def close(self):
        self._ctx.close()
