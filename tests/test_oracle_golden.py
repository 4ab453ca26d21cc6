"""CPU: the oracle (C restatement and Python port) against the committed golden vectors,
which were produced by the unmodified reference (tests/golden/make_golden.py)."""
import json
import os

import numpy as np
import pytest

from oracle import coracle, pyoracle

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def load(name):
    with open(os.path.join(GOLD, name)) as f:
        return json.load(f)


def fixture_records(name):
    d = load(name)
    if "records_hex" in d:
        return bytes.fromhex(d["records_hex"]), d["reference"]
    import chronos_b200  # noqa: F401  (host twin of the generator lives in the product library)
    from chronos_b200 import tracegen
    import hashlib

    spec = d["generator"]
    rows = tracegen.generate_host(spec, n=spec["n"])
    data = rows.tobytes()
    assert hashlib.sha256(data).hexdigest() == d["records_sha256"], "trace generator drifted: regenerate the fixtures"
    return data, d["reference"]


def windows_from_coracle(out, data):
    rows = np.frombuffer(data, dtype=np.uint8).reshape(-1, 288)
    res = []
    for w in range(len(out["win_pid"])):
        lo, hi = int(out["win_off"][w]), int(out["win_off"][w + 1])
        res.append([int(out["win_pid"][w]), [pyoracle.event_str_of(bytes(rows[int(i)])) for i in out["members"][lo:hi]]])
    return res


def state_from_coracle(eng, data):
    rows = np.frombuffer(data, dtype=np.uint8).reshape(-1, 288)
    pids, offs, mem = eng.state()
    return {str(int(p)): [pyoracle.event_str_of(bytes(rows[int(i)])) for i in mem[int(offs[k]):int(offs[k + 1])]]
            for k, p in enumerate(pids)}


@pytest.mark.parametrize("name", ["screenshot_kat.json", "adversarial.json", "config1_trace.json"])
def test_c_oracle_matches_reference_output(name):
    data, ref = fixture_records(name)
    eng = coracle.COracle()
    out = eng.feed(data, 0)
    assert out["survivors"].tolist() == ref["survivors"]
    assert windows_from_coracle(out, data) == ref["windows"]
    assert state_from_coracle(eng, data) == ref["process_memory"]  # includes the [] of flushed PIDs


@pytest.mark.parametrize("name", ["screenshot_kat.json", "adversarial.json", "config1_trace.json"])
def test_python_port_matches_reference_output(name):
    data, ref = fixture_records(name)
    s = pyoracle.PySensor().feed(data, 0)
    assert s.survivors == ref["survivors"]
    assert [[p, h] for p, h in s.windows] == ref["windows"]
    assert {str(k): v for k, v in s.process_memory.items()} == ref["process_memory"]


def test_screenshot_windows_are_the_transcribed_ones():
    import cases

    _, ref = fixture_records("screenshot_kat.json")
    assert [(p, h) for p, h in ref["windows"]] == cases.SCREENSHOT_WINDOWS


def test_prompt_format_is_byte_exact():
    for name in ("screenshot_kat.json", "adversarial.json"):
        _, ref = fixture_records(name)
        assert len(ref["prompts"]) == len(ref["windows"])
        for (pid, hist), prompt in zip(ref["windows"], ref["prompts"]):
            assert pyoracle.prompt_of(hist) == prompt


def test_streaming_equals_one_shot_in_the_oracle():
    data, ref = fixture_records("config1_trace.json")
    eng = coracle.COracle()
    wins, surv = [], []
    n = len(data) // 288
    base = 0
    for size in (1, 99, 2500, 3, 4000, n):
        size = min(size, n - base)
        out = eng.feed(data[base * 288:(base + size) * 288], base)
        surv += out["survivors"].tolist()
        rows = np.frombuffer(data, dtype=np.uint8).reshape(-1, 288)
        for w in range(len(out["win_pid"])):
            lo, hi = int(out["win_off"][w]), int(out["win_off"][w + 1])
            wins.append([int(out["win_pid"][w]),
                         [pyoracle.event_str_of(bytes(rows[int(i)])) for i in out["members"][lo:hi]]])
        base += size
        if base == n:
            break
    assert surv == ref["survivors"] and wins == ref["windows"]
