#!/usr/bin/env python
"""Regenerate tests/golden/*.json by running the UNMODIFIED reference
(/root/reference/chronos_sensor.py under oracle/ref_shim.py) in the build container.

  python tests/golden/make_golden.py

The fixtures hold the input records (hex, or the generator parameters plus a digest for
the synthetic trace) and what the reference produced: surviving arrival indices, the
(pid, history) windows in emission order, the final process_memory and, for the small
sets, the prompts it POSTed.
"""
import hashlib
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle import ref_shim  # noqa: E402
import cases  # noqa: E402


def dump(name, obj):
    with open(os.path.join(HERE, name), "w") as f:
        json.dump(obj, f, indent=0, sort_keys=True)
        f.write("\n")


def run(records, with_prompts):
    buf = b"".join(records)
    out = ref_shim.run_reference(buf)
    res = {
        "survivors": out["survivors"],
        "windows": [[pid, hist] for pid, hist in out["windows"]],
        "process_memory": {str(k): v for k, v in out["process_memory"].items()},
    }
    if with_prompts:
        res["prompts"] = out["prompts"]
        banner = "[+] CHRONOS: LIVE. Stateful Monitoring Active.\n"  # chronos_sensor.py:159
        so = out["stdout"]
        res["stdout_after_banner"] = so[so.index(banner) + len(banner):]
    return buf, res


def main():
    assert ref_shim.available(), "needs /root/reference"
    buf, res = run(cases.screenshot_trace(), True)
    assert [tuple(w) for w in map(lambda w: (w[0], w[1]), res["windows"])] == cases.SCREENSHOT_WINDOWS
    dump("screenshot_kat.json", {"records_hex": buf.hex(), "reference": res})

    buf, res = run(cases.adversarial_trace(), True)
    dump("adversarial.json", {"records_hex": buf.hex(), "reference": res})

    import chronos_b200  # noqa: F401
    from chronos_b200 import tracegen

    spec = tracegen.WORKLOADS["attack_replay_10k"]
    rows = tracegen.generate_host(spec, n=spec["n"])
    data = rows.tobytes()
    out = ref_shim.run_reference(data)
    dump("config1_trace.json", {
        "generator": {k: v for k, v in spec.items()},
        "records_sha256": hashlib.sha256(data).hexdigest(),
        "reference": {
            "survivors": out["survivors"],
            "windows": [[pid, hist] for pid, hist in out["windows"]],
            "process_memory": {str(k): v for k, v in out["process_memory"].items()},
        },
    })
    print("golden fixtures written to", HERE)


if __name__ == "__main__":
    main()
