"""One core, same sample: the UNMODIFIED reference script (oracle/ref_shim.py: runpy of
/root/reference/chronos_sensor.py under a fake bcc/requests, stdout swallowed, LLM answering at
once) against the port bench.py's CPU arm times (oracle/_ref C filter + oracle.pyoracle.PySensor).
Build container only (/root/reference does not exist on the GPU box). Prints one JSON line."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from chronos_b200 import tracegen  # noqa: E402
from oracle import pyoracle, ref_shim  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 300_000
rows = tracegen.generate_host_nocuda(tracegen.WORKLOADS["uniform_10m_4k"], first=0, n=n)
buf = rows.tobytes()
t0 = time.perf_counter()
out = ref_shim.run_reference(buf)
t_ref = time.perf_counter() - t0
t0 = time.perf_counter()
drop = ref_shim.ref_filter_drop(buf)
s = pyoracle.PySensor()
s.feed(buf, 0, drop=bytes(drop))
t_port = time.perf_counter() - t0
print(json.dumps({"events": n, "unmodified_reference_events_per_s": n / t_ref, "port_events_per_s": n / t_port,
                  "unmodified_seconds": t_ref, "port_seconds": t_port,
                  "windows_reference": len(out["windows"]) if isinstance(out, dict) and "windows" in out else None,
                  "windows_port": len(s.windows), "cpu": os.uname().machine, "cores_used": 1}))
