#!/bin/bash
# Are there slow bursts of the filter kernel? per-step filter times of a few bench runs.
show() { python - "$1" <<'PY'
import json, sys
d=[json.loads(l) for l in open("gpurun_out/s.json") if l.startswith("{")][-1]; p=d["roofline"]["pipeline"]
f=p.get("filter_ms_per_step")
slow = sum(1 for x in f if x > 0.6) if f else None
print(sys.argv[1], "ms/step %.4f lib mean %.4f median %.4f min %.4f filter mean %.4f slow steps %s/%s" % (d["ms_per_step"], p["library_event_ms"], p["library_event_ms_median"], p["library_event_ms_min"], d["roofline"]["kernel_ms"], slow, len(f) if f else None))
PY
}
for i in 1 2 3; do
  (cd _r01 && python bench.py --steps 40 --warmup 5 --no-cpu --no-stream > ../gpurun_out/s.json 2>../gpurun_out/s.err); show "r01      "
  python bench.py --steps 40 --warmup 5 --no-cpu --no-stream > gpurun_out/s.json 2>gpurun_out/s.err; show "new      "
  CHRONOS_NO_CARVEOUT=1 python bench.py --steps 40 --warmup 5 --no-cpu --no-stream > gpurun_out/s.json 2>gpurun_out/s.err; show "new nocrv"
done
