#!/bin/bash
show() { python - "$1" <<'PY'
import json, sys
d=[json.loads(l) for l in open("gpurun_out/s.json") if l.startswith("{")][-1]; p=d["roofline"]["pipeline"]
print(sys.argv[1], "ms/step %.4f lib mean %.4f median %.4f min %.4f filter mean %.4f" % (d["ms_per_step"], p["library_event_ms"], p["library_event_ms_median"], p["library_event_ms_min"], d["roofline"]["kernel_ms"]))
PY
}
for i in 1 2 3; do for k in 18 17; do
  CHRONOS_FILTER_CTAS=$k python bench.py --steps 1500 --warmup 5 --no-cpu --no-stream > gpurun_out/s.json 2>gpurun_out/s.err; show "ctas=$k"
done; done
