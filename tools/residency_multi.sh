#!/bin/bash
# multi-GPU: is the sharded step steady at 18 / 17 filter CTAs per SM?  usage: residency_multi.sh N
N=${1:-4}
for rep in 1 2; do for k in 18 17; do
CHRONOS_FILTER_CTAS=$k python bench.py --gpus $N --steps 300 --warmup 5 --no-parity --no-skew > gpurun_out/s.json 2> gpurun_out/s.err
python - $k <<'PY'
import json, sys
d=[json.loads(l) for l in open("gpurun_out/s.json") if l.startswith("{")][-1]
print("ctas", sys.argv[1], "ms/step %.4f value %.2f G" % (d["ms_per_step"], d["value"]/1e9), "per-rank filter", [round(r["stages_ms"]["filter"],3) for r in d["per_rank"]], "step", [round(r["step_ms"],3) for r in d["per_rank"]])
PY
done; done
