#!/bin/bash
python bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/bench_n8.json 2> gpurun_out/bench_n8.err; tail -c 500 gpurun_out/bench_n8.err
python - <<'PY'
import json
d=[json.loads(l) for l in open("gpurun_out/bench_n8.json") if l.startswith("{")][-1]
print("N=8 value %.2f G ev/s ms/step %.4f same-slice single %.4f e2e %.1f M ev/s (%.1f ms)" % (d["value"]/1e9, d["ms_per_step"], d["config"]["single_gpu_same_slice_ms"], d["e2e"]["value"]/1e6, d["e2e"]["ms_per_step"]))
print("parity", d["parity"]["ok"], "windows", d["parity"]["windows"], "sec", d["parity"]["seconds"], d["parity"]["checks"])
for r in d["per_rank"]: print({k: round(v,4) for k,v in r["stages_ms"].items()}, round(r["step_ms"],4), r["received"])
print("skew", d["skew"]); print("e2e", {k: v for k, v in d["e2e"].items() if k in ("plain_copy", "host_trim", "upload")})
PY
