#!/bin/bash
# multi-GPU visit: real-rank exchange test + bench.py --gpus N (with the parity step). usage: gpu_round_multi.sh N
N=${1:-2}
python -m pytest tests/test_gpu_multiprocess.py -x -q -s > gpurun_out/pytest_mp_n$N.log 2>&1; grep "MP_EXCHANGE\|passed\|failed\|skipped" gpurun_out/pytest_mp_n$N.log
python bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; tail -c 600 gpurun_out/bench_n$N.err
python - $N <<'PY'
import json, sys
N=sys.argv[1]
d=[json.loads(l) for l in open(f"gpurun_out/bench_n{N}.json") if l.startswith("{")][-1]
print("N=%s value %.2f G ev/s ms/step %.4f same-slice single %.4f e2e %.1f M ev/s" % (N, d["value"]/1e9, d["ms_per_step"], d["config"]["single_gpu_same_slice_ms"], d["e2e"]["value"]/1e6))
print("parity", d["parity"]["ok"], d["parity"]["checks"], "windows", d["parity"]["windows"], "sec", d["parity"]["seconds"])
for r in d["per_rank"]: print({k: round(v,4) for k,v in r["stages_ms"].items()}, round(r["step_ms"],4), r["received"])
print("skew", d["skew"])
PY
