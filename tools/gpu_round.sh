#!/bin/bash
# One GPU visit: filter probe, the -m gpu suite against the product library and against the
# checked build (CH_ASSERT bounds checks), then the N=1 bench. Logs under gpurun_out/.
P=$(ls -d project-chronos*/)
python tools/filter_probe.py 40 0 2>&1 | tail -1
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -3 gpurun_out/pytest_gpu.log
CHRONOS_LIB=$PWD/${P}libchronos_b200_checked.so python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_checked.log 2>&1; tail -3 gpurun_out/pytest_gpu_checked.log
python bench.py --steps 20 --warmup 5 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; tail -c 400 gpurun_out/bench_n1.err
python - <<'PY'
import json
d=[json.loads(l) for l in open("gpurun_out/bench_n1.json") if l.startswith("{")][-1]
p=d["roofline"]["pipeline"]
print("value %.3f G ev/s  ms/step %.4f  lib mean %.4f median %.4f  filter %.4f frac %.3f  pipeline frac %.3f  e2e %.1f M ev/s" % (d["value"]/1e9, d["ms_per_step"], p["library_event_ms"], p["library_event_ms_median"], d["roofline"]["kernel_ms"], d["roofline"]["frac"], p["frac"], d["e2e"]["value"]/1e6))
print("stages", d["stages_ms_rank0"]); print("pipelined", d["pipelined"])
s=d["streaming"]; print("stream p50", s["device_resident"]["p50_ms"], "host", s["host_buffers"]["p50_ms"], "2inflight", s["device_resident_two_in_flight"]["ms_per_batch"], "1M pids", s["stream_64k_1m_pids"]["first_100"]["p50_ms"], s["stream_64k_1m_pids"]["last_100"]["p50_ms"])
print("cpu", d["cpu_baseline"]["value"], d["config"]["warmup_extra_steps"])
PY
