#!/bin/bash
# variants: lib (wi8/wi4) x gather_first x ctas
python -m pytest tests -m gpu -x -q > gpurun_out/t1.log 2>&1; tail -4 gpurun_out/t1.log
P=$(ls -d project-chronos*/)
for lib in wi8 wi4; do
 for gf in 1 0; do
  for g in 1 2; do
   for low in 0 1; do
   L=""; [ $lib = wi4 ] && L="$PWD/${P}libchronos_b200_wi4.so"
   extra=""; [ $low = 1 ] && extra="CHRONOS_GATHER_LOW=1"
   env CHRONOS_LIB=$L CHRONOS_GATHER_FIRST=$gf CHRONOS_GATHER_CTAS=$g $extra python bench.py --steps 20 --warmup 5 --no-cpu --no-stream > gpurun_out/s.json 2> gpurun_out/s.err
   python - <<PY
import json
try:
    d=json.load(open("gpurun_out/s.json")); p=d["roofline"]["pipeline"]
    print("$lib first=$gf ctas=$g low=$low", "ms/step %.4f lib %.4f min %.4f filt %.4f"%(d["ms_per_step"], p["library_event_ms"], p["library_event_ms_min"], d["roofline"]["kernel_ms"]))
except Exception as e: print("$lib $gf $g ERR", e, open("gpurun_out/s.err").read()[-300:])
PY
   done
  done
 done
done
