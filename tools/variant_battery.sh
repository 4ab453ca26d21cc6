#!/bin/bash
# Run on the GPU box: step / stage times of every library under variants/ (tools/variant_build.sh),
# and the duration of selected kernels under ncu.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
OUT=gpurun_out/variant_battery.txt
: > $OUT
if [ -n "$PYTEST_LIB" ]; then CHRONOS_LIB=$PWD/variants/$PYTEST_LIB.so timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_robustness.py -x -q 2>&1 | tail -2 >> $OUT; fi
for v in variants/*.so; do
  CHRONOS_LIB=$PWD/$v timeout 300 python tools/variant_probe.py 100 >> $OUT 2>&1
  CHRONOS_LIB=$PWD/$v ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none -k regex:"${KERNELS:-group_forward}" -s 3 -c 3 --csv python tools/profile_step.py 10000000 3 2>&1 | grep "gpu__time" | cut -d, -f5,12- >> $OUT
done
cat $OUT
