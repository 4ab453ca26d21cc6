#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
OUT=gpurun_out/variant_battery.txt
: > $OUT
P=$(ls -d project-chronos*/)
CHRONOS_LIB=$PWD/${P}libchronos_b200_checked.so timeout 900 python -m pytest tests/test_gpu_exchange.py -x -q 2>&1 | tail -2 >> $OUT
CHRONOS_LIB=$PWD/variants/new.so timeout 900 python -m pytest tests/test_gpu_exchange.py -x -q 2>&1 | tail -2 >> $OUT
for v in base new; do
  CHRONOS_LIB=$PWD/variants/$v.so python tools/profile_exchange.py 12500000 2 3 2>&1 | tail -1 >> $OUT
  CHRONOS_LIB=$PWD/variants/$v.so python tools/profile_exchange.py 6000000 8 2 2>&1 | tail -1 >> $OUT
done
cat $OUT
