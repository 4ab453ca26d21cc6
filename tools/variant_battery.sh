#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
OUT=gpurun_out/variant_battery.txt
: > $OUT
for v in plain_nohint; do
  for k in "CHRONOS_FILTER_CTAS=19" "CHRONOS_FILTER_CTAS=18"; do
    echo "== $v $k" >> $OUT
    env $k CHRONOS_LIB=$PWD/variants/$v.so timeout 600 python tools/long_probe.py 2000 2>&1 | cut -c1-600 >> $OUT
    env $k SWEEP=read CHRONOS_LIB=$PWD/variants/$v.so timeout 600 python tools/long_probe.py 300 2>&1 | cut -c1-600 >> $OUT
    env $k SWEEP=write CHRONOS_LIB=$PWD/variants/$v.so timeout 600 python tools/long_probe.py 300 2>&1 | cut -c1-600 >> $OUT
  done
done
echo "== trig_nohint 19" >> $OUT
CHRONOS_FILTER_CTAS=19 CHRONOS_LIB=$PWD/variants/trig_nohint.so timeout 600 python tools/long_probe.py 1000 2>&1 | cut -c1-600 >> $OUT
echo "== sharded slice, plain_nohint 19 / plain 18" >> $OUT
CHRONOS_FILTER_CTAS=19 CHRONOS_LIB=$PWD/variants/plain_nohint.so timeout 300 python tools/variant_probe.py 200 12500000 sharded_100m_1m >> $OUT 2>&1
CHRONOS_LIB=$PWD/variants/plain.so timeout 300 python tools/variant_probe.py 200 12500000 sharded_100m_1m >> $OUT 2>&1
cat $OUT
