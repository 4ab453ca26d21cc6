#!/bin/bash
P=$(ls -d project-chronos*/)
for v in small nolineinfo; do
  echo -n "$v: "; CHRONOS_LIB=$PWD/${P}libvar_$v.so python tools/filter_probe.py 30 0 2>&1 | tail -1
done
echo -n "default: "; python tools/filter_probe.py 30 0 2>&1 | tail -1
