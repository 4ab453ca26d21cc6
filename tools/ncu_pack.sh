#!/bin/bash
# ncu --set full with source counters on xchg_pack (2 emulated ranks on one GPU; writes are local)
python tools/profile_exchange.py 12500000 2 2 > gpurun_out/profile_exchange_plain.log 2>&1 || { tail -5 gpurun_out/profile_exchange_plain.log; exit 1; }
tail -2 gpurun_out/profile_exchange_plain.log
ncu --set full --import-source on --clock-control none --cache-control none -k regex:xchg_pack -s 2 -c 1 -o gpurun_out/r02_pack -f python tools/profile_exchange.py 12500000 2 2 > gpurun_out/ncu_pack.log 2>&1
ls -la gpurun_out/r02_pack.ncu-rep
