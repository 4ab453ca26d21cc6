#!/bin/bash
# ncu evidence for profiles/ (run AFTER the same command exited 0 without ncu)
python tools/profile_step.py 10000000 3 > gpurun_out/profile_step_plain.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/r02_launches_n1.csv python tools/profile_step.py 10000000 3 > gpurun_out/ncu_a.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none -c 60 --csv --log-file gpurun_out/r02_launches_n1_warm.csv python tools/profile_step.py 10000000 3 > gpurun_out/ncu_b.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:filter_kernel -s 1 -c 1 -o gpurun_out/r02_filter -f python tools/profile_step.py 10000000 3 > gpurun_out/ncu_c.log 2>&1
ncu -i gpurun_out/r02_filter.ncu-rep --page raw --csv > gpurun_out/r02_filter_ncu_raw.csv
ncu -i gpurun_out/r02_filter.ncu-rep --page details --csv > gpurun_out/r02_filter_ncu_details.csv
ncu --set full --clock-control none --cache-control none -k regex:"compact_scan|sort_prepare|onesweep_pass|gather_records|group_forward|window_reverse|emit_ordered" -s 8 -c 9 -o gpurun_out/r02_tail -f python tools/profile_step.py 10000000 3 > gpurun_out/ncu_d.log 2>&1
ncu -i gpurun_out/r02_tail.ncu-rep --page raw --csv > gpurun_out/r02_tail_raw.csv
tail -3 gpurun_out/profile_step_plain.log; ls -la gpurun_out/r02_*
# the exchange pack (two emulated ranks on one GPU: writes are local)
python tools/profile_exchange.py 12500000 2 2 > gpurun_out/profile_exchange_plain.log 2>&1 && \
ncu --set full --import-source on --clock-control none --cache-control none -k regex:"xchg_pack|xchg_hist|xchg_unpack" -s 3 -c 3 -o gpurun_out/r02_exchange -f python tools/profile_exchange.py 12500000 2 2 > gpurun_out/ncu_e.log 2>&1
ncu -i gpurun_out/r02_exchange.ncu-rep --page raw --csv > gpurun_out/r02_exchange_raw.csv
