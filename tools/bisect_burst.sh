#!/bin/bash
for rep in 1 2; do
for d in _r01 _b_067860c _b_36d0cea _b_1624008 .; do
  (cd $d && python tools/filter_probe.py 60 0 2>&1 | tail -1)
done
done
(cd . && python tools/filter_probe.py 60 1 2>&1 | tail -1)
(cd _b_067860c && python tools/filter_probe.py 60 1 2>&1 | tail -1)
