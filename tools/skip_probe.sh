#!/bin/bash
python tools/long_probe.py 2000 2>&1 | tail -1 | cut -c1-300
for g in 1 2 4; do for gw in 0 2; do echo -n "gather_ctas=$g when=$gw "; CHRONOS_GATHER_CTAS=$g CHRONOS_GATHER_WHEN=$gw python tools/filter_probe.py 100 0 | tail -1; done; done
