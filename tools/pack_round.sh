#!/bin/bash
P=$(ls -d project-chronos*/)
CHRONOS_LIB=$PWD/${P}libchronos_b200_checked.so timeout 900 python -m pytest tests/test_gpu_exchange.py -x -q 2>&1 | tail -3
timeout 900 python -m pytest tests/test_gpu_exchange.py -x -q 2>&1 | tail -3
python tools/profile_exchange.py 12500000 2 3 2>&1 | tail -2
python tools/profile_exchange.py 6000000 8 2 2>&1 | tail -3
