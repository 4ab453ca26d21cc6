#!/bin/bash
# Library with every CH_ASSERT bounds check compiled in (csrc/common.cuh), next to the product one.
#   tools/checked_build.sh && CHRONOS_LIB=$PWD/<pkg>/libchronos_b200_checked.so python -m pytest tests -m gpu -q
set -e
cd "$(dirname "$0")/.."
P=$(ls -d project-chronos*/)
nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC -shared -I include \
     -DCHRONOS_CHECKED -o ${P}libchronos_b200_checked.so ${P}csrc/chronos.cu ${P}csrc/tracegen_tu.cu
echo ${P}libchronos_b200_checked.so
