#!/bin/bash
# Build the library with extra -D flags into variants/<name>.so (experiments; variants/ is git-ignored).
#   tools/variant_build.sh name "-DCHRONOS_RADIX_TILE=4096 ..."
set -e
cd "$(dirname "$0")/.."
P=$(ls -d project-chronos*/)
mkdir -p variants
nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -Xcompiler -fPIC -shared -I include \
     $2 -o variants/$1.so ${P}csrc/chronos.cu ${P}csrc/tracegen_tu.cu
echo variants/$1.so
