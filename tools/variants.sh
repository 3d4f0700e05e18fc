#!/bin/bash
# Development aid: build libhavac_cuda variants with different kernel geometry into tools/variants/.
#   tools/variants.sh name "-DSSV_ROW_CHUNK_V=1024" ...
set -e
cd "$(dirname "$0")/.."
mkdir -p tools/variants
name=$1; shift
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -shared -cudart static \
  -I include "$@" -o tools/variants/libhavac_cuda_$name.so havac_b200/csrc/*.cu
