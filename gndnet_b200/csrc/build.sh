#!/bin/bash
# Builds gndnet_b200/libgndnet_b200.so in-tree for sm_100a (the only target).
set -euo pipefail
cd "$(dirname "$0")"
OUT=../libgndnet_b200.so
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC,-fvisibility=hidden --fmad=true"
SRCS="common.cu voxelize.cu segment.cu pfn.cu segnet.cu conv_tc.cu conv_halo2.cu weights.cu abi.cu cloud_io.cu raster.cu"
mkdir -p build
pids=()
for s in $SRCS; do
  o=build/${s%.cu}.o
  if [ ! -f "$o" ] || [ "$s" -nt "$o" ] || [ -n "$(find . -maxdepth 1 -name "*.cuh" -newer "$o")" ] || [ ../../include/gndnet_b200.h -nt "$o" ]; then
    $NVCC $FLAGS ${EXTRA_NVCC_FLAGS:-} -c "$s" -o "$o" &
    pids+=($!)
  fi
done
for p in "${pids[@]:-}"; do [ -n "$p" ] && wait "$p"; done
$NVCC -gencode arch=compute_100a,code=sm_100a -shared -o "$OUT" build/*.o -lcudart_static -ldl -lrt -lpthread
echo "built $(realpath $OUT)"
