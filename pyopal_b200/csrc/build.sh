#!/bin/bash
# Build libpyopal_cuda.so in-tree for sm_100a (called by __graft_entry__.build()).
set -e
cd "$(dirname "$0")"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS="-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -I. -I../../include"
mkdir -p build
pids=()
for f in kernels_w16_l1 kernels_w16_l0 kernels_w16_l1_end kernels_w16_l0_end kernels_dual kernels_stream_l1 kernels_stream_l0 kernels_w32_l1 kernels_w32_l0 giant tgiant topk api pairdp multi microbench; do
  if [ ! -f build/$f.o ] || [ $f.cu -nt build/$f.o ] || [ kernels.cuh -nt build/$f.o ] || [ stream.cuh -nt build/$f.o ] || [ dual.cuh -nt build/$f.o ] || [ store.h -nt build/$f.o ] || [ ../../include/pyopal_cuda.h -nt build/$f.o ]; then
    $NVCC $FLAGS -c $f.cu -o build/$f.o &
    pids+=($!)
  fi
done
for p in "${pids[@]}"; do wait $p; done
$NVCC -gencode arch=compute_100a,code=sm_100a -shared -o ../libpyopal_cuda.so build/*.o -lcudart_static -lpthread -ldl -lrt
echo built $(pwd)/../libpyopal_cuda.so
bash "$(pwd)/build_host.sh"
