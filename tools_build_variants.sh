#!/bin/bash
# builds tuning variants of libnorib200.so into gpurun_out-independent dir variants/
set -e
cd /root/repo/optix-renderer_b200/csrc
mkdir -p ../variants
for cfg in "128 4" "128 5" "128 6" "256 2" "256 3" "64 8" "64 10"; do
  set -- $cfg
  out=../variants/libnorib200_t$1_b$2.so
  objs=""
  for f in nb_api nb_bvh nb_render nb_trace; do objs="$objs ../build/$f.o"; done
  nvcc -O3 -std=c++17 -lineinfo -Xcompiler -fPIC --expt-relaxed-constexpr -gencode arch=compute_100a,code=sm_100a -DNB_TILE_THREADS=$1 -DNB_TILE_MIN_BLOCKS=$2 -Xptxas -v -c nb_mega.cu -o /tmp/nb_mega_$1_$2.o 2>&1 | grep -A1 "ILi0EE" | grep -E "registers|spill" | tr '\n' ' '
  echo " <- $cfg"
  nvcc -gencode arch=compute_100a,code=sm_100a -shared -o $out $objs /tmp/nb_mega_$1_$2.o -cudart static
done
