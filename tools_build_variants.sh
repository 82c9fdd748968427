#!/bin/bash
# tuning builds of libnorib200.so with different K_tile_walk parameters -> optix-renderer_b200/variants/
set -e
cd /root/repo/optix-renderer_b200/csrc
rm -rf ../variants; mkdir -p ../variants
i=0
while read -r name flags; do
  [ -z "$name" ] && continue
  objs=""
  for f in nb_api nb_bvh nb_render nb_trace; do objs="$objs ../build/$f.o"; done
  nvcc -O3 -std=c++17 -lineinfo -Xcompiler -fPIC --expt-relaxed-constexpr -gencode arch=compute_100a,code=sm_100a -prec-div=false -prec-sqrt=false $flags -Xptxas -v -c nb_mega.cu -o /tmp/nb_mega_$name.o 2>&1 | grep -A2 "tile_walkILb1" | grep -E "registers|spill" | tr '\n' ' '
  echo " <- $name $flags"
  nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../variants/libnorib200_$name.so $objs /tmp/nb_mega_$name.o -cudart static
done <<LIST
base
b5 -DNB_WALK_MIN_BLOCKS=5
b6 -DNB_WALK_MIN_BLOCKS=6
leaf12 -DNB_WALK_T_LEAF=12
leaf16s6 -DNB_WALK_T_LEAF=16 -DNB_WALK_NODE_STEPS=6
steps8 -DNB_WALK_NODE_STEPS=8
shade16 -DNB_WALK_T_SHADE=16 -DNB_WALK_T_REGEN=12
b6leaf12 -DNB_WALK_MIN_BLOCKS=6 -DNB_WALK_T_LEAF=12 -DNB_WALK_T_SHADE=12
LIST
