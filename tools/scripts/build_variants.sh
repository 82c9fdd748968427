#!/bin/bash
# tuning builds of libnorib200.so with different K_tile_walk parameters -> optix-renderer_b200/variants/
# usage: bash tools/scripts/build_variants.sh [file with "name flags" lines]  (default: the min-blocks sweep)
set -e
cd /root/repo/optix-renderer_b200/csrc
mkdir -p ../variants
build_one() {
  name=$1; shift
  nvcc -O3 -std=c++17 -lineinfo -Xcompiler -fPIC --expt-relaxed-constexpr -gencode arch=compute_100a,code=sm_100a -prec-div=false -prec-sqrt=false "$@" -Xptxas -v -c nb_mega.cu -o /tmp/nb_mega_$name.o 2>&1 | grep -A2 "tile_walkILb1ELb0ELb1" | grep -E "registers|spill" | tr '\n' ' '
  echo " <- $name $*"
  objs=""
  for f in ../build/nb_*.o; do [ "$f" != "../build/nb_mega.o" ] && objs="$objs $f"; done
  nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../variants/libnorib200_$name.so $objs /tmp/nb_mega_$name.o -cudart static -ldl
}
if [ -n "$1" ]; then LIST=$(cat "$1"); else LIST=$'b4 -DNB_WALK_MIN_BLOCKS=4\nb5 -DNB_WALK_MIN_BLOCKS=5\nb6 -DNB_WALK_MIN_BLOCKS=6\nb8 -DNB_WALK_MIN_BLOCKS=8'; fi
while read -r name flags; do
  [ -z "$name" ] && continue
  build_one $name $flags &
done <<< "$LIST"
wait
