#!/bin/bash
# builds tile-kernel variants of libb200nb.so into build/variants/ (for GPU parameter sweeps): tools/build_variants.sh "W B U" ...
set -e
cd "$(dirname "$0")/.."
mkdir -p build/variants build/vobj
FLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr"
for v in "$@"; do
  set -- $v; W=$1; B=$2; U=$3
  name="w${W}_b${B}_u${U}"
  nvcc $FLAGS -DTILE_WARPS=$W -DTILE_MIN_BLOCKS=$B -DTILE_UNROLL=$U -c openmm_b200/csrc/b200_tile.cu -o build/vobj/tile_$name.o
  objs=$(ls build/obj/*.o | grep -v b200_tile.o)
  nvcc -shared -o build/variants/libb200nb_$name.so $objs build/vobj/tile_$name.o -gencode arch=compute_100a,code=sm_100a -ldl
  echo built $name
done
