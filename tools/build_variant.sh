#!/bin/bash
# builds a whole-library variant into build/variants/libb200nb_<name>.so with extra nvcc flags: tools/build_variant.sh <name> "<flags>"
set -e
cd "$(dirname "$0")/.."
name=$1; flags=$2
mkdir -p build/variants build/vobj/$name
FLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr"
for f in openmm_b200/csrc/*.cu; do
  b=$(basename $f .cu)
  nvcc $FLAGS $flags -c $f -o build/vobj/$name/$b.o &
done
wait
nvcc -shared -o build/variants/libb200nb_$name.so build/vobj/$name/*.o -gencode arch=compute_100a,code=sm_100a -ldl
echo built $name
