#!/bin/bash
# Builds experimental copies of the C-ABI library with different step-kernel tunables:
#   [VARIANT_FILES="kernels_neighbor"] [VARIANT_GREP=row_sweep2] scripts/build_variants.sh name "-DLG_TB=128 -DLG_TILE=512 ..." [name2 "flags2" ...]
# -> libgran_b200/variants/liblgb200_<name>.so, selected at run time with LGB200_LIB=<path>.
set -e
cd "$(dirname "$0")/../libgran_b200/csrc"
mkdir -p ../variants
while [ $# -ge 2 ]; do
  name=$1; flags=$2; shift 2
  d=build/var_$name; mkdir -p $d
  for f in api kernels_force kernels_force_tiled kernels_rows kernels_integrate kernels_neighbor kernels_halo kernels_bonds generators; do
    if [[ " ${VARIANT_FILES:-kernels_force_tiled kernels_rows} " == *" $f "* ]] || [ ! -f build/$f.o ]; then
      /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -ccbin /usr/bin/g++ -std=c++17 -O3 -lineinfo --fmad=false \
        -Xcompiler -fPIC,-ffp-contract=off -Xptxas -v $flags -c $f.cu -o $d/$f.o 2> $d/$f.ptxas.log &
    else cp build/$f.o $d/$f.o; fi
  done; wait
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -ccbin /usr/bin/g++ -shared -o ../variants/liblgb200_$name.so $d/*.o
  grep -A2 "${VARIANT_GREP:-step_rows2_kernelILi6ELi1ELb0}" $d/*.ptxas.log | grep -E "spill|Used" | tr '\n' ' '; echo " <- $name"
done
