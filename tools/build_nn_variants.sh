#!/bin/bash
# Builds variants of the library that differ only in the compile-time knobs of csrc/nn.cu, for tools/explore_nn_libs.py:
#   tools/build_nn_variants.sh name:"-DNN_F32X2=1" name2:"-DNN_MIN_BLOCKS=6" ...
# -> colmap-pcd_b200/build/var/libpcd_b200_<name>.so (git-ignored; travels to the GPU box with the snapshot).
set -e
cd "$(dirname "$0")/../colmap-pcd_b200"
make -s -j8 libpcd_b200.so
mkdir -p build/var
for v in "$@"; do
  name=${v%%:*}; defs=${v#*:}
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -fmad=true \
    -Xcompiler -fPIC,-ffp-contract=off,-fvisibility=hidden,-Wall,-Wno-unused-function $defs -Xptxas -v -c csrc/nn.cu -o build/var/nn_$name.o \
    2> build/var/nn_$name.ptxas.txt
  objs=$(ls build/*.o | grep -v '/nn.o$')
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o build/var/libpcd_b200_$name.so $objs build/var/nn_$name.o -cudart static -Xcompiler -pthread
  echo "$name ($defs): $(grep -A1 'k_nn_searchILi1' build/var/nn_$name.ptxas.txt | grep -o '[0-9]* bytes spill stores' | head -1), $(grep -A2 'k_nn_searchILi1' build/var/nn_$name.ptxas.txt | grep -o 'Used [0-9]* registers' | head -1)"
done
