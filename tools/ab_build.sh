#!/bin/bash
# Build variants of the batch-affine MSM translation unit and link each into its own library (build_ab/lib<name>.so) for
# tools/ab.sh:   tools/ab_build.sh name1:"-DFLAG_A" name2:"-DFLAG_B -DFLAG_C" ...
# engine.o and host_bls.o come from the in-tree build (run make first).
set -e
cd "$(dirname "$0")/../kateth_b200/csrc"
mkdir -p ../../build_ab
for spec in "$@"; do
  name=${spec%%:*}; flags=${spec#*:}
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -Xptxas -v $flags -c msm_ba.cu -o ../../build_ab/msm_ba_$name.o 2> ../../build_ab/ptxas_$name.log \
    || { cat ../../build_ab/ptxas_$name.log; exit 1; }
  nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../../build_ab/lib$name.so engine.o ../../build_ab/msm_ba_$name.o host_bls.o group.o -lcudart -ldl
  echo "$name: $(grep -A2 'k_msm_baE' ../../build_ab/ptxas_$name.log | grep -o 'Used [0-9]* registers' | head -1), $(grep -A1 'Function properties for _ZN5kb2008k_msm_baE' ../../build_ab/ptxas_$name.log | tail -1 | sed 's/^ *//')"
done
