#!/bin/bash
# like tools/ab.sh but ONE round per library (many variants on one box): tools/ab1.sh <tag> <libA.so> ...   ("-" = the in-tree build)
TAG=$1; shift
O=gpurun_out/$TAG; mkdir -p $O
for L in "$@"; do
  name=$(basename $L .so)
  if [ "$L" = "-" ]; then unset KB200_LIB; name=tree; else export KB200_LIB=$PWD/$L; fi
  timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > $O/${name}.json 2> $O/${name}.err
  python tools/jl.py $O/${name}.json
done
