#!/bin/bash
# like tools/ab.sh, with tools/e2e_probe.py (per-call host-buffer times): tools/ab_e2e.sh <tag> <lib...>   ("-" = in-tree build)
TAG=$1; shift
O=gpurun_out/$TAG; mkdir -p $O
for round in 1 2; do
  for L in "$@"; do
    name=$(basename $L .so)
    if [ "$L" = "-" ]; then unset KB200_LIB; name=tree; else export KB200_LIB=$PWD/$L; fi
    timeout 200 python tools/e2e_probe.py 2> $O/${name}_$round.err | python -c "
import json, sys
d = json.loads(sys.stdin.read())
print('$name', {k: round(v, 2) for k, v in d.items() if k.endswith('_ms')}, flush=True)"
  done
done
