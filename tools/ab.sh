#!/bin/bash
# A/B of library builds on ONE box: tools/ab.sh <tag> <libA.so> <libB.so> ...   ("-" = the in-tree build); two rounds each.
TAG=$1; shift
O=gpurun_out/$TAG; mkdir -p $O
for round in 1 2; do
  for L in "$@"; do
    name=$(basename $L .so)
    if [ "$L" = "-" ]; then unset KB200_LIB; name=tree; else export KB200_LIB=$PWD/$L; fi
    timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > $O/${name}_$round.json 2> $O/${name}_$round.err
    python - "$O/${name}_$round.json" "$name" <<'PY'
import json, sys
d = json.loads(open(sys.argv[1]).read())
print(sys.argv[2], round(d["value"]), {k: round(v, 2) for k, v in d["stages_ms"].items()}, "msm_ms", round(d["roofline"]["launch_ms"], 2), "e2e", round(d["e2e"]["value"]), flush=True)
PY
  done
done
