#!/bin/bash
# One multi-GPU bench run under torchrun: tools/scale_run.sh <N> <tag> [extra bench.py args]; output gpurun_out/<tag>/bench_n<N>[...].json
N=$1; TAG=$2; shift 2
O=gpurun_out/$TAG; mkdir -p $O
SUF=$(echo "$*" | tr -c 'a-zA-Z0-9\n' '_' | sed 's/_*$//')
if [ "$N" = "1" ]; then
  timeout 900 python bench.py --gpus 1 --steps 10 --warmup 3 --no-cpu-baseline "$@" > $O/bench_n1$SUF.json 2> $O/bench_n1$SUF.err
else
  NCCL_DEBUG=WARN timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29571 bench.py --gpus $N --steps 10 --warmup 3 "$@" > $O/bench_n$N$SUF.json 2> $O/bench_n$N$SUF.err
fi
echo "N=$N $* rc=$?"
python - "$O/bench_n$N$SUF.json" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read())
    print("  value", round(d["value"]), "ms/step", round(d["ms_per_step"], 2), "e2e", round(d["e2e"]["value"]), "stages", {k: round(v, 1) for k, v in d["stages_ms"].items()}, "msm_ms", round(d["roofline"]["launch_ms"], 2))
except Exception as e:
    print("  no line:", e)
PY
