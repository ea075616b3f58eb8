#!/bin/bash
# One GPU-box session: parity tests, bench, ncu launch list and one full capture of the dominant kernel.
# usage: tools/gpu_round.sh <tag> [stages...]   stages: tests bench ncu_list ncu_full   (default: all)
set -u
TAG=${1:-run}; shift || true
STAGES=${*:-"tests bench ncu_list ncu_full"}
O=gpurun_out/$TAG; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv,noheader > $O/gpu.txt 2>&1
for st in $STAGES; do
  case $st in
    tests)   timeout 1500 python -m pytest tests -x -q -m gpu > $O/pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc.txt; tail -3 $O/pytest.log;;
    bench)   timeout 900 python bench.py --steps 5 --warmup 3 > $O/bench.json 2> $O/bench.err; echo "bench rc=$?" | tee -a $O/rc.txt; cat $O/bench.json;;
    ncu_list) timeout 900 python bench.py --steps 1 --warmup 1 --blobs 1184 --no-cpu-baseline > $O/plain_small.log 2>&1 && \
             timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/launches.csv \
               python bench.py --steps 1 --warmup 1 --blobs 1184 --no-cpu-baseline > $O/ncu_list.log 2>&1; echo "ncu_list rc=$?" | tee -a $O/rc.txt;;
    ncu_full) timeout 900 python bench.py --steps 1 --warmup 1 --blobs 1184 --no-cpu-baseline > $O/plain_small2.log 2>&1 && \
             timeout 1200 ncu --set full --clock-control none --import-source on -k regex:${KREGEX:-k_msm} -s 2 -c 1 -o $O/k_msm \
               python bench.py --steps 1 --warmup 1 --blobs 1184 --no-cpu-baseline > $O/ncu_full.log 2>&1; echo "ncu_full rc=$?" | tee -a $O/rc.txt;;
  esac
done
