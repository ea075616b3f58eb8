#!/bin/bash
# ncu --set full over the kernels of the verification stage (tools/verify_probe.py); raw + source pages exported on the box,
# the .ncu-rep itself is dropped (gpurun_out/ is capped at 64 MiB).   usage: tools/ncu_verify.sh <tag> [n]
TAG=${1:-v}; N=${2:-4096}
O=gpurun_out/$TAG; mkdir -p $O
timeout 200 python tools/verify_probe.py $N > $O/plain.log 2>&1 || { tail -5 $O/plain.log; exit 1; }
tail -2 $O/plain.log
KB200_TRACE=1 timeout 200 python tools/verify_probe.py $N > $O/trace.log 2>&1
timeout 600 ncu --set full --clock-control none -k regex:"${KREGEX:-k_rlc|k_g1_decompress|k_eval_finish|k_eval_prod|k_challenge_w2}" -c ${COUNT:-10} -o /tmp/verify_$TAG \
  python tools/verify_probe.py $N > $O/ncu.log 2>&1; echo "ncu rc=$?"
ncu -i /tmp/verify_$TAG.ncu-rep --page raw --csv > $O/raw.csv 2>/dev/null
for k in ${KERNELS:-k_rlc k_g1_decompress k_eval_finish k_eval_prod k_challenge_w2 k_rlc_bucket}; do
  ncu -i /tmp/verify_$TAG.ncu-rep --page source --csv -k regex:"^$k\$" -c 1 2>/dev/null | gzip > $O/src_$k.csv.gz
done
ls -la $O
