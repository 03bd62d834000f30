#!/bin/bash
# L3 iteration: lpb parity tests, kbench, one ncu --set full capture of the kernel named by $2
tag=$1; regex=$2; shift 2
out=gpurun_out/$tag; mkdir -p $out
timeout 1200 python -m pytest tests -m gpu -x -q ${PYTEST_ARGS} > $out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $out/pytest_gpu.log
tail -5 $out/pytest_gpu.log
timeout 900 python tools/kbench.py "$@" > $out/kbench.jsonl 2> $out/kbench.err; echo "kbench exit $?"
cat $out/kbench.jsonl; tail -5 $out/kbench.err
if [ -n "$regex" ]; then
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:"$regex" -s 3 -c 1 -o $out/prof python tools/kbench.py --workloads ${NCU_WL:-fold1} --paths 2 --reps 1 > $out/ncu_full.log 2>&1
  tail -2 $out/ncu_full.log
fi
