#!/bin/bash
# sweep of L3 launch shapes: "cfg:rep:warps" in $SWEEP, workloads in $WL
out=gpurun_out/$1; mkdir -p $out
for cfg in $SWEEP; do
  IFS=: read k r w <<< "$cfg"
  echo "== cfg=$k rep<=$r warps<=$w" | tee -a $out/sweep.log
  B200ANS_L3_CFG=$k B200ANS_L3_REP=$r B200ANS_L3_WARPS=$w timeout 600 python tools/kbench.py --workloads ${WL:-fold1} --paths 2 --reps 3 2>&1 | cut -c1-200 | tee -a $out/sweep.log
done
