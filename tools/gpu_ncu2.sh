#!/bin/bash
# ncu --set full of one kernel for several workloads: gpu_ncu2.sh tag regex wl1 wl2 ...
tag=$1; regex=$2; shift 2
out=gpurun_out/$tag; mkdir -p $out
for wl in "$@"; do
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:"$regex" -s ${NCU_SKIP:-3} -c 1 -o $out/$wl python tools/kbench.py --workloads $wl --paths 2 --reps 1 > $out/ncu_$wl.log 2>&1
  tail -2 $out/ncu_$wl.log
done
