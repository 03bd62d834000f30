#!/bin/bash
# ncu --set full of selected kernels on a kbench command. usage: gpu_ncu.sh <tag> <kernel-regex> <skip> <count> <kbench args...>
tag=$1; regex=$2; skip=$3; count=$4; shift 4
out=gpurun_out/$tag; mkdir -p $out
CMD="python tools/kbench.py $*"
$CMD > $out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"$regex" -s $skip -c $count -o $out/prof $CMD > $out/ncu_full.log 2>&1
echo "exit $?"; cat $out/plain.log; tail -3 $out/ncu_full.log
