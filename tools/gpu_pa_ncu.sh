#!/bin/bash
tag=${1:-pancu}; out=gpurun_out/$tag; mkdir -p $out
timeout 900 ncu --set full --clock-control none --import-source on -k 'regex:pa_smem' -s 2 -c 2 -o $out/full_pa \
  python tools/pabench.py --model-syms 8192 --reps 1 > $out/ncu_full_pa.log 2>&1; echo "ncu exit $?"; tail -3 $out/ncu_full_pa.log
