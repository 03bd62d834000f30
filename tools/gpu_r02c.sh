#!/bin/bash
# r02c: baseline kernel timings of this session + ncu --set full of the L3 decoder (fold-1 2^30; the C5 model, M = 2^18, on 2^30 symbols)
tag=${1:-r02c}; out=gpurun_out/$tag; mkdir -p $out
timeout 600 python tools/kbench.py --workloads fold1,fold2,ans30 --paths 2 --reps 3 > $out/kbench.txt 2>&1; tail -4 $out/kbench.txt
timeout 300 python tools/kbench.py --workloads c5w --paths 2 --reps 3 --bins-scale 4 > $out/kbench_c5m.txt 2>&1; tail -2 $out/kbench_c5m.txt
timeout 600 ncu --set full --clock-control none --import-source on -k regex:lpb3_decode -s 3 -c 1 -o $out/dec_fold1 python tools/kbench.py --workloads fold1 --paths 2 --reps 1 > $out/ncu_fold1.log 2>&1; tail -2 $out/ncu_fold1.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:lpb3_decode -s 3 -c 1 -o $out/dec_c5m python tools/kbench.py --workloads c5w --paths 2 --reps 1 --bins-scale 4 > $out/ncu_c5m.log 2>&1; tail -2 $out/ncu_c5m.log
ls -la $out
