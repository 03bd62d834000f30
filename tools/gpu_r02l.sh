#!/bin/bash
# r02l: tasks dealt round-robin over all SMs: C2 / C1 / 2^30 timings + parity of the lane-per-block paths
tag=${1:-r02l}; out=gpurun_out/$tag; mkdir -p $out
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -k "lpb" > $out/pytest_parity.log 2>&1; echo "parity exit $?"; tail -2 $out/pytest_parity.log
timeout 600 python tools/kbench.py --workloads c2,c1 --paths 0 --reps 5 --block-syms 4096 > $out/kbench_small.txt 2>&1; cut -c1-420 $out/kbench_small.txt | tail -2
timeout 600 python tools/kbench.py --workloads fold1,ans30 --paths 2 --reps 3 > $out/kbench_big.txt 2>&1; cut -c1-420 $out/kbench_big.txt | tail -2
