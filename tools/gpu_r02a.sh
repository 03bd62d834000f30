#!/bin/bash
# r02a: exploratory. Host facts, the 22-warp encoder shape (parity + timing), C3 / C3r / 2^32-symbol C5 on one GPU.
tag=${1:-r02a}; out=gpurun_out/$tag; mkdir -p $out
{ nproc; free -g; lscpu | head -20; nvidia-smi; ulimit -l; } > $out/host.txt 2>&1
for wl in c3 c3r; do
  timeout 900 python tools/kbench.py --workloads $wl --paths 0 --reps 2 > $out/kbench_$wl.txt 2>&1; echo "$wl exit $?"; tail -1 $out/kbench_$wl.txt | cut -c1-400
done
timeout 900 python tools/kbench.py --workloads c5 --n 4294967296 --paths 0 --reps 2 > $out/kbench_c5_2p32.txt 2>&1; echo "c5 2^32 exit $?"; tail -1 $out/kbench_c5_2p32.txt | cut -c1-400
