#!/bin/bash
# launch list of the remapped pseudo-adaptive coder with per-block plain ANS
tag=${1:-r02ak}
out=gpurun_out/$tag
mkdir -p $out
timeout 200 python tools/pabench.py --remap --remap-method int --model-syms 8192 --reps 1 > $out/pabench_int.json 2> $out/pabench_int.err || { echo failed; tail -3 $out/pabench_int.err; exit 1; }
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file $out/launches.csv python tools/pabench.py --remap --remap-method int --model-syms 8192 --reps 1 > $out/ncu.log 2>&1
echo "ncu exit $?"
