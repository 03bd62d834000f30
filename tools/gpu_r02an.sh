#!/bin/bash
tag=${1:-r02an}
out=gpurun_out/$tag
mkdir -p $out
for path in 0 1; do
timeout 200 python tools/pabench.py --path $path --remap --remap-method int --model-syms 8192 --reps 2 2>/dev/null | tail -1 | cut -c1-330
done
