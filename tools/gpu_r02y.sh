#!/bin/bash
# ncu --set full of the PA3 coders (lane per block, per-lane models)
tag=${1:-r02y}
out=gpurun_out/$tag
mkdir -p $out
timeout 300 python tools/pabench.py --model-syms 8192 --reps 1 > $out/pabench.json 2> $out/pabench.err || { echo "pabench failed"; tail -5 $out/pabench.err; exit 1; }
timeout 900 ncu --set full --import-source on --clock-control none -k regex:'lpb3_encode_kernel|lpb3_decode_kernel|pa_fit' -c 6 -o $out/pa3 -f python tools/pabench.py --model-syms 8192 --reps 1 > $out/ncu.log 2>&1
echo "ncu exit $?"
ls -la $out
