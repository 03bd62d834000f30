#!/bin/bash
# ncu --set full with source counters: pa_fit_kernel, pa_prelude_decode_kernel, pa_smem_decode_kernel
tag=${1:-r02ab}
out=gpurun_out/$tag
mkdir -p $out
timeout 300 python tools/pabench.py --model-syms 8192 --reps 1 > $out/pabench.json 2> $out/pabench.err || { echo "pabench failed"; tail -5 $out/pabench.err; exit 1; }
timeout 900 ncu --set full --import-source on --clock-control none -k regex:'pa_fit_kernel|pa_prelude_decode_kernel|pa_smem_decode_kernel' -c 3 -o $out/pa -f python tools/pabench.py --model-syms 8192 --reps 1 > $out/ncu.log 2>&1
echo "ncu exit $?"
ls -la $out
