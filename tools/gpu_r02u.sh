#!/bin/bash
# ncu --set full of the two serial interpolative decoders (pseudo-adaptive preludes, remap alphabets) and the PA coders
tag=${1:-r02u}
out=gpurun_out/$tag
mkdir -p $out
timeout 300 python tools/pabench.py --remap --model-syms 8192 --reps 1 > $out/pabench.json 2> $out/pabench.err || { echo "pabench failed"; tail -5 $out/pabench.err; exit 1; }
tail -2 $out/pabench.json
timeout 900 ncu --set full --import-source on --clock-control none -k regex:'remap_alpha_decode_kernel|pa_prelude_decode_kernel|pa_prelude_encode_kernel|pa_sort_kernel' -c 8 -o $out/pa_serial -f python tools/pabench.py --remap --model-syms 8192 --reps 1 > $out/ncu.log 2>&1
echo "ncu exit $?"
ls -la $out
