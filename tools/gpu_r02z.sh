#!/bin/bash
# PA3 encoder without the inlined divide, pa_fit with prefetched keys: parity, stage times, launch list
tag=${1:-r02z}
out=gpurun_out/$tag
mkdir -p $out
timeout 900 python -m pytest tests/test_gpu_pseudo_adaptive.py tests/test_gpu_pa_remap.py -m gpu -x -q > $out/pytest.log 2>&1
echo "pytest exit $?"; tail -3 $out/pytest.log
timeout 300 python tools/pabench.py --remap --model-syms 8192 --reps 3 > $out/pabench.json 2> $out/pabench.err || { echo "pabench failed"; tail -5 $out/pabench.err; }
cat $out/pabench.json
B200ANS_NO_PA3=1 timeout 300 python tools/pabench.py --model-syms 8192 --reps 3 2>/dev/null | tee $out/pabench_nopa3.json

timeout 300 python tools/pabench.py --model-syms 16384,65536 --reps 3 > $out/pabench_sizes.json 2> $out/pabench_sizes.err
cat $out/pabench_sizes.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $out/launches.csv python tools/pabench.py --model-syms 8192 --reps 1 > $out/ncu.log 2>&1
echo "ncu exit $?"
