#!/bin/bash
# PA prelude coders / remap alphabet decoder rewritten: parity tests, then the stage times
tag=${1:-r02v}
out=gpurun_out/$tag
mkdir -p $out
timeout 900 python -m pytest tests/test_gpu_pseudo_adaptive.py tests/test_gpu_pa_remap.py -m gpu -x -q > $out/pytest.log 2>&1
echo "pytest exit $?"; tail -5 $out/pytest.log
timeout 300 python tools/pabench.py --remap --model-syms 8192 --reps 3 > $out/pabench.json 2> $out/pabench.err || { echo "pabench failed"; tail -5 $out/pabench.err; }
cat $out/pabench.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $out/launches.csv python tools/pabench.py --remap --model-syms 8192 --reps 1 > $out/ncu.log 2>&1
echo "ncu exit $?"
