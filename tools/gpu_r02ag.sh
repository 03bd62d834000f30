#!/bin/bash
# per-block plain ANS behind the remap (run<ANSint>): parity, then every PA test, then stage times
tag=${1:-r02ag}
out=gpurun_out/$tag
mkdir -p $out
timeout 240 python -m pytest tests/test_gpu_pa_remap.py -m gpu -x -q -s > $out/pytest_remap.log 2>&1
echo "pytest remap exit $?"; tail -8 $out/pytest_remap.log
timeout 240 python -m pytest tests/test_gpu_pseudo_adaptive.py -m gpu -x -q > $out/pytest_pa.log 2>&1
echo "pytest pa exit $?"; tail -3 $out/pytest_pa.log
timeout 200 python tools/pabench.py --remap --remap-method int --model-syms 8192 --reps 2 > $out/pabench_int.json 2> $out/pabench_int.err || { echo "pabench int failed"; tail -5 $out/pabench_int.err; }
tail -1 $out/pabench_int.json
timeout 200 python tools/pabench.py --remap --remap-method msb --model-syms 8192 --reps 2 > $out/pabench_msb.json 2> $out/pabench_msb.err
tail -1 $out/pabench_msb.json
