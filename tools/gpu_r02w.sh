#!/bin/bash
# (1) ceiling for a lane-per-block PA coder: the semi-static L3 coder on C4's geometry (2^27 symbols, 8192-symbol blocks)
# (2) counting sort in the PA normaliser: parity + stage times
tag=${1:-r02w}
out=gpurun_out/$tag
mkdir -p $out
timeout 300 python tools/kbench.py --n 134217728 --block-syms 8192 --workloads fold1 --paths 2 --reps 5 > $out/kbench_c4geom.txt 2>&1
tail -3 $out/kbench_c4geom.txt
timeout 900 python -m pytest tests/test_gpu_pseudo_adaptive.py tests/test_gpu_pa_remap.py -m gpu -x -q > $out/pytest.log 2>&1
echo "pytest exit $?"; tail -5 $out/pytest.log
B200ANS_PA_BITONIC=1 B200ANS_NO_PA3=1 timeout 600 python -m pytest tests/test_gpu_pseudo_adaptive.py -m gpu -x -q > $out/pytest_old.log 2>&1
echo "pytest (bitonic sort, 4-lane coders) exit $?"; tail -2 $out/pytest_old.log
B200ANS_NO_PA3=1 timeout 600 python -m pytest tests/test_gpu_pseudo_adaptive.py -m gpu -x -q > $out/pytest_nopa3.log 2>&1
echo "pytest (counting sort, 4-lane coders) exit $?"; tail -2 $out/pytest_nopa3.log
B200ANS_NO_PA3=1 timeout 300 python tools/pabench.py --model-syms 8192 --reps 3 > $out/pabench_nopa3.json 2> $out/pabench_nopa3.err
cat $out/pabench_nopa3.json
timeout 300 python tools/pabench.py --remap --model-syms 8192 --reps 3 > $out/pabench.json 2> $out/pabench.err || { echo "pabench failed"; tail -5 $out/pabench.err; }
cat $out/pabench.json
timeout 300 python tools/pabench.py --model-syms 8192,16384,65536 --reps 3 > $out/pabench_sizes.json 2> $out/pabench_sizes.err
cat $out/pabench_sizes.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $out/launches.csv python tools/pabench.py --model-syms 8192 --reps 1 > $out/ncu.log 2>&1
echo "ncu exit $?"
