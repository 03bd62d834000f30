#!/bin/bash
# PA3 coders (lane per block, per-lane models) after the load-ordering fix: parity, stage times, launch list
tag=${1:-r02x}
out=gpurun_out/$tag
mkdir -p $out
timeout 900 python -m pytest tests/test_gpu_pseudo_adaptive.py tests/test_gpu_pa_remap.py -m gpu -x -q > $out/pytest.log 2>&1
echo "pytest exit $?"; tail -3 $out/pytest.log
timeout 300 python tools/pabench.py --remap --model-syms 8192 --reps 3 > $out/pabench.json 2> $out/pabench.err || { echo "pabench failed"; tail -5 $out/pabench.err; }
cat $out/pabench.json
for blog in 8 7 6; do
B200ANS_PA3_BLOG=$blog timeout 300 python tools/pabench.py --model-syms 8192 --reps 3 2>/dev/null | tee $out/pabench_blog$blog.json
done
timeout 300 python tools/pabench.py --model-syms 16384,65536 --reps 3 > $out/pabench_sizes.json 2> $out/pabench_sizes.err
cat $out/pabench_sizes.json
timeout 300 python tools/kbench.py --workloads fold1,ans30,c2 --paths 2 --reps 5 > $out/kbench.txt 2>&1
tail -4 $out/kbench.txt
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $out/launches.csv python tools/pabench.py --model-syms 8192 --reps 1 > $out/ncu.log 2>&1
echo "ncu exit $?"
