#!/bin/bash
# Run on the B200 box through gpurun: GPU parity tests, bench lines, ncu launch list + full capture.
# Usage: gpurun --timeout 2400 -- 'bash tools/gpu_check.sh [tag]'
tag=${1:-r01}
out=gpurun_out/$tag
mkdir -p $out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $out/smi.txt 2>&1
nproc > $out/host.txt; grep -m1 'model name' /proc/cpuinfo >> $out/host.txt
python -m pytest tests -m gpu -x -q > $out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > $out/smoke.log 2>&1; echo "smoke exit $?" >> $out/smoke.log
python bench.py --steps 5 --warmup 3 > $out/bench_c2.json 2> $out/bench_c2.err
python bench.py --workload fold1 --steps 5 --warmup 3 --cpu-sample 20000000 > $out/bench_fold1.json 2> $out/bench_fold1.err
python bench.py --workload fold2 --steps 3 --warmup 3 --no-cpu > $out/bench_fold2.json 2> $out/bench_fold2.err
python bench.py --workload ans30 --steps 3 --warmup 3 --no-cpu > $out/bench_ans30.json 2> $out/bench_ans30.err
python bench.py --workload c1 --steps 5 --warmup 3 --no-cpu > $out/bench_c1.json 2> $out/bench_c1.err
python bench.py --impl reference --steps 2 --warmup 1 > $out/bench_ref_c2.json 2> $out/bench_ref_c2.err
CMD="python bench.py --workload fold1 --symbols 67108864 --steps 1 --warmup 3 --no-cpu"
$CMD > $out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $out/launches.csv $CMD > $out/ncu_launch.log 2>&1
$CMD > $out/plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'encode_kernel|decode_kernel|histogram_kernel' -s 9 -c 3 -o $out/prof $CMD > $out/ncu_full.log 2>&1
ls -la $out
tail -3 $out/pytest_gpu.log; cat $out/bench_c2.json | head -c 1500
