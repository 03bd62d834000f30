#!/bin/bash
# Round check after the pseudo-adaptive work: GPU parity tests, smoke, bench lines (c2, c4, fold1), c4 launch list + full capture.
tag=${1:-r01d}
out=gpurun_out/$tag; mkdir -p $out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $out/smi.txt 2>&1
nproc > $out/host.txt; grep -m1 'model name' /proc/cpuinfo >> $out/host.txt
timeout 1500 python -m pytest tests -m gpu -x -q > $out/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a $out/pytest_gpu.log
tail -3 $out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > $out/smoke.log 2>&1; echo "smoke exit $?"; tail -4 $out/smoke.log
timeout 600 python bench.py > $out/bench_c2.json 2> $out/bench_c2.err; echo "bench c2 exit $?"
for wl in fold1 c4; do
  timeout 600 python bench.py --workload $wl --no-targets > $out/bench_$wl.json 2> $out/bench_$wl.err; echo "bench $wl exit $?"
done
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/bench_c4_launches.csv \
  python bench.py --workload c4 --steps 2 --warmup 1 --no-targets --no-cpu > $out/ncu_launches_c4.log 2>&1; echo "launch list c4 exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k 'regex:^(pa_.*|encode_kernel|decode_kernel)$' -s 9 -c 9 -o $out/full_c4 \
  python tools/pabench.py --model-syms 8192 --reps 1 > $out/ncu_full_c4.log 2>&1; echo "ncu full c4 exit $?"
timeout 600 python tools/pabench.py --model-syms 1024,8192,65536 --reps 3 > $out/pabench.log 2>&1; echo "pabench exit $?"; tail -20 $out/pabench.log
ls -la $out
cat $out/bench_c2.json $out/bench_fold1.json $out/bench_c4.json | cut -c1-2500
