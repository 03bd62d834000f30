#!/bin/bash
# Round profile set: bench lines, the ncu launch list of the default bench command, and one
# `ncu --set full` capture of the four hot kernels per workload.  usage: gpu_profiles.sh tag
tag=${1:-r01e}
out=gpurun_out/$tag; mkdir -p $out
K="histogram_kernel|lpb3_encode_kernel|compact_kernel|lpb3_decode_kernel"
timeout 1500 python -m pytest tests -m gpu -x -q > $out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -3 $out/pytest_gpu.log
timeout 600 python bench.py > $out/bench_c2.json 2> $out/bench_c2.err; echo "bench c2 exit $?"
for wl in fold1 fold2 ans30 c4 c1 c5; do
  timeout 600 python bench.py --workload $wl --no-targets > $out/bench_$wl.json 2> $out/bench_$wl.err; echo "bench $wl exit $?"
done
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > $out/bench_ref_c2.json 2> $out/bench_ref_c2.err; echo "bench ref exit $?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/bench_c2_launches.csv \
  python bench.py --steps 2 --warmup 1 --no-targets > $out/ncu_launches.log 2>&1; echo "launch list exit $?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/bench_c4_launches.csv \
  python bench.py --workload c4 --steps 2 --warmup 1 --no-targets --no-cpu > $out/ncu_launches_c4.log 2>&1; echo "launch list c4 exit $?"
python -c "import __graft_entry__ as g; g.smoke()" > $out/smoke.log 2>&1; echo "smoke exit $?"; tail -4 $out/smoke.log
for wl in fold1 ans30; do
  timeout 1500 ncu --set full --clock-control none --import-source on -k regex:"$K" -s 12 -c 4 -o $out/full_$wl \
    python tools/kbench.py --workloads $wl --paths 0 --reps 1 > $out/ncu_full_$wl.log 2>&1; echo "ncu full $wl exit $?"
  python tools/ncu_summary.py $out/full_$wl.ncu-rep > $out/${wl}_ncu_full_summary.txt 2>/dev/null
  python tools/ncu_traffic.py $wl $out/full_$wl.ncu-rep > /dev/null
done
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"$K" -s 12 -c 4 -o $out/full_c2 \
  python tools/kbench.py --workloads c2 --paths 0 --reps 1 --block-syms 4096 > $out/ncu_full_c2.log 2>&1; echo "ncu full c2 exit $?"
python tools/ncu_summary.py $out/full_c2.ncu-rep > $out/c2_ncu_full_summary.txt 2>/dev/null
python tools/ncu_traffic.py c2 $out/full_c2.ncu-rep > /dev/null
timeout 900 ncu --set full --clock-control none --import-source on -k 'regex:^(pa_.*)$' -s 7 -c 7 -o $out/full_c4 \
  python tools/pabench.py --model-syms 8192 --reps 1 > $out/ncu_full_c4.log 2>&1; echo "ncu full c4 exit $?"
python tools/ncu_summary.py $out/full_c4.ncu-rep > $out/c4_ncu_full_summary.txt 2>/dev/null
python tools/ncu_traffic.py c4 $out/full_c4.ncu-rep > /dev/null
cp profiles/traffic.json $out/traffic.json
# gpurun brings back at most 64 MiB: the summaries travel, of the reports only the two smallest
rm -f $out/full_ans30.ncu-rep $out/full_c4.ncu-rep
ls -la $out
if [ -n "$WITH_2P32" ]; then
# maximum size: 2^32 symbols (16 GiB of input) on one GPU, ANSfold-1, round trip checked on the device
timeout 900 python tools/kbench.py --workloads c5 --n 4294967296 --paths 0 --reps 1 > $out/kbench_2p32.log 2>&1; echo "2^32 exit $?"; tail -2 $out/kbench_2p32.log | cut -c1-400
fi
du -sh $out
