#!/bin/bash
# r01g set: block sizes from b200ans_suggest_block_syms in the 2^30 / 2^32 lines; tests, bench lines, launch list,
# one full ncu capture of ANSfold-1.  usage: gpu_r01g.sh tag
tag=${1:-r01g}; out=gpurun_out/$tag; mkdir -p $out
K="histogram_kernel|lpb3_encode_kernel|compact_kernel|lpb3_decode_kernel"
timeout 1500 python -m pytest tests -m gpu -x -q > $out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -3 $out/pytest_gpu.log
timeout 600 python bench.py > $out/bench_c2.json 2> $out/bench_c2.err; echo "bench c2 exit $?"
for wl in ${WLS:-fold1 fold2 ans30 c5 msb}; do
  timeout 600 python bench.py --workload $wl --no-targets > $out/bench_$wl.json 2> $out/bench_$wl.err; echo "bench $wl exit $?"
done
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/bench_fold1_launches.csv \
  python bench.py --workload fold1 --steps 2 --warmup 1 --no-targets --no-cpu > $out/ncu_launches.log 2>&1; echo "launch list exit $?"
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:"$K" -s 12 -c 4 -o $out/full_fold1 \
  python tools/kbench.py --workloads fold1 --paths 0 --reps 1 > $out/ncu_full_fold1.log 2>&1; echo "ncu full fold1 exit $?"
python tools/ncu_summary.py $out/full_fold1.ncu-rep > $out/fold1_ncu_full_summary.txt 2>/dev/null
python tools/ncu_traffic.py fold1 $out/full_fold1.ncu-rep > /dev/null
cp profiles/traffic.json $out/traffic.json
rm -f $out/full_fold1.ncu-rep
for f in $out/bench_*.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); k=d["kernels"]
    print(sys.argv[1].split('/')[-1], d["config"]["block_syms"], round(d['value'],1), 'enc', round(d["encode_ms"],3), 'dec', round(d["decode_ms"],3), 'bps', round(d["bits_per_symbol"],4), {n:(round(v["ms"],3), round(v["frac"],3)) for n,v in k.items()}, 'e2e', round(d['e2e']['value'],2), d.get('roundtrip_ok'))
except Exception as e:
    print(sys.argv[1], 'ERR', e)
PY
done
du -sh $out
