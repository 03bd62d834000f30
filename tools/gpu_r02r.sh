#!/bin/bash
# r02r: full GPU suite + C4 line with the driver's remap beside it
tag=${1:-r02r}; out=gpurun_out/$tag; mkdir -p $out
t0=$(date +%s)
timeout 1800 python -m pytest tests -m gpu -x -q > $out/pytest_gpu.log 2>&1; echo "pytest exit $? ($(( $(date +%s) - t0 )) s)"; tail -4 $out/pytest_gpu.log
timeout 600 python bench.py --workload c4 --no-targets > $out/bench_c4.json 2> $out/bench_c4.err; echo "bench c4 exit $?"; tail -2 $out/bench_c4.err
python - <<PY
import json
try:
    d=json.loads(open("$out/bench_c4.json").read().strip().splitlines()[-1])
    print('c4 value', round(d['value'],1), 'step', round(d['ms_per_step'],3), 'bps', round(d['bits_per_symbol'],3), d.get('roundtrip_ok'), 'e2e', round(d['e2e']['value'],2))
    print(d.get('driver_remap'))
except Exception as e: print('ERR', e)
PY
