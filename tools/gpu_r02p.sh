#!/bin/bash
# r02p: pseudo-adaptive mode after the normaliser changes: tests, launch list, C4 line
tag=${1:-r02p}; out=gpurun_out/$tag; mkdir -p $out
timeout 900 python -m pytest tests/test_gpu_pseudo_adaptive.py -x -q > $out/pytest_pa.log 2>&1; echo "pa tests exit $?"; tail -3 $out/pytest_pa.log
timeout 600 python bench.py --workload c4 --no-targets > $out/bench_c4.json 2> $out/bench_c4.err; echo "bench c4 exit $?"; tail -2 $out/bench_c4.err
python - <<PY
import json
try:
    d=json.loads(open("$out/bench_c4.json").read().strip().splitlines()[-1]); k=d["kernels"]
    print('c4 value', round(d['value'],1), 'step', round(d['ms_per_step'],3), 'enc', round(d["encode_ms"],3), 'dec', round(d["decode_ms"],3), {n:(v["kernel"], round(v["ms"],3)) for n,v in k.items()}, d.get('roundtrip_ok'), 'parity', (d.get('parity') or {}).get('ok'), 'e2e', round(d['e2e']['value'],2))
except Exception as e: print('ERR', e)
PY
bash tools/gpu_launchlist.sh $tag c4 40 16 -- python bench.py --workload c4 --no-targets --steps 2 --warmup 3 --no-cpu --no-parity | grep -v "^void at::" | tail -18
