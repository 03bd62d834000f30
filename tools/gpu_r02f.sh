#!/bin/bash
# r02f: full GPU suite after the model cache / model structs / rfold changes, C3r line
tag=${1:-r02f}; out=gpurun_out/$tag; mkdir -p $out
t0=$(date +%s)
timeout 1800 python -m pytest tests -m gpu -x -q > $out/pytest_gpu.log 2>&1; echo "pytest exit $? ($(( $(date +%s) - t0 )) s)"; tail -8 $out/pytest_gpu.log
timeout 900 python bench.py --workload c3r --no-targets > $out/bench_c3r.json 2> $out/bench_c3r.err; echo "bench c3r exit $?"; tail -3 $out/bench_c3r.err
python - <<PY
import json
try:
    d=json.loads(open("$out/bench_c3r.json").read().strip().splitlines()[-1]); k=d["kernels"]
    print('c3r value', round(d['value'],1), 'enc', round(d["encode_ms"],3), 'dec', round(d["decode_ms"],3), {n:(v["kernel"], round(v["ms"],3), round(v["frac"],3)) for n,v in k.items()}, d.get('roundtrip_ok'), 'parity', (d.get('parity') or {}).get('ok'), 'e2e', d['e2e']['value'])
except Exception as e: print('ERR', e)
PY
